#include "scene_builder.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <fstream>
#include <stdexcept>

namespace pth {
namespace {

// ---- 4x4 column-major helpers in GLM 0.9.9.7 operation order (type_mat4x4.inl, ext/matrix_transform.inl)
struct M4 { float c[4][4]; };
M4 identity() { M4 m; std::memset(&m, 0, sizeof(m)); for (int i = 0; i < 4; ++i) m.c[i][i] = 1.f; return m; }
void axpy_col(float* dst, const float* col, float s, bool first) {
    for (int r = 0; r < 4; ++r) dst[r] = first ? col[r] * s : dst[r] + col[r] * s;
}
M4 mul(const M4& a, const M4& b) {      // ((a0*b0 + a1*b1) + a2*b2) + a3*b3 per column
    M4 r;
    for (int j = 0; j < 4; ++j)
        for (int k = 0; k < 4; ++k) axpy_col(r.c[j], a.c[k], b.c[j][k], k == 0);
    return r;
}
M4 translate(const float v[3]) {
    M4 m = identity(), r = identity();
    for (int k = 0; k < 3; ++k) axpy_col(r.c[3], m.c[k], v[k], k == 0);
    for (int row = 0; row < 4; ++row) r.c[3][row] = r.c[3][row] + m.c[3][row];
    return r;
}
M4 scale(const float v[3]) {
    M4 m = identity(), r = identity();
    for (int k = 0; k < 3; ++k) for (int row = 0; row < 4; ++row) r.c[k][row] = m.c[k][row] * v[k];
    return r;
}
M4 rotate(float angle, float ax, float ay, float az) {
    M4 m = identity();
    float c = std::cos(angle), s = std::sin(angle);
    float inv = 1.f / std::sqrt(ax * ax + ay * ay + az * az);
    float axis[3] = {ax * inv, ay * inv, az * inv};
    float temp[3] = {(1.f - c) * axis[0], (1.f - c) * axis[1], (1.f - c) * axis[2]};
    float R[3][3];
    R[0][0] = c + temp[0] * axis[0];
    R[0][1] = temp[0] * axis[1] + s * axis[2];
    R[0][2] = temp[0] * axis[2] - s * axis[1];
    R[1][0] = temp[1] * axis[0] - s * axis[2];
    R[1][1] = c + temp[1] * axis[1];
    R[1][2] = temp[1] * axis[2] + s * axis[0];
    R[2][0] = temp[2] * axis[0] + s * axis[1];
    R[2][1] = temp[2] * axis[1] - s * axis[0];
    R[2][2] = c + temp[2] * axis[2];
    M4 r = identity();
    for (int j = 0; j < 3; ++j)
        for (int k = 0; k < 3; ++k) axpy_col(r.c[j], m.c[k], R[j][k], k == 0);
    for (int row = 0; row < 4; ++row) r.c[3][row] = m.c[3][row];
    return r;
}
float radians(float d) { return d * 0.01745329251994329576923690768489f; }

// ---- BoundingBox of src/vk_engine.h:81-115 (vec4 bounds, sentinel +-1e30)
struct Box4 {
    float lo[4] = {1e30f, 1e30f, 1e30f, 1e30f}, hi[4] = {-1e30f, -1e30f, -1e30f, -1e30f};
    void add(const ptl::TrianglePoint& p) {               // iMin/iMax on xyz (:83-88)
        const float q[3] = {p.px, p.py, p.pz};
        for (int i = 0; i < 3; ++i) { lo[i] = q[i] < lo[i] ? q[i] : lo[i]; hi[i] = q[i] < hi[i] ? hi[i] : q[i]; }   // iMin(p, lo), iMax(p, hi): ties keep p
    }
    void add(const Box4& b) {                             // glm::min/max on all four lanes (:90-93)
        for (int i = 0; i < 4; ++i) { lo[i] = b.lo[i] < lo[i] ? b.lo[i] : lo[i]; hi[i] = hi[i] < b.hi[i] ? b.hi[i] : hi[i]; }
    }
    float half_area() const {                             // :102-107
        float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
        return x * y + y * z + z * x;
    }
};

const unsigned kBins = 20;                                // BINS, src/vk_engine.h:203

std::string dir_of(const std::string& path) { return path.substr(0, path.rfind("/") + 1); }

}  // namespace

void SceneBuilder::object_matrix(const Placement& p, float out16[16]) {
    M4 m = mul(mul(mul(mul(translate(p.position), rotate(radians(p.rotation[0]), 1.f, 0.f, 0.f)),
                       rotate(radians(p.rotation[1]), 0.f, 1.f, 0.f)),
                   rotate(radians(p.rotation[2]), 0.f, 0.f, 1.f)),
               scale(p.scale));
    std::memcpy(out16, m.c, 64);
}

ptl::Material SceneBuilder::default_material() {          // RayMaterial defaults, src/vk_engine.h:69-79
    ptl::Material m;
    std::memset(&m, 0, sizeof(m));
    m.albedo[0] = m.albedo[1] = m.albedo[2] = 1.f;
    m.ior = -1.f;
    m.albedoIndex = m.metalnessIndex = m.alphaIndex = m.bumpIndex = -1;
    return m;
}

void SceneBuilder::push_default_materials() {
    ptl::Material white = default_material(), red = white, green = white, light = white, mirror = white, glass = white;
    red.albedo[1] = red.albedo[2] = 0.f;
    green.albedo[0] = green.albedo[2] = 0.f;
    light.albedo[0] = light.albedo[1] = light.albedo[2] = 0.f;
    light.emissionColor[0] = light.emissionColor[1] = light.emissionColor[2] = 1.f;
    light.emissionStrength = 2.4f;
    mirror.reflectance = 1.f;
    glass.ior = 2.f;
    for (const ptl::Material& m : {white, red, green, light, mirror, glass}) materials.push_back(m);
}

// three private TrianglePoints per face + the centroid the BVH build sorts by (:919-958)
uint32_t SceneBuilder::push_face(const ptl::TrianglePoint* v, bool frontOnly) {
    uint32_t base = (uint32_t)points.size();
    C3 c = {0.f, 0.f, 0.f};
    for (int i = 0; i < 3; ++i) {
        points.push_back(v[i]);
        c.x += v[i].px; c.y += v[i].py; c.z += v[i].pz;
    }
    ptl::Triangle t;
    std::memset(&t, 0, sizeof(t));
    t.v0 = base; t.v1 = base + 1; t.v2 = base + 2; t.frontOnly = frontOnly ? 1u : 0u;
    triangles.push_back(t);
    centroids_.push_back({c.x / 3.f, c.y / 3.f, c.z / 3.f});
    return base;
}

void SceneBuilder::finish_object(size_t firstTri, const Placement& where, uint32_t material, bool smooth, uint32_t sampler) {
    ptl::RenderObject o;
    std::memset(&o, 0, sizeof(o));
    object_matrix(where, o.m);
    o.smoothShade = smooth ? 1u : 0u;
    o.bvhIndex = (uint32_t)nodes.size();
    o.materialIndex = material;
    o.samplerIndex = sampler;
    objects.push_back(o);
    build_bvh((uint32_t)firstTri, (uint32_t)(triangles.size() - firstTri));
}

void SceneBuilder::add_mesh(const ptl::TrianglePoint* fv, size_t n, const Placement& where, int material) {
    size_t first = triangles.size();
    for (size_t t = 0; t < n; ++t) push_face(fv + 3 * t, where.frontOnly);
    finish_object(first, where, (uint32_t)material, false, 0);
}

// ---- OBJ / MTL ingest. The reference's parser is position-based (find(' ') / substr / stof), and
// its behaviour on slightly unusual files is part of what decides triangle order and content, so the
// same positions are used here (A-Q17).
namespace {

bool parse_floats(const std::string& line, size_t start, int n, float* out) {
    size_t index = start;
    for (int i = 0; i < n; ++i) {
        size_t next = line.find(' ', index);
        out[i] = std::stof(line.substr(index, next - index));
        index = next + 1;
    }
    return true;
}

struct FaceRefs { std::vector<uint32_t> v, vt, vn; int corners = 0; };

// "f a/b/c d/e/f g/h/i": vertex / uv / normal indices (1-based in the file)
FaceRefs parse_face(const std::string& line, bool haveNormals, bool& sawUVs) {
    FaceRefs f;
    for (int i = 0; i < (int)line.size() - 1; ++i)
        if (line[i] == ' ') f.corners++;
    int index = 0;
    for (int i = 0; i < f.corners; ++i) {
        int space = (int)line.find(' ', index);
        int next = (int)line.find(' ', space + 1);
        std::string corner = line.substr(space + 1, next - space - (i == f.corners - 1 ? 0 : 1));
        int s1 = (int)corner.find('/');
        int s2 = (int)corner.find('/', s1 + 1);
        std::string a = corner.substr(0, s1);
        if (!a.empty()) f.v.push_back((uint32_t)(std::stoi(a) - 1));
        std::string b = corner.substr(s1 + 1, s2 - s1 - 1);
        if (!b.empty()) { f.vt.push_back((uint32_t)(std::stoi(b) - 1)); sawUVs = true; }
        if (!haveNormals) continue;                    // the cursor is not advanced on this path (:906)
        std::string c = corner.substr(s2 + 1, corner.size() - s2 - 1);
        if (!c.empty()) f.vn.push_back((uint32_t)(std::stoi(c) - 1));
        index = next;
    }
    return f;
}

}  // namespace

bool SceneBuilder::read_obj(const std::string& path, const Placement& where, int material) {
    auto known = loadedObjects_.find(path);
    if (known != loadedObjects_.end()) {                 // same file again: an instance sharing the BVH (:802-815)
        ptl::RenderObject o;
        std::memset(&o, 0, sizeof(o));
        object_matrix(where, o.m);
        o.bvhIndex = known->second;
        o.materialIndex = (uint32_t)material;
        objects.push_back(o);
        return true;
    }
    std::ifstream in(path);
    if (!in.is_open()) return false;                     // the reference returns silently (:834)

    struct P3 { float x, y, z; };
    struct P2 { float u, v; };
    std::vector<P3> pos, nrm;
    std::vector<P2> uvs;
    std::string line, mtlFile, group;
    bool smooth = false, sawUVs = false;
    size_t firstTri = triangles.size();
    const std::string base = dir_of(path);

    while (in) {
        std::getline(in, line);
        if (line.find("mtllib") != std::string::npos) {
            mtlFile = line.substr(7, line.size() - 7);
            read_mtl(base + mtlFile);
        }
        std::string tag = line.substr(0, line.find(' '));
        if (tag == "v") {
            float f[3]; parse_floats(line, 2, 3, f);
            pos.push_back({f[0], f[1], f[2]});
        } else if (tag == "vt") {
            size_t a = line.find(' ', 2), b = line.find(' ', a + 1);
            float u = std::stof(line.substr(a + 1, b - a - 1));
            float v = std::stof(line.substr(b, line.length() - b));
            uvs.push_back({u, v});
        } else if (tag == "vn") {
            float f[3]; parse_floats(line, 3, 3, f);
            nrm.push_back({f[0], f[1], f[2]});
        } else if (tag == "f") {
            FaceRefs f = parse_face(line, !nrm.empty(), sawUVs);
            if (f.corners > 4) throw std::runtime_error("OBJ face with more than 4 corners: " + path);
            ptl::TrianglePoint tp[4];
            std::memset(tp, 0, sizeof(tp));
            uint32_t firstPoint = (uint32_t)points.size();
            for (int i = 0; i < f.corners; ++i) {
                P3 n = nrm.empty() ? P3{0.f, 0.f, 0.f} : nrm.at(f.vn.at(i));
                P2 t = sawUVs ? uvs.at(f.vt.at(i)) : P2{0.f, 0.f};
                P3 p = pos.at(f.v.at(i));
                tp[i] = {p.x, p.y, p.z, t.u, n.x, n.y, n.z, t.v};
            }
            // every corner becomes a stored point (a 4th corner too, although only 3 are indexed, :919-958)
            uint32_t idx[4] = {0, 0, 0, 0};
            for (int i = 0; i < f.corners; ++i) { idx[i] = (uint32_t)points.size(); points.push_back(tp[i]); }
            (void)firstPoint;
            ptl::Triangle t;
            std::memset(&t, 0, sizeof(t));
            t.v0 = idx[0]; t.v1 = idx[1]; t.v2 = idx[2]; t.frontOnly = where.frontOnly ? 1u : 0u;
            C3 c = {0.f, 0.f, 0.f};
            for (int i = 0; i < 3; ++i) { const ptl::TrianglePoint& q = points[idx[i]]; c.x += q.px; c.y += q.py; c.z += q.pz; }
            triangles.push_back(t);
            centroids_.push_back({c.x / 3.f, c.y / 3.f, c.z / 3.f});
        } else if (tag == "usemtl") {
            size_t sp = line.find(' ');
            std::string name = line.substr(sp + 1, line.size() - sp - 1);
            if (group.empty()) { group = name; continue; }
            // a new group closes the previous one: one object + BVH per usemtl group (:967-1002)
            uint32_t mat = loadedMaterials_.at(base + mtlFile + "/" + group);
            loadedObjects_.emplace(path + "/" + group, (uint32_t)nodes.size());
            finish_object(firstTri, where, mat, smooth, where.samplerIndex);
            group = name;
            firstTri = triangles.size();
            smooth = false;
        } else if (tag == "s") {
            smooth = (line.at(2) - '0') == 1;
        }
    }
    uint32_t mat = group.empty() ? (uint32_t)material : loadedMaterials_.at(base + mtlFile + "/" + group);
    loadedObjects_.emplace(path, (uint32_t)nodes.size());
    finish_object(firstTri, where, mat, smooth, 0);       // the last object keeps sampler 0 (:1009-1019)
    return true;
}

void SceneBuilder::read_mtl(const std::string& path) {
    std::ifstream in(path);
    if (!in.is_open()) return;
    std::string line, name;
    ptl::Material cur = default_material();
    const std::string base = dir_of(path);
    auto slot = [&](const std::string& l) {
        size_t sp = l.find(' ');
        texturePaths.push_back(base + l.substr(sp + 1, l.size() - sp - 1));
        return (int32_t)(texturePaths.size() - 1);
    };
    while (in) {
        std::getline(in, line);
        if (line.find("newmtl") != std::string::npos) {
            if (!name.empty()) {
                loadedMaterials_.emplace(path + "/" + name, (uint32_t)materials.size());
                materials.push_back(cur);
                cur = default_material();
            }
            name = line.substr(7, line.size() - 7);
            continue;
        }
        line.erase(std::remove(line.begin(), line.end(), '\t'), line.end());
        std::string tag = line.substr(0, line.find(' '));
        if (tag == "Ka" || tag == "Kd") {                 // both multiply into albedo (A-Q18)
            size_t a = line.find(' '), b = line.find(' ', a + 1), c = line.find(' ', b + 1);
            cur.albedo[0] *= std::stof(line.substr(a + 1, b - a - 1));
            cur.albedo[1] *= std::stof(line.substr(b + 1, c - b - 1));
            cur.albedo[2] *= std::stof(line.substr(c + 1, line.size() - c - 1));
        } else if (tag == "map_Ka" || tag == "map_Kd") {
            cur.albedoIndex = slot(line);
        } else if (tag == "map_Ks") {
            cur.metalnessIndex = slot(line);
        } else if (tag == "map_d") {
            cur.alphaIndex = slot(line);
        } else if (tag == "map_bump") {                   // lower-case spelling only (A-Q18)
            cur.bumpIndex = slot(line);
        }                                                 // Ni and d are read and dropped by the reference
    }
    loadedMaterials_.emplace(path + "/" + name, (uint32_t)materials.size());
    materials.push_back(cur);
}

// ---- BVH2, binned SAH, 20 bins per axis (src/vk_engine.cpp:1169-1337)
void SceneBuilder::fit_bounds(uint32_t ni) {               // update_bvh_bounds :1196-1210
    ptl::BVHNode& n = nodes[ni];
    Box4 b;
    for (uint32_t i = 0; i < n.triCount; ++i) {
        const ptl::Triangle& t = triangles[n.index + i];
        b.add(points[t.v0]); b.add(points[t.v1]); b.add(points[t.v2]);
    }
    n.minX = b.lo[0]; n.maxX = b.hi[0]; n.minY = b.lo[1]; n.maxY = b.hi[1]; n.minZ = b.lo[2]; n.maxZ = b.hi[2];
}

float SceneBuilder::best_split(const ptl::BVHNode& node, int& axis, float& pos) const {   // find_bvh_split_plane :1279-1337
    float best = 1e30f;
    for (int a = 0; a < 3; ++a) {
        auto cen = [&](uint32_t i) { const C3& c = centroids_[node.index + i]; return a == 0 ? c.x : (a == 1 ? c.y : c.z); };
        float lo = 1e30f, hi = -1e30f;
        for (uint32_t i = 0; i < node.triCount; ++i) {
            float c = cen(i);
            lo = lo < c ? lo : c;
            hi = hi < c ? c : hi;
        }
        if (lo == hi) continue;

        Box4 binBox[kBins];
        uint32_t binCount[kBins] = {0};
        float scale = kBins / (hi - lo);
        for (uint32_t i = 0; i < node.triCount; ++i) {
            const ptl::Triangle& t = triangles[node.index + i];
            float f = std::floor((cen(i) - lo) * scale);
            float lim = (float)(kBins - 1);
            int b = (int)(lim < f ? lim : f);
            binCount[b]++;
            binBox[b].add(points[t.v0]); binBox[b].add(points[t.v1]); binBox[b].add(points[t.v2]);
        }

        // sweep from both ends. The reference stores the right-hand area of step i at BOTH [i] and
        // [kBins-2-i] (A-Q14); the later write wins, which is what the loop below reproduces.
        float leftArea[kBins - 1], rightArea[kBins - 1], leftN[kBins - 1], rightN[kBins - 1];
        Box4 lbox, rbox;
        int lsum = 0, rsum = 0;
        for (unsigned i = 0; i < kBins - 1; ++i) {
            lsum += binCount[i];
            leftN[i] = (float)lsum;
            lbox.add(binBox[i]);
            leftArea[i] = lbox.half_area();
            rsum += binCount[kBins - 1 - i];
            rightN[kBins - 2 - i] = (float)rsum;
            rbox.add(binBox[kBins - 1 - i]);
            float ra = rbox.half_area();
            rightArea[i] = ra;
            rightArea[kBins - 2 - i] = ra;
        }
        scale = (hi - lo) / kBins;
        for (unsigned i = 0; i < kBins - 1; ++i) {
            float cost = leftN[i] * leftArea[i] + rightN[i] * rightArea[i];   // empty side: 0 * inf = NaN, never < best (A-Q15)
            if (cost < best) { best = cost; axis = a; pos = lo + scale * (float)(i + 1); }
        }
    }
    return best;
}

void SceneBuilder::build_bvh(uint32_t firstTri, uint32_t count) {   // build_bvh :1169-1194 + subdivide_bvh :1212-1277
    if (count == 0) throw std::runtime_error("object without triangles");
    if (bvhFn_ && count >= bvhMinTris_) {                   // plugged-in builder (b200pt_build_bvh): same bytes
        uint32_t root = (uint32_t)nodes.size(), made = 0;
        double ms = 0.0;
        std::vector<ptl::BVHNode> out((size_t)count * 2 - 1);
        if (bvhFn_(bvhDevice_, points.data(), (uint32_t)points.size(), triangles.data(), firstTri, count, root, out.data(), &made, &ms))
            throw std::runtime_error("plugged-in BVH builder failed");
        nodes.insert(nodes.end(), out.begin(), out.begin() + made);
        nodesUsed_ = (uint32_t)nodes.size();
        bvhDeviceMs += ms; bvhPluggedBuilds++;
        return;                                             // centroids_ keep their original order: only used by build_bvh of this range
    }
    nodesUsed_++;
    uint32_t root = (uint32_t)nodes.size();
    ptl::BVHNode zero;
    std::memset(&zero, 0, sizeof(zero));
    nodes.resize(nodes.size() + (size_t)count * 2 - 1, zero);
    nodes[root].index = firstTri;
    nodes[root].triCount = count;
    fit_bounds(root);

    // depth-first, left subtree first: the order in which the recursive reference hands out node pairs
    struct Todo { uint32_t node, depth; };
    std::vector<Todo> stack;
    stack.push_back({root, 0});
    while (!stack.empty()) {
        Todo t = stack.back(); stack.pop_back();
        ptl::BVHNode& n = nodes[t.node];
        if (n.triCount <= 2 || t.depth >= 64) continue;
        int axis = 0; float pos = 0.f;
        float cost = best_split(n, axis, pos);
        Box4 pb;
        pb.lo[0] = n.minX; pb.lo[1] = n.minY; pb.lo[2] = n.minZ; pb.hi[0] = n.maxX; pb.hi[1] = n.maxY; pb.hi[2] = n.maxZ;
        float stay = (float)n.triCount * pb.half_area();
        if (cost >= stay) continue;

        int i = (int)n.index, j = i + (int)n.triCount - 1;      // in-place partition, swapping from the back (:1233-1244)
        while (i <= j) {
            const C3& c = centroids_[i];
            float v = axis == 0 ? c.x : (axis == 1 ? c.y : c.z);
            if (v < pos) { i++; }
            else { std::swap(triangles[i], triangles[j]); std::swap(centroids_[i], centroids_[j]); j--; }
        }
        uint32_t left = (uint32_t)i - n.index;
        if (left == 0 || left == n.triCount) continue;

        uint32_t kids = nodesUsed_;
        nodesUsed_ += 2;
        nodes[kids].index = n.index; nodes[kids].triCount = left;
        nodes[kids + 1].index = (uint32_t)i; nodes[kids + 1].triCount = n.triCount - left;
        n.index = kids;
        n.triCount = 0;
        fit_bounds(kids);
        fit_bounds(kids + 1);
        stack.push_back({kids + 1, t.depth + 1});
        stack.push_back({kids, t.depth + 1});
    }
    nodes.resize(nodesUsed_);
}

// ---- scene presets (src/vk_engine.cpp:638-758)
void SceneBuilder::cornell_box(const std::string& dir) {
    auto place = [](float px, float py, float pz, float rx, float ry, float rz) {
        Placement p; p.frontOnly = true;
        p.position[0] = px; p.position[1] = py; p.position[2] = pz;
        p.rotation[0] = rx; p.rotation[1] = ry; p.rotation[2] = rz;
        return p;
    };
    read_obj(dir + "light2.obj", place(0.f, -1.5f, 0.f, 0.f, 0.f, 0.f), 3);
    read_obj(dir + "plane.obj", place(0.f, 0.5f, 0.f, 0.f, 0.f, 0.f), 0);         // bottom
    read_obj(dir + "plane.obj", place(-1.f, -0.5f, 0.f, 90.f, 0.f, 90.f), 2);     // left, green
    read_obj(dir + "plane.obj", place(1.f, -0.5f, 0.f, 90.f, 0.f, -90.f), 1);     // right, red
    read_obj(dir + "ceiling.obj", place(0.f, -1.5f, 0.f, 0.f, 0.f, 0.f), 0);      // top
    read_obj(dir + "plane.obj", place(0.f, -0.5f, 1.f, 90.f, 0.f, 0.f), 0);       // back
    read_obj(dir + "plane.obj", place(0.f, -0.5f, -1.f, -90.f, 0.f, 0.f), 0);     // front
}

void SceneBuilder::default_scene(const std::string& dir) {        // prepare_storage_buffers, HEAD state
    ptl::Sphere none;
    std::memset(&none, 0, sizeof(none));
    spheres.assign(10, none);
    push_default_materials();
    Placement cube;
    cube.samplerIndex = 1;
    cube.scale[0] = cube.scale[1] = cube.scale[2] = 0.25f;
    cube.rotation[1] = -30.f;
    cube.position[0] = -0.4f; cube.position[1] = 0.25f; cube.position[2] = -0.45f;
    read_obj(dir + "cube.obj", cube, 0);
    cube.scale[0] = 0.3f; cube.scale[1] = 0.7f; cube.scale[2] = 0.3f;
    cube.rotation[1] = 30.f;
    cube.position[0] = 0.4f; cube.position[1] = -0.2f; cube.position[2] = 0.45f;
    read_obj(dir + "cube.obj", cube, 0);
    cornell_box(dir);
}

}  // namespace pth
