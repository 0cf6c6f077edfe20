// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Plain-C entry points so tests/ (ctypes) and bench.py's cpu_baseline leg can drive the oracle.
#include <cstring>
#include <string>

#include "oracle_scene.h"
#include "oracle_shader.h"

using namespace orc;

static om::vec3 v3p(const float* p) { return om::vec3(p[0], p[1], p[2]); }

static void work_out(const Work& w, uint64_t* o) {
    if (!o) return;
    o[0] += w.n_box; o[1] += w.n_tri; o[2] += w.n_obj; o[3] += w.n_sph;
    o[4] += w.n_hit; o[5] += w.n_ray; o[6] += w.n_bounce; o[7] += w.n_path;
}

// ---- math known-answer hooks (checked against the real GLM in tests/golden/glm_kat.json) ----
static om::mat4 m4(const float* p) {
    om::mat4 m;
    for (int c = 0; c < 4; ++c) m[c] = om::vec4(p[4 * c], p[4 * c + 1], p[4 * c + 2], p[4 * c + 3]);
    return m;
}
static void put(float* o, const om::mat4& m) { for (int c = 0; c < 4; ++c) for (int r = 0; r < 4; ++r) o[4 * c + r] = m[c][r]; }
static void put(float* o, const om::vec3& v) { o[0] = v.x; o[1] = v.y; o[2] = v.z; }
static void put(float* o, const om::vec4& v) { o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }


extern "C" {

void* orc_scene_create(const void* spheres, uint32_t nS, const void* materials, uint32_t nM, const void* points,
                       uint32_t nP, const void* tris, uint32_t nT, const void* objects, uint32_t nO,
                       const void* nodes, uint32_t nN) {
    Scene* s = new Scene();
    s->spheres.assign((const ol::Sphere*)spheres, (const ol::Sphere*)spheres + nS);
    s->materials.assign((const ol::RayMaterial*)materials, (const ol::RayMaterial*)materials + nM);
    s->points.assign((const ol::TrianglePoint*)points, (const ol::TrianglePoint*)points + nP);
    s->triangles.assign((const ol::Triangle*)tris, (const ol::Triangle*)tris + nT);
    s->objects.assign((const ol::RenderObject*)objects, (const ol::RenderObject*)objects + nO);
    s->nodes.assign((const ol::BVHNode*)nodes, (const ol::BVHNode*)nodes + nN);
    s->finalize();
    return s;
}
void orc_scene_destroy(void* s) { delete (Scene*)s; }
// texture extension (see Scene::useTextures)
int orc_scene_set_texture(void* scene, uint32_t slot, const void* rgba8, uint32_t w, uint32_t h) {
    if (slot >= 64 || !rgba8 || !w || !h) return 1;
    Scene::Texture& t = ((Scene*)scene)->textures[slot];
    t.rgba.assign((const uint8_t*)rgba8, (const uint8_t*)rgba8 + size_t(w) * h * 4);
    t.w = w; t.h = h;
    return 0;
}
void orc_scene_use_textures(void* scene, int on) { ((Scene*)scene)->useTextures = on != 0; }

int orc_render_dispatch(void* scene, const void* pc208, int W, int H, float* image, uint8_t* rgba8, float* outcol,
                        float* stats, uint64_t* work8, int threads) {
    ol::PushConstants pc;
    std::memcpy(&pc, pc208, sizeof(pc));
    Work w;
    render_dispatch(*(Scene*)scene, pc, W, H, image, rgba8, outcol, stats, &w, threads);
    work_out(w, work8);
    return 0;
}

int orc_render_frames(void* scene, const void* pc208, int W, int H, uint32_t first, uint32_t n, float* accum,
                      uint64_t* work8, int threads) {
    ol::PushConstants pc;
    std::memcpy(&pc, pc208, sizeof(pc));
    Work w;
    render_frames(*(Scene*)scene, pc, W, H, first, n, accum, &w, threads);
    work_out(w, work8);
    return 0;
}

// as orc_render_frames; work16 = 8 counters over all rays followed by 8 over extension rays only
int orc_render_frames2(void* scene, const void* pc208, int W, int H, uint32_t first, uint32_t n, float* accum,
                       uint64_t* work16, int threads) {
    ol::PushConstants pc;
    std::memcpy(&pc, pc208, sizeof(pc));
    Work w, e;
    render_frames(*(Scene*)scene, pc, W, H, first, n, accum, &w, threads, &e);
    work_out(w, work16);
    work_out(e, work16 ? work16 + 8 : nullptr);
    return 0;
}

// window / pixel lattice of the full W x H image (see render_window); work16 as orc_render_frames2
int orc_render_window(void* scene, const void* pc208, int W, int H, uint32_t first, uint32_t n, int x0, int y0, int w, int h,
                      int sx, int sy, float* accum, uint64_t* work16, int threads) {
    if (w < 0 || h < 0 || sx < 1 || sy < 1 || x0 < 0 || y0 < 0) return 1;
    ol::PushConstants pc;
    std::memcpy(&pc, pc208, sizeof(pc));
    Work wk, e;
    render_window(*(Scene*)scene, pc, W, H, first, n, x0, y0, w, h, sx, sy, accum, &wk, threads, &e);
    work_out(wk, work16);
    work_out(e, work16 ? work16 + 8 : nullptr);
    return 0;
}

int orc_trace_batch(void* scene, const void* pc208, const float* rays, uint64_t n, void* hits48, uint64_t* work8,
                    int threads) {
    ol::PushConstants pc;
    std::memcpy(&pc, pc208, sizeof(pc));
    Work w;
    trace_batch(*(Scene*)scene, pc, rays, (size_t)n, (HitOut*)hits48, &w, threads);
    work_out(w, work8);
    return 0;
}

int orc_pixel_ties(void* scene, const void* pc208, int W, int H, uint32_t frame, const uint32_t* xy, uint64_t n, uint32_t* ties) {
    ol::PushConstants pc;
    std::memcpy(&pc, pc208, sizeof(pc));
    pixel_ties(*(Scene*)scene, pc, W, H, frame, xy, (size_t)n, ties);
    return 0;
}

uint64_t orc_pixel_rays(void* scene, const void* pc208, int W, int H, uint32_t frame, uint32_t x, uint32_t y, uint64_t cap, float* rays8,
                        void* hits48) {
    ol::PushConstants pc;
    std::memcpy(&pc, pc208, sizeof(pc));
    return pixel_rays(*(Scene*)scene, pc, W, H, frame, x, y, (size_t)cap, rays8, (HitOut*)hits48);
}

void orc_set_math_mode(int mode) { orc::set_math_mode(mode); }
int orc_math_mode() { return orc::math_mode(); }
float orc_random(uint32_t* state) { return orc::random(*state); }
uint32_t orc_starting_seed(uint32_t f) { return starting_seed(f); }

// ---- scene builder ----
void* orc_builder_create() { return new SceneBuilder(); }
void orc_builder_destroy(void* b) { delete (SceneBuilder*)b; }

static ImGuiObject mk_obj(const float* pos, const float* rot, const float* scl, uint32_t samplerIndex, int frontOnly) {
    ImGuiObject o;
    o.position = v3p(pos); o.rotation = v3p(rot); o.scale = v3p(scl);
    o.samplerIndex = samplerIndex; o.frontOnly = frontOnly != 0;
    return o;
}

int orc_builder_read_obj(void* b, const char* path, const float* pos, const float* rot, const float* scl,
                         uint32_t samplerIndex, int frontOnly, int material) {
    try {
        return ((SceneBuilder*)b)->read_obj(path, mk_obj(pos, rot, scl, samplerIndex, frontOnly), material) ? 0 : 1;
    } catch (const std::exception&) { return 2; }
}
int orc_builder_add_mesh(void* b, const void* pts, uint64_t nTris, const float* pos, const float* rot,
                         const float* scl, int frontOnly, int material) {
    try {
        ((SceneBuilder*)b)->add_mesh((const ol::TrianglePoint*)pts, (size_t)nTris, mk_obj(pos, rot, scl, 0, frontOnly), material);
        return 0;
    } catch (const std::exception&) { return 2; }
}
int orc_builder_prepare_default_scene(void* b, const char* assetDir) {
    try { ((SceneBuilder*)b)->prepare_default_scene(assetDir); return 0; } catch (const std::exception&) { return 2; }
}
int orc_builder_cornell_box(void* b, const char* assetDir) {
    try { ((SceneBuilder*)b)->cornell_box(assetDir); return 0; } catch (const std::exception&) { return 2; }
}
void orc_builder_push_default_materials(void* b) { ((SceneBuilder*)b)->push_default_materials(); }
void orc_builder_add_material(void* b, const void* mat64) {
    ol::RayMaterial m; std::memcpy(&m, mat64, 64); ((SceneBuilder*)b)->rayMaterials.push_back(m);
}
void orc_builder_set_spheres(void* b, const void* sph, uint32_t n) {
    ((SceneBuilder*)b)->spheres.assign((const ol::Sphere*)sph, (const ol::Sphere*)sph + n);
}
// counts: spheres, materials, points, triangles, objects, nodes, texturesUsed
void orc_builder_counts(void* b, uint32_t* out7) {
    SceneBuilder* s = (SceneBuilder*)b;
    out7[0] = s->spheres.size(); out7[1] = s->rayMaterials.size(); out7[2] = s->triPoints.size();
    out7[3] = s->triangles.size(); out7[4] = s->objects.size(); out7[5] = s->bvhNodes.size();
    out7[6] = s->texturesUsed;
}
void orc_builder_copy(void* b, int which, void* dst) {
    SceneBuilder* s = (SceneBuilder*)b;
    switch (which) {
        case 0: std::memcpy(dst, s->spheres.data(), s->spheres.size() * 32); break;
        case 1: std::memcpy(dst, s->rayMaterials.data(), s->rayMaterials.size() * 64); break;
        case 2: std::memcpy(dst, s->triPoints.data(), s->triPoints.size() * 32); break;
        case 3: std::memcpy(dst, s->triangles.data(), s->triangles.size() * 48); break;
        case 4: std::memcpy(dst, s->objects.data(), s->objects.size() * 80); break;
        case 5: std::memcpy(dst, s->bvhNodes.data(), s->bvhNodes.size() * 32); break;
    }
}
// texture slot -> file path recorded by read_mtl; returns length, copies up to cap bytes
int orc_builder_texture_path(void* b, uint32_t slot, char* dst, int cap) {
    SceneBuilder* s = (SceneBuilder*)b;
    if (slot >= s->textureFiles.size()) return -1;
    const std::string& p = s->textureFiles[slot];
    if (dst && cap > 0) { std::strncpy(dst, p.c_str(), cap - 1); dst[cap - 1] = 0; }
    return (int)p.size();
}

void orc_default_push_constants(int W, int H, uint32_t sphereCount, uint32_t objectCount, void* out208) {
    ol::PushConstants pc = default_push_constants(W, H, sphereCount, objectCount);
    std::memcpy(out208, &pc, sizeof(pc));
}
void orc_set_camera_angles(void* pc208, float ax, float ay, float az) {
    ol::PushConstants pc;
    std::memcpy(&pc, pc208, sizeof(pc));
    set_camera_angles(pc, ax, ay, az);
    std::memcpy(pc208, &pc, sizeof(pc));
}

// returns number of outputs written, -1 for an unknown function name
int orc_math_eval(const char* fn, const float* in, float* out) {
    using namespace om;
    std::string f(fn);
    if (f == "dot3") { out[0] = dot(v3p(in), v3p(in + 3)); return 1; }
    if (f == "dot4") { out[0] = dot(vec4(in[0], in[1], in[2], in[3]), vec4(in[4], in[5], in[6], in[7])); return 1; }
    if (f == "cross") { put(out, cross(v3p(in), v3p(in + 3))); return 3; }
    if (f == "normalize3") { put(out, normalize(v3p(in))); return 3; }
    if (f == "normalize4") { put(out, normalize(vec4(in[0], in[1], in[2], in[3]))); return 4; }
    if (f == "reflect") { put(out, reflect(v3p(in), v3p(in + 3))); return 3; }
    if (f == "refract") { put(out, refract(v3p(in), v3p(in + 3), in[6])); return 3; }
    if (f == "mix1") { out[0] = mix(in[0], in[1], in[2]); return 1; }
    if (f == "mix3") { put(out, mix(v3p(in), v3p(in + 3), in[6])); return 3; }
    if (f == "smoothstep") { out[0] = smoothstep(in[0], in[1], in[2]); return 1; }
    if (f == "sign") { out[0] = sign(in[0]); return 1; }
    if (f == "radians") { out[0] = radians(in[0]); return 1; }
    if (f == "min") { out[0] = gmin(in[0], in[1]); return 1; }
    if (f == "max") { out[0] = gmax(in[0], in[1]); return 1; }
    if (f == "inverse") { put(out, inverse(m4(in))); return 16; }
    if (f == "mat4vec4") { put(out, m4(in) * vec4(in[16], in[17], in[18], in[19])); return 4; }
    if (f == "mat4mat4") { put(out, m4(in) * m4(in + 16)); return 16; }
    if (f == "translate") { put(out, translate(v3p(in))); return 16; }
    if (f == "rotate") { put(out, rotate(in[0], v3p(in + 1))); return 16; }
    if (f == "scale") { put(out, scale(v3p(in))); return 16; }
    if (f == "object_matrix") {
        ImGuiObject o; o.position = v3p(in); o.rotation = v3p(in + 3); o.scale = v3p(in + 6);
        put(out, object_matrix(o)); return 16;
    }
    if (f == "camera_rotation") {
        ol::PushConstants pc; std::memset(&pc, 0, sizeof(pc));
        set_camera_angles(pc, in[0], in[1], in[2]);
        std::memcpy(out, pc.camInfo.cameraRotation, 64); return 16;
    }
    if (f == "mat3vec3") {
        mat3 m; m[0] = v3p(in); m[1] = v3p(in + 3); m[2] = v3p(in + 6);
        put(out, m * v3p(in + 9)); return 3;
    }
    return -1;
}

}  // extern "C"
