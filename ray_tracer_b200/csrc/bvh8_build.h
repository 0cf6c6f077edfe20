// Host-side builder of the internal wide BVH (pt_bvh8.h) from the reference-layout buffers:
// world-space boxes per (object, triangle) -> binned-SAH binary tree -> greedy collapse to 8
// children -> octant-ordered slot assignment -> 8-bit quantisation (rounded outward).
// This is upload-time work replacing nothing in the shader; the reference's own per-object BVH2
// (src/vk_engine.cpp:1169-1337) is still accepted and used by the reference-layout kernel.
#pragma once
#include <algorithm>
#include <cmath>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "pt_bvh8.h"

namespace pt {

struct Bvh8Host {
    std::vector<Bvh8Node> nodes;
    std::vector<PrimRecord> prims;
    std::vector<ShadeRecord> shade;
    std::vector<UvRecord> uvs;
    uint32_t maxDepth = 0;
};

struct WideSceneHost {
    Bvh8Host all, emis;
    std::vector<Xform> xforms;
    std::vector<uint32_t> objXform;
    std::vector<UpNode> up;
    bool anyEmissive = false;
    uint32_t uniqSphereMask = 0, emisSphereMask = 0, plainSphereMask = 0;
    bool emissionFinite = true;
    float emisLo[3] = {0, 0, 0}, emisHi[3] = {0, 0, 0};
    // for the device build of `all` (wide_build.cu): every (object, triangle) instance and the BVH2 leaf of each triangle
    std::vector<uint32_t> instObject, instTri, leafOf;
    // (object, triangle) -> record index in all.prims (WideScene::primOf): per object the first slot and its lowest triangle
    // index (an object's triangles are one contiguous range of the reference's `triangles`); primOf itself is filled from
    // the finished tree (host: fill_prim_of below; device build: k_wide_prim_of)
    std::vector<uint32_t> instBase, triFirst, primOf;
    uint32_t nInstSlots = 0;
    float extent = 0.f;        // WideScene::extent
};

// Largest |coordinate| the scene reaches in world space: the corners of every object's BVH2 root box through its matrix,
// and the spheres. Scale of the contest window (pt_traverse.h); non-finite values are skipped.
inline float wide_extent(const std::vector<ptl::Sphere>& spheres, const std::vector<ptl::RenderObject>& objects, const std::vector<ptl::BVHNode>& nodes) {
    float m = 0.f;
    auto grow = [&](float v) { v = std::fabs(v); if (v < 1e30f && v > m) m = v; };
    for (const auto& s : spheres) { grow(std::fabs(s.px) + std::fabs(s.radius)); grow(std::fabs(s.py) + std::fabs(s.radius)); grow(std::fabs(s.pz) + std::fabs(s.radius)); }
    for (const auto& ob : objects) {
        if (ob.bvhIndex >= nodes.size()) continue;
        const ptl::BVHNode& r = nodes[ob.bvhIndex];
        for (int c = 0; c < 8; ++c) {
            v3 p = mul_point(ob.m, mk3((c & 1) ? r.maxX : r.minX, (c & 2) ? r.maxY : r.minY, (c & 4) ? r.maxZ : r.minZ));
            grow(p.x); grow(p.y); grow(p.z);
        }
    }
    return m;
}

inline void fill_prim_of(WideSceneHost& out) {
    out.primOf.assign(out.nInstSlots, 0xffffffffu);
    for (size_t i = 0; i < out.all.prims.size(); ++i) {
        const PrimRecord& r = out.all.prims[i];
        out.primOf[out.instBase[r.objectIndex] + (r.triIndexFlags & PT_PRIM_TRIMASK) - out.triFirst[r.objectIndex]] = (uint32_t)i;
    }
}

struct WideBox { float lo[3], hi[3]; };

// World-space box of one triangle instance, padded: the hit the object-space test accepts lies within rounding of it.
// Shared by the builders and by the refit after a transform edit (wide_build.cu: k_wide_refit). Returns false when a
// world-space vertex is not finite.
PT_HD bool wide_world_box(const ptl::RenderObject& ob, v3 pa, v3 pb, v3 pc, WideBox& bx) {
    for (int ax = 0; ax < 3; ++ax) { bx.lo[ax] = 1e30f; bx.hi[ax] = -1e30f; }
    v3 w[3] = {mul_point(ob.m, pa), mul_point(ob.m, pb), mul_point(ob.m, pc)};
    bool finite = true;
    for (int k = 0; k < 3; ++k) {
        float q[3] = {w[k].x, w[k].y, w[k].z};
        for (int ax = 0; ax < 3; ++ax) {
            if (!(fabsf(q[ax]) < 1e30f)) finite = false;
            bx.lo[ax] = fminf(bx.lo[ax], q[ax]); bx.hi[ax] = fmaxf(bx.hi[ax], q[ax]);
        }
    }
    float mag = 0.f;
    for (int ax = 0; ax < 3; ++ax) mag = fmaxf(mag, fmaxf(fabsf(bx.lo[ax]), fabsf(bx.hi[ax])));
    float pad = 1e-5f * mag + 1e-7f;
    for (int ax = 0; ax < 3; ++ax) { bx.lo[ax] -= pad; bx.hi[ax] += pad; }
    return finite;
}

// Records of one (object, triangle) instance; shared by the host builder and the device build so both produce the
// same bytes. Returns false when a world-space vertex is not finite (such a triangle can never be hit: dropped).
PT_HD bool wide_make_records(uint32_t o, uint32_t t, const ptl::RenderObject& ob, const ptl::Triangle& tr, const ptl::TrianglePoint& a,
                             const ptl::TrianglePoint& b, const ptl::TrianglePoint& c, uint32_t leaf, const UpNode* up, uint32_t xformId,
                             bool em, PrimRecord& r, ShadeRecord& s, UvRecord& uvr, WideBox& bx) {
    v3 pa = mk3(a.px, a.py, a.pz), pb = mk3(b.px, b.py, b.pz), pc = mk3(c.px, c.py, c.pz);
    TriEdges te = tri_edges(pa, pb, pc);
    r.p0x = te.p0.x; r.p0y = te.p0.y; r.p0z = te.p0.z; r.objectIndex = o;
    r.e1x = te.e1.x; r.e1y = te.e1.y; r.e1z = te.e1.z;
    r.triIndexFlags = (t & PT_PRIM_TRIMASK) | (tr.frontOnly ? PT_PRIM_FRONTONLY : 0u) | (em ? PT_PRIM_EMISSIVE : 0u);
    if (leaf != 0xffffffffu) {
        const UpNode lf = up[leaf];
        if (lf.minX == lf.maxX) r.triIndexFlags |= PT_PRIM_LEAF_FLAT_X;
        if (lf.minY == lf.maxY) r.triIndexFlags |= PT_PRIM_LEAF_FLAT_Y;
        if (lf.minZ == lf.maxZ) r.triIndexFlags |= PT_PRIM_LEAF_FLAT_Z;
        if (lf.parent == 0xffffffffu) r.triIndexFlags |= PT_PRIM_LEAF_IS_ROOT;
    }
    r.e2x = te.e2.x; r.e2y = te.e2.y; r.e2z = te.e2.z; r.xformId = xformId;
    r.nx = te.n.x; r.ny = te.n.y; r.nz = te.n.z; r.leafNode = leaf;
    s.n0x = a.nx; s.n0y = a.ny; s.n0z = a.nz; s.materialIndex = ob.materialIndex;
    s.n1x = b.nx; s.n1y = b.ny; s.n1z = b.nz; s.objectIndex = o;
    s.n2x = c.nx; s.n2y = c.ny; s.n2z = c.nz; s.triIndex = t;
    const bool finite = wide_world_box(ob, pa, pb, pc, bx);
    uvr.u0 = a.u; uvr.v0 = a.v; uvr.u1 = b.u; uvr.v1 = b.v; uvr.u2 = c.u; uvr.v2 = c.v;
    uvr.samplerIndex = ob.samplerIndex;
    uvr.degenerate = ((a.u == b.u && a.v == b.v) || (b.u == c.u && b.v == c.v) || (c.u == a.u && c.v == a.v)) ? 1u : 0u;
    return finite;
}

namespace b8 {

struct Box { float lo[3], hi[3]; };
inline void box_reset(Box& b) { for (int a = 0; a < 3; ++a) { b.lo[a] = 1e30f; b.hi[a] = -1e30f; } }
inline void box_grow(Box& b, const Box& o) { for (int a = 0; a < 3; ++a) { b.lo[a] = std::min(b.lo[a], o.lo[a]); b.hi[a] = std::max(b.hi[a], o.hi[a]); } }
inline float box_half_area(const Box& b) {
    float x = b.hi[0] - b.lo[0], y = b.hi[1] - b.lo[1], z = b.hi[2] - b.lo[2];
    return x * y + y * z + z * x;
}

struct Instance { uint32_t object, tri; };     // one leaf primitive

struct Node2 {                                 // binary SAH tree
    Box box;
    uint32_t left = 0, right = 0;              // children (inner) ...
    uint32_t first = 0, count = 0;             // ... or index range (leaf, count > 0)
};

struct Builder2 {
    const std::vector<Box>& boxes;
    std::vector<float> cx, cy, cz;
    std::vector<uint32_t> idx;
    std::vector<Node2> nodes;
    static const int NB = 16;
    int maxLeaf = 3;            // triangles per leaf slot, 1..3 (tuning: B200PT_BVH_LEAF)
    float travCost = 0.5f;      // cost of one more box level relative to one triangle test (tuning: B200PT_BVH_CT)

    explicit Builder2(const std::vector<Box>& b) : boxes(b) {
        size_t n = b.size();
        cx.resize(n); cy.resize(n); cz.resize(n); idx.resize(n);
        for (size_t i = 0; i < n; ++i) {
            idx[i] = (uint32_t)i;
            cx[i] = 0.5f * (b[i].lo[0] + b[i].hi[0]);
            cy[i] = 0.5f * (b[i].lo[1] + b[i].hi[1]);
            cz[i] = 0.5f * (b[i].lo[2] + b[i].hi[2]);
        }
        nodes.reserve(2 * n + 1);
        if (const char* e = std::getenv("B200PT_BVH_LEAF")) maxLeaf = std::max(1, std::min(3, std::atoi(e)));
        if (const char* e = std::getenv("B200PT_BVH_CT")) travCost = (float)std::atof(e);
        if (const char* e = std::getenv("B200PT_BVH_COLLAPSE")) dpOn = std::string(e) == "dp";
        if (const char* e = std::getenv("B200PT_BVH_CN")) cNode = (float)std::atof(e);
    }
    float cen(uint32_t i, int a) const { return a == 0 ? cx[i] : (a == 1 ? cy[i] : cz[i]); }

    // SAH-optimal collapse into 8-wide nodes (Ylitie, Karras, Laine 2017, section 3.1), leaves being the binary leaves:
    //   C(n, 1) = A(n) * cNode + D(n, 8)                       n becomes a wide node over at most 8 children
    //   D(n, j) = min over 0 < k < j of C(left, k) + C(right, j - k)
    //   C(n, i) = min(D(n, i), C(n, i - 1))                    the subtree of n as at most i children of some wide node
    // dpK[n][j] = the k of D(n, j); dpUse[n] bit i = C(n, i) takes D(n, i). Tuning: B200PT_BVH_COLLAPSE=dp, B200PT_BVH_CN.
    std::vector<float> dpC; std::vector<uint8_t> dpK; std::vector<uint16_t> dpUse;
    bool dpOn = false;
    float cNode = 1.f, cTri = 1.f;
    void collapse_dp() {
        const size_t n = nodes.size();
        dpC.assign(n * 9, 0.f); dpK.assign(n * 9, 0); dpUse.assign(n, 0);
        for (size_t m = n; m-- > 0;) {                        // children have larger indices than their parent
            const Node2& nd = nodes[m];
            float* C = &dpC[m * 9];
            const float A = box_half_area(nd.box);
            if (nd.count) { for (int i = 1; i <= 8; ++i) C[i] = A * nd.count * cTri; continue; }
            const float* L = &dpC[(size_t)nd.left * 9]; const float* R = &dpC[(size_t)nd.right * 9];
            float D[9];
            for (int j = 2; j <= 8; ++j) {
                float best = 1e30f; int bk = 1;
                for (int k = 1; k < j; ++k) { float c = L[k] + R[j - k]; if (c < best) { best = c; bk = k; } }
                D[j] = best; dpK[m * 9 + j] = (uint8_t)bk;
            }
            C[1] = A * cNode + D[8];
            for (int i = 2; i <= 8; ++i) {
                if (D[i] < C[i - 1]) { C[i] = D[i]; dpUse[m] |= (uint16_t)(1u << i); }
                else C[i] = C[i - 1];
            }
        }
    }
    void dp_gather(uint32_t m, int i, uint32_t* ch, int& nch) const {
        const Node2& nd = nodes[m];
        if (nd.count) { ch[nch++] = m; return; }
        while (i > 1 && !(dpUse[m] & (1u << i))) i--;
        if (i == 1) { ch[nch++] = m; return; }
        const int k = dpK[(size_t)m * 9 + i];
        dp_gather(nd.left, k, ch, nch); dp_gather(nd.right, i - k, ch, nch);
    }

    uint32_t build(uint32_t first, uint32_t count) {
        uint32_t me = (uint32_t)nodes.size();
        nodes.emplace_back();
        Box box, cbox;
        box_reset(box); box_reset(cbox);
        for (uint32_t i = first; i < first + count; ++i) {
            uint32_t p = idx[i];
            box_grow(box, boxes[p]);
            for (int a = 0; a < 3; ++a) { float c = cen(p, a); cbox.lo[a] = std::min(cbox.lo[a], c); cbox.hi[a] = std::max(cbox.hi[a], c); }
        }
        nodes[me].box = box;
        if (count == 1) { nodes[me].first = first; nodes[me].count = 1; return me; }

        // binned SAH over the centroid box
        int bestAxis = -1, bestBin = -1;
        float bestCost = 1e30f;
        for (int a = 0; a < 3; ++a) {
            float lo = cbox.lo[a], ext = cbox.hi[a] - lo;
            if (!(ext > 0.f)) continue;
            Box bb[NB]; uint32_t bc[NB];
            for (int k = 0; k < NB; ++k) { box_reset(bb[k]); bc[k] = 0; }
            float scale = NB / ext;
            for (uint32_t i = first; i < first + count; ++i) {
                uint32_t p = idx[i];
                int k = std::min(NB - 1, (int)((cen(p, a) - lo) * scale));
                bc[k]++; box_grow(bb[k], boxes[p]);
            }
            float rightArea[NB]; uint32_t rightCnt[NB];
            Box acc; box_reset(acc); uint32_t cnt = 0;
            for (int k = NB - 1; k > 0; --k) { box_grow(acc, bb[k]); cnt += bc[k]; rightArea[k] = cnt ? box_half_area(acc) : 0.f; rightCnt[k] = cnt; }
            box_reset(acc); cnt = 0;
            for (int k = 0; k < NB - 1; ++k) {
                box_grow(acc, bb[k]); cnt += bc[k];
                if (cnt == 0 || rightCnt[k + 1] == 0) continue;
                float cost = box_half_area(acc) * cnt + rightArea[k + 1] * rightCnt[k + 1];
                if (cost < bestCost) { bestCost = cost; bestAxis = a; bestBin = k; }
            }
        }
        // leaves hold at most 3 triangles (the unary count field of a node slot)
        if (count <= (uint32_t)maxLeaf) {
            float leafCost = box_half_area(box) * count;
            float splitCost = bestAxis < 0 ? 1e30f : box_half_area(box) * travCost + bestCost;
            if (!(splitCost < leafCost)) { nodes[me].first = first; nodes[me].count = count; return me; }
        }
        uint32_t mid;
        if (bestAxis >= 0) {
            float lo = cbox.lo[bestAxis], scale = NB / (cbox.hi[bestAxis] - lo);
            uint32_t* b = idx.data() + first;
            uint32_t* e = b + count;
            uint32_t* m = std::partition(b, e, [&](uint32_t p) {
                return std::min(NB - 1, (int)((cen(p, bestAxis) - lo) * scale)) <= bestBin;
            });
            mid = (uint32_t)(m - idx.data());
        } else {
            mid = first + count / 2;          // all centroids coincide: split by position in the list
        }
        if (mid == first || mid == first + count) mid = first + count / 2;
        uint32_t l = build(first, mid - first);
        uint32_t r = build(mid, first + count - mid);
        nodes[me].left = l; nodes[me].right = r; nodes[me].count = 0;
        return me;
    }
};

inline uint8_t exponent_for(float extent) {
    // smallest e with 2^e * 255 >= extent, as a biased exponent byte (2^(b-127))
    if (!(extent > 0.f)) return 0;
    int e;
    float m = std::frexp(extent / 255.f, &e);      // extent/255 = m * 2^e, m in [0.5, 1)
    (void)m;
    int b = e + 127;                               // 2^e >= extent/255
    if (b < 1) b = 1;
    if (b > 254) b = 254;
    return (uint8_t)b;
}
inline float pow2_from_byte(uint8_t b) { uint32_t u = (uint32_t)b << 23; float f; std::memcpy(&f, &u, 4); return f; }

}  // namespace b8

// Encode the subtree rooted at binary node `n2` as wide node `me` (already allocated in out.nodes).
inline void bvh8_emit(const b8::Builder2& B, const std::vector<b8::Instance>& inst,
                      const std::vector<PrimRecord>& primOf, const std::vector<ShadeRecord>& shadeOf, const std::vector<UvRecord>& uvOf,
                      uint32_t n2, uint32_t me, uint32_t depth, Bvh8Host& out) {
    using namespace b8;
    if (depth > out.maxDepth) out.maxDepth = depth;
    // 1. gather up to 8 children: repeatedly open the inner child with the largest area
    uint32_t ch[8]; int nch = 0;
    const Node2& root = B.nodes[n2];
    if (root.count) { ch[nch++] = n2; }
    else if (B.dpOn) { const int k = B.dpK[(size_t)n2 * 9 + 8]; B.dp_gather(root.left, k, ch, nch); B.dp_gather(root.right, 8 - k, ch, nch); }
    else { ch[nch++] = root.left; ch[nch++] = root.right; }
    while (!B.dpOn && nch < 8) {
        int best = -1; float bestA = -1.f;
        for (int i = 0; i < nch; ++i) {
            const Node2& c = B.nodes[ch[i]];
            if (c.count) continue;
            float a = box_half_area(c.box);
            if (a > bestA) { bestA = a; best = i; }
        }
        if (best < 0) break;
        uint32_t open = ch[best];
        ch[best] = B.nodes[open].left;
        ch[nch++] = B.nodes[open].right;
    }
    // 2. slots: greedy assignment minimising (child centre - node centre) . d_s, where d_s is the
    //    direction whose rays visit slot s first (traversal priority is slot ^ (7 - octant))
    const Box& nb = root.box;
    float ncx[3] = {0.5f * (nb.lo[0] + nb.hi[0]), 0.5f * (nb.lo[1] + nb.hi[1]), 0.5f * (nb.lo[2] + nb.hi[2])};
    float cost[8][8];
    for (int c = 0; c < nch; ++c) {
        const Box& cb = B.nodes[ch[c]].box;
        float d[3] = {0.5f * (cb.lo[0] + cb.hi[0]) - ncx[0], 0.5f * (cb.lo[1] + cb.hi[1]) - ncx[1], 0.5f * (cb.lo[2] + cb.hi[2]) - ncx[2]};
        for (int s = 0; s < 8; ++s)
            cost[c][s] = d[0] * ((s & 4) ? -1.f : 1.f) + d[1] * ((s & 2) ? -1.f : 1.f) + d[2] * ((s & 1) ? -1.f : 1.f);
    }
    int slotOf[8], childIn[8];
    for (int i = 0; i < 8; ++i) { slotOf[i] = -1; childIn[i] = -1; }
    for (int k = 0; k < nch; ++k) {
        int bc = -1, bs = -1; float bv = 1e30f;
        for (int c = 0; c < nch; ++c) if (slotOf[c] < 0)
            for (int s = 0; s < 8; ++s) if (childIn[s] < 0 && cost[c][s] < bv) { bv = cost[c][s]; bc = c; bs = s; }
        slotOf[bc] = bs; childIn[bs] = bc;
    }
    // 3. quantisation frame
    Bvh8Node node;
    std::memset(&node, 0, sizeof(node));
    node.px = nb.lo[0]; node.py = nb.lo[1]; node.pz = nb.lo[2];
    uint8_t eb[3]; float sc[3];
    for (int a = 0; a < 3; ++a) {
        eb[a] = exponent_for(nb.hi[a] - nb.lo[a]);
        for (;;) {      // make sure the largest plane is representable after float rounding
            sc[a] = pow2_from_byte(eb[a]);
            if (eb[a] == 0 || eb[a] >= 254 || nb.lo[a] + 255.f * sc[a] >= nb.hi[a]) break;
            eb[a]++;
        }
    }
    node.ex = eb[0]; node.ey = eb[1]; node.ez = eb[2];
    // 4. children in slot order: inner nodes get contiguous indices, leaf triangles contiguous records
    uint32_t nInner = 0;
    for (int s = 0; s < 8; ++s) if (childIn[s] >= 0 && B.nodes[ch[childIn[s]]].count == 0) nInner++;
    node.childBase = (uint32_t)out.nodes.size();
    out.nodes.resize(out.nodes.size() + nInner);
    node.primBase = (uint32_t)out.prims.size();
    uint32_t innerSeen = 0, primOff = 0;
    uint32_t innerNode2[8], innerIdx[8]; int nRec = 0;
    for (int s = 0; s < 8; ++s) {
        int c = childIn[s];
        if (c < 0) continue;
        const Node2& cn = B.nodes[ch[c]];
        const float* lo = cn.box.lo; const float* hi = cn.box.hi;
        uint8_t* qlo[3] = {node.qlox, node.qloy, node.qloz};
        uint8_t* qhi[3] = {node.qhix, node.qhiy, node.qhiz};
        const float p[3] = {node.px, node.py, node.pz};
        for (int a = 0; a < 3; ++a) {
            int l = 0, h = 0;
            if (sc[a] > 0.f) {
                l = (int)std::floor((lo[a] - p[a]) / sc[a]);
                h = (int)std::ceil((hi[a] - p[a]) / sc[a]);
                l = std::max(0, std::min(255, l)); h = std::max(0, std::min(255, h));
                while (l > 0 && p[a] + l * sc[a] > lo[a]) l--;       // outward after rounding
                while (h < 255 && p[a] + h * sc[a] < hi[a]) h++;
            }
            qlo[a][s] = (uint8_t)l; qhi[a][s] = (uint8_t)h;
        }
        if (cn.count == 0) {
            node.imask |= (uint8_t)(1u << s);
            node.meta[s] = (uint8_t)((1u << 5) | (24u + s));
            innerNode2[nRec] = ch[c]; innerIdx[nRec] = node.childBase + innerSeen; nRec++;
            innerSeen++;
        } else {
            uint32_t unary = cn.count == 1 ? 1u : (cn.count == 2 ? 3u : 7u);
            node.meta[s] = (uint8_t)((unary << 5) | primOff);
            for (uint32_t k = 0; k < cn.count; ++k) {
                uint32_t id = B.idx[cn.first + k];
                out.prims.push_back(primOf[id]);
                out.shade.push_back(shadeOf[id]);
                out.uvs.push_back(uvOf[id]);
            }
            primOff += cn.count;
        }
    }
    (void)inst;
    out.nodes[me] = node;
    for (int i = 0; i < nRec; ++i) bvh8_emit(B, inst, primOf, shadeOf, uvOf, innerNode2[i], innerIdx[i], depth + 1, out);
}

inline void bvh8_build(const std::vector<b8::Box>& boxes, const std::vector<b8::Instance>& inst,
                       const std::vector<PrimRecord>& primOf, const std::vector<ShadeRecord>& shadeOf, const std::vector<UvRecord>& uvOf,
                       Bvh8Host& out) {
    out.nodes.clear(); out.prims.clear(); out.shade.clear(); out.uvs.clear(); out.maxDepth = 0;
    out.nodes.resize(1);
    if (boxes.empty()) { std::memset(&out.nodes[0], 0, sizeof(Bvh8Node)); return; }
    auto t0 = std::chrono::steady_clock::now();
    b8::Builder2 B(boxes);
    uint32_t root = B.build(0, (uint32_t)boxes.size());
    if (B.dpOn) B.collapse_dp();
    auto t1 = std::chrono::steady_clock::now();
    out.prims.reserve(boxes.size()); out.shade.reserve(boxes.size()); out.uvs.reserve(boxes.size());
    bvh8_emit(B, inst, primOf, shadeOf, uvOf, root, 0, 1, out);
    if (std::getenv("B200PT_BUILD_TIMING"))
        fprintf(stderr, "[bvh8] %zu prims: binary build %.3f s, collapse+emit %.3f s\n", boxes.size(),
                std::chrono::duration<double>(t1 - t0).count(), std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count());
}

// Sphere slot masks (WideScene) and the flags that depend only on spheres and materials; also used alone when only
// those two buffers changed and no triangle moved between the emissive and the non-emissive set.
inline void wide_sphere_masks(const std::vector<ptl::Sphere>& spheres, const std::vector<ptl::Material>& materials, WideSceneHost& out) {
    auto emissive = [&](uint32_t m) { return !(materials[m].emissionStrength == 0.f); };
    out.uniqSphereMask = out.emisSphereMask = out.plainSphereMask = 0;
    out.anyEmissive = false;
    for (size_t i = 0; i < spheres.size(); ++i) {
        bool em = emissive(spheres[i].materialIndex);
        if (em) out.anyEmissive = true;
        if (i >= 32) continue;
        bool firstAll = true, firstClass = true;
        for (size_t k = 0; k < i; ++k) {
            if (std::memcmp(&spheres[k], &spheres[i], 16) != 0) continue;      // px, py, pz, radius
            firstAll = false;
            if (emissive(spheres[k].materialIndex) == em) firstClass = false;
        }
        if (firstAll) out.uniqSphereMask |= 1u << i;
        if (firstClass) (em ? out.emisSphereMask : out.plainSphereMask) |= 1u << i;
    }
    out.emissionFinite = true;
    for (const auto& m : materials)
        for (int k = 0; k < 3; ++k) if (!std::isfinite(m.emissionColor[k] * m.emissionStrength)) out.emissionFinite = false;
}

// Reference-layout buffers -> WideSceneHost. Returns nullptr on success or a message.
// distinct object matrices (bit pattern): objects that share one share the transformed ray
inline void wide_xforms(const std::vector<ptl::RenderObject>& objects, WideSceneHost& out) {
    out.xforms.clear(); out.objXform.assign(objects.size(), 0);
    for (size_t o = 0; o < objects.size(); ++o) {
        uint32_t id = (uint32_t)out.xforms.size();
        for (uint32_t k = 0; k < out.xforms.size(); ++k)
            if (std::memcmp(out.xforms[k].M, objects[o].m, 64) == 0) { id = k; break; }
        if (id == out.xforms.size()) {
            Xform x;
            std::memcpy(x.M, objects[o].m, 64);
            inverse4(objects[o].m, x.Minv);
            out.xforms.push_back(x);
        }
        out.objXform[o] = id;
    }
}

// After a transform edit (b200pt_update_objects): the transform table and the small tree over the triangles of EMISSIVE
// objects only (world-space boxes move with the objects), from the triangle -> leaf map and parent links kept from the
// last full re-layout. The tree over all triangles is refitted on the device (wide_build.cu: wide_refit_device).
inline const char* build_emissive_scene(const std::vector<ptl::Material>& materials, const std::vector<ptl::TrianglePoint>& points,
                                        const std::vector<ptl::Triangle>& tris, const std::vector<ptl::RenderObject>& objects,
                                        const std::vector<ptl::BVHNode>& nodes, const std::vector<uint32_t>& leafOf, const std::vector<UpNode>& up,
                                        WideSceneHost& out) {
    using namespace b8;
    wide_xforms(objects, out);
    std::vector<Box> boxesE; std::vector<Instance> instE; std::vector<PrimRecord> primE; std::vector<ShadeRecord> shadeE; std::vector<UvRecord> uvsE;
    std::vector<uint32_t> stack;
    for (size_t o = 0; o < objects.size(); ++o) {
        const ptl::RenderObject& ob = objects[o];
        if (materials[ob.materialIndex].emissionStrength == 0.f) continue;
        stack.clear(); stack.push_back(ob.bvhIndex);
        size_t guard = 0;
        while (!stack.empty()) {
            if (++guard > 4 * nodes.size() + 8) return "BVH of an object is cyclic";
            const ptl::BVHNode& nd = nodes[stack.back()]; stack.pop_back();
            if (!nd.triCount) { stack.push_back(nd.index + 1); stack.push_back(nd.index); continue; }
            for (uint32_t t = nd.index; t < nd.index + nd.triCount; ++t) {
                const ptl::Triangle& tr = tris[t];
                PrimRecord r; ShadeRecord s; UvRecord uvr; WideBox wb;
                if (!wide_make_records((uint32_t)o, t, ob, tr, points[tr.v0], points[tr.v1], points[tr.v2], leafOf[t], up.data(), out.objXform[o], true,
                                       r, s, uvr, wb)) continue;
                Box bx;
                for (int ax = 0; ax < 3; ++ax) { bx.lo[ax] = wb.lo[ax]; bx.hi[ax] = wb.hi[ax]; }
                boxesE.push_back(bx); instE.push_back({(uint32_t)o, t}); primE.push_back(r); shadeE.push_back(s); uvsE.push_back(uvr);
            }
        }
    }
    b8::Box u; b8::box_reset(u);
    for (const auto& bx : boxesE) b8::box_grow(u, bx);
    for (int a = 0; a < 3; ++a) {
        float mag = std::max(std::fabs(u.lo[a]), std::fabs(u.hi[a]));
        out.emisLo[a] = u.lo[a] - (1e-4f * mag + 1e-6f); out.emisHi[a] = u.hi[a] + (1e-4f * mag + 1e-6f);
    }
    out.emis = Bvh8Host();
    bvh8_build(boxesE, instE, primE, shadeE, uvsE, out.emis);
    if (out.emis.maxDepth > PT_BVH8_MAX_DEPTH) return "wide BVH deeper than the traversal stack";
    return nullptr;
}

inline const char* build_wide_scene(const std::vector<ptl::Sphere>& spheres, const std::vector<ptl::Material>& materials,
                                    const std::vector<ptl::TrianglePoint>& points, const std::vector<ptl::Triangle>& tris,
                                    const std::vector<ptl::RenderObject>& objects, const std::vector<ptl::BVHNode>& nodes,
                                    WideSceneHost& out, bool buildAll = true) {
    using namespace b8;
    out.anyEmissive = false;
    out.extent = wide_extent(spheres, objects, nodes);
    out.instObject.clear(); out.instTri.clear(); out.leafOf.clear();
    auto emissive = [&](uint32_t m) { return !(materials[m].emissionStrength == 0.f); };
    wide_sphere_masks(spheres, materials, out);

    wide_xforms(objects, out);

    // parent links and triangle -> leaf map of the reference BVH2
    out.up.resize(nodes.size());
    std::vector<uint32_t> leafOf(tris.size(), 0xffffffffu);
    for (size_t i = 0; i < nodes.size(); ++i) {
        UpNode& u = out.up[i];
        u.minX = nodes[i].minX; u.maxX = nodes[i].maxX; u.minY = nodes[i].minY; u.maxY = nodes[i].maxY;
        u.minZ = nodes[i].minZ; u.maxZ = nodes[i].maxZ; u.parent = 0xffffffffu; u.pad = 0;
    }
    for (size_t i = 0; i < nodes.size(); ++i) {
        if (nodes[i].triCount) { for (uint32_t t = nodes[i].index; t < nodes[i].index + nodes[i].triCount; ++t) leafOf[t] = (uint32_t)i; }
        else { out.up[nodes[i].index].parent = (uint32_t)i; out.up[nodes[i].index + 1].parent = (uint32_t)i; }
    }

    auto tPrep = std::chrono::steady_clock::now();
    std::vector<Box> boxes, boxesE;
    std::vector<Instance> inst, instE;
    std::vector<PrimRecord> prim, primE;
    std::vector<ShadeRecord> shade, shadeE;
    std::vector<UvRecord> uvs, uvsE;
    std::vector<uint32_t> stack, triList;
    out.instBase.assign(objects.size(), 0); out.triFirst.assign(objects.size(), 0); out.nInstSlots = 0;
    for (size_t o = 0; o < objects.size(); ++o) {
        const ptl::RenderObject& ob = objects[o];
        // triangles of this object = leaves under its BVH root (instances share bvhIndex: src/vk_engine.cpp:802-815)
        triList.clear(); stack.clear(); stack.push_back(ob.bvhIndex);
        size_t guard = 0;
        while (!stack.empty()) {
            if (++guard > 4 * nodes.size() + 8) return "BVH of an object is cyclic";
            const ptl::BVHNode& nd = nodes[stack.back()]; stack.pop_back();
            if (nd.triCount) for (uint32_t t = nd.index; t < nd.index + nd.triCount; ++t) triList.push_back(t);
            else { stack.push_back(nd.index + 1); stack.push_back(nd.index); }
        }
        if (!triList.empty()) {
            uint32_t lo = triList[0], hi = triList[0];
            for (uint32_t t : triList) { lo = std::min(lo, t); hi = std::max(hi, t); }
            out.instBase[o] = out.nInstSlots; out.triFirst[o] = lo;
            out.nInstSlots += hi - lo + 1;
        }
        bool em = emissive(ob.materialIndex);
        if (em && !triList.empty()) out.anyEmissive = true;
        for (uint32_t t : triList) {
            if (!buildAll) { out.instObject.push_back((uint32_t)o); out.instTri.push_back(t); if (!em) continue; }
            const ptl::Triangle& tr = tris[t];
            PrimRecord r; ShadeRecord s; UvRecord uvr; WideBox wb;
            if (!wide_make_records((uint32_t)o, t, ob, tr, points[tr.v0], points[tr.v1], points[tr.v2], leafOf[t], out.up.data(),
                                   out.objXform[o], em, r, s, uvr, wb)) continue;   // NaN/Inf vertices can never produce an accepted hit
            Box bx;
            for (int ax = 0; ax < 3; ++ax) { bx.lo[ax] = wb.lo[ax]; bx.hi[ax] = wb.hi[ax]; }
            if (buildAll) { boxes.push_back(bx); inst.push_back({(uint32_t)o, t}); prim.push_back(r); shade.push_back(s); uvs.push_back(uvr); }
            if (em) { boxesE.push_back(bx); instE.push_back({(uint32_t)o, t}); primE.push_back(r); shadeE.push_back(s); uvsE.push_back(uvr); }
        }
    }
    {
        b8::Box u; b8::box_reset(u);
        for (const auto& bx : boxesE) b8::box_grow(u, bx);
        for (int a = 0; a < 3; ++a) {
            float mag = std::max(std::fabs(u.lo[a]), std::fabs(u.hi[a]));
            out.emisLo[a] = u.lo[a] - (1e-4f * mag + 1e-6f); out.emisHi[a] = u.hi[a] + (1e-4f * mag + 1e-6f);
        }
    }
    if (std::getenv("B200PT_BUILD_TIMING"))
        fprintf(stderr, "[bvh8] records of %zu triangles: %.3f s\n", boxes.size(), std::chrono::duration<double>(std::chrono::steady_clock::now() - tPrep).count());
    if (buildAll) { bvh8_build(boxes, inst, prim, shade, uvs, out.all); fill_prim_of(out); }
    else out.leafOf = leafOf;      // `all` is built on the device from instObject / instTri / leafOf
    bvh8_build(boxesE, instE, primE, shadeE, uvsE, out.emis);
    if (out.all.maxDepth > PT_BVH8_MAX_DEPTH || out.emis.maxDepth > PT_BVH8_MAX_DEPTH) return "wide BVH deeper than the traversal stack";
    return nullptr;
}

}  // namespace pt
