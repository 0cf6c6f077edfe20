// TEST INFRASTRUCTURE — host emulation of the wavefront pipeline (see emu_ref.cpp for the rules).
// Runs the stage bodies of ray_tracer_b200/csrc/pt_wavefront.h in CPU loops with the same queue
// discipline as wavefront.cu, over the same re-laid-out scene (bvh8_build.h).
#include <cstring>
#include <vector>

#include "../../ray_tracer_b200/csrc/bvh8_build.h"
#include "../../ray_tracer_b200/csrc/pt_wavefront.h"

using namespace pt;

namespace {
struct EmuScene {
    std::vector<ptl::Sphere> spheres; std::vector<ptl::Material> materials; std::vector<ptl::TrianglePoint> points;
    std::vector<ptl::Triangle> tris; std::vector<ptl::RenderObject> objects; std::vector<ptl::BVHNode> nodes;
    WideSceneHost wide;
    WideScene view;
    std::vector<float> minv;      // inverse(transformMatrix) per object, for the reference-order re-trace of contested rays
    RefScene ref;
};
}  // namespace

extern "C" {

void* emu_scene_create(const void* spheres, uint32_t nS, const void* materials, uint32_t nM, const void* points,
                       uint32_t nP, const void* tris, uint32_t nT, const void* objects, uint32_t nO, const void* nodes,
                       uint32_t nN) {
    EmuScene* s = new EmuScene();
    s->spheres.assign((const ptl::Sphere*)spheres, (const ptl::Sphere*)spheres + nS);
    s->materials.assign((const ptl::Material*)materials, (const ptl::Material*)materials + nM);
    s->points.assign((const ptl::TrianglePoint*)points, (const ptl::TrianglePoint*)points + nP);
    s->tris.assign((const ptl::Triangle*)tris, (const ptl::Triangle*)tris + nT);
    s->objects.assign((const ptl::RenderObject*)objects, (const ptl::RenderObject*)objects + nO);
    s->nodes.assign((const ptl::BVHNode*)nodes, (const ptl::BVHNode*)nodes + nN);
    const char* err = build_wide_scene(s->spheres, s->materials, s->points, s->tris, s->objects, s->nodes, s->wide);
    if (err) { delete s; return nullptr; }
    WideScene& v = s->view;
    v.all = {s->wide.all.nodes.data(), s->wide.all.prims.data(), (uint32_t)s->wide.all.nodes.size(), (uint32_t)s->wide.all.prims.size()};
    v.emis = {s->wide.emis.nodes.data(), s->wide.emis.prims.data(), (uint32_t)s->wide.emis.nodes.size(), (uint32_t)s->wide.emis.prims.size()};
    v.shade = s->wide.all.shade.data(); v.shadeEmis = s->wide.emis.shade.data();
    v.uv = s->wide.all.uvs.data(); v.tex = nullptr; v.srgbLut = nullptr;
    v.xforms = s->wide.xforms.data(); v.up = s->wide.up.data(); v.spheres = s->spheres.data(); v.materials = s->materials.data();
    v.objXform = s->wide.objXform.data();
    v.nXforms = (uint32_t)s->wide.xforms.size(); v.nSpheres = nS; v.nMaterials = nM; v.nObjects = nO;
    v.anyEmissive = s->wide.anyEmissive ? 1u : 0u;
    v.emissionFinite = s->wide.emissionFinite ? 1u : 0u;
    v.extent = s->wide.extent;
    for (int a = 0; a < 3; ++a) {     v.emisLo[a] = s->wide.emisLo[a];     v.emisHi[a] = s->wide.emisHi[a]; }
    v.uniqSphereMask = s->wide.uniqSphereMask; v.emisSphereMask = s->wide.emisSphereMask; v.plainSphereMask = s->wide.plainSphereMask;
    v.primOf = s->wide.primOf.data(); v.instBase = s->wide.instBase.data(); v.triFirst = s->wide.triFirst.data();
    s->minv.resize(16 * (size_t)nO);
    for (uint32_t o = 0; o < nO; ++o) inverse4(s->objects[o].m, &s->minv[16 * (size_t)o]);
    s->ref = RefScene{s->spheres.data(), s->materials.data(), s->points.data(), s->tris.data(), s->objects.data(), s->nodes.data(), s->minv.data(),
                      nS, nM, nP, nT, nO, nN};
    return s;
}
void emu_scene_destroy(void* s) { delete (EmuScene*)s; }

// stats: nodes, prims, emissive prims, max depth, xforms
void emu_scene_stats(void* sp, uint32_t* out5) {
    EmuScene* s = (EmuScene*)sp;
    out5[0] = (uint32_t)s->wide.all.nodes.size(); out5[1] = (uint32_t)s->wide.all.prims.size();
    out5[2] = (uint32_t)s->wide.emis.prims.size(); out5[3] = s->wide.all.maxDepth; out5[4] = (uint32_t)s->wide.xforms.size();
}

int emu_wf_render_frames(void* sp, const void* pc208, int W, int H, uint32_t first, uint32_t n, uint32_t stride,
                         uint32_t slots, float* accum, uint64_t* rays2) {
    EmuScene* s = (EmuScene*)sp;
    const WideScene& sc = s->view;
    WfParams P;
    std::memcpy(&P.pc, pc208, sizeof(P.pc));
    P.cp = camera_plane(P.pc.cam);
    P.triVote32 = 8; P.fetchMinIdle = 8; P.sharePrimary = 0; P.connectAfterExtend = 0; P.useTextures = 0; P.continueSample = 0;
    P.W = W; P.H = H; P.row0 = 0; P.rows = (uint32_t)H; P.rowStride = 1; P.stride = stride ? stride : 1;
    size_t cap = (size_t)slots * W * H;
    std::vector<float4> rayO(cap), rayD(cap), atten(cap), total(cap), direct(cap), qO(cap), qL(cap), qA(cap), qE(cap);
    std::vector<Hit> hit(cap);
    PathArrays A{rayO.data(), rayD.data(), atten.data(), total.data(), direct.data(), hit.data(), qO.data(), qL.data(), qA.data(), qE.data()};
    std::vector<uint32_t> q0(cap), q1(cap), qs(cap);
    for (uint32_t done = 0; done < n; done += slots) {
        P.nSlots = std::min(slots, n - done);
        P.firstFrame = first + done * P.stride;
        P.nPaths = P.nSlots * (uint32_t)H * (uint32_t)W;
#pragma omp parallel for
        for (long long p = 0; p < (long long)P.nPaths; ++p) { stage_generate(P, A, (uint32_t)p); q0[p] = (uint32_t)p; }
        uint32_t count = P.nPaths;
        uint32_t* cur = q0.data(); uint32_t* nxt = q1.data();
        for (uint32_t j = 0; j <= P.pc.params.bounceLimit && count; ++j) {
#pragma omp parallel for schedule(dynamic, 256)
            for (long long i = 0; i < (long long)count; ++i) stage_extend(P, sc, s->ref, A, cur[i]);
            if (rays2) rays2[0] += count;
            uint32_t nNext = 0, nShadow = 0;
            std::vector<uint32_t> flags(count);
#pragma omp parallel for schedule(dynamic, 256)
            for (long long i = 0; i < (long long)count; ++i) flags[i] = stage_shade(P, sc, A, cur[i], j);
            for (uint32_t i = 0; i < count; ++i) {
                if (flags[i] & PT_SHADE_CONTINUE) nxt[nNext++] = cur[i];
                if ((flags[i] & PT_SHADE_CONTINUE) && (flags[i] & PT_SHADE_CONNECT)) qs[nShadow++] = cur[i];
            }
#pragma omp parallel for schedule(dynamic, 256)
            for (long long i = 0; i < (long long)nShadow; ++i) stage_connect(P, sc, s->ref, A, qs[i]);
            if (rays2) rays2[1] += 2ull * nShadow;
            std::swap(cur, nxt);
            count = nNext;
        }
#pragma omp parallel for
        for (long long pix = 0; pix < (long long)W * H; ++pix) stage_accumulate(P, A, (float4*)accum, (uint32_t)pix);
    }
    return 0;
}

// closest hits of a ray batch through the wide scene; 48-byte records as b200pt_trace_batch
int emu_wf_trace_batch2(void* sp, const void* pc208, const float* rays, uint64_t n, void* hits48, uint64_t* nContested);
int emu_wf_trace_batch(void* sp, const void* pc208, const float* rays, uint64_t n, void* hits48) { return emu_wf_trace_batch2(sp, pc208, rays, n, hits48, nullptr); }
// nContested: optional counter of rays that took the reference-order re-trace
int emu_wf_trace_batch2(void* sp, const void* pc208, const float* rays, uint64_t n, void* hits48, uint64_t* nContested) {
    EmuScene* s = (EmuScene*)sp;
    const WideScene& sc = s->view;
    ptl::PushConstants pc;
    std::memcpy(&pc, pc208, sizeof(pc));
    struct Rec { float dst; uint32_t kind, obj, tri, mat, front; float n[3], hp[3]; };
    Rec* out = (Rec*)hits48;
#pragma omp parallel for schedule(dynamic, 256)
    for (long long i = 0; i < (long long)n; ++i) {
        Ray r; r.o = mk3(rays[8 * i], rays[8 * i + 1], rays[8 * i + 2]); r.d = mk3(rays[8 * i + 4], rays[8 * i + 5], rays[8 * i + 6]);
        uint32_t contested = 0u;
        Hit h = wide_closest_hit(sc, r, pc.params.sphereCount, pc.params.objectCount, contested);
        if (contested) { h = exact_closest_hit(s->ref, sc, pc.params, r, h.dst); if (nContested) {
#pragma omp atomic
            nContested[0]++; } }
        Rec o; std::memset(&o, 0, sizeof(o));
        o.dst = h.dst; o.obj = 0xffffffffu;
        if (h.id != PT_HIT_MISS) {
            v3 hp, nn; uint32_t mat, obj, prim; bool front;
            hit_surface(sc, r, h, hp, nn, mat, front, obj, prim);
            o.kind = (h.id & PT_HIT_SPHERE) ? 1u : 2u; o.obj = obj; o.tri = prim; o.mat = mat; o.front = front ? 1u : 0u;
            o.n[0] = nn.x; o.n[1] = nn.y; o.n[2] = nn.z; o.hp[0] = hp.x; o.hp[1] = hp.y; o.hp[2] = hp.z;
        }
        out[i] = o;
    }
    return 0;
}

}  // extern "C"
