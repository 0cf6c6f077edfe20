// Reference-layout kernels: one thread per pixel evaluating pt_reftrace.h (a literal restatement
// of shaders/raytrace.comp main()/trace()/calculateIntersections()). Serves b200pt_render (every
// switch of RayTracerData incl. the debug heat-maps, whose box/triangle counts are defined by the
// reference's own BVH2) and is the on-device cross-check of the wavefront pipeline
// (option "pipeline" = 0). Compiled with -fmad=false: see pt_math.h.
#include "pt_context.h"

namespace pt {

__global__ void __launch_bounds__(64) k_ref_dispatch(RefScene sc, const __grid_constant__ ptl::PushConstants c_pc, CameraPlane cp, int W, int H, float4* image, float2* stats,
                                                     int storeUnorm8) {
    // 8x8 tiles as the reference dispatches them (local_size 8x8, vkCmdDispatch(ceil(W/8), ceil(H/8)));
    // the reference has no bounds guard although it over-covers (A-Q7) — guarded here.
    uint32_t gx = blockIdx.x * 8 + (threadIdx.x & 7);
    uint32_t gy = blockIdx.y * 8 + (threadIdx.x >> 3);
    if (gx >= (uint32_t)W || gy >= (uint32_t)H) return;
    size_t p = (size_t)gy * W + gx;
    RefPixel px = ref_pixel(sc, c_pc, cp, gx, gy, W, H);
    float4 old = image[p];
    v3 fin = ref_resolve(c_pc, px, mk3(old.x, old.y, old.z));
    if (storeUnorm8) fin = mk3(unorm8_round(fin.x), unorm8_round(fin.y), unorm8_round(fin.z));
    image[p] = make_float4(fin.x, fin.y, fin.z, 1.f);
    stats[p] = make_float2(px.stats0, px.stats1);
}

__global__ void __launch_bounds__(64) k_ref_frames(RefScene sc, const __grid_constant__ ptl::PushConstants c_pc, CameraPlane cp, int W, int H, uint32_t first, uint32_t n,
                                                   uint32_t stride, uint32_t row0, uint32_t row1, float4* accum) {
    uint32_t gx = blockIdx.x * 8 + (threadIdx.x & 7);
    uint32_t gy = row0 + blockIdx.y * 8 + (threadIdx.x >> 3);
    if (gx >= (uint32_t)W || gy >= row1) return;
    size_t p = (size_t)gy * W + gx;
    float4 a = accum[p];
    ptl::PushConstants pc = c_pc;
    for (uint32_t k = 0; k < n; ++k) {
        pc.frameCount = first + k * stride;
        RefPixel px = ref_pixel(sc, pc, cp, gx, gy, W, H);
        v3 c = px.outColor;
        if (isnan3(c) || isinf3(c)) c = mk3(1.f, 0.f, 1.f);      // raytrace.comp:579-581
        a.x += c.x; a.y += c.y; a.z += c.z; a.w += 1.f;
    }
    accum[p] = a;
}

__global__ void k_ref_trace_batch(RefScene sc, const __grid_constant__ ptl::PushConstants c_pc, const float4* rays, uint64_t n, HitRecord48* out) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 o = rays[2 * i], d = rays[2 * i + 1];
    Ray r; r.o = mk3(o.x, o.y, o.z); r.d = mk3(d.x, d.y, d.z);
    RefHit h = ref_closest_hit(sc, c_pc.params, r, nullptr);
    HitRecord48 rec;
    rec.dst = h.dst;
    rec.kind = h.kind;
    rec.objectHitIndex = h.kind == 2 ? h.objectHitIndex : 0xffffffffu;
    rec.triHitIndex = h.kind ? h.triHitIndex : 0u;
    rec.materialIndex = h.materialIndex;
    rec.frontFace = (h.kind && h.front) ? 1u : 0u;
    rec.normal[0] = h.kind ? h.n.x : 0.f; rec.normal[1] = h.kind ? h.n.y : 0.f; rec.normal[2] = h.kind ? h.n.z : 0.f;
    rec.hitPoint[0] = h.kind ? h.hp.x : 0.f; rec.hitPoint[1] = h.kind ? h.hp.y : 0.f; rec.hitPoint[2] = h.kind ? h.hp.z : 0.f;
    out[i] = rec;
}

cudaError_t launch_ref_dispatch(b200pt_ctx* c, const ptl::PushConstants& pc) {
    dim3 grid((c->W + 7) / 8, (c->H + 7) / 8);
    k_ref_dispatch<<<grid, 64, 0, c->stream>>>(c->ref, pc, camera_plane(pc.cam), c->W, c->H, (float4*)c->dImage.p,
                                               (float2*)c->dStats.p, c->storageRgba8);
    c->counters[0] += 1;
    uint64_t samples = pc.params.singleRender ? pc.params.sampleLimit : pc.params.raysPerPixel;
    c->counters[1] += (uint64_t)c->W * c->H * samples;
    return cudaGetLastError();
}

cudaError_t launch_ref_frames(b200pt_ctx* c, const ptl::PushConstants& pc, uint32_t first, uint32_t n, uint32_t stride,
                              uint32_t row0, uint32_t row1) {
    if (row1 <= row0 || n == 0) return cudaSuccess;
    dim3 grid((c->W + 7) / 8, (row1 - row0 + 7) / 8);
    k_ref_frames<<<grid, 64, 0, c->stream>>>(c->ref, pc, camera_plane(pc.cam), c->W, c->H, first, n, stride, row0, row1, c->accum);
    c->counters[0] += 1;
    uint64_t samples = pc.params.singleRender ? pc.params.sampleLimit : pc.params.raysPerPixel;
    c->counters[1] += (uint64_t)c->W * (row1 - row0) * n * samples;
    return cudaGetLastError();
}

cudaError_t launch_ref_trace_batch(b200pt_ctx* c, const ptl::PushConstants& pc, const float* dRays, uint64_t n,
                                   HitRecord48* dHits) {
    if (n == 0) return cudaSuccess;
    k_ref_trace_batch<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(c->ref, pc, (const float4*)dRays, n, dHits);
    c->counters[0] += 1;
    c->counters[2] += n;
    return cudaGetLastError();
}

}  // namespace pt
