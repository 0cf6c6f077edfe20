// Internal state behind the opaque b200pt_ctx of include/b200pt.h.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "pt_reftrace.h"

namespace pt {

struct HitRecord48 {          // b200pt_trace_batch output record
    float dst;
    uint32_t kind, objectHitIndex, triHitIndex, materialIndex, frontFace;
    float normal[3];
    float hitPoint[3];
};
static_assert(sizeof(HitRecord48) == 48, "HitRecord48");

struct WavefrontState;        // wavefront.cu

struct DeviceBuffer {
    void* p = nullptr;
    size_t bytes = 0;
};

}  // namespace pt

struct b200pt_comm;           // comm.cu: NCCL communicator joining the contexts of a multi-GPU job

struct b200pt_ctx {
    int device = 0;
    int W = 0, H = 0;
    cudaStream_t ownStream = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t evStart = nullptr, evStop = nullptr;
    std::string error;

    // reference-layout scene (device) + host copies kept for re-layout on update_*
    pt::RefScene ref{};
    pt::DeviceBuffer dSpheres, dMaterials, dPoints, dTris, dObjects, dNodes, dMinv;
    std::vector<ptl::Sphere> hSpheres;
    std::vector<ptl::Material> hMaterials;
    std::vector<ptl::TrianglePoint> hPoints;
    std::vector<ptl::Triangle> hTris;
    std::vector<ptl::RenderObject> hObjects;
    std::vector<ptl::BVHNode> hNodes;
    bool sceneLoaded = false;

    // outputs
    pt::DeviceBuffer dAccumOwn, dImage, dStats;
    float4* accum = nullptr;      // own buffer or caller-attached
    // TextureBuffer[64] (texture extension): texels in a cudaArray behind a point-filtered texture object
    struct Texture { cudaArray_t array = nullptr; cudaTextureObject_t object = 0; uint32_t w = 0, h = 0; } textures[64];
    pt::DeviceBuffer dTexTable, dSrgbLut;      // 64 x TexEntry; 256 floats
    int useTextures = 0;
    int storageRgba8 = 0;         // b200pt_render keeps the storage image at 8-bit precision like the reference's rgba8 image

    // options
    int pipeline = 1;
    int framesInFlight = 4;
    int profileStages = 0;
    int countWork = 0;            // traversal kernels count their node visits / triangle tests (counters [7..10])
    int triVote32 = 12;           // see WfParams; swept on the C4 stand-in: 8 / 10 / 11 / 12 / 13 / 16 -> 300.9 / 305.2 / 305.7 / 305.6 / 304.6 / 298.3 Mpaths/s
    int fetchMinIdle = 7;         // 4 / 6 / 7 / 8 / 10 / 12 -> 298.2 / 305.5 / 306.2 / 305.6 / 303.7 / 294.2 (with triVote32 = 12)
    int triVoteOccl32 = 12, fetchMinIdleOccl = 7;     // k_occlude's own pair
    int overlapConnect = 1;       // connect chain of bounce j beside extend(j+1) on a second stream: 0 never, 1 for batches of up to 8 frames, 2 always
    int sharePrimary = 1;         // trace the primary ray once per pixel and share the hit between frame slots (A-Q4)
    int wideBuild = 1;            // 1: the wide BVH of all triangles is built on the device (wide_build.cu), 0: on the host
    int exactOrder = 1;           // contested rays are re-traced in the reference's order (k_exact_*); 0 skips that (results then depend on the traversal order for ~1e-4 of paths)
    int refitOnUpdate = 1;        // b200pt_update_objects with only matrices changed refits the device-built tree instead of rebuilding it
    int lastRefit = 0;            // the last re-layout was such a refit
    double lastSceneBuildMs = 0.0, lastWideDeviceMs = 0.0;   // wall time of the last re-layout; device part of it

    // counters (host side; device-side ray counters are folded in by the wavefront)
    uint64_t counters[16] = {0};
    float lastMs = 0.f;

    pt::WavefrontState* wf = nullptr;
    b200pt_comm* comm = nullptr;
};

namespace pt {
void set_global_error(const char* what);      // api.cu: text behind b200pt_global_error()
// megakernel.cu
cudaError_t launch_ref_dispatch(b200pt_ctx* c, const ptl::PushConstants& pc);
cudaError_t launch_ref_frames(b200pt_ctx* c, const ptl::PushConstants& pc, uint32_t first, uint32_t n, uint32_t stride,
                              uint32_t row0, uint32_t row1);
cudaError_t launch_ref_trace_batch(b200pt_ctx* c, const ptl::PushConstants& pc, const float* dRays, uint64_t n,
                                   HitRecord48* dHits);
}  // namespace pt
