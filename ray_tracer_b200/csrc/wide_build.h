// Device build of the wide BVH of all instance triangles (wide_build.cu); the host keeps the small parts
// (bvh8_build.h: transforms, parent links of the reference BVH2, the instance list, the emissive-only BVH).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <vector>

#include "bvh8_build.h"

namespace pt {

struct WideDeviceInputs {                    // reference-layout buffers already on the device (b200pt_upload_scene)
    const ptl::TrianglePoint* points; const ptl::Triangle* tris; const ptl::RenderObject* objects; const ptl::Material* materials;
};
struct WideDeviceResult {                    // cudaMalloc'ed, owned by the caller. In/out: buffers of an earlier build with room for
    void* nodes = nullptr; void* prims = nullptr; void* shade = nullptr; void* uv = nullptr;   // as many instances are written again,
    size_t capacityInstances = 0;                                                               // otherwise freed and replaced
    uint32_t nNodes = 0, nPrims = 0, maxDepth = 0;
    uint32_t nDropped = 0;                   // instances left out because a world-space vertex was not finite
    std::vector<uint32_t> levelCount;        // wide nodes per level, root first: level k is the index range after levels < k
    void* nodeBox = nullptr;                 // refit scratch: 6 floats per wide node (owned like the other buffers)
    double deviceMs = 0.0;
};
struct WideBuildCache {                      // scratch arena kept between builds (released by the owner's destructor)
    static constexpr size_t kKeepBytes = (size_t)1 << 30;      // larger arenas are freed after the build
    void* p = nullptr;
    size_t bytes = 0;
    WideBuildCache() = default;
    WideBuildCache(const WideBuildCache&) = delete;
    WideBuildCache& operator=(const WideBuildCache&) = delete;
    ~WideBuildCache() { if (p) cudaFree(p); }
};
// host: result of build_wide_scene(..., buildAll = false) (instObject / instTri / leafOf / up / objXform filled).
// Returns nullptr or an error text.
const char* wide_build_device(cudaStream_t stream, const WideDeviceInputs& in, const WideSceneHost& host, WideDeviceResult& out,
                              WideBuildCache* cache = nullptr);

// Refit after a transform edit (b200pt_update_objects): same tree, same triangle order; every wide node's child boxes are
// recomputed bottom-up from the triangles' new world-space boxes (level by level, deepest first) and re-quantised, and the
// triangle records get their objects' current transform ids. `objects` / `objXform`: device arrays already updated.
// Only valid for a tree built by wide_build_device with nDropped == 0. Returns nullptr or an error text.
const char* wide_refit_device(cudaStream_t stream, const WideDeviceInputs& in, const uint32_t* objXform, WideDeviceResult& tree, float* deviceMs);

}  // namespace pt
