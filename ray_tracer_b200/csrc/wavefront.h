// Host-side interface of the wavefront pipeline (wavefront.cu) used by api.cu.
#pragma once
#include "pt_context.h"

namespace pt {
enum { WF_CHANGED_ALL = 0, WF_CHANGED_MATERIALS = 1, WF_CHANGED_SPHERES = 2, WF_CHANGED_TRANSFORMS = 3 };
// re-layout of the uploaded scene (or, when only materials / spheres changed and no object switched between emitting and
// not emitting, just the flags that depend on them; when only object MATRICES changed, a refit of the device-built tree:
// same topology, child boxes recomputed bottom-up, the small emissive tree rebuilt); 0 or a B200PT_ERR_* with c->error set
int wavefront_build_scene(b200pt_ctx* c, int changed = WF_CHANGED_ALL);
void wavefront_destroy(b200pt_ctx* c);
void wavefront_set_textures(b200pt_ctx* c);      // texture table / sRGB table pointers changed
void wavefront_invalidate_frame_state(b200pt_ctx* c);
// dispatchImage != NULL (n == 1, whole image): instead of adding the frame to the accumulator, finish it as ONE literal
// dispatch into that storage image (progressive blend etc., raytrace.comp:568-593)
// rowStride > 1: rows row0, row0 + rowStride, ... below row1 (interleaved split of one frame over several GPUs)
int wavefront_render_frames(b200pt_ctx* c, const ptl::PushConstants& pc, uint32_t first, uint32_t n, uint32_t stride,
                            uint32_t row0, uint32_t row1, float4* dispatchImage = nullptr, uint32_t rowStride = 1);
int wavefront_trace_batch(b200pt_ctx* c, const ptl::PushConstants& pc, const float* dRays, uint64_t n, HitRecord48* dHits);
// the same launches with the scratch in a static device array (no allocation; n <= B200PT_STATIC_MAX_RAYS)
int wavefront_trace_batch_static(b200pt_ctx* c, const ptl::PushConstants& pc, const float* dRays, uint64_t n, HitRecord48* dHits);
void wavefront_fold_counters(b200pt_ctx* c);
void wavefront_resolve_stage_times(b200pt_ctx* c);     // stream idle
void wavefront_stage_times(b200pt_ctx* c, double* ms8, uint64_t* launches8, bool reset);
cudaError_t launch_image_to_rgba8(b200pt_ctx* c, const float4* image, uchar4* out, size_t n);
}  // namespace pt
