/*
 * b200pt.h — C ABI of the B200 path-tracing backend (libb200pt.so).
 *
 * Drop-in for the compute pass of CobaltCrabby/ray_tracer: everything the reference binds and
 * dispatches for shaders/raytrace.comp has one entry point here. Buffers are taken in the
 * reference's own layouts, byte for byte; re-layout for the GPU is internal.
 *
 *   reference interface (file:line)                               entry point
 *   -----------------------------------------------------------   --------------------------------
 *   init_image: W x H storage image            vk_engine.cpp:1375-1399   b200pt_create / b200pt_resize
 *   prepare_storage_buffers + copy_buffer x6   vk_engine.cpp:680-758,1401-1444   b200pt_upload_scene
 *   update_buffer(materials)  (UI button)      vk_engine.cpp:1545,1446-1475      b200pt_update_materials
 *   update_buffer(spheres)                     vk_engine.cpp:1572                b200pt_update_spheres
 *   update_buffer(objects)                     vk_engine.cpp:1594-1603           b200pt_update_objects
 *   load_images_from_file -> TextureBuffer[64] vk_engine.cpp:1155, vk_textures.cpp:103-200  b200pt_upload_texture
 *   run_compute: push constants + vkCmdDispatch  vk_engine.cpp:1623-1676         b200pt_render
 *   draw(): progressive frame loop             vk_engine.cpp:1774-1815           b200pt_render_frames
 *   run_graphics: sampling the compute image   vk_engine.cpp:1678-1737           b200pt_read_* / b200pt_device_*_ptr
 *   VK_CHECK (print + abort)                   vk_engine.cpp:20-27               int status + b200pt_last_error
 *
 * Conventions: every call returns 0 on success, non-zero on failure (never aborts); the text of
 * the last failure is b200pt_last_error(ctx). Host input buffers are copied during the call and
 * may be freed afterwards. One context per GPU; calls on one context must be serialised by the
 * caller (the reference is single-threaded with one queue). Work is asynchronous on the
 * context's CUDA stream; b200pt_sync or any b200pt_read_* waits for it.
 * There is no CPU fallback: without a CUDA device b200pt_create fails.
 */
#ifndef B200PT_H
#define B200PT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200pt_ctx b200pt_ctx;

/* sizes of the reference's buffer elements (std140; src/vk_engine.h:49-189) */
#define B200PT_SPHERE_BYTES 32
#define B200PT_MATERIAL_BYTES 64
#define B200PT_TRIPOINT_BYTES 32
#define B200PT_TRIANGLE_BYTES 48
#define B200PT_OBJECT_BYTES 80
#define B200PT_BVHNODE_BYTES 32
#define B200PT_PUSHCONSTANT_BYTES 208
#define B200PT_MAX_TEXTURES 64 /* src/vk_engine.h:204 */

/* status codes */
#define B200PT_OK 0
#define B200PT_ERR_ARG 1
#define B200PT_ERR_CUDA 2
#define B200PT_ERR_STATE 3

int b200pt_version(void);
/* non-context failures (e.g. b200pt_create) */
const char* b200pt_global_error(void);

/* Context = device buffers + W x H outputs on CUDA device `device`. */
int b200pt_create(int device, int width, int height, b200pt_ctx** out);
int b200pt_destroy(b200pt_ctx* ctx);
int b200pt_resize(b200pt_ctx* ctx, int width, int height);
const char* b200pt_last_error(b200pt_ctx* ctx);

/* Run on a caller-owned CUDA stream (cudaStream_t as void*; NULL = the context's own stream). */
int b200pt_set_stream(b200pt_ctx* ctx, void* cuda_stream);
int b200pt_sync(b200pt_ctx* ctx);

/* The six storage buffers, reference layout, element counts (not bytes). */
int b200pt_upload_scene(b200pt_ctx* ctx, const void* spheres, uint32_t n_spheres, const void* materials,
                        uint32_t n_materials, const void* tri_points, uint32_t n_points, const void* triangles,
                        uint32_t n_triangles, const void* objects, uint32_t n_objects, const void* bvh_nodes,
                        uint32_t n_nodes);
/* In-place updates, same element counts as uploaded (the reference's update_buffer never resizes).
 * b200pt_update_objects with nothing but transform matrices changed REFITS the internal tree on the device (same
 * topology, boxes recomputed; option "refit_on_update", default 1; "last_update_was_refit" tells); any other change, or a
 * tree built on the host, re-lays the scene out. Images equal those of a fresh upload either way. */
int b200pt_update_materials(b200pt_ctx* ctx, const void* materials, uint32_t n);
int b200pt_update_spheres(b200pt_ctx* ctx, const void* spheres, uint32_t n);
int b200pt_update_objects(b200pt_ctx* ctx, const void* objects, uint32_t n);
/* RGBA8 (sRGB-encoded) texels for TextureBuffer[slot]; stored for the texture extension (the HEAD
 * shader never samples: SURVEY.md §0). */
int b200pt_upload_texture(b200pt_ctx* ctx, uint32_t slot, const void* rgba8, uint32_t width, uint32_t height);

/* One reference dispatch: main() of raytrace.comp for every pixel with the 208-byte push-constant
 * block; read-modify-writes the float storage image (the reference's image is rgba8; see
 * b200pt_read_rgba8). Honours progressive, singleRender, raysPerPixel, sampleLimit, bounceLimit,
 * debug, frameCount. Without a debug heat-map the frame is traced by the wavefront kernels (several samples per
 * dispatch as consecutive passes that hand the pixel's RNG state over) and finished by the same main() tail; the
 * debug modes and option "pipeline" = 0 use the one-thread-per-pixel kernel in the reference's traversal order,
 * which is also the only one that fills b200pt_read_stats. */
int b200pt_render(b200pt_ctx* ctx, const void* push_constants_208);

/* Throughput entry: progressive frames first_frame .. first_frame+n_frames-1 (frameCount = f,
 * raysPerPixel samples each); accum.rgb += outColor_f (non-finite -> magenta, raytrace.comp:579-581),
 * accum.a += 1. Frames are summed in frame order, so the result does not depend on how many frames
 * are in flight. frame_stride > 1 renders first_frame, first_frame+stride, ... (multi-GPU sample split). */
int b200pt_render_frames(b200pt_ctx* ctx, const void* push_constants_208, uint32_t first_frame, uint32_t n_frames,
                         uint32_t frame_stride);
/* Row band [row_begin, row_end) only (interactive multi-GPU tile split); other rows untouched. */
int b200pt_render_frames_rows(b200pt_ctx* ctx, const void* push_constants_208, uint32_t first_frame,
                              uint32_t n_frames, uint32_t frame_stride, uint32_t row_begin, uint32_t row_end);
/* Rows row_begin, row_begin + row_stride, ... below row_end (interleaved split of a frame: rows of different cost are
 * spread evenly over the GPUs); other rows untouched. */
int b200pt_render_frames_rows_strided(b200pt_ctx* ctx, const void* push_constants_208, uint32_t first_frame,
                                      uint32_t n_frames, uint32_t frame_stride, uint32_t row_begin, uint32_t row_end,
                                      uint32_t row_stride);
int b200pt_clear(b200pt_ctx* ctx); /* zero accumulator, storage image and counters */

/* Outputs. Accumulator and image are W*H float4, row 0 first (row 0 is the top of the screen). */
int b200pt_read_accum_f32(b200pt_ctx* ctx, float* host_rgba);
int b200pt_read_image_f32(b200pt_ctx* ctx, float* host_rgba);
int b200pt_write_image_f32(b200pt_ctx* ctx, const float* host_rgba);
int b200pt_read_rgba8(b200pt_ctx* ctx, uint8_t* host_rgba8); /* unorm8 round-to-nearest of the image */
/* Per-pixel stats[0], stats[1] of the last b200pt_render (box tests, triangle tests; raytrace.comp:310,338). */
int b200pt_read_stats(b200pt_ctx* ctx, float* host_w_h_2);
/* Headless device buffers (where Vulkan external-memory interop would attach). */
int b200pt_device_accum_ptr(b200pt_ctx* ctx, void** device_ptr);
int b200pt_device_image_ptr(b200pt_ctx* ctx, void** device_ptr);
/* Accumulate into a caller-owned device buffer of W*H float4 (e.g. a torch tensor handed to NCCL);
 * NULL returns to the context's own buffer. */
int b200pt_attach_accum(b200pt_ctx* ctx, void* device_ptr);

/* Closest hits of caller-supplied rays (8 floats each: origin.xyz, pad, dir.xyz, pad) through
 * calculateIntersections (raytrace.comp:276-353); 48-byte records:
 * {f32 dst; u32 kind (0 miss,1 sphere,2 triangle); u32 objectHitIndex; u32 triHitIndex;
 *  u32 materialIndex; u32 frontFace; f32 normal[3]; f32 hitPoint[3]}. */
int b200pt_trace_batch(b200pt_ctx* ctx, const void* push_constants_208, const float* host_rays, uint64_t n,
                       void* host_hits48);

/* The same with rays and records already in device memory (n * 32 and n * 48 bytes); asynchronous on the context's
 * stream. The throughput form: the soup roofline case of the bench times this call. */
int b200pt_trace_batch_device(b200pt_ctx* ctx, const void* push_constants_208, const void* device_rays, uint64_t n,
                              void* device_hits48);

/* b200pt_trace_batch_device for callers that cannot allocate. No counterpart in the reference: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.
 * Same inputs, same records, same kernels; the scratch memory (ray slots, hits, queues, counters) is a file-scope
 * static device array of B200PT_STATIC_MAX_RAYS * 60 + 256 bytes (3.75 MiB, held by every GPU that loads the library
 * until exit) instead of buffers grown with cudaMalloc. The call makes no cudaMalloc / cudaFree and calls no allocating
 * library routine; the scene must have been uploaded before (that is where allocation happens). n above
 * B200PT_STATIC_MAX_RAYS is rejected with B200PT_ERR_ARG and a message. Asynchronous on the context's stream.
 * All calls on one GPU share the array: calls that overlap on the GPU (two streams, two contexts, two threads) race. */
#define B200PT_STATIC_MAX_RAYS 65536u
int b200pt_trace_batch_device_static(b200pt_ctx* ctx, const void* push_constants_208, const void* device_rays, uint64_t n,
                                     void* device_hits48);

/* GPU build of one object's BVH2 — replaces VulkanEngine::build_bvh(size, triIndex) with update_bvh_bounds,
 * subdivide_bvh and find_bvh_split_plane (src/vk_engine.cpp:1169-1337): 20-bin SAH, back-swapping partition,
 * depth-first node numbering. Bit-identical to the reference's sequential build for finite coordinates: same node
 * bounds, same node order, same final triangle order.
 *   tri_points      the engine's triPoints (32-byte TrianglePoint), n_points entries
 *   triangles       the engine's triangles (48-byte Triangle); [first_tri, first_tri + count) is this object's
 *                   range and is REORDERED IN PLACE as the reference's partition does
 *   root_index      bvhNodes.size() before the call (= RenderObject.bvhIndex); children indices are absolute
 *   nodes_out       room for 2 * count - 1 32-byte BVHNodes; *n_nodes_out receives the number written, which the
 *                   caller appends to bvhNodes
 *   device_ms       optional: device time of the build without the host<->device copies
 * No context is needed (the scene is not uploaded yet). Errors: b200pt_global_error(). */
int b200pt_build_bvh(int device, const void* tri_points, uint32_t n_points, void* triangles, uint32_t first_tri,
                     uint32_t count, uint32_t root_index, void* nodes_out, uint32_t* n_nodes_out, double* device_ms);

/* ---- multi-GPU. The reference drives ONE GPU through one queue (src/vk_engine.cpp:115-117); a
 * maintainer who wants N GPUs behind the same dispatch point (run_compute, src/vk_engine.cpp:1623-1676) creates one
 * context per GPU (one thread or one process each, scene uploaded to every one) and joins them with a communicator.
 * NCCL is bound at run time (libnccl.so.2); these entries fail with a message when it is absent. Collective calls
 * (reduce_accum, render_frame_interleaved) must be made by every rank; they are asynchronous on each context's stream. */
#define B200PT_UNIQUE_ID_BYTES 128
int b200pt_comm_unique_id(void* out128);   /* rank 0 makes it and hands it to the other ranks (shared memory, a file, MPI ...) */
int b200pt_comm_init_rank(b200pt_ctx* ctx, int nranks, int rank, const void* unique_id128);
int b200pt_comm_destroy(b200pt_ctx* ctx);
int b200pt_comm_info(b200pt_ctx* ctx, int* nranks, int* rank);      /* 1, 0 without a communicator */
/* Sample split of draw()'s progressive loop: of the frames first_frame .. first_frame + n_frames - 1 this rank renders
 * first_frame + rank, + nranks, ... (independent seeds: raytrace.comp:562-564) into its own accumulator. */
int b200pt_render_frames_split(b200pt_ctx* ctx, const void* push_constants_208, uint32_t first_frame, uint32_t n_frames);
/* Sum of all ranks' accumulators (W*H float4: rgb sums and the frame count in a) onto `root`, in place, over NVLink;
 * zero_others != 0 clears the accumulators of the other ranks for the next batch. */
int b200pt_reduce_accum(b200pt_ctx* ctx, int root, int zero_others);
/* One progressive frame split by INTERLEAVED rows (rank r renders rows r, r + nranks, ...: a pixel's seed depends only
 * on x, y, W and the frame, so any subset of pixels is exact, and interleaving balances rows of unequal cost). Every
 * rank accumulates its rows into its own accumulator; `root` then receives the other ranks' rows into its accumulator. */
int b200pt_render_frame_interleaved(b200pt_ctx* ctx, const void* push_constants_208, uint32_t frame, int root);

/* Options: "pipeline" 0 = reference-layout per-pixel kernel, 1 = wavefront (default for
 * render_frames/trace_batch); "frames_in_flight" (1..64); "profile_stages" 0/1; "count_work" 0/1: the traversal
 * kernels count their node visits and triangle tests (counters [7..10]; a few percent slower);
 * "textures" 0/1: texture extension (albedo / metalness maps; off = the HEAD shader, which never samples);
 * "storage_rgba8" 0/1: b200pt_render stores the image at unorm8 precision, as the reference's rgba8 image does;
 * "overlap_connect" 0/1/2: the NEE (connect) kernels of a bounce run on a second stream beside the next extension —
 * never / for batches of up to 8 frames (default) / always; "share_primary" 0/1, "trav_tri_vote" 1..32, "trav_fetch_idle" 1..32 (k_extend), "occl_tri_vote", "occl_fetch_idle" (k_occlude): scheduling of the wavefront kernels
 * (results do not depend on them); "wide_build" 1 (default) / 0: the internal wide BVH of all triangles is built on
 * the device / on the host at b200pt_upload_scene and b200pt_update_* (read before the upload; images do not depend
 * on it); "refit_on_update" 1 (default) / 0; "exact_order" 1 (default) / 0: rays on which two candidates lie
 * closer together than the rounding error of their distances (coincident surfaces; about 6 per thousand in the C4 stand-in)
 * are re-traced in the reference's own traversal order, which makes every result the reference's bit for bit; with it the
 * traversal kernels also accept candidate hits optimistically and separate kernels confirm each ray's final hit against the
 * reference's BVH2 (failures join the re-trace). 0 skips the re-trace and keeps that confirmation inside the traversal
 * loop (1.5 ms less latency per 1-spp frame on one GPU, 2.1 ms less on eight, about 1e-4 of the paths then depend on the traversal order). Read-only: "last_scene_build_us", "last_wide_device_us",
 * "last_update_was_refit". Unknown keys fail. */
int b200pt_set_option(b200pt_ctx* ctx, const char* key, int64_t value);
int b200pt_get_option(b200pt_ctx* ctx, const char* key, int64_t* value);

/* Counters since the last b200pt_clear / b200pt_reset_counters, 16 x u64:
 * [0] kernel launches, [1] paths traced;
 * algorithmic work — what the reference's trace() does for those paths: [2] extension rays (one
 * calculateIntersections per path per bounce), [3] connect items (diffuse bounces whose next-event query reached the
 * connect stage; the reference traces three rays for each, raytrace.comp:443,447,453);
 * traced work — what the kernels walked: [4] rays of the closest-hit launches (one primary ray per PIXEL when a batch
 * shares the primary hit between its frame slots), [5] occlusion walks, [6] exact pdf rays (evaluated lazily);
 * with option "count_work": [7] node visits and [8] triangle tests of the closest-hit kernel, [9] and [10] the same
 * for the occlusion kernel (wide layout: 80 B per node, 64 B per triangle record); [11] extension rays and [12]
 * light-sample queries that were CONTESTED (two candidates closer together than the rounding slack of their distances:
 * coincident surfaces), or whose optimistically accepted hit the reference would not have tested, and were re-traced in
 * the reference's own order over the reference's own buffers, so that the result is the reference's bit for bit;
 * [13..15] reserved. */
int b200pt_read_counters(b200pt_ctx* ctx, uint64_t* out16);
int b200pt_reset_counters(b200pt_ctx* ctx);
/* With option "profile_stages" = 1 the wavefront records a CUDA event between consecutive launches; this
 * returns the accumulated device milliseconds and launch counts per stage, 8 entries each
 * (0 generate, 1 extend, 2 shade, 3 connect-prep, 4 accumulate, 5 occlude, 6 connect-finish, 7 confirmation of the
 * traversal kernels' hits + reference-order re-trace of contested rays)
 * and optionally resets them. */
int b200pt_read_stage_times(b200pt_ctx* ctx, double* ms8, uint64_t* launches8, int reset);
/* Device time of the last render call in milliseconds (CUDA events on the context's stream). */
int b200pt_last_render_ms(b200pt_ctx* ctx, float* ms);

#ifdef __cplusplus
}
#endif
#endif /* B200PT_H */
