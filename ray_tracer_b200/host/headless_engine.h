// The part of VulkanEngine that surrounds the compute pass, without a window: the host code a user of the reference
// keeps when the Vulkan dispatch is replaced by libb200pt.so (C ABI, include/b200pt.h). Member names and the order
// of operations follow the reference so that the two can be read side by side:
//   init_image()                src/vk_engine.cpp:1375-1399   -> b200pt_create
//   prepare_storage_buffers()   :680-758 (+ copy_buffer :1401-1444) -> b200pt_upload_scene
//   update_buffer(...)          :1545, :1572, :1603 (ImGui buttons)  -> b200pt_update_*
//   run_compute()               :1623-1676 (camera matrix, counts, push constants, dispatch) -> b200pt_render
//   draw()                      :1782-1812 (sampleLimit gate, _frameNumber / totalSamples bookkeeping)
//   cleanup()                   DeletionQueue, src/vk_engine.h:25-39 -> b200pt_destroy
// Added (not in the reference): render_batch() = b200pt_render_frames, the throughput entry point, and image files
// (PPM of the rgba8 view, PFM of the float image) in place of the swapchain — no graphics stack on this box.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../csrc/pt_layout.h"
#include "scene_builder.h"

struct b200pt_ctx;

namespace pth {

class HeadlessEngine {
   public:
    // defaults of src/vk_engine.h:145-171, 325, 338
    uint32_t width = 1728, height = 1117;
    float cameraAngles[3] = {4.f, 0.f, 0.f};
    ptl::CameraInfo cameraInfo;
    ptl::EnvironmentData environment;
    ptl::RayTracerData rayTracerParams;
    ptl::PushConstants constants;
    int frameNumber = 0;                  // _frameNumber
    uint32_t totalSamples = 0;
    SceneBuilder scene;                   // spheres, rayMaterials, triPoints, triangles, objects, bvhNodes

    HeadlessEngine();
    ~HeadlessEngine() { cleanup(); }

    bool init_image(int device);                          // creates the context (the storage image lives on the device)
    bool load_scene_file(const std::string& path);        // the six buffers from an on-disk scene cache (host_api.cpp)
    bool prepare_storage_buffers();                       // uploads what `scene` holds
    enum Buffer { MATERIALS, SPHERES, OBJECTS };
    bool update_buffer(Buffer which);
    void fill_constants();                                // what run_compute() does before vkCmdPushConstants
    bool run_compute();                                   // one dispatch with `constants.frameCount = frameNumber`
    bool draw();                                          // one iteration of the engine's loop body
    bool render_batch(uint32_t firstFrame, uint32_t nFrames);   // progressive frames firstFrame.. in flight together
    // multi-GPU (one engine per GPU, one host thread each): the engines of a job share an NCCL communicator inside the
    // library; render_batch_split renders this rank's frames of the batch, reduce_accum sums the accumulators onto rank 0,
    // render_frame_interleaved splits ONE frame by interleaved rows and gathers it on rank 0
    bool comm_init(int nranks, int rank, const void* uniqueId128);
    bool render_batch_split(uint32_t firstFrame, uint32_t nFrames);
    bool reduce_accum(int root);
    bool render_frame_interleaved(uint32_t frame, int root);
    bool set_option(const char* key, long long value);
    bool read_image(std::vector<float>& rgba);            // the storage image (progressive blend), float4
    bool read_accum_mean(std::vector<float>& rgba);       // sum of outColor / frames of render_batch
    bool read_rgba8(std::vector<uint8_t>& rgba8);
    bool last_render_ms(float& ms);
    void cleanup();
    const std::string& error() const { return error_; }
    b200pt_ctx* context() { return ctx_; }

    static bool write_ppm(const std::string& path, const uint8_t* rgba8, uint32_t w, uint32_t h);
    static bool write_pfm(const std::string& path, const float* rgba, uint32_t w, uint32_t h);   // rgb, bottom row first

   private:
    b200pt_ctx* ctx_ = nullptr;
    std::string error_;
    bool fail(const char* what);
};

}  // namespace pth
