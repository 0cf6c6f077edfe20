// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// Harness around the REFERENCE'S OWN HOST CODE for the scene side of the hot path: OBJ/MTL ingest, the binned-SAH
// BVH2 builder, the scene presets and the push-constant block of run_compute(). oracle/ref_from_engine.py slices
// those definitions verbatim out of /root/reference/src/vk_engine.{h,cpp} into oracle/_ref/vk_engine_*.inc
// (git-ignored). The application cannot be built here (Vulkan, SDL2, ImGui headers absent), so this file declares
// a `VulkanEngine` holding only the members the sliced functions touch, and inert stand-ins for the Vulkan / VMA
// entry points they call (the stand-ins record what the GPU side would have received: the 208 bytes given to
// vkCmdPushConstants, the group counts of vkCmdDispatch, the texture files handed to load_images_from_file).
// GLM is the reference's vendored 0.9.9.7, used in place. Built by `make -C oracle ref` into
// oracle/_ref/libvk_engine_scene.so. No arithmetic is restated here.
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <deque>
#include <fstream>
#include <functional>
#include <iostream>
#include <sstream>
#include <string>
#include <unordered_map>
#include <vector>

#include <glm/glm.hpp>
#include <glm/gtx/transform.hpp>

// ---- inert Vulkan / VMA surface (only what the sliced text names) ----
typedef void* VkFence; typedef void* VkCommandPool; typedef void* VkCommandBuffer; typedef void* VkImageView;
typedef void* VkImage; typedef void* VmaAllocation; typedef void* VkBuffer; typedef void* VkDevice;
typedef void* VmaAllocator; typedef void* VkPipeline; typedef void* VkPipelineLayout; typedef void* VkDescriptorSet;
typedef void* VkSemaphore; typedef void* VkQueue;
typedef uint32_t VkPipelineStageFlags; typedef uint32_t VkBufferUsageFlags; typedef int VkResult;
struct VkExtent2D { uint32_t width, height; };
struct VkCommandBufferBeginInfo {};
struct VkImageViewCreateInfo {};
struct VkSubmitInfo {
    uint32_t waitSemaphoreCount; const VkSemaphore* pWaitSemaphores; const VkPipelineStageFlags* pWaitDstStageMask;
    uint32_t signalSemaphoreCount; const VkSemaphore* pSignalSemaphores;
};
enum { VK_PIPELINE_STAGE_COMPUTE_SHADER_BIT = 0x800, VK_PIPELINE_BIND_POINT_COMPUTE = 1, VK_SHADER_STAGE_COMPUTE_BIT = 0x20,
       VK_BUFFER_USAGE_STORAGE_BUFFER_BIT = 0x20, VK_FORMAT_R8G8B8A8_SRGB = 43, VK_IMAGE_ASPECT_COLOR_BIT = 1 };
#define VK_NULL_HANDLE nullptr
struct AllocatedBuffer { VkBuffer buffer; VmaAllocation allocation; };   // src/vk_types.h:10-13
struct AllocatedImage { VkImage image; VmaAllocation allocation; };     // src/vk_types.h:15-18

namespace capture {
static uint8_t pushConstants[256];
static uint32_t pushSize = 0;
static uint32_t groups[3] = {0, 0, 0};
static std::vector<std::string> textureFiles;
}  // namespace capture

inline VkResult vkBeginCommandBuffer(VkCommandBuffer, const VkCommandBufferBeginInfo*) { return 0; }
inline void vkCmdBindPipeline(VkCommandBuffer, int, VkPipeline) {}
inline void vkCmdBindDescriptorSets(VkCommandBuffer, int, VkPipelineLayout, uint32_t, uint32_t, const VkDescriptorSet*, uint32_t, const uint32_t*) {}
inline void vkCmdPushConstants(VkCommandBuffer, VkPipelineLayout, int, uint32_t, uint32_t size, const void* p) {
    capture::pushSize = size;
    std::memcpy(capture::pushConstants, p, size < 256 ? size : 256);
}
inline void vkCmdDispatch(VkCommandBuffer, uint32_t x, uint32_t y, uint32_t z) { capture::groups[0] = x; capture::groups[1] = y; capture::groups[2] = z; }
inline VkResult vkEndCommandBuffer(VkCommandBuffer) { return 0; }
inline VkResult vkQueueSubmit(VkQueue, uint32_t, const VkSubmitInfo*, void*) { return 0; }
inline void vkDestroyImageView(VkDevice, VkImageView, const void*) {}
inline VkResult vkCreateImageView(VkDevice, const VkImageViewCreateInfo*, const void*, VkImageView*) { return 0; }
inline void vmaDestroyImage(VmaAllocator, VkImage, VmaAllocation) {}
namespace vkinit {
inline VkCommandBufferBeginInfo commandBufferBeginInfo() { return {}; }
inline VkSubmitInfo submitInfo(VkCommandBuffer*) { return {}; }
inline VkImageViewCreateInfo imageViewCreateInfo(int, VkImage, int) { return {}; }
}  // namespace vkinit

#include "_ref/vk_engine_types.inc"

// The members of class VulkanEngine (src/vk_engine.h:205-352) that the sliced functions use, same names and types.
class VulkanEngine {
public:
    void read_obj(std::string filePath, ImGuiObject imGui, int material);
    void calculate_binormal(int v1, int v2, int v3, glm::vec3& tangent, glm::vec3& binormal);
    void read_mtl(std::string filePath);
    void build_bvh(int size, int triIndex, BoundingBox scene);
    void update_bvh_bounds(uint index);
    void subdivide_bvh(uint intex, uint depth, BVHStats& stats, BoundingBox scene);
    float find_bvh_split_plane(BVHNode& node, int& axis, float& splitPos, BoundingBox scene);
    float scene_interior_cost(BoundingBox node, BoundingBox scene);
    void prepare_storage_buffers();
    void copy_buffer(size_t, AllocatedBuffer&, VkBufferUsageFlags, void*) {}   // upload: nothing to do on the host
    void run_compute();
    void cornell_box();

    DeletionQueue deletionQueue;
    VkDevice device = nullptr;
    VkCommandBuffer computeCmdBuffer = nullptr;
    std::vector<Sphere> spheres;
    std::vector<RayMaterial> rayMaterials;
    std::vector<Texture> textures = std::vector<Texture>(MAX_TEXTURES);   // init_image creates MAX_TEXTURES slots (:1390-1398)
    std::vector<TrianglePoint> triPoints;
    std::vector<Triangle> triangles;
    std::vector<RenderObject> objects;
    std::vector<ImGuiObject> imGuiObjects;
    std::vector<glm::vec3> centroids;
    uint texturesUsed = 0;
    std::vector<BVHNode> bvhNodes;
    BoundingBox scene;
    uint nodesUsed = 0;
    std::unordered_map<std::string, int> loadedObjects;
    std::unordered_map<std::string, int> loadedMaterials;
    VkDescriptorSet computeSet = nullptr;
    AllocatedBuffer sphereBuffer{}, triPointBuffer{}, materialBuffer{}, triangleBuffer{}, objectBuffer{}, bvhBuffer{};
    VkPipelineLayout computePipeLayout = nullptr;
    VkPipeline computePipeline = nullptr;
    VkSemaphore computeSemaphore = nullptr, graphicsSemaphore = nullptr;
    VkQueue computeQueue = nullptr;
    VmaAllocator allocator = nullptr;
    int _frameNumber{0};
    float cameraAngles[3] = {4.f, 0.f, 0.f};
    RayTracerData rayTracerParams;
    CameraInfo cameraInfo;
    EnvironmentData environment;
    PushConstants constants;
    VkExtent2D _windowExtent{1728, 1117};
};

namespace vkutil {
inline bool load_images_from_file(VulkanEngine&, const char* files[], AllocatedImage*[], int size) {
    for (int i = 0; i < size; ++i) capture::textureFiles.push_back(files[i]);
    return true;
}
}  // namespace vkutil

using namespace std;
#define VK_CHECK(x) do { VkResult err = x; if (err) { abort(); } } while (0)   // src/vk_engine.cpp:20-27

#include "_ref/vk_engine_funcs.inc"

// ------------------------------------------------------------------------------------------------------
namespace {
struct Quiet {   // the sliced functions print build statistics to std::cout
    std::streambuf* old;
    std::ostringstream sink;
    Quiet() : old(std::cout.rdbuf(sink.rdbuf())) {}
    ~Quiet() { std::cout.rdbuf(old); }
};
struct InDir {   // scene presets name their assets relative to the build directory ("../assets/...")
    char prev[4096];
    bool ok;
    explicit InDir(const char* d) { ok = getcwd(prev, sizeof prev) && (!d || !*d || chdir(d) == 0); }
    ~InDir() { if (ok) { int r = chdir(prev); (void)r; } }
};
ImGuiObject mk(const float* pos, const float* rot, const float* scl, uint32_t sampler, int frontOnly) {
    ImGuiObject o;
    o.position = glm::vec3(pos[0], pos[1], pos[2]);
    o.rotation = glm::vec3(rot[0], rot[1], rot[2]);
    o.scale = glm::vec3(scl[0], scl[1], scl[2]);
    o.samplerIndex = sampler;
    o.frontOnly = frontOnly != 0;
    return o;
}
}  // namespace

extern "C" {

void* refeng_create() { capture::textureFiles.clear(); return new VulkanEngine(); }
void refeng_destroy(void* e) { delete (VulkanEngine*)e; }

// VulkanEngine::prepare_storage_buffers() run from `cwd` (a directory from which "../assets" is the asset folder)
int refeng_prepare_storage_buffers(void* e, const char* cwd) {
    Quiet q; InDir d(cwd);
    if (!d.ok) return 1;
    try { ((VulkanEngine*)e)->prepare_storage_buffers(); } catch (const std::exception&) { return 2; }
    return 0;
}
int refeng_cornell_box(void* e, const char* cwd) {
    Quiet q; InDir d(cwd);
    if (!d.ok) return 1;
    try { ((VulkanEngine*)e)->cornell_box(); } catch (const std::exception&) { return 2; }
    return 0;
}
int refeng_read_obj(void* e, const char* path, const float* pos, const float* rot, const float* scl, uint32_t sampler,
                    int frontOnly, int material) {
    Quiet q;
    try { ((VulkanEngine*)e)->read_obj(path, mk(pos, rot, scl, sampler, frontOnly), material); } catch (const std::exception&) { return 2; }
    return 0;
}
void refeng_add_material(void* e, const void* mat64) {
    RayMaterial m; std::memcpy((void*)&m, mat64, 64); ((VulkanEngine*)e)->rayMaterials.push_back(m);
}
// counts: spheres, materials, points, triangles, objects, nodes, texturesUsed
void refeng_counts(void* e, uint32_t* out7) {
    VulkanEngine* s = (VulkanEngine*)e;
    out7[0] = s->spheres.size(); out7[1] = s->rayMaterials.size(); out7[2] = s->triPoints.size();
    out7[3] = s->triangles.size(); out7[4] = s->objects.size(); out7[5] = s->bvhNodes.size(); out7[6] = s->texturesUsed;
}
void refeng_copy(void* e, int which, void* dst) {
    VulkanEngine* s = (VulkanEngine*)e;
    static_assert(sizeof(Sphere) == 32 && sizeof(RayMaterial) == 64 && sizeof(TrianglePoint) == 32 && sizeof(Triangle) == 48 &&
                  sizeof(RenderObject) == 80 && sizeof(BVHNode) == 32 && sizeof(PushConstants) == 208, "reference layouts (A-Q1)");
    switch (which) {
        case 0: std::memcpy(dst, s->spheres.data(), s->spheres.size() * 32); break;
        case 1: std::memcpy(dst, (void*)s->rayMaterials.data(), s->rayMaterials.size() * 64); break;
        case 2: std::memcpy(dst, s->triPoints.data(), s->triPoints.size() * 32); break;
        case 3: std::memcpy(dst, s->triangles.data(), s->triangles.size() * 48); break;
        case 4: std::memcpy(dst, s->objects.data(), s->objects.size() * 80); break;
        case 5: std::memcpy(dst, s->bvhNodes.data(), s->bvhNodes.size() * 32); break;
    }
}
int refeng_texture_path(uint32_t slot, char* dst, int cap) {
    if (slot >= capture::textureFiles.size()) return -1;
    const std::string& p = capture::textureFiles[slot];
    if (dst && cap > 0) { std::strncpy(dst, p.c_str(), cap - 1); dst[cap - 1] = 0; }
    return (int)p.size();
}
// VulkanEngine::run_compute() with the engine's default camera / environment / parameters, the given window extent,
// camera angles and frame number; returns what vkCmdPushConstants and vkCmdDispatch received.
int refeng_run_compute(void* e, uint32_t W, uint32_t H, const float* angles3, int frameNumber, void* out208, uint32_t* groups3) {
    VulkanEngine* s = (VulkanEngine*)e;
    s->_windowExtent = VkExtent2D{W, H};
    for (int i = 0; i < 3; ++i) s->cameraAngles[i] = angles3[i];
    s->_frameNumber = frameNumber;
    std::memset((void*)&s->constants, 0, sizeof(s->constants));
    s->run_compute();
    if (capture::pushSize != 208) return 1;
    std::memcpy(out208, capture::pushConstants, 208);
    for (int i = 0; i < 3; ++i) groups3[i] = capture::groups[i];
    return 0;
}

}  // extern "C"
