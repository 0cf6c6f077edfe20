// Host-side scene construction for the compute pass: produces the six storage buffers in the
// reference's layout, with the reference's triangle ORDER and BVH2 node array, because those are
// part of the data contract ("BVH primitive ordering must be bit-exact").
//   read_obj / read_mtl         src/vk_engine.cpp:800-1037, 1060-1167
//   build_bvh and helpers       src/vk_engine.cpp:1169-1337, BoundingBox src/vk_engine.h:81-115
//   cornell_box, default scene  src/vk_engine.cpp:638-758
// The float evaluation order of the split search is the reference's (incl. its rightArea
// indexing quirk, A-Q14) — compile with -ffp-contract=off.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

#include "../csrc/pt_layout.h"

namespace pth {

struct Placement {                  // ImGuiObject, src/vk_engine.h:125-132
    float position[3] = {0.f, 0.f, 0.f};
    float rotation[3] = {0.f, 0.f, 0.f};     // Euler degrees
    float scale[3] = {1.f, 1.f, 1.f};
    uint32_t samplerIndex = 0;
    bool frontOnly = false;
};

class SceneBuilder {
   public:
    std::vector<ptl::Sphere> spheres;
    std::vector<ptl::Material> materials;
    std::vector<ptl::TrianglePoint> points;
    std::vector<ptl::Triangle> triangles;
    std::vector<ptl::RenderObject> objects;
    std::vector<ptl::BVHNode> nodes;
    std::vector<std::string> texturePaths;   // slot -> file, in read_mtl order

    static ptl::Material default_material();
    void push_default_materials();           // white, red, green, light, mirror, dielectric (:689-722)
    bool read_obj(const std::string& path, const Placement& where, int material);
    void read_mtl(const std::string& path);
    // one object from caller-supplied face vertices (3 TrianglePoints per face): the tail of read_obj
    void add_mesh(const ptl::TrianglePoint* faceVertices, size_t nTriangles, const Placement& where, int material);
    void cornell_box(const std::string& assetDir);
    void default_scene(const std::string& assetDir);
    static void object_matrix(const Placement& p, float out16[16]);   // translate*rotX*rotY*rotZ*scale (:807-811)

    // Optional replacement for build_bvh: the signature of b200pt_build_bvh (include/b200pt.h), so the CUDA library's
    // entry point can be plugged in without this library linking against it. Objects with fewer triangles than
    // `minTriangles` keep the host build. Both produce the same bytes.
    typedef int (*BvhBuildFn)(int device, const void* triPoints, uint32_t nPoints, void* triangles, uint32_t firstTri, uint32_t count,
                              uint32_t rootIndex, void* nodesOut, uint32_t* nNodesOut, double* deviceMs);
    void set_bvh_builder(BvhBuildFn fn, int device, uint32_t minTriangles) { bvhFn_ = fn; bvhDevice_ = device; bvhMinTris_ = minTriangles; }
    double bvhDeviceMs = 0.0;                // accumulated device time of plugged-in builds
    uint32_t bvhPluggedBuilds = 0;

   private:
    struct C3 { float x, y, z; };
    std::vector<C3> centroids_;
    std::map<std::string, uint32_t> loadedObjects_;
    std::map<std::string, uint32_t> loadedMaterials_;
    uint32_t nodesUsed_ = 0;
    BvhBuildFn bvhFn_ = nullptr;
    int bvhDevice_ = 0;
    uint32_t bvhMinTris_ = 0;

    uint32_t push_face(const ptl::TrianglePoint* v, bool frontOnly);
    void finish_object(size_t firstTri, const Placement& where, uint32_t material, bool smooth, uint32_t sampler);
    void build_bvh(uint32_t firstTri, uint32_t count);
    void fit_bounds(uint32_t node);
    float best_split(const ptl::BVHNode& node, int& axis, float& pos) const;
};

}  // namespace pth
