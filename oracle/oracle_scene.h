// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// CPU restatement of the reference's host-side scene construction:
//   read_obj            src/vk_engine.cpp:800-1037
//   read_mtl            src/vk_engine.cpp:1060-1167   (data only; no Vulkan image upload)
//   build_bvh …         src/vk_engine.cpp:1169-1337   + BoundingBox src/vk_engine.h:81-115
//   cornell_box         src/vk_engine.cpp:638-678
//   prepare_storage_buffers (materials + default objects) src/vk_engine.cpp:680-758
//   run_compute camera/params block  src/vk_engine.cpp:1631-1661
#pragma once
#include <string>
#include <unordered_map>
#include <vector>
#include "oracle_layout.h"
#include "oracle_math.h"

namespace orc {

struct ImGuiObject {  // src/vk_engine.h:125-132
    std::string name;
    om::vec3 position = om::vec3(0.f);
    om::vec3 rotation = om::vec3(0.f);
    om::vec3 scale = om::vec3(1.f);
    uint32_t samplerIndex = 0;
    bool frontOnly = false;
};

struct BoundingBox {  // src/vk_engine.h:81-115
    om::vec4 bounds[2] = {om::vec4(1e30f), om::vec4(-1e30f)};
    void grow(const ol::TrianglePoint& p);
    void grow(const BoundingBox& b);
    float surfaceArea() const;
};

class SceneBuilder {
   public:
    std::vector<ol::Sphere> spheres;
    std::vector<ol::RayMaterial> rayMaterials;
    std::vector<ol::TrianglePoint> triPoints;
    std::vector<ol::Triangle> triangles;
    std::vector<ol::RenderObject> objects;
    std::vector<ol::BVHNode> bvhNodes;
    std::vector<om::vec3> centroids;
    std::vector<std::string> textureFiles;  // slot i -> path, in read_mtl order
    uint32_t texturesUsed = 0;
    uint32_t nodesUsed = 0;
    std::unordered_map<std::string, int> loadedObjects;
    std::unordered_map<std::string, int> loadedMaterials;

    static ol::RayMaterial default_material();
    // returns false if the file cannot be opened (the reference silently returns, :834)
    bool read_obj(const std::string& filePath, const ImGuiObject& imGuiObj, int material);
    void read_mtl(const std::string& filePath);
    // the part of read_obj after parsing: one object over caller-supplied face-vertex data
    // (3 TrianglePoints per face), used for procedural scenes
    void add_mesh(const ol::TrianglePoint* pts, size_t nTris, const ImGuiObject& imGuiObj, int material);
    void build_bvh(int size, int triIndex);
    void push_default_materials();   // src/vk_engine.cpp:689-722
    void cornell_box(const std::string& assetDir);             // :638-678
    void prepare_default_scene(const std::string& assetDir);   // :680-758 (HEAD state: two cubes + cornell box)

   private:
    void update_bvh_bounds(uint32_t index);
    void subdivide_bvh(uint32_t index, uint32_t depth);
    float find_bvh_split_plane(const ol::BVHNode& node, int& axis, float& splitPos);
};

om::mat4 object_matrix(const ImGuiObject& o);  // translate*rotX*rotY*rotZ*scale, src/vk_engine.cpp:807-811

// run_compute's parameter block (src/vk_engine.cpp:1631-1661) with the reference's
// default member initialisers (src/vk_engine.h:145-171,325).
ol::PushConstants default_push_constants(int W, int H, uint32_t sphereCount, uint32_t objectCount);
void set_camera_angles(ol::PushConstants& pc, float ax, float ay, float az);

}  // namespace orc
