// Data contract of the compute pass: the six storage buffers and the 208-byte parameter
// block exactly as the reference uploads them (host structs src/vk_engine.h:49-123,145-189;
// std140 twins shaders/raytrace.comp:9-15,25-110,150-155). The backend takes these bytes
// verbatim through the C ABI (include/b200pt.h) and re-lays them out internally.
#pragma once
#include <stddef.h>
#include <stdint.h>

namespace ptl {

struct Sphere {            // 32 B  (src/vk_engine.h:49-53)
    float px, py, pz, radius;
    uint32_t materialIndex, pad0, pad1, pad2;
};
struct Triangle {          // 48 B  (src/vk_engine.h:55-62); binormal/tangent are never written
    uint32_t v0, v1, v2, frontOnly;
    float binormal[4];
    float tangent[4];
};
struct TrianglePoint {     // 32 B  (src/vk_engine.h:64-67): position.w = u, normal.w = v
    float px, py, pz, u;
    float nx, ny, nz, v;
};
struct Material {          // 64 B  (src/vk_engine.h:69-79)
    float albedo[3]; uint32_t pad0;
    float emissionColor[3];
    float emissionStrength;
    float reflectance;
    float ior;
    int32_t albedoIndex, metalnessIndex, alphaIndex, bumpIndex;
    uint32_t pad1[2];
};
struct RenderObject {      // 80 B  (src/vk_engine.h:117-123), column-major mat4
    float m[16];
    uint32_t smoothShade, bvhIndex, materialIndex, samplerIndex;
};
struct BVHNode {           // 32 B  (src/vk_engine.h:185-189)
    float minX, maxX, minY, maxY, minZ, maxZ;
    uint32_t index, triCount;
};
struct CameraInfo {        // 96 B  (src/vk_engine.h:145-151)
    float rot[16];
    float pos[3];
    float nearPlane, aspectRatio, fov;
    uint32_t pad[2];
};
struct EnvironmentData {   // 64 B  (src/vk_engine.h:153-158)
    float horizon[4];      // w = sun focus
    float zenith[4];       // w = sun intensity
    float ground[3]; uint32_t pad;
    float lightDir[4];     // w = environment on
};
struct RayTracerData {     // 40 B  (src/vk_engine.h:160-171): C++ bools are 1 byte in a 4-byte slot
    uint8_t progressive, p0[3];
    uint8_t singleRender, p1[3];
    int32_t debug;
    uint32_t raysPerPixel, bounceLimit, sphereCount, objectCount, triangleCap, boxCap, sampleLimit;
};
struct PushConstants {     // 208 B (src/vk_engine.h:173-178)
    CameraInfo cam;
    EnvironmentData env;
    RayTracerData params;
    uint32_t frameCount;
    uint32_t pad;
};

static_assert(sizeof(Sphere) == 32 && offsetof(Sphere, materialIndex) == 16, "Sphere");
static_assert(sizeof(Triangle) == 48 && offsetof(Triangle, frontOnly) == 12, "Triangle");
static_assert(sizeof(TrianglePoint) == 32, "TrianglePoint");
static_assert(sizeof(Material) == 64 && offsetof(Material, emissionStrength) == 28 &&
              offsetof(Material, reflectance) == 32 && offsetof(Material, ior) == 36, "Material");
static_assert(sizeof(RenderObject) == 80 && offsetof(RenderObject, bvhIndex) == 68, "RenderObject");
static_assert(sizeof(BVHNode) == 32 && offsetof(BVHNode, index) == 24, "BVHNode");
static_assert(sizeof(CameraInfo) == 96 && offsetof(CameraInfo, nearPlane) == 76, "CameraInfo");
static_assert(sizeof(EnvironmentData) == 64 && offsetof(EnvironmentData, lightDir) == 48, "EnvironmentData");
static_assert(sizeof(RayTracerData) == 40 && offsetof(RayTracerData, sampleLimit) == 36, "RayTracerData");
static_assert(sizeof(PushConstants) == 208 && offsetof(PushConstants, env) == 96 &&
              offsetof(PushConstants, params) == 160 && offsetof(PushConstants, frameCount) == 200, "PushConstants");

}  // namespace ptl
