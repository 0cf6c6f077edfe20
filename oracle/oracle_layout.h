// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// Byte layouts of the six storage buffers and the push-constant block, restated from
// the reference's host structs (src/vk_engine.h:49-123,145-189) and their std140 twins
// in the shader (shaders/raytrace.comp:9-15,25-110,150-155). Sizes/offsets are the ones
// measured in SURVEY.md Appendix A-Q1. Padding bytes are pinned to zero.
#pragma once
#include <cstddef>
#include <cstdint>

namespace ol {

struct Sphere {           // src/vk_engine.h:49-53, raytrace.comp:52-56
    float position[3];    // @0
    float radius;         // @12
    uint32_t materialIndex;  // @16
    uint32_t _pad[3];
};
static_assert(sizeof(Sphere) == 32 && offsetof(Sphere, radius) == 12 && offsetof(Sphere, materialIndex) == 16, "Sphere");

struct Triangle {         // src/vk_engine.h:55-62, raytrace.comp:58-65
    uint32_t v0, v1, v2;  // @0,4,8
    uint32_t frontOnly;   // @12
    float binormal[3];    // @16 (never written by the reference, calculate_binormal is a no-op)
    uint32_t _pad0;
    float tangent[3];     // @32
    uint32_t _pad1;
};
static_assert(sizeof(Triangle) == 48 && offsetof(Triangle, binormal) == 16 && offsetof(Triangle, tangent) == 32, "Triangle");

struct TrianglePoint {    // src/vk_engine.h:64-67, raytrace.comp:67-70
    float position[4];    // xyz + uv.x
    float normal[4];      // xyz + uv.y
};
static_assert(sizeof(TrianglePoint) == 32, "TrianglePoint");

struct RayMaterial {      // src/vk_engine.h:69-79, raytrace.comp:25-35
    float albedo[3];      // @0
    uint32_t _pad0;
    float emissionColor[3];  // @16
    float emissionStrength;  // @28
    float reflectance;    // @32
    float ior;            // @36
    int32_t albedoIndex;  // @40
    int32_t metalnessIndex;  // @44
    int32_t alphaIndex;   // @48
    int32_t bumpIndex;    // @52
    uint32_t _pad1[2];
};
static_assert(sizeof(RayMaterial) == 64 && offsetof(RayMaterial, emissionColor) == 16 &&
              offsetof(RayMaterial, emissionStrength) == 28 && offsetof(RayMaterial, reflectance) == 32 &&
              offsetof(RayMaterial, ior) == 36 && offsetof(RayMaterial, albedoIndex) == 40 &&
              offsetof(RayMaterial, bumpIndex) == 52, "RayMaterial");

struct RenderObject {     // src/vk_engine.h:117-123, raytrace.comp:43-49
    float transformMatrix[16];  // column-major @0
    uint32_t smoothShade;  // @64
    uint32_t bvhIndex;     // @68
    uint32_t materialIndex;  // @72
    uint32_t samplerIndex;   // @76
};
static_assert(sizeof(RenderObject) == 80 && offsetof(RenderObject, bvhIndex) == 68, "RenderObject");

struct BVHNode {          // src/vk_engine.h:185-189, raytrace.comp:85-89
    float boundsX[2], boundsY[2], boundsZ[2];  // (min,max) pairs @0,8,16
    uint32_t index;       // @24
    uint32_t triCount;    // @28
};
static_assert(sizeof(BVHNode) == 32 && offsetof(BVHNode, index) == 24, "BVHNode");

struct CameraInfo {       // src/vk_engine.h:145-151
    float cameraRotation[16];  // @0
    float pos[3];         // @64
    float nearPlane;      // @76
    float aspectRatio;    // @80
    float fov;            // @84
    uint32_t _pad[2];
};
static_assert(sizeof(CameraInfo) == 96 && offsetof(CameraInfo, nearPlane) == 76 && offsetof(CameraInfo, fov) == 84, "CameraInfo");

struct EnvironmentData {  // src/vk_engine.h:153-158
    float horizonColor[4];  // @0  w = sun focus
    float zenithColor[4];   // @16 w = sun intensity
    float groundColor[3];   // @32
    uint32_t _pad;
    float lightDir[4];      // @48 w = environment on
};
static_assert(sizeof(EnvironmentData) == 64 && offsetof(EnvironmentData, lightDir) == 48, "EnvironmentData");

struct RayTracerData {    // src/vk_engine.h:160-171 — bools are 1 byte in a 4-byte slot
    uint8_t progressive;  uint8_t _p0[3];   // @0
    uint8_t singleRender; uint8_t _p1[3];   // @4
    int32_t debug;        // @8
    uint32_t raysPerPixel;  // @12
    uint32_t bounceLimit;   // @16
    uint32_t sphereCount;   // @20
    uint32_t objectCount;   // @24
    uint32_t triangleCap;   // @28
    uint32_t boxCap;        // @32
    uint32_t sampleLimit;   // @36
};
static_assert(sizeof(RayTracerData) == 40 && offsetof(RayTracerData, sampleLimit) == 36, "RayTracerData");

struct PushConstants {    // src/vk_engine.h:173-178
    CameraInfo camInfo;           // @0
    EnvironmentData environment;  // @96
    RayTracerData rayTraceParams; // @160
    uint32_t frameCount;          // @200
    uint32_t _pad;
};
static_assert(sizeof(PushConstants) == 208 && offsetof(PushConstants, environment) == 96 &&
              offsetof(PushConstants, rayTraceParams) == 160 && offsetof(PushConstants, frameCount) == 200, "PushConstants");

}  // namespace ol
