// Internal acceleration layout: an 8-wide BVH over WORLD-space boxes of every (object instance,
// triangle) pair, with child boxes quantised to 8 bits per plane relative to the node
// (compressed-wide-BVH encoding after Ylitie, Karras, Laine 2017): 80-byte nodes read as five
// 16-byte vectors, 64-byte triangle records read as four. It replaces the reference's
// "linear loop over objects, one BVH2 per object, 32-byte nodes + 144 bytes of vertices per test"
// (shaders/raytrace.comp:289-350) — the boxes only decide WHICH triangles are tested; the
// triangle test itself is the reference's arithmetic in object space (pt_shade.h: tri_test), so
// accepted hits are bit-identical. Boxes are conservative (padded, rounded outward).
#pragma once
#include "pt_shade.h"

namespace pt {

struct Bvh8Node {              // 80 B
    float px, py, pz;          // quantisation origin (node box minimum)
    uint8_t ex, ey, ez;        // biased exponents: plane = p + q * 2^(e-127)
    uint8_t imask;             // bit s set: slot s holds an inner node
    uint32_t childBase;        // first inner child (children contiguous, ordered by slot)
    uint32_t primBase;         // first triangle record of this node's leaf children
    uint8_t meta[8];           // inner: 001|24+slot ; leaf: unary count (1,3,7)<<5 | offset ; empty: 0
    uint8_t qlox[8], qloy[8], qloz[8];
    uint8_t qhix[8], qhiy[8], qhiz[8];
};
static_assert(sizeof(Bvh8Node) == 80, "Bvh8Node");

// Triangle record for intersection: edges and face normal are those of triangleIntersection()
// (raytrace.comp:229-232), evaluated once in object space.
struct PrimRecord {            // 64 B
    float p0x, p0y, p0z; uint32_t objectIndex;
    float e1x, e1y, e1z; uint32_t triIndexFlags;   // bit31 frontOnly, bit30 emissive material, bits 29..26 leaf flags, low 26 bits triangle index
    float e2x, e2y, e2z; uint32_t xformId;
    float nx, ny, nz;    uint32_t leafNode;        // reference BVH2 leaf holding this triangle (index into UpNode[])
};
static_assert(sizeof(PrimRecord) == 64, "PrimRecord");
#define PT_PRIM_FRONTONLY 0x80000000u
#define PT_PRIM_EMISSIVE 0x40000000u
#define PT_PRIM_LEAF_FLAT_X 0x20000000u    /* the reference BVH2 leaf holding the triangle has minX == maxX */
#define PT_PRIM_LEAF_FLAT_Y 0x10000000u
#define PT_PRIM_LEAF_FLAT_Z 0x08000000u
#define PT_PRIM_LEAF_IS_ROOT 0x04000000u   /* that leaf is the object's root: its box is never tested */
#define PT_PRIM_TRIMASK 0x03ffffffu

// Shading record of the same triangle (vertex normals; raytrace.comp:259)
struct ShadeRecord {           // 48 B
    float n0x, n0y, n0z; uint32_t materialIndex;
    float n1x, n1y, n1z; uint32_t objectIndex;
    float n2x, n2y, n2z; uint32_t triIndex;
};
static_assert(sizeof(ShadeRecord) == 48, "ShadeRecord");

// TEXTURE EXTENSION (the HEAD shader never samples — semantics are this project's).
// Per-triangle texture coordinates (TrianglePoint.position.w / normal.w, raytrace.comp:249-251), parallel to the
// triangle records; read only for hits on materials that name a texture while option "textures" is on.
struct UvRecord {              // 32 B
    float u0, v0, u1, v1, u2, v2;
    uint32_t samplerIndex;     // RenderObject.samplerIndex: 0 NEAREST/REPEAT, else NEAREST/CLAMP_TO_EDGE
    uint32_t degenerate;       // two vertices share their uv: the shader substitutes (0.5, 0.5) (raytrace.comp:254-256)
};
static_assert(sizeof(UvRecord) == 32, "UvRecord");
struct TexEntry {              // 16 B: TextureBuffer[slot]
    unsigned long long object; // cudaTextureObject_t over a cudaArray of uchar4 (point filter, unnormalised coordinates); 0 = empty
    uint32_t w, h;
};
static_assert(sizeof(TexEntry) == 16, "TexEntry");

struct Xform { float M[16]; float Minv[16]; };   // per distinct object matrix

// The reference's own BVH2 (src/vk_engine.h:185-189), re-linked child -> parent. The reference
// only tests a triangle when its slab test (raytrace.comp:263-274) passed for every node between
// the object's root and the triangle's leaf; a ray that grazes one of those boxes can miss a
// triangle it geometrically hits. To return the reference's hit and not merely the geometrically
// closest one, a candidate found through the wide BVH is confirmed by re-running those slab tests
// leaf-to-root (pt_traverse.h: ref_reachable).
struct UpNode {                // 32 B
    float minX, maxX, minY, maxY, minZ, maxZ;
    uint32_t parent;           // 0xffffffff for an object's root (whose own box the shader never tests)
    uint32_t pad;
};
static_assert(sizeof(UpNode) == 32, "UpNode");

struct Bvh8View {
    const Bvh8Node* nodes;
    const PrimRecord* prims;
    uint32_t nNodes, nPrims;
};

struct WideScene {
    Bvh8View all;              // every instance triangle
    Bvh8View emis;             // triangles of objects whose material has emissionStrength != 0
    const ShadeRecord* shade;  // parallel to all.prims
    const ShadeRecord* shadeEmis;   // parallel to emis.prims
    const UvRecord* uv;        // parallel to all.prims (texture extension)
    const TexEntry* tex;       // 64 slots
    const float* srgbLut;      // 256 entries: sRGB-encoded byte -> linear float
    const Xform* xforms;
    const UpNode* up;          // parallel to the reference's bvhNodes
    const ptl::Sphere* spheres;
    const ptl::Material* materials;
    const uint32_t* objXform;  // object -> xformId
    // (object, reference triangle index) -> index of its record in all.prims, for hits found by the reference-order
    // re-trace of contested rays: primOf[instBase[object] + triangle - triFirst[object]]
    const uint32_t* primOf; const uint32_t* instBase; const uint32_t* triFirst;
    uint32_t nXforms, nSpheres, nMaterials, nObjects;
    uint32_t anyEmissive;      // any emissive sphere or triangle in the scene
    uint32_t emissionFinite;   // every material's emissionColor * emissionStrength is finite
    float extent;              // largest |coordinate| of the scene's world-space boxes and spheres (contest window, pt_traverse.h)
    float emisLo[3], emisHi[3];   // bounding box of all emissive triangles (grown; see ray_meets_box)
    // Sphere slot masks (valid when nSpheres <= 32). Slots with bit-identical (position, radius) give identical
    // test results, and every loop of the reference keeps the FIRST of equal candidates, so only the lowest
    // slot of each identical class is kept (the ten zero spheres of spheres.resize(10) collapse to one test).
    uint32_t uniqSphereMask;   // closest-hit loop: first slot of each class
    uint32_t emisSphereMask;   // first slot of each class among slots with an emissive material
    uint32_t plainSphereMask;  // ... among slots with a non-emissive material
};

// Traversal stack: at most one sibling group per level of the wide tree (<= PT_BVH8_MAX_DEPTH,
// checked at upload) plus postponed triangle groups, which are only pushed while the stack is
// below PT_BVH8_POSTPONE_LIMIT entries: LIMIT + 1 + MAX_DEPTH < STACK.
#define PT_BVH8_STACK 48
#define PT_BVH8_MAX_DEPTH 30
#ifndef PT_BVH8_POSTPONE_LIMIT
#define PT_BVH8_POSTPONE_LIMIT 16
#endif
#define PT_BVH8_SHARED_STACK 20     /* entries per ray kept in shared memory by the persistent kernels */

}  // namespace pt
