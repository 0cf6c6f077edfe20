// Traversal of the wide BVH (pt_bvh8.h): closest hit and bounded occlusion. Replaces the node
// loop of calculateIntersections() (shaders/raytrace.comp:289-350). Boxes are tested in world
// space with fused multiply-adds (they only have to be conservative); every triangle whose box is
// reached is tested with the reference's own arithmetic on the ray transformed into the
// triangle's object space (raytrace.comp:292-294, 227-248), so dst/u/v of an accepted hit are the
// reference's values bit for bit.
#pragma once
#include "pt_bvh8.h"

namespace pt {

#define PT_HIT_MISS 0xffffffffu
#define PT_HIT_SPHERE 0x80000000u
#define PT_HIT_FRONT 0x40000000u
#define PT_HIT_IDMASK 0x3fffffffu

struct Hit {
    float dst;
    uint32_t id;        // PT_HIT_MISS | PT_HIT_SPHERE|front|sphere index | front|prim index
    float u, v;
};

struct BoxRay {         // world-space ray prepared for slab tests
    v3 o, rd;           // origin, 1/direction (zero components nudged: boxes need not see exact axis-parallel rays)
    uint32_t octinv4;   // (7 - octant) replicated in four bytes
};
PT_HD float nudge(float d) { return fabsf(d) > 1e-20f ? d : (as_uint(d) >> 31 ? -1e-20f : 1e-20f); }
PT_HD BoxRay make_box_ray(const Ray& r) {
    BoxRay b;
    b.o = r.o;
    float dx = nudge(r.d.x), dy = nudge(r.d.y), dz = nudge(r.d.z);
    b.rd = mk3(1.f / dx, 1.f / dy, 1.f / dz);
    uint32_t oct = (dx < 0.f ? 4u : 0u) | (dy < 0.f ? 2u : 0u) | (dz < 0.f ? 1u : 0u);
    b.octinv4 = (7u - oct) * 0x01010101u;
    return b;
}

PT_HD uint32_t ld_u32(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
PT_HD int highest_bit(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return 31 - __clz(x);
#else
    return 31 - __builtin_clz(x);
#endif
}
PT_HD int popcount32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}

// Slab-test the 8 children of a node; returns hit bits: inner children in bits 24..31 ordered by
// traversal priority for this ray's octant, leaf triangles in bits 0..23 (offset from primBase).
PT_HD uint32_t bvh8_node_hits(const Bvh8Node& n, const BoxRay& br, float tmax) {
    float sx = as_float((uint32_t)n.ex << 23) * br.rd.x;
    float sy = as_float((uint32_t)n.ey << 23) * br.rd.y;
    float sz = as_float((uint32_t)n.ez << 23) * br.rd.z;
    float ox = (n.px - br.o.x) * br.rd.x;
    float oy = (n.py - br.o.y) * br.rd.y;
    float oz = (n.pz - br.o.z) * br.rd.z;
    const uint8_t* nx = br.rd.x < 0.f ? n.qhix : n.qlox;
    const uint8_t* fx = br.rd.x < 0.f ? n.qlox : n.qhix;
    const uint8_t* ny = br.rd.y < 0.f ? n.qhiy : n.qloy;
    const uint8_t* fy = br.rd.y < 0.f ? n.qloy : n.qhiy;
    const uint8_t* nz = br.rd.z < 0.f ? n.qhiz : n.qloz;
    const uint8_t* fz = br.rd.z < 0.f ? n.qloz : n.qhiz;
    uint32_t oct = br.octinv4 & 7u;
    uint32_t hits = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        uint32_t meta = n.meta[i];
        if (meta == 0) continue;
        float t0x = pt_fma((float)nx[i], sx, ox), t1x = pt_fma((float)fx[i], sx, ox);
        float t0y = pt_fma((float)ny[i], sy, oy), t1y = pt_fma((float)fy[i], sy, oy);
        float t0z = pt_fma((float)nz[i], sz, oz), t1z = pt_fma((float)fz[i], sz, oz);
        float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.f));
        float tf = fminf(fminf(t1x, t1y), fminf(t1z, tmax));
        if (tn <= tf) {
            bool inner = (meta & 0x18u) == 0x18u && (meta >> 5) == 1u;
            uint32_t bit = inner ? ((meta ^ oct) & 0x1fu) : (meta & 0x1fu);
            hits |= (meta >> 5) << bit;
        }
    }
    return hits;
}

struct ObjRayCache { uint32_t xform; Ray r; v3 inv; };

PT_HD const Ray& object_ray(ObjRayCache& c, const Xform* xforms, uint32_t xformId, const Ray& world) {
    if (c.xform != xformId) {
        const float* Minv = xforms[xformId].Minv;
        c.r.d = mul_dir(Minv, world.d);       // raytrace.comp:292 (direction not renormalised: t stays world-parametric)
        c.r.o = mul_point(Minv, world.o);     // raytrace.comp:293
        c.inv = mk3(1.f / c.r.d.x, 1.f / c.r.d.y, 1.f / c.r.d.z);   // raytrace.comp:294
        c.xform = xformId;
    }
    return c.r;
}

// Would the reference have tested the triangle in BVH2 leaf `leaf`? It does iff boxIntersection()
// (raytrace.comp:263-274) reports a hit for every node from the leaf up to, not including, the
// object's root (:324-348; the root's own box is never tested, :301-307). The walk stops early once
// the object-space hit point is inside a box with a margin far above the slab test's rounding
// error: every ancestor contains that box, so their tests pass a fortiori. Axes on which the leaf is
// flat (axis-aligned triangles) are exempt from the margin: there tMin == tMax is the plane distance
// itself and rounding is monotone in the box coordinate.
PT_HD bool ref_reachable(const UpNode* __restrict__ up, uint32_t leaf, const ObjRayCache& c, float dst) {
    UpNode n = up[leaf];
    if (n.parent == 0xffffffffu) return true;
    v3 h = c.r.o + c.r.d * dst;
    float m = 2e-5f * (fmaxf(fmaxf(fabsf(h.x), fabsf(h.y)), fabsf(h.z)) + fmaxf(fmaxf(fabsf(c.r.o.x), fabsf(c.r.o.y)), fabsf(c.r.o.z))) + 1e-30f;
    v3 dh = h - c.r.o;
    bool farEnough = fmaxf(fmaxf(fabsf(dh.x), fabsf(dh.y)), fabsf(dh.z)) > m;
    bool flatX = n.minX == n.maxX, flatY = n.minY == n.maxY, flatZ = n.minZ == n.maxZ;
    for (;;) {
        if (box_test_ref(mk3(n.minX, n.minY, n.minZ), mk3(n.maxX, n.maxY, n.maxZ), c.r.o, c.inv) == PT_FAR) return false;
        bool inX = flatX || (h.x >= n.minX + m && h.x <= n.maxX - m);
        bool inY = flatY || (h.y >= n.minY + m && h.y <= n.maxY - m);
        bool inZ = flatZ || (h.z >= n.minZ + m && h.z <= n.maxZ - m);
        if (inX && inY && inZ && farEnough) return true;
        n = up[n.parent];
        if (n.parent == 0xffffffffu) return true;
    }
}

PT_HD TriEdges prim_edges(const PrimRecord& p) {
    TriEdges t;
    t.p0 = mk3(p.p0x, p.p0y, p.p0z); t.e1 = mk3(p.e1x, p.e1y, p.e1z);
    t.e2 = mk3(p.e2x, p.e2y, p.e2z); t.n = mk3(p.nx, p.ny, p.nz);
    return t;
}

// Closest triangle hit strictly closer than hit.dst (the reference's `hitInfo.dst < closestHit.dst`).
// `key` is the position of the current best hit in the reference's loop order (sphere i -> i, object o
// -> nSpheres + o; 0xffffffff for none): the reference keeps the FIRST of equal-distance hits.
PT_HD_NOINLINE void bvh8_closest(const Bvh8View& bvh, const Xform* xforms, const UpNode* up, uint32_t nSpheres, const Ray& ray,
                                 uint32_t objectCount, Hit& hit, uint32_t& key) {
    BoxRay br = make_box_ray(ray);
    ObjRayCache oc; oc.xform = 0xffffffffu;
    uint32_t oct = br.octinv4 & 7u;
    uint32_t stackBase[PT_BVH8_STACK], stackBits[PT_BVH8_STACK];
    int sp = 0;
    uint32_t gBase = 0, gBits = 0x80000000u;          // node group: base index | hit bits(31..24) + imask(7..0)
    for (;;) {
        uint32_t tBase = 0, tBits = 0;
        if (gBits & 0xff000000u) {
            int bit = highest_bit(gBits);
            gBits &= ~(1u << bit);
            uint32_t slot = ((uint32_t)bit - 24u) ^ oct;
            uint32_t rel = (uint32_t)popcount32(gBits & 0xffu & ~(0xffffffffu << slot));
            uint32_t nodeIdx = gBase + rel;
            if (gBits & 0xff000000u) { stackBase[sp] = gBase; stackBits[sp] = gBits; sp++; }
            const Bvh8Node n = bvh.nodes[nodeIdx];
            uint32_t h = bvh8_node_hits(n, br, hit.dst);
            gBase = n.childBase; gBits = (h & 0xff000000u) | n.imask;
            tBase = n.primBase; tBits = h & 0x00ffffffu;
        } else {
            gBits = 0;
        }
        while (tBits) {
            int k = highest_bit(tBits);
            tBits &= ~(1u << k);
            uint32_t pi = tBase + (uint32_t)k;
            const PrimRecord p = bvh.prims[pi];
            if (p.objectIndex >= objectCount) continue;
            const Ray& tr = object_ray(oc, xforms, p.xformId, ray);
            float dst, u, v; bool front;
            if (tri_test(prim_edges(p), tr, (p.triIndexFlags & PT_PRIM_FRONTONLY) != 0, dst, u, v, front) &&
                (dst < hit.dst || (dst == hit.dst && nSpheres + p.objectIndex < key)) && ref_reachable(up, p.leafNode, oc, dst)) {
                hit.dst = dst; hit.u = u; hit.v = v;
                hit.id = pi | (front ? PT_HIT_FRONT : 0u);
                key = nSpheres + p.objectIndex;
            }
        }
        if ((gBits & 0xff000000u) == 0) {
            if (sp == 0) break;
            sp--; gBase = stackBase[sp]; gBits = stackBits[sp];
        }
    }
}

// Is there a NON-emissive triangle hit that the reference's closest-hit search would have
// preferred to an emissive hit at (tE, keyE)? (strictly closer, or equal distance and earlier in
// the reference's loop order)
PT_HD_NOINLINE bool bvh8_occluded(const Bvh8View& bvh, const Xform* xforms, const UpNode* up, uint32_t nSpheres, const Ray& ray,
                                  uint32_t objectCount, float tE, uint32_t keyE) {
    BoxRay br = make_box_ray(ray);
    ObjRayCache oc; oc.xform = 0xffffffffu;
    uint32_t oct = br.octinv4 & 7u;
    uint32_t stackBase[PT_BVH8_STACK], stackBits[PT_BVH8_STACK];
    int sp = 0;
    uint32_t gBase = 0, gBits = 0x80000000u;
    for (;;) {
        uint32_t tBase = 0, tBits = 0;
        if (gBits & 0xff000000u) {
            int bit = highest_bit(gBits);
            gBits &= ~(1u << bit);
            uint32_t slot = ((uint32_t)bit - 24u) ^ oct;
            uint32_t rel = (uint32_t)popcount32(gBits & 0xffu & ~(0xffffffffu << slot));
            uint32_t nodeIdx = gBase + rel;
            if (gBits & 0xff000000u) { stackBase[sp] = gBase; stackBits[sp] = gBits; sp++; }
            const Bvh8Node n = bvh.nodes[nodeIdx];
            uint32_t h = bvh8_node_hits(n, br, tE);
            gBase = n.childBase; gBits = (h & 0xff000000u) | n.imask;
            tBase = n.primBase; tBits = h & 0x00ffffffu;
        } else {
            gBits = 0;
        }
        while (tBits) {
            int k = highest_bit(tBits);
            tBits &= ~(1u << k);
            const PrimRecord p = bvh.prims[tBase + (uint32_t)k];
            if (p.objectIndex >= objectCount || (p.triIndexFlags & PT_PRIM_EMISSIVE)) continue;
            const Ray& tr = object_ray(oc, xforms, p.xformId, ray);
            float dst, u, v; bool front;
            if (tri_test(prim_edges(p), tr, (p.triIndexFlags & PT_PRIM_FRONTONLY) != 0, dst, u, v, front) &&
                (dst < tE || (dst == tE && nSpheres + p.objectIndex < keyE)) && ref_reachable(up, p.leafNode, oc, dst))
                return true;
        }
        if ((gBits & 0xff000000u) == 0) {
            if (sp == 0) break;
            sp--; gBase = stackBase[sp]; gBits = stackBits[sp];
        }
    }
    return false;
}

}  // namespace pt
