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
#define PT_HIT_CONTESTED 0x20000000u   /* only between k_extend and k_verify_extend / k_exact_extend (wavefront.cu): the ray goes to the reference-order re-trace; record indices stay below 2^29 */

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

// A node held in registers: the 80 bytes as twenty 32-bit words (five 16-byte loads).
//  w0..2 px,py,pz | w3 ex,ey,ez,imask | w4 childBase | w5 primBase | w6,7 meta[8]
//  w8,9 qlox | w10,11 qloy | w12,13 qloz | w14,15 qhix | w16,17 qhiy | w18,19 qhiz
struct NodeWords { uint32_t w[20]; };
PT_HD NodeWords load_node(const Bvh8Node* nodes, uint32_t idx) {
    NodeWords n;
#if defined(__CUDA_ARCH__)
    const uint4* p = reinterpret_cast<const uint4*>(nodes + idx);
    uint4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2), d = __ldg(p + 3), e = __ldg(p + 4);
    n.w[0] = a.x; n.w[1] = a.y; n.w[2] = a.z; n.w[3] = a.w; n.w[4] = b.x; n.w[5] = b.y; n.w[6] = b.z; n.w[7] = b.w;
    n.w[8] = c.x; n.w[9] = c.y; n.w[10] = c.z; n.w[11] = c.w; n.w[12] = d.x; n.w[13] = d.y; n.w[14] = d.z; n.w[15] = d.w;
    n.w[16] = e.x; n.w[17] = e.y; n.w[18] = e.z; n.w[19] = e.w;
#else
    memcpy(n.w, nodes + idx, sizeof(Bvh8Node));
#endif
    return n;
}
// Byte k of `word` dropped into the mantissa of 1.0f (bits 8..15): 1 + q * 2^-15, exact, one PRMT
// on the device — no integer->float conversion (quarter-rate pipe) and no subtraction.
#ifndef PT_UNIT_FROM_CONSTANT
#define PT_UNIT_FROM_CONSTANT 1
#endif
#if defined(__CUDACC__) && PT_UNIT_FROM_CONSTANT
// The bits of 1.0f read from constant memory: with BOTH PRMT inputs compile-time constants the compiler keeps 0x3F800000
// as the immediate and re-materialises the four byte selectors into registers around every use (20 extra moves per
// node); with this operand coming from the constant bank the selector is the immediate.
static __constant__ uint32_t c_unit_float_bits = 0x3F800000u;
#endif
PT_HD float byte_as_unit_float(uint32_t word, int k) {
#if defined(__CUDA_ARCH__) && PT_UNIT_FROM_CONSTANT
    return __uint_as_float(__byte_perm(word, c_unit_float_bits, 0x7604u + ((uint32_t)k << 4)));
#elif defined(__CUDA_ARCH__)
    return __uint_as_float(__byte_perm(word, 0x3F800000u, 0x7604u + ((uint32_t)k << 4)));
#else
    return as_float(0x3F800000u | (((word >> (8 * k)) & 0xffu) << 8));
#endif
}

// Slab-test the 8 children of a node; returns hit bits: inner children in bits 24..31 ordered by
// traversal priority for this ray's octant, leaf triangles in bits 0..23 (offset from primBase).
// A child plane is at t = (p + q * 2^(e-127) - o) * rd = q * s + o with s = 2^(e-127) * rd, o = (p - o) * rd;
// in mantissa form = Q * S + O with Q = 1 + q * 2^-15, S = 2^15 * s and O = o - S: one FMA per plane
// either way. O carries extra rounding of at most |S| * 2^-23 (a 1/256 of a quantisation step); planes
// evaluated in that form (the entry planes) are moved down by twice that, so the test stays conservative.
PT_HD float byte_to_float(uint32_t word, int k) {
#if defined(__CUDA_ARCH__)
    return __uint2float_rn((word >> (8 * k)) & 0xffu);        // I2F.U8 with a byte selector: one instruction on the XU pipe
#else
    return (float)((word >> (8 * k)) & 0xffu);
#endif
}
#ifndef PT_DEBUG_COUNT
#define PT_DEBUG_COUNT(i)      /* host-side instrumentation hook (profiling experiments only) */
#endif
#ifndef PT_TRAV_PREFETCH
#define PT_TRAV_PREFETCH 0
#endif
#ifndef PT_NODE_I2F_NEAR
#define PT_NODE_I2F_NEAR 2        /* how many of the three entry-plane axes also go through I2F: 0 / 1 / 2 / 3 -> 305.9 / 307.7 / 309.4 / 303.9 Mpaths/s */
#endif
#ifndef PT_NODE_IMAD_SELECT
#define PT_NODE_IMAD_SELECT 0     /* pick near/far words with integer multiply-adds (FMA pipe) instead of SEL (ALU pipe) */
#endif
PT_HD uint32_t pick_word(bool neg, uint32_t ifNeg, uint32_t ifPos) {
#if PT_NODE_IMAD_SELECT && defined(__CUDA_ARCH__)
    uint32_t d, r, f = neg ? 1u : 0u;
    asm("mad.lo.u32 %0, %1, 0xffffffff, %2;" : "=r"(d) : "r"(ifPos), "r"(ifNeg));      // ifNeg - ifPos
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(d), "r"(f), "r"(ifPos));
    return r;
#else
    return neg ? ifNeg : ifPos;
#endif
}
#ifndef PT_NODE_HITBITS_PRMT
#define PT_NODE_HITBITS_PRMT 1     /* hit-bit assembly: one PRMT per child extracts (count, position) */
#endif
#ifndef PT_NODE_FFMA2
#define PT_NODE_FFMA2 0           /* node test: child planes two at a time through FFMA2 (device only). A/B on B200: 306.2 vs 311.7 Mpaths/s without —
                                     an FFMA2 takes one issue slot (peaks_micro: FFMA2 + LOP3 interleaved 1108 G/s) but holds the FMA pipe for two
                                     cycles and wants aligned register pairs: 24 fewer instructions per node visit, a few spills at the 72-register cap */
#endif
#if defined(__CUDA_ARCH__)
// (a0, a1) * s + o with one packed FMA (fma.rn.f32x2, sm_100): each half rounds exactly like fma.rn.f32
static __device__ __forceinline__ float2 fma2_bcast(float a0, float a1, float s, float o) {
    unsigned long long a, ss, oo, r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
    asm("mov.b64 %0, {%1, %1};" : "=l"(ss) : "f"(s));
    asm("mov.b64 %0, {%1, %1};" : "=l"(oo) : "f"(o));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(ss), "l"(oo));
    float2 w;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(w.x), "=f"(w.y) : "l"(r));
    return w;
}
#endif
PT_HD uint32_t bvh8_node_hits(const NodeWords& n, const BoxRay& br, float tmax) {
    // entry planes: mantissa form Q * S + O (PRMT on the ALU pipe + FFMA); exit planes: q * s + o with q converted by
    // I2F (XU pipe + FFMA) — the ALU pipe is this kernel's busiest unit (ncu: 68 %), the XU pipe is nearly idle.
    const float sx = as_float((n.w[3] & 0xffu) << 23) * br.rd.x;
    const float sy = as_float(((n.w[3] >> 8) & 0xffu) << 23) * br.rd.y;
    const float sz = as_float(((n.w[3] >> 16) & 0xffu) << 23) * br.rd.z;
    const float Sx = sx * 32768.f, Sy = sy * 32768.f, Sz = sz * 32768.f;
    const float ofx = (as_float(n.w[0]) - br.o.x) * br.rd.x;
    const float ofy = (as_float(n.w[1]) - br.o.y) * br.rd.y;
    const float ofz = (as_float(n.w[2]) - br.o.z) * br.rd.z;
    const float onx = pt_fma(fabsf(Sx), -2.384185791015625e-07f, ofx - Sx);      // O - |S| * 2^-22
    const float ony = pt_fma(fabsf(Sy), -2.384185791015625e-07f, ofy - Sy);
    const float onz = pt_fma(fabsf(Sz), -2.384185791015625e-07f, ofz - Sz);
    const bool negx = br.rd.x < 0.f, negy = br.rd.y < 0.f, negz = br.rd.z < 0.f;
    const uint32_t nx[2] = {pick_word(negx, n.w[14], n.w[8]), pick_word(negx, n.w[15], n.w[9])}, fx[2] = {pick_word(negx, n.w[8], n.w[14]), pick_word(negx, n.w[9], n.w[15])};
    const uint32_t ny[2] = {pick_word(negy, n.w[16], n.w[10]), pick_word(negy, n.w[17], n.w[11])}, fy[2] = {pick_word(negy, n.w[10], n.w[16]), pick_word(negy, n.w[11], n.w[17])};
    const uint32_t nz[2] = {pick_word(negz, n.w[18], n.w[12]), pick_word(negz, n.w[19], n.w[13])}, fz[2] = {pick_word(negz, n.w[12], n.w[18]), pick_word(negz, n.w[13], n.w[19])};
    uint32_t hits = 0;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        // four children at a time (Ylitie et al. 2017): which are inner nodes, their bit position
        // (inner: 24 + (slot ^ (7 - octant)), leaf: offset of its first triangle) and unary count
        const uint32_t meta4 = n.w[6 + h];
        const uint32_t isInner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t innerMask4 = (isInner4 >> 4) * 0xffu;
        const uint32_t bitIndex4 = (meta4 ^ (br.octinv4 & innerMask4)) & 0x1f1f1f1fu;
        const uint32_t childBits4 = (meta4 >> 5) & 0x07070707u;
#if defined(__CUDA_ARCH__) && PT_NODE_HITBITS_PRMT
        // per child one byte: unary count in bits 5..7, final bit position in bits 0..4 — one PRMT extracts it, the shift
        // count of a 32-bit shift takes its low five bits by itself
        const uint32_t posCount4 = (meta4 & 0xe0e0e0e0u) | bitIndex4;
#endif
#if defined(__CUDA_ARCH__) && PT_NODE_FFMA2
        // two children per FFMA2 (sm_100 packed FP32 FMA: one issued instruction, the per-axis scale and offset broadcast):
        // the same fma.rn per plane as the scalar form, half the FMA issue slots of a node visit (48 -> 24)
#pragma unroll
        for (int k = 0; k < 4; k += 2) {
            const float2 t0x = PT_NODE_I2F_NEAR > 0 ? fma2_bcast(byte_to_float(nx[h], k), byte_to_float(nx[h], k + 1), sx, ofx)
                                                    : fma2_bcast(byte_as_unit_float(nx[h], k), byte_as_unit_float(nx[h], k + 1), Sx, onx);
            const float2 t0y = PT_NODE_I2F_NEAR > 1 ? fma2_bcast(byte_to_float(ny[h], k), byte_to_float(ny[h], k + 1), sy, ofy)
                                                    : fma2_bcast(byte_as_unit_float(ny[h], k), byte_as_unit_float(ny[h], k + 1), Sy, ony);
            const float2 t0z = PT_NODE_I2F_NEAR > 2 ? fma2_bcast(byte_to_float(nz[h], k), byte_to_float(nz[h], k + 1), sz, ofz)
                                                    : fma2_bcast(byte_as_unit_float(nz[h], k), byte_as_unit_float(nz[h], k + 1), Sz, onz);
            const float2 t1x = fma2_bcast(byte_to_float(fx[h], k), byte_to_float(fx[h], k + 1), sx, ofx);
            const float2 t1y = fma2_bcast(byte_to_float(fy[h], k), byte_to_float(fy[h], k + 1), sy, ofy);
            const float2 t1z = fma2_bcast(byte_to_float(fz[h], k), byte_to_float(fz[h], k + 1), sz, ofz);
            const float tn0 = fmaxf(fmaxf(t0x.x, t0y.x), fmaxf(t0z.x, 0.f)), tf0 = fminf(fminf(t1x.x, t1y.x), fminf(t1z.x, tmax));
            const float tn1 = fmaxf(fmaxf(t0x.y, t0y.y), fmaxf(t0z.y, 0.f)), tf1 = fminf(fminf(t1x.y, t1y.y), fminf(t1z.y, tmax));
            if (tn0 <= tf0) hits |= ((childBits4 >> (8 * k)) & 0xffu) << ((bitIndex4 >> (8 * k)) & 0xffu);
            if (tn1 <= tf1) hits |= ((childBits4 >> (8 * (k + 1))) & 0xffu) << ((bitIndex4 >> (8 * (k + 1))) & 0xffu);
        }
#else
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float t0x = PT_NODE_I2F_NEAR > 0 ? pt_fma(byte_to_float(nx[h], k), sx, ofx) : pt_fma(byte_as_unit_float(nx[h], k), Sx, onx);
            float t0y = PT_NODE_I2F_NEAR > 1 ? pt_fma(byte_to_float(ny[h], k), sy, ofy) : pt_fma(byte_as_unit_float(ny[h], k), Sy, ony);
            float t0z = PT_NODE_I2F_NEAR > 2 ? pt_fma(byte_to_float(nz[h], k), sz, ofz) : pt_fma(byte_as_unit_float(nz[h], k), Sz, onz);
            float t1x = pt_fma(byte_to_float(fx[h], k), sx, ofx);
            float t1y = pt_fma(byte_to_float(fy[h], k), sy, ofy);
            float t1z = pt_fma(byte_to_float(fz[h], k), sz, ofz);
            float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.f));
            float tf = fminf(fminf(t1x, t1y), fminf(t1z, tmax));
#if defined(__CUDA_ARCH__) && PT_NODE_HITBITS_PRMT
            const uint32_t pc = __byte_perm(posCount4, 0u, 0x4440u + (uint32_t)k);
            if (tn <= tf) hits |= (pc >> 5) << (pc & 31u);
#else
            if (tn <= tf) hits |= ((childBits4 >> (8 * k)) & 0xffu) << ((bitIndex4 >> (8 * k)) & 0xffu);
#endif
        }
#endif
    }
    return hits;
}

struct ObjRayCache { uint32_t xform; Ray r; };

PT_HD const Ray& object_ray(ObjRayCache& c, const Xform* xforms, uint32_t xformId, const Ray& world) {
    if (c.xform != xformId) {
#if defined(__CUDA_ARCH__)
        float Minv[16];                       // four 16-byte loads (read element by element through a float* they were 12 scalar loads)
        const float4* q = reinterpret_cast<const float4*>(xforms[xformId].Minv);
#pragma unroll
        for (int j = 0; j < 4; ++j) { const float4 col = __ldg(q + j); Minv[4 * j] = col.x; Minv[4 * j + 1] = col.y; Minv[4 * j + 2] = col.z; Minv[4 * j + 3] = col.w; }
#else
        const float* Minv = xforms[xformId].Minv;
#endif
        c.r.d = mul_dir(Minv, world.d);       // raytrace.comp:292 (direction not renormalised: t stays world-parametric)
        c.r.o = mul_point(Minv, world.o);     // raytrace.comp:293
        c.xform = xformId;
    }
    return c.r;
}

PT_HD PrimRecord load_prim(const PrimRecord* prims, uint32_t i) {
#if defined(__CUDA_ARCH__)
    const uint4* q = reinterpret_cast<const uint4*>(prims + i);
    uint4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3);
    PrimRecord p;
    p.p0x = __uint_as_float(a.x); p.p0y = __uint_as_float(a.y); p.p0z = __uint_as_float(a.z); p.objectIndex = a.w;
    p.e1x = __uint_as_float(b.x); p.e1y = __uint_as_float(b.y); p.e1z = __uint_as_float(b.z); p.triIndexFlags = b.w;
    p.e2x = __uint_as_float(c.x); p.e2y = __uint_as_float(c.y); p.e2z = __uint_as_float(c.z); p.xformId = c.w;
    p.nx = __uint_as_float(d.x); p.ny = __uint_as_float(d.y); p.nz = __uint_as_float(d.z); p.leafNode = d.w;
    return p;
#else
    return prims[i];
#endif
}

// Would the reference have tested the triangle in BVH2 leaf `leaf`? It does iff boxIntersection()
// (raytrace.comp:263-274) reports a hit for every node from the leaf up to, not including, the
// object's root (:324-348; the root's own box is never tested, :301-307). A box that contains the
// object-space hit point h with a margin m far above the slab test's rounding error passes, and so
// does every box around it.
//   quick accept (no memory access), from the triangle's own bounding box T rebuilt from the record
//     (p0, p0+e1, p0+e2) — the leaf box and all its ancestors contain T:
//     (a) h lies inside T with margin m on every axis on which the LEAF is not flat (flags
//         precomputed at upload): every enclosing box passes a fortiori;
//     (b) exactly one axis a fails (a): typically a wall triangle lying ON a face of its leaf box.
//         On the other two axes the slab intervals of every enclosing box contain dst with a
//         parametric margin m/|d_b|; on axis a the interval end is within (m/16 + rounding)/|d_a| of
//         dst as long as h_a is within m/16 of T, and t1_a <= t2_a always. With
//         |d_a| >= max(|d_b|, |d_c|)/4 that slack is below the others' margin, so
//         tNear <= tFar and tFar > 0 hold for every enclosing box. When the TRIANGLE itself is flat
//         on axis a (all three vertices share p0_a, so e1_a == e2_a == 0 exactly) no bound on d_a is
//         needed: dst = (o_a - p0_a) n_a / -(d_a n_a) and the slab value of the plane p0_a,
//         (p0_a - o_a) * (1/d_a), start from the same rounded difference and agree to 4e-7 * dst,
//         rounding is monotone in the box coordinate so t1_a <= that value <= t2_a for every box
//         around the plane, and the other axes' margins are >= 2e-5 * dst.
//   otherwise: walk leaf -> root re-running the reference's slab test, stopping at the first box
//     that contains h with the margin. Axes on which the leaf is flat (axis-aligned triangles) are
//     exempt from the margin: there tMin == tMax is the plane distance itself and rounding is
//     monotone in the box coordinate.
// quick accept: true = reachable for certain; false = the leaf -> root walk has to decide
PT_HD bool ref_reachable_quick(const PrimRecord& p, const ObjRayCache& c, float dst) {
    const v3 h = c.r.o + c.r.d * dst;
    const float m = 2e-5f * (fmaxf(fmaxf(fabsf(h.x), fabsf(h.y)), fabsf(h.z)) + fmaxf(fmaxf(fabsf(c.r.o.x), fabsf(c.r.o.y)), fabsf(c.r.o.z))) + 1e-30f;
    const v3 dh = h - c.r.o;
    const bool farEnough = fmaxf(fmaxf(fabsf(dh.x), fabsf(dh.y)), fabsf(dh.z)) > m;
    const bool flatX = (p.triIndexFlags & PT_PRIM_LEAF_FLAT_X) != 0, flatY = (p.triIndexFlags & PT_PRIM_LEAF_FLAT_Y) != 0,
               flatZ = (p.triIndexFlags & PT_PRIM_LEAF_FLAT_Z) != 0;
    PT_DEBUG_COUNT(2);
    if (p.triIndexFlags & PT_PRIM_LEAF_IS_ROOT) return true;
    if (farEnough) {
        const float lx = p.p0x + fminf(fminf(p.e1x, p.e2x), 0.f), hx = p.p0x + fmaxf(fmaxf(p.e1x, p.e2x), 0.f);
        const float ly = p.p0y + fminf(fminf(p.e1y, p.e2y), 0.f), hy = p.p0y + fmaxf(fmaxf(p.e1y, p.e2y), 0.f);
        const float lz = p.p0z + fminf(fminf(p.e1z, p.e2z), 0.f), hz = p.p0z + fmaxf(fmaxf(p.e1z, p.e2z), 0.f);
        const bool inX = flatX || (h.x >= lx + m && h.x <= hx - m);
        const bool inY = flatY || (h.y >= ly + m && h.y <= hy - m);
        const bool inZ = flatZ || (h.z >= lz + m && h.z <= hz - m);
        const int nOut = (inX ? 0 : 1) + (inY ? 0 : 1) + (inZ ? 0 : 1);
        if (nOut == 0) { PT_DEBUG_COUNT(3); return true; }
        // (b) needs real margins on the other two axes: not when the leaf is flat on one of them (its slab interval there
        // is the single value of the plane distance, and a hit point m/16 outside the triangle's box on axis a means the
        // ray has left the zero-thickness box: found on the C4 stand-in, frame 402 pixel (1740, 780), a trim-box face)
        if (nOut == 1 && !(flatX || flatY || flatZ)) {
            const float m16 = m * 0.0625f;
            const float ax = fabsf(c.r.d.x), ay = fabsf(c.r.d.y), az = fabsf(c.r.d.z);
            bool ok;
            if (!inX) ok = h.x >= lx - m16 && h.x <= hx + m16 && ((lx == hx && ax >= 1e-30f) || 4.f * ax >= fmaxf(ay, az));
            else if (!inY) ok = h.y >= ly - m16 && h.y <= hy + m16 && ((ly == hy && ay >= 1e-30f) || 4.f * ay >= fmaxf(ax, az));
            else ok = h.z >= lz - m16 && h.z <= hz + m16 && ((lz == hz && az >= 1e-30f) || 4.f * az >= fmaxf(ax, ay));
            if (ok) { PT_DEBUG_COUNT(4); return true; }
        }
    }
    return false;
}
// the walk: the reference's slab tests leaf -> root
PT_HD bool ref_reachable_walk(const UpNode* __restrict__ up, const PrimRecord& p, const ObjRayCache& c, float dst) {
    const v3 h = c.r.o + c.r.d * dst;
    const float m = 2e-5f * (fmaxf(fmaxf(fabsf(h.x), fabsf(h.y)), fabsf(h.z)) + fmaxf(fmaxf(fabsf(c.r.o.x), fabsf(c.r.o.y)), fabsf(c.r.o.z))) + 1e-30f;
    const v3 dh = h - c.r.o;
    const bool farEnough = fmaxf(fmaxf(fabsf(dh.x), fabsf(dh.y)), fabsf(dh.z)) > m;
    const bool flatX = (p.triIndexFlags & PT_PRIM_LEAF_FLAT_X) != 0, flatY = (p.triIndexFlags & PT_PRIM_LEAF_FLAT_Y) != 0,
               flatZ = (p.triIndexFlags & PT_PRIM_LEAF_FLAT_Z) != 0;
    PT_DEBUG_COUNT(5);
    const v3 inv = mk3(1.f / c.r.d.x, 1.f / c.r.d.y, 1.f / c.r.d.z);   // raytrace.comp:294
    UpNode n = up[p.leafNode];
    for (;;) {
        if (box_test_ref(mk3(n.minX, n.minY, n.minZ), mk3(n.maxX, n.maxY, n.maxZ), c.r.o, inv) == PT_FAR) { PT_DEBUG_COUNT(6); return false; }
        bool inX = flatX || (h.x >= n.minX + m && h.x <= n.maxX - m);
        bool inY = flatY || (h.y >= n.minY + m && h.y <= n.maxY - m);
        bool inZ = flatZ || (h.z >= n.minZ + m && h.z <= n.maxZ - m);
        if (inX && inY && inZ && farEnough) return true;
        n = up[n.parent];
        if (n.parent == 0xffffffffu) return true;
    }
}
PT_HD bool ref_reachable(const UpNode* __restrict__ up, const PrimRecord& p, const ObjRayCache& c, float dst) {
    return ref_reachable_quick(p, c, dst) || ref_reachable_walk(up, p, c, dst);
}

PT_HD TriEdges prim_edges(const PrimRecord& p) {
    TriEdges t;
    t.p0 = mk3(p.p0x, p.p0y, p.p0z); t.e1 = mk3(p.e1x, p.e1y, p.e1z);
    t.e2 = mk3(p.e2x, p.e2y, p.e2z); t.n = mk3(p.nx, p.ny, p.nz);
    return t;
}

// ---- exactness against the reference's traversal ORDER -----------------------------------------------------------
// The reference prunes a BVH2 node when its slab distance is not below the closest hit found SO FAR in its own loop
// order (raytrace.comp:346-347), so which triangles it tests depends on history. Taking the minimum over all reachable
// triangles is the same answer EXCEPT when a triangle's hit distance dst_T is not above the slab distance b_T of one of
// its enclosing boxes — possible only by rounding (triangle on a box face: b_T - dst_T is a few ulps) — and another
// candidate lies in between: the reference then never tests T. Analysis: the two answers can
// differ only if two reachable candidates are closer together than delta = b_T - dst_T <= the rounding error of the
// Moeller-Trumbore distance, about 4 eps * (scene extent) / cos(incidence) in ray-parameter units (the slab distance
// itself is accurate to 3 ulp: lo - o is a difference of two floats).
// Rule: a ray is CONTESTED when (i) two reachable candidates lie within PT_TAU * (distance + scene extent) of each other
// — 60 times that error at normal incidence — or (ii) its best hit is grazing (1 / cos above PT_GRAZE: the error bound
// then approaches (i)'s window). Contested rays are re-traced in the reference's own order (pt_reftrace.h:
// ref_closest_hit — k_exact_* in wavefront.cu). That also settles hits at exactly equal distance on two triangles of one
// object, the former "documented tie class": there is none left. Nodes are culled against best + window, so a competitor
// inside the window beyond the best hit is always tested. State: two bits of br.octinv4 (bits 5..7 of its bytes are free:
// the node test masks with 0x1f) — separate registers for it cost 17 % of k_extend at its 72-register cap.
#ifndef PT_CONTEST
#define PT_CONTEST 1          /* 0: measurement only — no contest detection (results then depend on the traversal order) */
#endif
#define PT_TAU 1.52587890625e-5f       /* 2^-16 */
#define PT_GRAZE 64.f
#define PT_TRAV_CONTESTED 0x80000000u
#define PT_TRAV_GRAZING 0x40000000u   /* the current best hit is grazing */
PT_HD float maxabs3(v3 a) { return fmaxf(fmaxf(fabsf(a.x), fabsf(a.y)), fabsf(a.z)); }
PT_HD float contest_window(float dst, float extent) { return PT_TAU * (dst + extent); }
PT_HD float cull_dst(float best, float extent) { return best + contest_window(best, extent); }
// 1 / cos(incidence) from what tri_test already formed: invD0 = 1 / -dot(d, n) (n = cross(e1, e2), not normalised)
PT_HD float inv_cos_incidence(const TriEdges& t, const Ray& r, float invD0) { return fabsf(invD0) * maxabs3(t.n) * maxabs3(r.d); }
PT_HD bool hit_is_grazing(const TriEdges& t, const Ray& r, float invD0) { return inv_cos_incidence(t, r, invD0) > PT_GRAZE; }

// ---- traversal as per-ray steps ---------------------------------------------------------------
// The state of one ray's walk through the wide BVH. Work items are "groups": a node group
// (base = first inner child, bits 31..24 = children still to visit in priority order, bits 7..0 =
// imask) or a triangle group (base = first record, bits 23..0 = records still to test). The
// stack holds groups of either kind (a triangle group has no bit above 23). Steps are separate so
// that a warp can run one KIND of step for all its lanes at a time (wavefront.cu: the persistent
// kernels vote on which) — and the same functions, looped per ray, are the host emulation.
struct Trav {
    Ray ray;            // world space
    BoxRay br;
    ObjRayCache oc;
    uint32_t gBase, gBits;      // current node group
    uint32_t tBase, tBits;      // current triangle group
    int sp;
};
PT_HD bool trav_contested(const Trav& t) { return (t.br.octinv4 & 0xc0000000u) != 0u; }      // contested, or the best hit is grazing (see below)
PT_HD void trav_begin(Trav& t, const Ray& ray) {
    t.ray = ray; t.br = make_box_ray(ray); t.oc.xform = 0xffffffffu;
    t.gBase = 0; t.gBits = 0x80000000u;               // the root as the only child of a virtual group
    t.tBase = 0; t.tBits = 0; t.sp = 0;
}
// The per-ray stack of groups. PrivateStack: two arrays in thread-private (local) memory. The persistent kernels use
// a SplitStack (wavefront.cu): the first PT_BVH8_SHARED_STACK entries live in shared memory, [entry][thread] so that a
// warp's accesses are conflict-free, the rest (rarely reached) in local memory.
struct PrivateStack {
    uint32_t base[PT_BVH8_STACK], bits[PT_BVH8_STACK];
    PT_HD void put(int i, uint32_t b, uint32_t v) { base[i] = b; bits[i] = v; }
    PT_HD void get(int i, uint32_t& b, uint32_t& v) { b = base[i]; v = bits[i]; }
};
// The same stack with its top entry held in two registers: a pop hands out registers and refills them with a load
// nobody waits for until the next pop or push (the persistent kernels' pop was a load-to-use stall for the whole
// converged warp, profiles/r01/k_extend_final_by_line.txt: pt_traverse.h trav_pop 7.4 % of the stall samples), and an
// entry that is pushed and popped again before anything displaces it never goes through memory.
// Invariant: with sp entries on the stack, entry sp-1 is in (topB, topV) and entries 0..sp-2 are in memory.
struct TopCachedStack {
    uint32_t base[PT_BVH8_STACK], bits[PT_BVH8_STACK];
    uint32_t topB, topV;
    PT_HD void put(int i, uint32_t b, uint32_t v) {
        if (i > 0) { base[i - 1] = topB; bits[i - 1] = topV; }
        topB = b; topV = v;
    }
    PT_HD void get(int i, uint32_t& b, uint32_t& v) {
        b = topB; v = topV;
        if (i > 0) { topB = base[i - 1]; topV = bits[i - 1]; }
    }
};
PT_HD bool trav_has_node(const Trav& t) { return (t.gBits & 0xff000000u) != 0; }
PT_HD bool trav_has_tri(const Trav& t) { return t.tBits != 0; }
// Nothing in hand: take the next group from the stack. Returns false when the walk is over.
template <class Stack>
PT_HD bool trav_pop(Trav& t, Stack& stack) {
    if (t.sp == 0) return false;
    t.sp--;
    uint32_t b, bits;
    stack.get(t.sp, b, bits);
    if (bits & 0xff000000u) { t.gBase = b; t.gBits = bits; }
    else { t.tBase = b; t.tBits = bits; t.gBits = 0; }
    return true;
}
// Visit the nearest pending child of the current node group (precondition: trav_has_node).
// A triangle group still in hand is postponed: it goes on top of the stack, so it is the next
// thing popped once this subtree is done.
template <class Stack>
PT_HD void trav_node_step(const Bvh8View& bvh, Trav& t, Stack& stack, float tmax) {
    const uint32_t oct = t.br.octinv4 & 7u;
    int bit = highest_bit(t.gBits);
    t.gBits &= ~(1u << bit);
    uint32_t slot = ((uint32_t)bit - 24u) ^ oct;
    uint32_t rel = (uint32_t)popcount32(t.gBits & 0xffu & ~(0xffffffffu << slot));
    uint32_t nodeIdx = t.gBase + rel;
    if (t.gBits & 0xff000000u) { stack.put(t.sp, t.gBase, t.gBits); t.sp++; }
    if (t.tBits) { stack.put(t.sp, t.tBase, t.tBits); t.sp++; }
    PT_DEBUG_COUNT(0);
    const NodeWords n = load_node(bvh.nodes, nodeIdx);
    uint32_t h = bvh8_node_hits(n, t.br, tmax);
    t.gBase = n.w[4]; t.gBits = (h & 0xff000000u) | (n.w[3] >> 24);
    t.tBase = n.w[5]; t.tBits = h & 0x00ffffffu;
#if defined(__CUDA_ARCH__) && PT_TRAV_PREFETCH
    // the records the next steps of this lane will read: first triangle of the new group, nearest child hit
    if ((PT_TRAV_PREFETCH & 1) && t.tBits) asm volatile("prefetch.global.L1 [%0];" ::"l"(bvh.prims + t.tBase + (uint32_t)highest_bit(t.tBits)));
    if ((PT_TRAV_PREFETCH & 2) && (t.gBits & 0xff000000u)) {
        uint32_t slot2 = ((uint32_t)highest_bit(t.gBits) - 24u) ^ oct;
        const Bvh8Node* nn = bvh.nodes + t.gBase + (uint32_t)popcount32(t.gBits & 0xffu & ~(0xffffffffu << slot2));
        asm volatile("prefetch.global.L1 [%0];" ::"l"(nn));
        asm volatile("prefetch.global.L1 [%0];" ::"l"((const char*)nn + 64));
    }
#endif
}

// Test the next triangle of the current group for a closest hit strictly closer than hit.dst (the
// reference's `hitInfo.dst < closestHit.dst`). `key` is the position of the current best hit in the
// reference's loop order (sphere i -> i, object o -> nSpheres + o; 0xffffffff for none): the
// reference keeps the FIRST of equal-distance hits.
// REACH = false (the persistent k_extend): candidates are accepted WITHOUT the reachability confirmation — ref_reachable ran
// at 2-3 lanes per warp and was 15 % of that kernel's instructions — and the ray's final hit is confirmed afterwards by
// k_verify_extend at full lane count; a ray whose final hit fails goes to the reference-order re-trace, which is right by
// construction. Until then an unreachable candidate can only shadow farther ones, and every candidate within the contest
// window of the final hit is still seen (nodes are culled against best + window), flagged contested whether reachable or not.
template <bool REACH = true>
PT_HD void trav_tri_step_closest(const Bvh8View& bvh, const Xform* xforms, const UpNode* up, uint32_t nSpheres, uint32_t objectCount,
                                 float extent, Trav& t, Hit& hit, uint32_t& key, float* bestInvCos = nullptr) {
    int k = highest_bit(t.tBits);
    t.tBits &= ~(1u << k);
    uint32_t pi = t.tBase + (uint32_t)k;
    PT_DEBUG_COUNT(1);
    const PrimRecord p = load_prim(bvh.prims, pi);
    if (p.objectIndex >= objectCount) return;
    const Ray& tr = object_ray(t.oc, xforms, p.xformId, t.ray);
    float dst, u, v, invD0; bool front;
    const TriEdges te = prim_edges(p);
    if (!tri_test_r(te, tr, (p.triIndexFlags & PT_PRIM_FRONTONLY) != 0, dst, u, v, front, invD0)) return;
    const bool wins = dst < hit.dst || (dst == hit.dst && nSpheres + p.objectIndex < key);
#if PT_CONTEST
    // only a hit that wins, or one that comes within tau of the best, can change the answer
    const bool near = fabsf(dst - hit.dst) <= contest_window(fmaxf(dst, hit.dst), extent) && hit.id != PT_HIT_MISS;
    if (!(wins || near) || (REACH && !ref_reachable(up, p, t.oc, dst))) return;
    if (near && (hit.id & ~PT_HIT_FRONT) != pi) t.br.octinv4 |= PT_TRAV_CONTESTED;
    if (wins) {
        const float g = inv_cos_incidence(te, tr, invD0);
        t.br.octinv4 = (t.br.octinv4 & ~PT_TRAV_GRAZING) | (g > PT_GRAZE ? PT_TRAV_GRAZING : 0u);
        if (bestInvCos) *bestInvCos = g;
    }
#else
    if (!wins || (REACH && !ref_reachable(up, p, t.oc, dst))) return;
#endif
    if (wins) {
        hit.dst = dst; hit.u = u; hit.v = v;
        hit.id = pi | (front ? PT_HIT_FRONT : 0u);
        key = nSpheres + p.objectIndex;
    }
}
// Occlusion form: is the next triangle a NON-emissive hit that the reference's closest-hit search
// would have preferred to an emissive hit at (tE, keyE)? (strictly closer, or equal distance and
// earlier in the reference's loop order) -> PT_OCCL_YES. A reachable hit within the slacks of tE on either side makes
// the query CONTESTED (see above): PT_OCCL_CONTESTED, answered by the reference-order search.
#define PT_OCCL_NO 0
#define PT_OCCL_YES 1
#define PT_OCCL_CONTESTED 2
// REACH = false (the persistent k_occlude): as in trav_tri_step_closest<false>, the occluder is taken without the
// reachability confirmation and reported through `occluder` / `occDst`; k_verify_occluded confirms it afterwards.
template <bool REACH = true>
PT_HD int trav_tri_step_occluded(const Bvh8View& bvh, const Xform* xforms, const UpNode* up, uint32_t nSpheres, uint32_t objectCount,
                                 float extent, float windowScale, Trav& t, float tE, uint32_t keyE, uint32_t* occluder = nullptr, float* occDst = nullptr) {
    int k = highest_bit(t.tBits);
    t.tBits &= ~(1u << k);
    const uint32_t pi = t.tBase + (uint32_t)k;
    const PrimRecord p = load_prim(bvh.prims, pi);
    if (p.objectIndex >= objectCount || (p.triIndexFlags & PT_PRIM_EMISSIVE)) return PT_OCCL_NO;
    const Ray& tr = object_ray(t.oc, xforms, p.xformId, t.ray);
    float dst, u, v, invD0; bool front;
    const TriEdges te = prim_edges(p);
    if (!tri_test_r(te, tr, (p.triIndexFlags & PT_PRIM_FRONTONLY) != 0, dst, u, v, front, invD0)) return PT_OCCL_NO;
    const bool closer = dst < tE || (dst == tE && nSpheres + p.objectIndex < keyE);
    const float gap = fabsf(dst - tE), window = contest_window(fmaxf(dst, tE), extent);
    const bool near = gap <= window * windowScale;          // windowScale >= 1: the emissive hit's own error grows with its 1 / cos
    if (!(closer || near) || (REACH && !ref_reachable(up, p, t.oc, dst))) return PT_OCCL_NO;
#if PT_CONTEST
    // an occluder far in front of tE decides unless its own distance error (which grows with 1 / cos) reaches back to tE
    if (near || gap <= window * inv_cos_incidence(te, tr, invD0) * (1.f / 16.f)) return PT_OCCL_CONTESTED;
#endif
    if (closer && occluder) { *occluder = pi; *occDst = dst; }
    return closer ? PT_OCCL_YES : PT_OCCL_NO;
}

// Per-ray loops over the steps (host emulation, the small emissive BVH). `contested` is OR-ed: with bestInvCos == NULL
// it includes "the best hit is grazing"; otherwise 1 / cos(incidence) of the best triangle hit is returned instead.
PT_HD_NOINLINE void bvh8_closest(const Bvh8View& bvh, const Xform* xforms, const UpNode* up, uint32_t nSpheres, const Ray& ray,
                                 uint32_t objectCount, float extent, Hit& hit, uint32_t& key, uint32_t& contested, float* bestInvCos = nullptr) {
    PrivateStack stack;
    Trav t;
    trav_begin(t, ray);
    for (;;) {
        if (trav_has_tri(t)) trav_tri_step_closest(bvh, xforms, up, nSpheres, objectCount, extent, t, hit, key, bestInvCos);
        else if (trav_has_node(t)) trav_node_step(bvh, t, stack, cull_dst(hit.dst, extent));
        else if (!trav_pop(t, stack)) break;
    }
    contested |= (bestInvCos ? (t.br.octinv4 & PT_TRAV_CONTESTED) != 0u : trav_contested(t)) ? 1u : 0u;
}
// -> PT_OCCL_NO / PT_OCCL_YES / PT_OCCL_CONTESTED
PT_HD_NOINLINE int bvh8_occluded(const Bvh8View& bvh, const Xform* xforms, const UpNode* up, uint32_t nSpheres, const Ray& ray,
                                 uint32_t objectCount, float extent, float windowScale, float tE, uint32_t keyE) {
    PrivateStack stack;
    Trav t;
    trav_begin(t, ray);
    int contested = PT_OCCL_NO;
    for (;;) {
        if (trav_has_tri(t)) {
            const int r = trav_tri_step_occluded(bvh, xforms, up, nSpheres, objectCount, extent, windowScale, t, tE, keyE);
            if (r == PT_OCCL_YES) return PT_OCCL_YES;
            if (r == PT_OCCL_CONTESTED) contested = PT_OCCL_CONTESTED;
        }
        else if (trav_has_node(t)) trav_node_step(bvh, t, stack, tE + contest_window(tE, extent) * windowScale);
        else if (!trav_pop(t, stack)) break;
    }
    return contested;
}

}  // namespace pt
