// Device build of the internal 8-wide BVH over every (object, triangle) instance (pt_bvh8.h) — the upload-time work
// of b200pt_upload_scene / b200pt_update_objects, which the host builder (bvh8_build.h) does in 13.5 s for 10 M
// triangles. Same algorithm as the host builder: world boxes -> binned-SAH binary tree (16 bins, leaves of <= 3
// triangles) -> greedy collapse to 8 children -> octant-ordered slots -> 8-bit planes rounded outward. Traversal
// results do not depend on the tree (every accepted hit is the reference's own triangle test, confirmed against
// the reference's BVH2: pt_traverse.h), so there is no bit-exactness requirement against the host tree — only
// that boxes are conservative; tests/test_gpu_parity.py runs every parity case through this build.
//   k_wide_records   one thread per instance: PrimRecord / ShadeRecord / UvRecord / padded world box
//   k_wide_split     one block per binary node and level: bounds, bins, SAH, stable partition, children
//   k_wide_emit      one thread per wide node and level: collapse, slots, quantisation, record copies
#include <cuda_runtime.h>
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "bvh8_build.h"
#include "wide_build.h"

namespace pt {
namespace wb {

#ifndef PT_WIDE_BINS
#define PT_WIDE_BINS 64
#endif
#ifndef PT_WIDE_BINS_SMALL
#define PT_WIDE_BINS_SMALL 16
#endif
// SAH bins per axis. Tree quality is worth a few per cent of tracing throughput (C4 stand-in, 24 frames in flight:
// 16 bins 285.1 Mpaths/s, 32: 293.3, 64: 295.1, 128: 297.0, 256: 295.4; on the statue scene 64 and 128 are level), the sweep
// over the bins is serial per node and axis, and most nodes are tiny: nodes of up to 64 triangles use NB_SMALL bins
// (16, 32 or 64 there measure the same). Tiers measured on the final traversal kernels (Mpaths/s, device build ms):
// 64 everywhere above 64 triangles 310.3 / 5.7; 128 above 2048 triangles and 64 below 311.5 / 7.2; 128 everywhere 312.4 / 9.0;
// 256 above 2048: 309.7 / 41.6 — differences at the level of scene-specific noise, 64 stays.
#ifndef PT_WIDE_BINS_MID
#define PT_WIDE_BINS_MID PT_WIDE_BINS       /* bins for nodes of 65 .. PT_WIDE_MID_MAX triangles */
#endif
#ifndef PT_WIDE_MID_MAX
#define PT_WIDE_MID_MAX 2048
#endif
constexpr int NB = PT_WIDE_BINS;
constexpr int NB_MID = PT_WIDE_BINS_MID;
constexpr int NB_SMALL = PT_WIDE_BINS_SMALL;
constexpr uint32_t HUGE_MIN = 32768;         // nodes with more triangles are split by many blocks (k_huge_*)

struct Node2 {                               // binary SAH tree (as b8::Node2)
    float lo[3], hi[3];
    uint32_t left, right;                    // children (count == 0)
    uint32_t first, count;                   // index range (leaf)
};

struct EmitItem { uint32_t n2, me, depth; };

struct Dev {
    // inputs
    const ptl::TrianglePoint* points; const ptl::Triangle* tris; const ptl::RenderObject* objects; const ptl::Material* materials;
    const UpNode* up; const uint32_t* leafOf; const uint32_t* objXform; const uint32_t* instObject; const uint32_t* instTri;
    uint32_t nInst;
    // per-instance records
    PrimRecord* prim; ShadeRecord* shade; UvRecord* uv; WideBox* box;
    // binary build
    uint32_t* idx; uint32_t* idxTmp; Node2* nodes2; uint8_t* valid;
    uint32_t* counters;                      // [0] valid instances [1] binary nodes [10..13] next split lists [5] wide nodes [6] prims out
                                             // [7] next emit list [8] max depth [9] instances dropped (non-finite vertices)
    uint32_t* nextList[4];                   // by triangle count: <= 64, <= 4096, <= HUGE_MIN, above
    int maxLeaf; float travCost;
    // SAH-optimal collapse (k_dp_*): per binary node the cost of its subtree as at most i = 1..8 children of a wide node
    float* dpC; uint8_t* dpK; uint8_t* dpUse; uint32_t* parent; uint32_t* arrived; int dp;
    // output
    Bvh8Node* outNodes; PrimRecord* outPrim; ShadeRecord* outShade; UvRecord* outUv;
    EmitItem* emitNext;
};

__device__ __forceinline__ uint32_t f2o(float f) { uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__device__ __forceinline__ float o2f(uint32_t o) { return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o); }

__global__ void k_wide_records(Dev d) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= d.nInst) return;
    const uint32_t o = d.instObject[i], t = d.instTri[i];
    const ptl::RenderObject ob = d.objects[o];
    const ptl::Triangle tr = d.tris[t];
    const bool em = !(d.materials[ob.materialIndex].emissionStrength == 0.f);
    PrimRecord r; ShadeRecord s; UvRecord u; WideBox b;
    bool ok = wide_make_records(o, t, ob, tr, d.points[tr.v0], d.points[tr.v1], d.points[tr.v2], d.leafOf[t], d.up, d.objXform[o], em, r, s, u, b);
    d.prim[i] = r; d.shade[i] = s; d.uv[i] = u; d.box[i] = b;
    d.idx[i] = i;                                               // instance order: the same tree on every build
    d.valid[i] = ok ? 1 : 0;
    if (!ok) atomicAdd(&d.counters[9], 1u);
}

struct SplitShared {
    uint32_t lo[3], hi[3], cLo[3], cHi[3];
    uint32_t binCount[3][NB];
    uint32_t binLo[3][NB][3], binHi[3][NB][3];
    uint32_t warpSum[32];
    int axis, bin, leaf;
    int axisBin[3]; float axisCost[3];
    uint32_t kids;
};

__device__ __forceinline__ uint32_t block_exclusive(SplitShared& s, bool flag, uint32_t& total) {
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nWarps = blockDim.x >> 5;
    const unsigned ballot = __ballot_sync(0xffffffffu, flag);
    if (lane == 0) s.warpSum[warp] = (uint32_t)__popc(ballot);
    __syncthreads();
    uint32_t before = 0, all = 0;
    for (unsigned w = 0; w < nWarps; ++w) { uint32_t c = s.warpSum[w]; all += c; if (w < warp) before += c; }
    __syncthreads();
    total = all;
    return before + (uint32_t)__popc(ballot & ((1u << lane) - 1u));
}

__device__ __forceinline__ float half_area(const float* lo, const float* hi) {
    float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
    return x * y + y * z + z * x;
}
__device__ __forceinline__ int bin_of(float c, float lo, float scale, int nb = NB) { return min(nb - 1, (int)((c - lo) * scale)); }

// binned SAH along one axis (b8::Builder2::build): best last bin of the left side, first of equal costs
__device__ void sah_axis(const uint32_t* binCount, const uint32_t (*binLo)[3], const uint32_t (*binHi)[3], int nb, int& bestBin, float& bestCost) {
    float rightArea[NB]; uint32_t rightCnt[NB];
    float alo[3] = {1e30f, 1e30f, 1e30f}, ahi[3] = {-1e30f, -1e30f, -1e30f};
    uint32_t cnt = 0;
    for (int k = nb - 1; k > 0; --k) {
        for (int c = 0; c < 3; ++c) { alo[c] = fminf(alo[c], o2f(binLo[k][c])); ahi[c] = fmaxf(ahi[c], o2f(binHi[k][c])); }
        cnt += binCount[k];
        rightArea[k] = cnt ? half_area(alo, ahi) : 0.f; rightCnt[k] = cnt;
    }
    for (int c = 0; c < 3; ++c) { alo[c] = 1e30f; ahi[c] = -1e30f; }
    cnt = 0;
    for (int k = 0; k < nb - 1; ++k) {
        for (int c = 0; c < 3; ++c) { alo[c] = fminf(alo[c], o2f(binLo[k][c])); ahi[c] = fmaxf(ahi[c], o2f(binHi[k][c])); }
        cnt += binCount[k];
        if (cnt == 0 || rightCnt[k + 1] == 0) continue;
        float cost = half_area(alo, ahi) * (float)cnt + rightArea[k + 1] * (float)rightCnt[k + 1];
        if (cost < bestCost) { bestCost = cost; bestBin = k; }
    }
}
// ... over the three axes, the first axis of equal costs
__device__ void sah_best(const uint32_t (*binCount)[NB], const uint32_t (*binLo)[NB][3], const uint32_t (*binHi)[NB][3], const float* ext, int nb,
                         int& bestAxis, int& bestBin, float& bestCost) {
    for (int a = 0; a < 3; ++a) {
        if (!(ext[a] > 0.f)) continue;
        int bin = -1; float cost = bestCost;
        sah_axis(binCount[a], binLo[a], binHi[a], nb, bin, cost);
        if (bin >= 0) { bestCost = cost; bestAxis = a; bestBin = bin; }
    }
}

// b8::Builder2::build for one node (bvh8_build.h), one block per node
__global__ void __launch_bounds__(1024) k_wide_split(Dev d, const uint32_t* __restrict__ list, uint32_t nList) {
    __shared__ SplitShared s;
    const uint32_t tid = threadIdx.x, nT = blockDim.x;
    for (uint32_t item = blockIdx.x; item < nList; item += gridDim.x) {
        const uint32_t me = list[item];
        const uint32_t first = d.nodes2[me].first, n = d.nodes2[me].count;
        const int nb = n > PT_WIDE_MID_MAX ? NB : n > 64 ? NB_MID : NB_SMALL;
        uint32_t* idx = d.idx + first;
        if (tid < 3) { s.lo[tid] = s.cLo[tid] = 0xffffffffu; s.hi[tid] = s.cHi[tid] = 0u; }
        for (uint32_t i = tid; i < 3u * (uint32_t)nb; i += nT) {          // only the bins this node uses
            const uint32_t a = i / (uint32_t)nb, k = i % (uint32_t)nb;
            s.binCount[a][k] = 0u;
            for (int c = 0; c < 3; ++c) { s.binLo[a][k][c] = 0xffffffffu; s.binHi[a][k][c] = 0u; }
        }
        __syncthreads();
        {
            uint32_t lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u}, clo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, chi[3] = {0u, 0u, 0u};
            for (uint32_t k = tid; k < n; k += nT) {
                const WideBox b = d.box[idx[k]];
                for (int a = 0; a < 3; ++a) {
                    uint32_t oc = f2o(0.5f * (b.lo[a] + b.hi[a]));
                    lo[a] = min(lo[a], f2o(b.lo[a])); hi[a] = max(hi[a], f2o(b.hi[a]));
                    clo[a] = min(clo[a], oc); chi[a] = max(chi[a], oc);
                }
            }
            for (int a = 0; a < 3; ++a) {
                uint32_t l = __reduce_min_sync(0xffffffffu, lo[a]), h = __reduce_max_sync(0xffffffffu, hi[a]);
                uint32_t cl = __reduce_min_sync(0xffffffffu, clo[a]), ch = __reduce_max_sync(0xffffffffu, chi[a]);
                if ((tid & 31u) == 0) { atomicMin(&s.lo[a], l); atomicMax(&s.hi[a], h); atomicMin(&s.cLo[a], cl); atomicMax(&s.cHi[a], ch); }
            }
        }
        __syncthreads();
        float lo[3], hi[3], cLo[3], ext[3], scale[3];
        for (int a = 0; a < 3; ++a) {
            lo[a] = o2f(s.lo[a]); hi[a] = o2f(s.hi[a]); cLo[a] = o2f(s.cLo[a]); ext[a] = o2f(s.cHi[a]) - cLo[a];
            scale[a] = (float)nb / ext[a];
        }
        if (tid < 3) { d.nodes2[me].lo[tid] = lo[tid]; d.nodes2[me].hi[tid] = hi[tid]; }
        if (n == 1) { __syncthreads(); continue; }            // leaf of one triangle (first / count already set)
        for (uint32_t k = tid; k < n; k += nT) {
            const WideBox b = d.box[idx[k]];
            for (int a = 0; a < 3; ++a) {
                if (!(ext[a] > 0.f)) continue;
                int bin = bin_of(0.5f * (b.lo[a] + b.hi[a]), cLo[a], scale[a], nb);
                atomicAdd(&s.binCount[a][bin], 1u);
                for (int c = 0; c < 3; ++c) { atomicMin(&s.binLo[a][bin][c], f2o(b.lo[c])); atomicMax(&s.binHi[a][bin][c], f2o(b.hi[c])); }
            }
        }
        __syncthreads();
        if (tid < 3) {                                        // one thread per axis; merged below in axis order
            s.axisBin[tid] = -1; s.axisCost[tid] = 1e30f;
            if (ext[tid] > 0.f) sah_axis(s.binCount[tid], s.binLo[tid], s.binHi[tid], nb, s.axisBin[tid], s.axisCost[tid]);
        }
        __syncthreads();
        if (tid == 0) {
            int bestAxis = -1, bestBin = -1;
            float bestCost = 1e30f;
            for (int a = 0; a < 3; ++a) if (s.axisBin[a] >= 0 && s.axisCost[a] < bestCost) { bestCost = s.axisCost[a]; bestAxis = a; bestBin = s.axisBin[a]; }
            int leaf = 0;
            if (n <= (uint32_t)d.maxLeaf) {                   // leaves hold at most 3 triangles (unary count of a node slot)
                float area = half_area(lo, hi);
                float leafCost = area * (float)n;
                float splitCost = bestAxis < 0 ? 1e30f : area * d.travCost + bestCost;
                if (!(splitCost < leafCost)) leaf = 1;
            }
            s.axis = bestAxis; s.bin = bestBin; s.leaf = leaf;
            if (!leaf) s.kids = atomicAdd(&d.counters[1], 2u);
        }
        __syncthreads();
        if (s.leaf) { __syncthreads(); continue; }
        uint32_t mid = n / 2;
        if (s.axis >= 0) {                                    // stable partition: bin <= bestBin first
            const int a = s.axis, bb = s.bin;
            uint32_t nLeft = 0;
            for (uint32_t k = tid; k < n; k += nT) {
                const WideBox b = d.box[idx[k]];
                if (bin_of(0.5f * (b.lo[a] + b.hi[a]), cLo[a], scale[a], nb) <= bb) nLeft++;
            }
            nLeft = __reduce_add_sync(0xffffffffu, nLeft);
            if (tid == 0) s.warpSum[0] = 0;
            __syncthreads();
            if ((tid & 31u) == 0 && nLeft) atomicAdd(&s.warpSum[0], nLeft);
            __syncthreads();
            const uint32_t totalLeft = s.warpSum[0];
            __syncthreads();
            if (totalLeft != 0 && totalLeft != n) {
                uint32_t* tmp = d.idxTmp + first;
                uint32_t runL = 0, runR = 0;
                for (uint32_t base = 0; base < n; base += nT) {
                    const uint32_t k = base + tid;
                    uint32_t p = 0; bool isLeft = false;
                    if (k < n) { p = idx[k]; const WideBox b = d.box[p]; isLeft = bin_of(0.5f * (b.lo[a] + b.hi[a]), cLo[a], scale[a], nb) <= bb; }
                    uint32_t tot;
                    uint32_t ex = block_exclusive(s, isLeft, tot);
                    if (k < n) tmp[isLeft ? runL + ex : totalLeft + runR + (k - base - ex)] = p;
                    runL += tot; runR += min(nT, n - base) - tot;
                }
                __syncthreads();
                for (uint32_t k = tid; k < n; k += nT) idx[k] = tmp[k];
                mid = totalLeft;
            }
        }
        __syncthreads();                                      // idx of this range is final
        if (tid < 2) {
            Node2 c;
            for (int a = 0; a < 3; ++a) { c.lo[a] = 0.f; c.hi[a] = 0.f; }
            c.left = c.right = 0;
            c.first = tid ? first + mid : first;
            c.count = tid ? n - mid : mid;
            const uint32_t kid = s.kids + tid;
            if (c.count == 1) {                               // single triangle: its box is the node's box, nothing to split
                const WideBox b = d.box[d.idx[c.first]];
                for (int a = 0; a < 3; ++a) { c.lo[a] = b.lo[a]; c.hi[a] = b.hi[a]; }
            }
            d.nodes2[kid] = c;
            if (c.count > 1) {
                const int cls = c.count <= 64 ? 0 : (c.count <= 4096 ? 1 : (c.count <= HUGE_MIN ? 2 : 3));
                d.nextList[cls][atomicAdd(&d.counters[10 + cls], 1u)] = kid;
            }
            if (tid == 0) { d.nodes2[me].left = s.kids; d.nodes2[me].right = s.kids + 1; d.nodes2[me].count = 0; }
        }
        __syncthreads();
    }
}

// ---- nodes above HUGE_MIN triangles: the same steps as k_wide_split, spread over all SMs ----------------------------
// One block per node would leave the first levels of a large scene on a handful of SMs (10 M triangles: 282 of
// 338 ms in the top seven levels). Such nodes are cut into chunks of HUGE_CHUNK positions; each step is one launch
// over all chunks of all huge nodes of the level, with per-node accumulators in global memory.
constexpr uint32_t HUGE_CHUNK = 4096;
constexpr uint32_t HUGE_THREADS = 256;

struct HugeNode {
    uint32_t id, first, count, chunk0, nChunks;
    uint32_t lo[3], hi[3], cLo[3], cHi[3];   // ordered uints
    uint32_t binCount[3][NB];
    uint32_t binLo[3][NB][3], binHi[3][NB][3];
    int axis, bin;
    uint32_t totalLeft, doPartition, kids;
};
struct HugeChunk { uint32_t slot, begin, len; };

__global__ void k_huge_init(Dev d, HugeNode* hn, uint32_t nHuge, const uint32_t* __restrict__ list) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nHuge) return;
    HugeNode& h = hn[i];
    h.id = list[i]; h.first = d.nodes2[h.id].first; h.count = d.nodes2[h.id].count;
    for (int a = 0; a < 3; ++a) { h.lo[a] = h.cLo[a] = 0xffffffffu; h.hi[a] = h.cHi[a] = 0u; }
    for (int k = 0; k < 3 * NB; ++k) {
        (&h.binCount[0][0])[k] = 0u;
        for (int c = 0; c < 3; ++c) { (&h.binLo[0][0][0])[k * 3 + c] = 0xffffffffu; (&h.binHi[0][0][0])[k * 3 + c] = 0u; }
    }
    h.axis = -1; h.bin = -1; h.totalLeft = 0; h.doPartition = 0; h.kids = 0;
}

__global__ void __launch_bounds__(HUGE_THREADS) k_huge_bounds(Dev d, HugeNode* hn, const HugeChunk* __restrict__ chunks) {
    const HugeChunk ck = chunks[blockIdx.x];
    HugeNode& h = hn[ck.slot];
    const uint32_t* idx = d.idx + h.first + ck.begin;
    uint32_t lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u}, clo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, chi[3] = {0u, 0u, 0u};
    for (uint32_t k = threadIdx.x; k < ck.len; k += HUGE_THREADS) {
        const WideBox b = d.box[idx[k]];
        for (int a = 0; a < 3; ++a) {
            uint32_t oc = f2o(0.5f * (b.lo[a] + b.hi[a]));
            lo[a] = min(lo[a], f2o(b.lo[a])); hi[a] = max(hi[a], f2o(b.hi[a]));
            clo[a] = min(clo[a], oc); chi[a] = max(chi[a], oc);
        }
    }
    __shared__ uint32_t sLo[3], sHi[3], sCLo[3], sCHi[3];
    if (threadIdx.x < 3) { sLo[threadIdx.x] = sCLo[threadIdx.x] = 0xffffffffu; sHi[threadIdx.x] = sCHi[threadIdx.x] = 0u; }
    __syncthreads();
    for (int a = 0; a < 3; ++a) {
        uint32_t l = __reduce_min_sync(0xffffffffu, lo[a]), hh = __reduce_max_sync(0xffffffffu, hi[a]);
        uint32_t cl = __reduce_min_sync(0xffffffffu, clo[a]), ch = __reduce_max_sync(0xffffffffu, chi[a]);
        if ((threadIdx.x & 31u) == 0) { atomicMin(&sLo[a], l); atomicMax(&sHi[a], hh); atomicMin(&sCLo[a], cl); atomicMax(&sCHi[a], ch); }
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        atomicMin(&h.lo[threadIdx.x], sLo[threadIdx.x]); atomicMax(&h.hi[threadIdx.x], sHi[threadIdx.x]);
        atomicMin(&h.cLo[threadIdx.x], sCLo[threadIdx.x]); atomicMax(&h.cHi[threadIdx.x], sCHi[threadIdx.x]);
    }
}

__global__ void __launch_bounds__(HUGE_THREADS) k_huge_bins(Dev d, HugeNode* hn, const HugeChunk* __restrict__ chunks) {
    __shared__ uint32_t binCount[3][NB], binLo[3][NB][3], binHi[3][NB][3];
    const HugeChunk ck = chunks[blockIdx.x];
    HugeNode& h = hn[ck.slot];
    const uint32_t* idx = d.idx + h.first + ck.begin;
    for (uint32_t i = threadIdx.x; i < 3 * NB; i += HUGE_THREADS) {
        (&binCount[0][0])[i] = 0u;
        for (int c = 0; c < 3; ++c) { (&binLo[0][0][0])[i * 3 + c] = 0xffffffffu; (&binHi[0][0][0])[i * 3 + c] = 0u; }
    }
    float cLo[3], ext[3], scale[3];
    for (int a = 0; a < 3; ++a) { cLo[a] = o2f(h.cLo[a]); ext[a] = o2f(h.cHi[a]) - cLo[a]; scale[a] = (float)NB / ext[a]; }
    __syncthreads();
    for (uint32_t k = threadIdx.x; k < ck.len; k += HUGE_THREADS) {
        const WideBox b = d.box[idx[k]];
        for (int a = 0; a < 3; ++a) {
            if (!(ext[a] > 0.f)) continue;
            int bin = bin_of(0.5f * (b.lo[a] + b.hi[a]), cLo[a], scale[a]);
            atomicAdd(&binCount[a][bin], 1u);
            for (int c = 0; c < 3; ++c) { atomicMin(&binLo[a][bin][c], f2o(b.lo[c])); atomicMax(&binHi[a][bin][c], f2o(b.hi[c])); }
        }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < 3 * NB; i += HUGE_THREADS) {
        uint32_t cnt = (&binCount[0][0])[i];
        if (!cnt) continue;
        atomicAdd(&(&h.binCount[0][0])[i], cnt);
        for (int c = 0; c < 3; ++c) { atomicMin(&(&h.binLo[0][0][0])[i * 3 + c], (&binLo[0][0][0])[i * 3 + c]); atomicMax(&(&h.binHi[0][0][0])[i * 3 + c], (&binHi[0][0][0])[i * 3 + c]); }
    }
}

__global__ void k_huge_sah(Dev d, HugeNode* hn, uint32_t nHuge) {       // huge nodes are never leaves (count > maxLeaf)
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nHuge) return;
    HugeNode& h = hn[i];
    float ext[3];
    for (int a = 0; a < 3; ++a) { d.nodes2[h.id].lo[a] = o2f(h.lo[a]); d.nodes2[h.id].hi[a] = o2f(h.hi[a]); ext[a] = o2f(h.cHi[a]) - o2f(h.cLo[a]); }
    int bestAxis = -1, bestBin = -1;
    float bestCost = 1e30f;
    sah_best(h.binCount, h.binLo, h.binHi, ext, NB, bestAxis, bestBin, bestCost);
    h.axis = bestAxis; h.bin = bestBin;
    h.kids = atomicAdd(&d.counters[1], 2u);
}

__device__ __forceinline__ bool huge_is_left(const Dev& d, const HugeNode& h, uint32_t p, float cLo, float scale) {
    const WideBox b = d.box[p];
    return bin_of(0.5f * (b.lo[h.axis] + b.hi[h.axis]), cLo, scale) <= h.bin;
}

__global__ void __launch_bounds__(HUGE_THREADS) k_huge_count(Dev d, HugeNode* hn, const HugeChunk* __restrict__ chunks, uint32_t* chunkLeft) {
    const HugeChunk ck = chunks[blockIdx.x];
    const HugeNode& h = hn[ck.slot];
    uint32_t nLeft = 0;
    if (h.axis >= 0) {
        const float cLo = o2f(h.cLo[h.axis]), scale = (float)NB / (o2f(h.cHi[h.axis]) - cLo);
        const uint32_t* idx = d.idx + h.first + ck.begin;
        for (uint32_t k = threadIdx.x; k < ck.len; k += HUGE_THREADS) nLeft += huge_is_left(d, h, idx[k], cLo, scale) ? 1u : 0u;
    }
    __shared__ uint32_t total;
    if (threadIdx.x == 0) total = 0;
    __syncthreads();
    nLeft = __reduce_add_sync(0xffffffffu, nLeft);
    if ((threadIdx.x & 31u) == 0 && nLeft) atomicAdd(&total, nLeft);
    __syncthreads();
    if (threadIdx.x == 0) chunkLeft[blockIdx.x] = total;
}

// per node: exclusive prefix of its chunks' left counts (in place) and the decision whether the partition is proper
__global__ void k_huge_scan(HugeNode* hn, uint32_t nHuge, uint32_t* chunkLeft) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nHuge) return;
    HugeNode& h = hn[i];
    uint32_t run = 0;
    for (uint32_t c = h.chunk0; c < h.chunk0 + h.nChunks; ++c) { uint32_t v = chunkLeft[c]; chunkLeft[c] = run; run += v; }
    h.totalLeft = run;
    h.doPartition = (h.axis >= 0 && run != 0 && run != h.count) ? 1u : 0u;
}

__global__ void __launch_bounds__(HUGE_THREADS) k_huge_scatter(Dev d, HugeNode* hn, const HugeChunk* __restrict__ chunks, const uint32_t* __restrict__ chunkLeft) {
    __shared__ SplitShared s;
    const HugeChunk ck = chunks[blockIdx.x];
    const HugeNode& h = hn[ck.slot];
    if (!h.doPartition) return;
    const float cLo = o2f(h.cLo[h.axis]), scale = (float)NB / (o2f(h.cHi[h.axis]) - cLo);
    const uint32_t* idx = d.idx + h.first + ck.begin;
    uint32_t* tmp = d.idxTmp + h.first;
    uint32_t runL = chunkLeft[blockIdx.x], runR = ck.begin - runL;           // lefts / rights of the node before this chunk
    for (uint32_t base = 0; base < ck.len; base += HUGE_THREADS) {
        const uint32_t k = base + threadIdx.x;
        uint32_t p = 0; bool isLeft = false;
        if (k < ck.len) { p = idx[k]; isLeft = huge_is_left(d, h, p, cLo, scale); }
        uint32_t tot;
        uint32_t ex = block_exclusive(s, isLeft, tot);
        if (k < ck.len) tmp[isLeft ? runL + ex : h.totalLeft + runR + (threadIdx.x - ex)] = p;
        runL += tot; runR += min(HUGE_THREADS, ck.len - base) - tot;
    }
}

__global__ void __launch_bounds__(HUGE_THREADS) k_huge_copyback(Dev d, HugeNode* hn, const HugeChunk* __restrict__ chunks) {
    const HugeChunk ck = chunks[blockIdx.x];
    const HugeNode& h = hn[ck.slot];
    if (!h.doPartition) return;
    const uint32_t o = h.first + ck.begin;
    for (uint32_t k = threadIdx.x; k < ck.len; k += HUGE_THREADS) d.idx[o + k] = d.idxTmp[o + k];
}

__global__ void k_huge_children(Dev d, HugeNode* hn, uint32_t nHuge) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * nHuge) return;
    const HugeNode& h = hn[i >> 1];
    const uint32_t side = i & 1u;
    const uint32_t mid = h.doPartition ? h.totalLeft : h.count / 2;
    Node2 c;
    for (int a = 0; a < 3; ++a) { c.lo[a] = 0.f; c.hi[a] = 0.f; }
    c.left = c.right = 0;
    c.first = side ? h.first + mid : h.first;
    c.count = side ? h.count - mid : mid;
    const uint32_t kid = h.kids + side;
    if (c.count == 1) {
        const WideBox b = d.box[d.idx[c.first]];
        for (int a = 0; a < 3; ++a) { c.lo[a] = b.lo[a]; c.hi[a] = b.hi[a]; }
    }
    d.nodes2[kid] = c;
    if (c.count > 1) {
        const int cls = c.count <= 64 ? 0 : (c.count <= 4096 ? 1 : (c.count <= HUGE_MIN ? 2 : 3));
        d.nextList[cls][atomicAdd(&d.counters[10 + cls], 1u)] = kid;
    }
    if (side == 0) { d.nodes2[h.id].left = h.kids; d.nodes2[h.id].right = h.kids + 1; d.nodes2[h.id].count = 0; }
}

__device__ __forceinline__ uint8_t exponent_for(float extent) {     // smallest e with 2^e * 255 >= extent, biased
    if (!(extent > 0.f)) return 0;
    int e;
    (void)frexpf(extent / 255.f, &e);
    int b = e + 127;
    return (uint8_t)max(1, min(254, b));
}

// ---- SAH-optimal collapse of the binary tree into 8-wide nodes (Ylitie, Karras, Laine 2017, section 3.1; b8::Builder2::
// collapse_dp is the host form), with the binary leaves as the leaves:
//   C(n, 1) = A(n) + D(n, 8)                              n becomes a wide node over at most 8 children
//   D(n, j) = min over 0 < k < j of C(left, k) + C(right, j - k)
//   C(n, i) = min(D(n, i), C(n, i - 1))                   the subtree of n as at most i children of some wide node
// dpK[n][j] = the k of D(n, j); bit i - 1 of dpUse[n] = C(n, i) takes D(n, i). Bottom-up: one thread per binary node starts
// at the leaves, and the second child to arrive at a parent computes it (children are read around L1: volatile).
__global__ void k_dp_parent(Dev d, uint32_t nNodes2) {
    const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= nNodes2) return;
    if (m == 0) d.parent[0] = 0xffffffffu;
    const Node2 nd = d.nodes2[m];
    if (nd.count == 0) { d.parent[nd.left] = m; d.parent[nd.right] = m; }
}
__global__ void k_dp_up(Dev d, uint32_t nNodes2) {
    uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= nNodes2) return;
    {
        const Node2 nd = d.nodes2[m];
        if (nd.count == 0) return;
        const float c = half_area(nd.lo, nd.hi) * (float)nd.count;
        for (int i = 1; i <= 8; ++i) d.dpC[(size_t)m * 8 + i - 1] = c;
    }
    for (;;) {
        const uint32_t p = d.parent[m];
        if (p == 0xffffffffu) return;
        __threadfence();
        if (atomicAdd(&d.arrived[p], 1u) == 0u) return;           // the sibling's subtree is not finished: its thread goes on
        __threadfence();
        const Node2 nd = d.nodes2[p];
        const volatile float* L = d.dpC + (size_t)nd.left * 8;
        const volatile float* R = d.dpC + (size_t)nd.right * 8;
        float l[8], r[8];
        for (int i = 0; i < 8; ++i) { l[i] = L[i]; r[i] = R[i]; }
        float D[9];
        for (int j = 2; j <= 8; ++j) {
            float best = 3.0e38f; int bk = 1;
            for (int k = 1; k < j; ++k) { const float c = l[k - 1] + r[j - k - 1]; if (c < best) { best = c; bk = k; } }
            D[j] = best; d.dpK[(size_t)p * 8 + j - 1] = (uint8_t)bk;
        }
        float C = half_area(nd.lo, nd.hi) + D[8];
        uint32_t use = 0;
        d.dpC[(size_t)p * 8] = C;
        for (int i = 2; i <= 8; ++i) {
            if (D[i] < C) { C = D[i]; use |= 1u << (i - 1); }
            d.dpC[(size_t)p * 8 + i - 1] = C;
        }
        d.dpUse[p] = (uint8_t)use;
        m = p;
    }
}

// bvh8_emit (bvh8_build.h) for one wide node: gather up to 8 children by opening the largest inner child, assign
// octant-ordered slots, quantise, copy the leaf triangles' records, queue the inner children for the next level.
__global__ void k_wide_emit(Dev d, const EmitItem* __restrict__ items, uint32_t nItems) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nItems) return;
    const EmitItem it = items[i];
    atomicMax(&d.counters[8], it.depth);
    uint32_t ch[8]; Node2 cn[8]; int nch = 0;
    const Node2 root = d.nodes2[it.n2];
    if (root.count) { ch[0] = it.n2; cn[0] = root; nch = 1; }
    else if (d.dp) {
        // the children the DP chose: D(root, 8) splits its 8 places between the two subtrees, and so on down
        uint32_t stN[16]; int stI[16]; int sp = 0;
        const int k8 = d.dpK[(size_t)it.n2 * 8 + 7];
        stN[sp] = root.right; stI[sp++] = 8 - k8;
        stN[sp] = root.left; stI[sp++] = k8;
        while (sp > 0) {
            const uint32_t m = stN[--sp]; int i = stI[sp];
            const Node2 nd = d.nodes2[m];
            if (nd.count == 0) {
                const uint32_t use = d.dpUse[m];
                while (i > 1 && !(use & (1u << (i - 1)))) i--;
            }
            if (nd.count != 0 || i == 1 || nch + sp >= 7) { ch[nch] = m; cn[nch] = nd; nch++; continue; }    // (the last test cannot trigger: the places add up to 8)
            const int k = d.dpK[(size_t)m * 8 + i - 1];
            stN[sp] = nd.right; stI[sp++] = i - k;
            stN[sp] = nd.left; stI[sp++] = k;
        }
    }
    else { ch[0] = root.left; ch[1] = root.right; cn[0] = d.nodes2[ch[0]]; cn[1] = d.nodes2[ch[1]]; nch = 2; }
    while (!d.dp && nch < 8) {
        int best = -1; float bestA = -1.f;
        for (int k = 0; k < nch; ++k) {
            if (cn[k].count) continue;
            float a = half_area(cn[k].lo, cn[k].hi);
            if (a > bestA) { bestA = a; best = k; }
        }
        if (best < 0) break;
        const uint32_t l = cn[best].left, r = cn[best].right;
        ch[best] = l; cn[best] = d.nodes2[l];
        ch[nch] = r; cn[nch] = d.nodes2[r]; nch++;
    }
    // slots: greedy assignment minimising (child centre - node centre) . d_s (traversal priority is slot ^ (7 - octant))
    float ncx[3];
    for (int a = 0; a < 3; ++a) ncx[a] = 0.5f * (root.lo[a] + root.hi[a]);
    float cost[8][8];
    for (int c = 0; c < nch; ++c) {
        float dd[3];
        for (int a = 0; a < 3; ++a) dd[a] = 0.5f * (cn[c].lo[a] + cn[c].hi[a]) - ncx[a];
        for (int sl = 0; sl < 8; ++sl)
            cost[c][sl] = dd[0] * ((sl & 4) ? -1.f : 1.f) + dd[1] * ((sl & 2) ? -1.f : 1.f) + dd[2] * ((sl & 1) ? -1.f : 1.f);
    }
    int slotOf[8], childIn[8];
    for (int k = 0; k < 8; ++k) { slotOf[k] = -1; childIn[k] = -1; }
    for (int k = 0; k < nch; ++k) {
        int bc = -1, bs = -1; float bv = 1e30f;
        for (int c = 0; c < nch; ++c) if (slotOf[c] < 0)
            for (int sl = 0; sl < 8; ++sl) if (childIn[sl] < 0 && cost[c][sl] < bv) { bv = cost[c][sl]; bc = c; bs = sl; }
        if (bc < 0) {                                          // non-finite costs: first free pair
            for (int c = 0; c < nch && bc < 0; ++c) if (slotOf[c] < 0) bc = c;
            for (int sl = 0; sl < 8 && bs < 0; ++sl) if (childIn[sl] < 0) bs = sl;
        }
        slotOf[bc] = bs; childIn[bs] = bc;
    }
    Bvh8Node node;
    memset(&node, 0, sizeof(node));
    node.px = root.lo[0]; node.py = root.lo[1]; node.pz = root.lo[2];
    uint8_t eb[3]; float sc[3];
    for (int a = 0; a < 3; ++a) {
        eb[a] = exponent_for(root.hi[a] - root.lo[a]);
        for (;;) {                                             // the largest plane must be representable after rounding
            sc[a] = __uint_as_float((uint32_t)eb[a] << 23);
            if (eb[a] == 0 || eb[a] >= 254 || root.lo[a] + 255.f * sc[a] >= root.hi[a]) break;
            eb[a]++;
        }
    }
    node.ex = eb[0]; node.ey = eb[1]; node.ez = eb[2];
    uint32_t nInner = 0, nLeafTris = 0;
    for (int sl = 0; sl < 8; ++sl) if (childIn[sl] >= 0) { if (cn[childIn[sl]].count == 0) nInner++; else nLeafTris += cn[childIn[sl]].count; }
    node.childBase = nInner ? atomicAdd(&d.counters[5], nInner) : 0u;
    node.primBase = nLeafTris ? atomicAdd(&d.counters[6], nLeafTris) : 0u;
    const uint32_t emitBase = nInner ? atomicAdd(&d.counters[7], nInner) : 0u;
    uint32_t innerSeen = 0, primOff = 0;
    for (int sl = 0; sl < 8; ++sl) {
        const int c = childIn[sl];
        if (c < 0) continue;
        const Node2& k = cn[c];
        uint8_t* qlo[3] = {node.qlox, node.qloy, node.qloz};
        uint8_t* qhi[3] = {node.qhix, node.qhiy, node.qhiz};
        const float p[3] = {node.px, node.py, node.pz};
        for (int a = 0; a < 3; ++a) {
            int l = 0, h = 0;
            if (sc[a] > 0.f) {
                l = (int)floorf((k.lo[a] - p[a]) / sc[a]);
                h = (int)ceilf((k.hi[a] - p[a]) / sc[a]);
                l = max(0, min(255, l)); h = max(0, min(255, h));
                while (l > 0 && p[a] + (float)l * sc[a] > k.lo[a]) l--;        // outward after rounding
                while (h < 255 && p[a] + (float)h * sc[a] < k.hi[a]) h++;
            }
            qlo[a][sl] = (uint8_t)l; qhi[a][sl] = (uint8_t)h;
        }
        if (k.count == 0) {
            node.imask |= (uint8_t)(1u << sl);
            node.meta[sl] = (uint8_t)((1u << 5) | (24u + sl));
            EmitItem nx; nx.n2 = ch[c]; nx.me = node.childBase + innerSeen; nx.depth = it.depth + 1;
            d.emitNext[emitBase + innerSeen] = nx;
            innerSeen++;
        } else {
            const uint32_t unary = k.count == 1 ? 1u : (k.count == 2 ? 3u : 7u);
            node.meta[sl] = (uint8_t)((unary << 5) | primOff);
            for (uint32_t q = 0; q < k.count; ++q) {
                const uint32_t id = d.idx[k.first + q], dst = node.primBase + primOff + q;
                d.outPrim[dst] = d.prim[id]; d.outShade[dst] = d.shade[id]; d.outUv[dst] = d.uv[id];
            }
            primOff += k.count;
        }
    }
    d.outNodes[it.me] = node;
}

}  // namespace wb

namespace {
thread_local char g_msg[256];
}  // namespace

#define WBC(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { snprintf(g_msg, sizeof g_msg, "wide BVH device build: %s: %s", #x, cudaGetErrorString(e_)); return g_msg; } } while (0)

const char* wide_build_device(cudaStream_t st, const WideDeviceInputs& in, const WideSceneHost& host, WideDeviceResult& out, WideBuildCache* cache) {
    using namespace wb;
    const size_t n = host.instTri.size();
    Dev d{};
    d.points = in.points; d.tris = in.tris; d.objects = in.objects; d.materials = in.materials;
    d.nInst = (uint32_t)n;
    d.maxLeaf = 3; d.travCost = 0.5f;        // measured: 0.5 / 0.25 are ~0.6 % faster to trace than 1.0 on the C4 stand-in
    if (const char* e = std::getenv("B200PT_BVH_LEAF")) d.maxLeaf = std::max(1, std::min(3, std::atoi(e)));
    if (const char* e = std::getenv("B200PT_BVH_CT")) d.travCost = (float)std::atof(e);
    UpNode* up; uint32_t *leafOf, *objXform, *instObject, *instTri, *lists[8], *chunkLeft;
    EmitItem* emit[2];
    HugeNode* hugeNodes; HugeChunk* hugeChunks;
    // One arena for all scratch (cudaMalloc / cudaFree of two dozen buffers cost as much as the kernels for a
    // 265 k-triangle scene); kept in `cache` between builds while it is small enough to be worth holding on to.
    WideBuildCache local;
    WideBuildCache& arena = cache ? *cache : local;
    size_t need = 0;
    auto plan = [&](size_t count, size_t elem) { size_t off = need; need += (std::max<size_t>(count, 1) * elem + 255) & ~(size_t)255; return off; };
    const size_t maxHuge = n / HUGE_MIN + 1, maxChunks = n / HUGE_CHUNK + maxHuge + 1;
    const size_t oUp = plan(host.up.size(), sizeof(UpNode)), oLeaf = plan(host.leafOf.size(), 4), oObjX = plan(host.objXform.size(), 4),
                 oInstO = plan(n, 4), oInstT = plan(n, 4), oPrim = plan(n, sizeof(PrimRecord)), oShade = plan(n, sizeof(ShadeRecord)),
                 oUv = plan(n, sizeof(UvRecord)), oBox = plan(n, sizeof(WideBox)), oIdx = plan(n, 4), oIdxT = plan(n, 4), oValid = plan(n, 1),
                 oNodes2 = plan(2 * n + 1, sizeof(Node2)), oCnt = plan(16, 4), oEmit0 = plan(n + 1, sizeof(EmitItem)), oEmit1 = plan(n + 1, sizeof(EmitItem)),
                 oHuge = plan(maxHuge, sizeof(HugeNode)), oChunks = plan(maxChunks, sizeof(HugeChunk)), oChunkLeft = plan(maxChunks, 4);
    d.dp = 1;                                 // B200PT_BVH_COLLAPSE=greedy: open the largest child instead (the round-1 collapse)
    if (const char* e = std::getenv("B200PT_BVH_COLLAPSE")) d.dp = std::string(e) == "greedy" ? 0 : 1;
    const size_t oDpC = plan(d.dp ? 2 * n + 1 : 1, 32), oDpK = plan(d.dp ? 2 * n + 1 : 1, 8), oDpUse = plan(d.dp ? 2 * n + 1 : 1, 1),
                 oParent = plan(d.dp ? 2 * n + 1 : 1, 4), oArrived = plan(d.dp ? 2 * n + 1 : 1, 4);
    size_t oList[8];
    for (auto& o : oList) o = plan(n / 2 + 1, 4);              // a listed node holds at least 2 triangles
    if (arena.bytes < need) {
        if (arena.p) cudaFree(arena.p);
        arena.p = nullptr; arena.bytes = 0;
        WBC(cudaMalloc(&arena.p, need));
        arena.bytes = need;
    }
    char* base = (char*)arena.p;
    up = (UpNode*)(base + oUp); leafOf = (uint32_t*)(base + oLeaf); objXform = (uint32_t*)(base + oObjX);
    instObject = (uint32_t*)(base + oInstO); instTri = (uint32_t*)(base + oInstT);
    d.prim = (PrimRecord*)(base + oPrim); d.shade = (ShadeRecord*)(base + oShade); d.uv = (UvRecord*)(base + oUv); d.box = (WideBox*)(base + oBox);
    d.idx = (uint32_t*)(base + oIdx); d.idxTmp = (uint32_t*)(base + oIdxT); d.valid = (uint8_t*)(base + oValid);
    d.nodes2 = (Node2*)(base + oNodes2); d.counters = (uint32_t*)(base + oCnt);
    emit[0] = (EmitItem*)(base + oEmit0); emit[1] = (EmitItem*)(base + oEmit1);
    hugeNodes = (HugeNode*)(base + oHuge); hugeChunks = (HugeChunk*)(base + oChunks); chunkLeft = (uint32_t*)(base + oChunkLeft);
    for (int i = 0; i < 8; ++i) lists[i] = (uint32_t*)(base + oList[i]);
    d.dpC = (float*)(base + oDpC); d.dpK = (uint8_t*)(base + oDpK); d.dpUse = (uint8_t*)(base + oDpUse);
    d.parent = (uint32_t*)(base + oParent); d.arrived = (uint32_t*)(base + oArrived);
    // the four result buffers: the caller's previous ones are written again when they are large enough
    struct Fresh { std::vector<void*> p; ~Fresh() { for (void* q : p) cudaFree(q); } } fresh;      // freed unless handed over
    const size_t capNodes = (n + 1) * sizeof(Bvh8Node), capPrim = std::max<size_t>(n, 1) * sizeof(PrimRecord),
                 capShade = std::max<size_t>(n, 1) * sizeof(ShadeRecord), capUv = std::max<size_t>(n, 1) * sizeof(UvRecord);
    const bool recycle = out.nodes && out.prims && out.shade && out.uv && out.capacityInstances >= n;
    if (recycle) {
        d.outNodes = (Bvh8Node*)out.nodes; d.outPrim = (PrimRecord*)out.prims; d.outShade = (ShadeRecord*)out.shade; d.outUv = (UvRecord*)out.uv;
    } else {
        void* q;
        WBC(cudaMalloc(&q, capNodes)); fresh.p.push_back(q); d.outNodes = (Bvh8Node*)q;
        WBC(cudaMalloc(&q, capPrim)); fresh.p.push_back(q); d.outPrim = (PrimRecord*)q;
        WBC(cudaMalloc(&q, capShade)); fresh.p.push_back(q); d.outShade = (ShadeRecord*)q;
        WBC(cudaMalloc(&q, capUv)); fresh.p.push_back(q); d.outUv = (UvRecord*)q;
    }
    auto h2d = [&](void* dst, const void* src, size_t bytes) { return bytes ? cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st) : cudaSuccess; };
    WBC(h2d(up, host.up.data(), host.up.size() * sizeof(UpNode)));
    WBC(h2d(leafOf, host.leafOf.data(), host.leafOf.size() * 4));
    WBC(h2d(objXform, host.objXform.data(), host.objXform.size() * 4));
    WBC(h2d(instObject, host.instObject.data(), n * 4));
    WBC(h2d(instTri, host.instTri.data(), n * 4));
    d.up = up; d.leafOf = leafOf; d.objXform = objXform; d.instObject = instObject; d.instTri = instTri;
    struct Events {
        cudaEvent_t e[2] = {nullptr, nullptr};
        ~Events() { for (cudaEvent_t x : e) if (x) cudaEventDestroy(x); }
    } ev;                                                   // destroyed on every exit path
    WBC(cudaEventCreate(&ev.e[0])); WBC(cudaEventCreate(&ev.e[1]));
    cudaEvent_t e0 = ev.e[0], e1 = ev.e[1];
    WBC(cudaMemsetAsync(d.counters, 0, 64, st));
    WBC(cudaEventRecord(e0, st));
    uint32_t h[16] = {0};
    if (n) k_wide_records<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d);
    WBC(cudaMemcpyAsync(h, d.counters, 64, cudaMemcpyDeviceToHost, st));
    WBC(cudaStreamSynchronize(st));
    uint32_t nValid = (uint32_t)n;
    if (h[9]) {                               // rare: drop the instances with non-finite vertices, keeping the order
        std::vector<uint8_t> valid(n);
        WBC(cudaMemcpyAsync(valid.data(), d.valid, n, cudaMemcpyDeviceToHost, st));
        WBC(cudaStreamSynchronize(st));
        std::vector<uint32_t> keep;
        keep.reserve(n);
        for (size_t i = 0; i < n; ++i) if (valid[i]) keep.push_back((uint32_t)i);
        nValid = (uint32_t)keep.size();
        WBC(h2d(d.idx, keep.data(), keep.size() * 4));
        WBC(cudaStreamSynchronize(st));
    }
    if (nValid == 0) {                        // no triangles: one empty node, as the host builder
        Bvh8Node zero; memset(&zero, 0, sizeof zero);
        WBC(cudaMemcpyAsync(d.outNodes, &zero, sizeof zero, cudaMemcpyHostToDevice, st));
        out.nNodes = 1; out.nPrims = 0; out.maxDepth = 0;
    } else {
        Node2 root{};
        root.first = 0; root.count = nValid;
        WBC(cudaMemcpyAsync(d.nodes2, &root, sizeof root, cudaMemcpyHostToDevice, st));
        uint32_t init[16] = {nValid, 1, 0, 0, 0, 1, 0, 0, 0};
        const int cls0 = nValid <= 64 ? 0 : (nValid <= 4096 ? 1 : (nValid <= HUGE_MIN ? 2 : 3));
        init[10 + cls0] = 1;
        uint32_t zeroId = 0;
        WBC(cudaMemcpyAsync(lists[cls0], &zeroId, 4, cudaMemcpyHostToDevice, st));
        WBC(cudaMemcpyAsync(d.counters, init, 64, cudaMemcpyHostToDevice, st));
        const unsigned threadsOf[3] = {32, 256, 1024};
        std::vector<HugeNode> hHuge;
        std::vector<HugeChunk> hChunks;
        int cur = 0;
        const bool timing = std::getenv("B200PT_BUILD_TIMING") != nullptr;
        auto tLevel = std::chrono::steady_clock::now();
        for (int level = 0;; ++level) {
            WBC(cudaMemcpyAsync(h, d.counters, 64, cudaMemcpyDeviceToHost, st));
            WBC(cudaStreamSynchronize(st));
            if (timing) {
                auto now = std::chrono::steady_clock::now();
                fprintf(stderr, "[wide] level %d done in %.3f ms; next lists %u / %u / %u / %u nodes\n", level - 1,
                        std::chrono::duration<double, std::milli>(now - tLevel).count(), h[10], h[11], h[12], h[13]);
                tLevel = now;
            }
            if (h[10] + h[11] + h[12] + h[13] == 0) break;
            if (level > 4096) return "wide BVH device build: binary tree does not terminate";
            const uint32_t* curList[4] = {lists[cur * 4], lists[cur * 4 + 1], lists[cur * 4 + 2], lists[cur * 4 + 3]};
            cur ^= 1;
            for (int c = 0; c < 4; ++c) d.nextList[c] = lists[cur * 4 + c];
            WBC(cudaMemsetAsync(d.counters + 10, 0, 16, st));
            for (int c = 0; c < 3; ++c)
                if (h[10 + c]) k_wide_split<<<std::min<uint32_t>(h[10 + c], 1u << 20), threadsOf[c], 0, st>>>(d, curList[c], h[10 + c]);
            if (const uint32_t nHuge = h[13]) {
                k_huge_init<<<(nHuge + 63) / 64, 64, 0, st>>>(d, hugeNodes, nHuge, curList[3]);
                hHuge.resize(nHuge);
                WBC(cudaMemcpyAsync(hHuge.data(), hugeNodes, nHuge * sizeof(HugeNode), cudaMemcpyDeviceToHost, st));
                WBC(cudaStreamSynchronize(st));
                hChunks.clear();
                for (uint32_t i = 0; i < nHuge; ++i) {
                    hHuge[i].chunk0 = (uint32_t)hChunks.size();
                    hHuge[i].nChunks = (hHuge[i].count + HUGE_CHUNK - 1) / HUGE_CHUNK;
                    for (uint32_t b = 0; b < hHuge[i].count; b += HUGE_CHUNK) hChunks.push_back({i, b, std::min(HUGE_CHUNK, hHuge[i].count - b)});
                }
                const uint32_t nChunks = (uint32_t)hChunks.size();
                if (nHuge > maxHuge || nChunks > maxChunks) return "wide BVH device build: chunk table overflow";
                WBC(cudaMemcpyAsync(hugeNodes, hHuge.data(), nHuge * sizeof(HugeNode), cudaMemcpyHostToDevice, st));
                WBC(cudaMemcpyAsync(hugeChunks, hChunks.data(), nChunks * sizeof(HugeChunk), cudaMemcpyHostToDevice, st));
                const unsigned gN = (nHuge + 63) / 64;
                k_huge_bounds<<<nChunks, HUGE_THREADS, 0, st>>>(d, hugeNodes, hugeChunks);
                k_huge_bins<<<nChunks, HUGE_THREADS, 0, st>>>(d, hugeNodes, hugeChunks);
                k_huge_sah<<<gN, 64, 0, st>>>(d, hugeNodes, nHuge);
                k_huge_count<<<nChunks, HUGE_THREADS, 0, st>>>(d, hugeNodes, hugeChunks, chunkLeft);
                k_huge_scan<<<gN, 64, 0, st>>>(hugeNodes, nHuge, chunkLeft);
                k_huge_scatter<<<nChunks, HUGE_THREADS, 0, st>>>(d, hugeNodes, hugeChunks, chunkLeft);
                k_huge_copyback<<<nChunks, HUGE_THREADS, 0, st>>>(d, hugeNodes, hugeChunks);
                k_huge_children<<<(2 * nHuge + 63) / 64, 64, 0, st>>>(d, hugeNodes, nHuge);
            }
            WBC(cudaGetLastError());
        }
        if (d.dp) {
            const uint32_t nNodes2 = h[1];
            WBC(cudaMemsetAsync(d.arrived, 0, (size_t)nNodes2 * 4, st));
            k_dp_parent<<<(nNodes2 + 255) / 256, 256, 0, st>>>(d, nNodes2);
            k_dp_up<<<(nNodes2 + 127) / 128, 128, 0, st>>>(d, nNodes2);
            WBC(cudaGetLastError());
        }
        // collapse + emit, level by level from the root (wide node 0 <- binary node 0)
        EmitItem first{0u, 0u, 1u};
        WBC(cudaMemcpyAsync(emit[0], &first, sizeof first, cudaMemcpyHostToDevice, st));
        uint32_t nItems = 1;
        int ecur = 0;
        out.levelCount.clear();
        while (nItems) {
            out.levelCount.push_back(nItems);
            if (timing) {
                auto now = std::chrono::steady_clock::now();
                fprintf(stderr, "[wide] emit level: %u nodes (previous step %.3f ms)\n", nItems, std::chrono::duration<double, std::milli>(now - tLevel).count());
                tLevel = now;
            }
            d.emitNext = emit[ecur ^ 1];
            WBC(cudaMemsetAsync(d.counters + 7, 0, 4, st));
            k_wide_emit<<<(nItems + 63) / 64, 64, 0, st>>>(d, emit[ecur], nItems);
            WBC(cudaGetLastError());
            WBC(cudaMemcpyAsync(h, d.counters, 64, cudaMemcpyDeviceToHost, st));
            WBC(cudaStreamSynchronize(st));
            nItems = h[7];
            ecur ^= 1;
        }
        out.nNodes = h[5]; out.nPrims = h[6]; out.maxDepth = h[8]; out.nDropped = h[9];
        if (out.nPrims != nValid) return "wide BVH device build: triangle count mismatch";
    }
    WBC(cudaEventRecord(e1, st));
    WBC(cudaStreamSynchronize(st));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    out.deviceMs = ms;
    if (!recycle) {
        if (out.nodes) cudaFree(out.nodes);
        if (out.prims) cudaFree(out.prims);
        if (out.shade) cudaFree(out.shade);
        if (out.uv) cudaFree(out.uv);
        if (out.nodeBox) { cudaFree(out.nodeBox); out.nodeBox = nullptr; }
        out.capacityInstances = n;
        fresh.p.clear();                      // handed over
    }
    out.nodes = d.outNodes; out.prims = d.outPrim; out.shade = d.outShade; out.uv = d.outUv;
    if (arena.bytes > WideBuildCache::kKeepBytes && cache) { cudaFree(cache->p); cache->p = nullptr; cache->bytes = 0; }
    return nullptr;
}

// ---- refit -----------------------------------------------------------------------------------------------------
namespace wb {

__global__ void k_wide_set_xform(PrimRecord* prims, uint32_t n, const uint32_t* __restrict__ objXform) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) prims[i].xformId = objXform[prims[i].objectIndex];
}

// one thread per wide node of one level: child boxes from the triangles (leaf children) or from the boxes the level
// below stored (inner children), then origin, exponents and the 8-bit planes exactly as k_wide_emit derives them
__global__ void k_wide_refit(Bvh8Node* nodes, const PrimRecord* __restrict__ prims, const ptl::TrianglePoint* __restrict__ points,
                             const ptl::Triangle* __restrict__ tris, const ptl::RenderObject* __restrict__ objects, float* nodeBox,
                             uint32_t first, uint32_t count, uint32_t* notFinite) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    const uint32_t me = first + k;
    Bvh8Node node = nodes[me];
    float clo[8][3], chi[8][3];
    float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
    uint32_t innerSeen = 0;
    for (int sl = 0; sl < 8; ++sl) {
        const uint32_t m = node.meta[sl];
        if (!m) continue;
        if (node.imask & (1u << sl)) {
            const float* b = nodeBox + 6 * (size_t)(node.childBase + innerSeen);
            for (int a = 0; a < 3; ++a) { clo[sl][a] = b[a]; chi[sl][a] = b[3 + a]; }
            innerSeen++;
        } else {
            const uint32_t unary = m >> 5, off = m & 31u, n = unary == 1u ? 1u : (unary == 3u ? 2u : 3u);
            for (int a = 0; a < 3; ++a) { clo[sl][a] = 1e30f; chi[sl][a] = -1e30f; }
            for (uint32_t q = 0; q < n; ++q) {
                const PrimRecord r = prims[node.primBase + off + q];
                const ptl::Triangle tr = tris[r.triIndexFlags & PT_PRIM_TRIMASK];
                const ptl::TrianglePoint a = points[tr.v0], b = points[tr.v1], c = points[tr.v2];
                WideBox bx;
                if (!wide_world_box(objects[r.objectIndex], mk3(a.px, a.py, a.pz), mk3(b.px, b.py, b.pz), mk3(c.px, c.py, c.pz), bx)) atomicAdd(notFinite, 1u);
                for (int ax = 0; ax < 3; ++ax) { clo[sl][ax] = fminf(clo[sl][ax], bx.lo[ax]); chi[sl][ax] = fmaxf(chi[sl][ax], bx.hi[ax]); }
            }
        }
        for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], clo[sl][a]); hi[a] = fmaxf(hi[a], chi[sl][a]); }
    }
    float* mine = nodeBox + 6 * (size_t)me;
    for (int a = 0; a < 3; ++a) { mine[a] = lo[a]; mine[3 + a] = hi[a]; }
    node.px = lo[0]; node.py = lo[1]; node.pz = lo[2];
    uint8_t eb[3]; float sc[3];
    for (int a = 0; a < 3; ++a) {
        eb[a] = exponent_for(hi[a] - lo[a]);
        for (;;) {
            sc[a] = __uint_as_float((uint32_t)eb[a] << 23);
            if (eb[a] == 0 || eb[a] >= 254 || lo[a] + 255.f * sc[a] >= hi[a]) break;
            eb[a]++;
        }
    }
    node.ex = eb[0]; node.ey = eb[1]; node.ez = eb[2];
    uint8_t* qlo[3] = {node.qlox, node.qloy, node.qloz};
    uint8_t* qhi[3] = {node.qhix, node.qhiy, node.qhiz};
    for (int sl = 0; sl < 8; ++sl) {
        if (!node.meta[sl]) continue;
        for (int a = 0; a < 3; ++a) {
            int l = 0, h = 0;
            if (sc[a] > 0.f) {
                l = (int)floorf((clo[sl][a] - lo[a]) / sc[a]);
                h = (int)ceilf((chi[sl][a] - lo[a]) / sc[a]);
                l = max(0, min(255, l)); h = max(0, min(255, h));
                while (l > 0 && lo[a] + (float)l * sc[a] > clo[sl][a]) l--;        // outward after rounding
                while (h < 255 && lo[a] + (float)h * sc[a] < chi[sl][a]) h++;
            }
            qlo[a][sl] = (uint8_t)l; qhi[a][sl] = (uint8_t)h;
        }
    }
    nodes[me] = node;
}

}  // namespace wb

const char* wide_refit_device(cudaStream_t st, const WideDeviceInputs& in, const uint32_t* objXform, WideDeviceResult& tree, float* deviceMs) {
    using namespace wb;
    if (!tree.nodes || tree.levelCount.empty() || tree.nDropped) return "wide BVH refit: no refittable tree";
    if (!tree.nodeBox) WBC(cudaMalloc(&tree.nodeBox, (size_t)tree.nNodes * 24 + 256));
    uint32_t* flag = (uint32_t*)((char*)tree.nodeBox + (size_t)tree.nNodes * 24);
    cudaEvent_t e0, e1;
    WBC(cudaEventCreate(&e0));
    if (cudaEventCreate(&e1) != cudaSuccess) { cudaEventDestroy(e0); return "wide BVH refit: cudaEventCreate"; }
    cudaEventRecord(e0, st);
    cudaMemsetAsync(flag, 0, 4, st);
    k_wide_set_xform<<<(tree.nPrims + 255) / 256, 256, 0, st>>>((PrimRecord*)tree.prims, tree.nPrims, objXform);
    std::vector<uint32_t> start(tree.levelCount.size());
    uint32_t acc = 0;
    for (size_t l = 0; l < tree.levelCount.size(); ++l) { start[l] = acc; acc += tree.levelCount[l]; }
    for (size_t l = tree.levelCount.size(); l-- > 0;)
        k_wide_refit<<<(tree.levelCount[l] + 63) / 64, 64, 0, st>>>((Bvh8Node*)tree.nodes, (const PrimRecord*)tree.prims, in.points, in.tris, in.objects,
                                                                   (float*)tree.nodeBox, start[l], tree.levelCount[l], flag);
    uint32_t bad = 0;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(&bad, flag, 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaEventRecord(e1, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    float ms = 0.f;
    if (e == cudaSuccess) cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (e != cudaSuccess) return cudaGetErrorString(e);
    if (deviceMs) *deviceMs = ms;
    if (acc != tree.nNodes) return "wide BVH refit: level table does not cover the tree";
    if (bad) return "wide BVH refit: a transformed vertex is not finite";      // the caller rebuilds (such triangles are dropped)
    return nullptr;
}

}  // namespace pt
