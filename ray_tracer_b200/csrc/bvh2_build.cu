// GPU build of the REFERENCE's BVH2: VulkanEngine::build_bvh / update_bvh_bounds / subdivide_bvh /
// find_bvh_split_plane (src/vk_engine.cpp:1169-1337), reproduced bit for bit — node bounds, node numbering and the
// order in which the triangles end up — so that scenes prepared on the GPU are interchangeable with scenes prepared
// by the engine (and by ray_tracer_b200/host/scene_builder.cpp, which tests compare against).
//
// What has to be reproduced, and how it is made parallel:
//  * 20-bin SAH per axis with the reference's sweep quirks (right-hand areas mirrored, 0 * inf = NaN never wins):
//    bins are filled with shared-memory atomics (integer counts; min/max are order independent), the 3 x 19 sweep
//    is evaluated by one thread in the reference's order.
//  * the in-place partition that swaps from the back (:1233-1244). Its result is a fixed permutation of the
//    segment that depends only on which elements are "left"; dest() below gives the closed form (front lefts stay,
//    the s-th hole takes the s-th left from the end, rights are laid down from the end in the order the sequential
//    loop meets them — including the rotation by one it performs when it runs past the last left element).
//  * node bounds: float min/max, where the only order dependence is the sign of a zero extreme (iMin keeps the
//    first of equal values, iMax the last: src/vk_engine.h:83-88). Reductions therefore run on 64-bit keys
//    (ordered value, position in the node, sign bit).
//  * node numbering is the depth-first hand-out of node pairs (nodesUsed += 2 before the left subtree is built).
//    The tree is built level by level into temporary nodes; subtree sizes (bottom-up) give every node its
//    depth-first index (top-down).
// One block per node and level, three block sizes by triangle count; nodes above 32 k triangles (the first levels of a
// large mesh) are instead processed by all SMs: each step of the node's split is one launch over 4096-position chunks. Arithmetic: f32, compiled with -fmad=false
// (the reference's host compiler does not contract); divisions are IEEE.
// Inputs must be finite (the reference's behaviour with NaN coordinates is undefined: float -> int of NaN).
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>
#include <algorithm>

#include "pt_layout.h"

namespace pt { void set_global_error(const char* what); }

namespace bvh2 {

constexpr int BINS = 20;                     // src/vk_engine.h:203
constexpr uint32_t MAX_DEPTH = 64;           // subdivide_bvh :1214
constexpr uint32_t HUGE_MIN = 32768;         // nodes with more triangles are split by many blocks (k_huge_*)
constexpr uint32_t HUGE_CHUNK = 4096;
constexpr uint32_t HUGE_THREADS = 256;
__device__ __forceinline__ int list_class(uint32_t count) { return count <= 64 ? 0 : (count <= 4096 ? 1 : (count <= HUGE_MIN ? 2 : 3)); }

struct TNode {                               // temporary node, breadth-first creation order
    uint32_t first, count;                   // triangle range, relative to the object's first triangle
    int32_t left;                            // temp id of the left child (right = left + 1), -1: leaf
    uint32_t depth;
    float lo[3], hi[3];
    uint32_t size;                           // nodes in this subtree
    uint32_t finalIdx;                       // index in the reference's bvhNodes
    uint32_t base;                           // inner: index of its node pair (= nodesUsed when it was subdivided)
};

struct Dev {
    const float* cen[3];                     // centroid per ORIGINAL triangle id
    const float* tlo[3];                     // triangle box per original id, with the reference's tie rules inside the triangle
    const float* thi[3];
    uint32_t* perm;                          // position -> original id
    uint32_t* permTmp;
    uint32_t* leftBefore;                    // scratch per position: lefts at earlier positions of the node
    uint32_t* hole;                          // scratch: hole[first + s - 1] = position of the s-th right element of the front region
    uint32_t* backLeft;                      // scratch: backLeft[first + s - 1] = position of the s-th left element from the end
    TNode* nodes;
    uint32_t* counters;                      // [0] temp nodes used, [1..4] next-level list sizes (by triangle-count class)
    uint32_t* nextList[4];                   // <= 64, <= 4096, <= HUGE_MIN, above
};

__device__ __forceinline__ uint32_t f2o(float f) {          // order-preserving float -> uint
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float o2f(uint32_t o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}
// 64-bit reduction keys for node bounds: equal values are bit-identical except +-0, whose sign follows the
// sequence order (first for the minimum, last for the maximum); `order` = position in the node.
__device__ __forceinline__ unsigned long long bound_key(float v, uint32_t order) {
    uint32_t sign = __float_as_uint(v) >> 31;
    float mag = v == 0.f ? 0.f : v;
    return ((unsigned long long)f2o(mag) << 32) | ((unsigned long long)order << 1) | (v == 0.f ? sign : 0u);
}
__device__ __forceinline__ float bound_value(unsigned long long key) {
    float v = o2f((uint32_t)(key >> 32));
    return (v == 0.f && (key & 1ull)) ? -0.f : v;
}

// one thread per triangle of the object: centroid (push_face :919-958 order of additions) and its box
__global__ void k_prepare(const ptl::TrianglePoint* __restrict__ pts, uint32_t pointBase, const ptl::Triangle* __restrict__ tris, uint32_t n,
                          float* cx, float* cy, float* cz, float* lox, float* loy, float* loz, float* hix, float* hiy, float* hiz, uint32_t* perm) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const ptl::Triangle t = tris[i];
    const uint32_t v[3] = {t.v0 - pointBase, t.v1 - pointBase, t.v2 - pointBase};
    float c[3] = {0.f, 0.f, 0.f}, lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
    for (int k = 0; k < 3; ++k) {
        const ptl::TrianglePoint p = pts[v[k]];
        const float q[3] = {p.px, p.py, p.pz};
        for (int a = 0; a < 3; ++a) {
            c[a] += q[a];
            lo[a] = q[a] < lo[a] ? q[a] : lo[a];            // iMin(p, lo): ties keep lo
            hi[a] = q[a] < hi[a] ? hi[a] : q[a];            // iMax(p, hi): ties take p
        }
    }
    cx[i] = c[0] / 3.f; cy[i] = c[1] / 3.f; cz[i] = c[2] / 3.f;
    lox[i] = lo[0]; loy[i] = lo[1]; loz[i] = lo[2];
    hix[i] = hi[0]; hiy[i] = hi[1]; hiz[i] = hi[2];
    perm[i] = i;
}

// ---- block-wide helpers (blockDim.x = 32 ... 1024, a multiple of 32) ---------------------------------------
struct SplitShared {
    uint32_t cLo[3], cHi[3];                 // centroid range (ordered uints)
    uint32_t binCount[3][BINS];
    uint32_t binLo[3][BINS][3], binHi[3][BINS][3];
    unsigned long long kidLo[2][3], kidHi[2][3];
    uint32_t warpSum[32];
    uint32_t running, nLeft, kids;
    int axis, doSplit;
    float pos;
};

// exclusive prefix of `flag` over the block, plus the block total; two barriers
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

struct SahBox {                              // BoundingBox of the sweep (glm::min / glm::max on values: order independent)
    float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
    __device__ void add(const float* blo, const float* bhi) {
        for (int i = 0; i < 3; ++i) { lo[i] = blo[i] < lo[i] ? blo[i] : lo[i]; hi[i] = hi[i] < bhi[i] ? bhi[i] : hi[i]; }
    }
    __device__ float half_area() const {     // src/vk_engine.h:102-107
        float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
        return x * y + y * z + z * x;
    }
};

// Where the sequential back-swapping partition (:1233-1244) leaves the element that starts at position k of an
// n-element segment. isLeft: centroid < split; L = lefts at positions < k; nL = all lefts; m = rights among the
// first nL positions (= lefts behind them). hole/backLeft as in Dev, relative to the segment.
__device__ __forceinline__ uint32_t dest(uint32_t k, uint32_t n, bool isLeft, uint32_t L, uint32_t nL, uint32_t m,
                                         const uint32_t* hole, const uint32_t* backLeft) {
    if (k < nL) {
        if (isLeft) return k;                                            // never touched
        uint32_t s = k - L + 1;                                          // s-th hole; met by the loop after s-1 earlier holes
        uint32_t x = (s - 1) + (s == 1 ? 0u : (n - 1 - backLeft[s - 2]) - (s - 2));   // + the rights pulled from the back so far
        return n - 1 - x;
    }
    uint32_t above = nL - L - (isLeft ? 1u : 0u);                        // lefts at later positions
    if (isLeft) return hole[above];                                      // (above+1)-th left from the end fills that hole
    uint32_t rightsAbove = (n - 1 - k) - above;
    uint32_t x;
    if (above == m && k == nL) x = m == 0 ? 0u : m + ((n - 1 - backLeft[m - 1]) - (m - 1));   // met when the loop runs past the last left
    else x = above + 1 + rightsAbove;
    return n - 1 - x;
}

// find_bvh_split_plane's sweep over the filled bins (:1304-1336), in the reference's order and with its quirks
__device__ void sah_sweep(const uint32_t (*binCount)[BINS], const uint32_t (*binLo)[BINS][3], const uint32_t (*binHi)[BINS][3],
                          const float* cLo, const float* cHi, float& best, int& axis, float& pos) {
    for (int a = 0; a < 3; ++a) {
        if (cLo[a] == cHi[a]) continue;
        float leftArea[BINS - 1], rightArea[BINS - 1], leftN[BINS - 1], rightN[BINS - 1];
        SahBox lbox, rbox;
        int lsum = 0, rsum = 0;
        for (int i = 0; i < BINS - 1; ++i) {
            float blo[3], bhi[3];
            lsum += (int)binCount[a][i];
            leftN[i] = (float)lsum;
            for (int c = 0; c < 3; ++c) { blo[c] = o2f(binLo[a][i][c]); bhi[c] = o2f(binHi[a][i][c]); }
            lbox.add(blo, bhi);
            leftArea[i] = lbox.half_area();
            rsum += (int)binCount[a][BINS - 1 - i];
            rightN[BINS - 2 - i] = (float)rsum;
            for (int c = 0; c < 3; ++c) { blo[c] = o2f(binLo[a][BINS - 1 - i][c]); bhi[c] = o2f(binHi[a][BINS - 1 - i][c]); }
            rbox.add(blo, bhi);
            float ra = rbox.half_area();
            rightArea[i] = ra;                      // the reference writes both entries; the later write wins
            rightArea[BINS - 2 - i] = ra;
        }
        const float step = (cHi[a] - cLo[a]) / (float)BINS;
        for (int i = 0; i < BINS - 1; ++i) {
            float cost = leftN[i] * leftArea[i] + rightN[i] * rightArea[i];
            if (cost < best) { best = cost; axis = a; pos = cLo[a] + step * (float)(i + 1); }
        }
    }
}

// One block per node of the level list: best split, partition, children.
__global__ void __launch_bounds__(1024) k_split(Dev d, const uint32_t* __restrict__ list, uint32_t nList) {
    __shared__ SplitShared s;
    const uint32_t tid = threadIdx.x, nT = blockDim.x;
    for (uint32_t item = blockIdx.x; item < nList; item += gridDim.x) {
        const uint32_t id = list[item];
        const TNode node = d.nodes[id];
        const uint32_t n = node.count, first = node.first;
        uint32_t* perm = d.perm + first;
        // ---- find_bvh_split_plane (:1279-1337) ----
        if (tid < 3) { s.cLo[tid] = f2o(1e30f); s.cHi[tid] = f2o(-1e30f); }
        for (uint32_t i = tid; i < 3 * BINS; i += nT) {
            (&s.binCount[0][0])[i] = 0u;
            for (int c = 0; c < 3; ++c) { (&s.binLo[0][0][0])[i * 3 + c] = f2o(1e30f); (&s.binHi[0][0][0])[i * 3 + c] = f2o(-1e30f); }
        }
        if (tid < 6) { (&s.kidLo[0][0])[tid] = ~0ull; (&s.kidHi[0][0])[tid] = 0ull; }
        __syncthreads();
        {
            uint32_t lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
            for (uint32_t k = tid; k < n; k += nT) {
                const uint32_t t = perm[k];
                for (int a = 0; a < 3; ++a) { uint32_t o = f2o(d.cen[a][t]); lo[a] = min(lo[a], o); hi[a] = max(hi[a], o); }
            }
            for (int a = 0; a < 3; ++a) {
                uint32_t l = __reduce_min_sync(0xffffffffu, lo[a]), h = __reduce_max_sync(0xffffffffu, hi[a]);
                if ((tid & 31u) == 0) { atomicMin(&s.cLo[a], l); atomicMax(&s.cHi[a], h); }
            }
        }
        __syncthreads();
        float cLo[3], cHi[3], scale[3];
        for (int a = 0; a < 3; ++a) { cLo[a] = o2f(s.cLo[a]); cHi[a] = o2f(s.cHi[a]); scale[a] = (float)BINS / (cHi[a] - cLo[a]); }
        for (uint32_t k = tid; k < n; k += nT) {
            const uint32_t t = perm[k];
            uint32_t blo[3], bhi[3];
            for (int c = 0; c < 3; ++c) { blo[c] = f2o(d.tlo[c][t]); bhi[c] = f2o(d.thi[c][t]); }
            for (int a = 0; a < 3; ++a) {
                if (cLo[a] == cHi[a]) continue;
                float f = floorf((d.cen[a][t] - cLo[a]) * scale[a]);
                const float lim = (float)(BINS - 1);
                int b = (int)(lim < f ? lim : f);
                atomicAdd(&s.binCount[a][b], 1u);
                for (int c = 0; c < 3; ++c) { atomicMin(&s.binLo[a][b][c], blo[c]); atomicMax(&s.binHi[a][b][c], bhi[c]); }
            }
        }
        __syncthreads();
        if (tid == 0) {
            float best = 1e30f, pos = 0.f;
            int axis = 0;
            sah_sweep(s.binCount, s.binLo, s.binHi, cLo, cHi, best, axis, pos);
            float ex = node.hi[0] - node.lo[0], ey = node.hi[1] - node.lo[1], ez = node.hi[2] - node.lo[2];
            float stay = (float)n * (ex * ey + ey * ez + ez * ex);
            s.doSplit = best >= stay ? 0 : 1;                // :1226-1229
            s.axis = axis; s.pos = pos; s.nLeft = 0; s.running = 0;
        }
        __syncthreads();
        if (!s.doSplit) continue;                            // stays a leaf, triangles untouched (uniform for the block)
        // ---- partition (:1233-1244) ----
        const float* cen = d.cen[s.axis];
        const float pos = s.pos;
        uint32_t* leftBefore = d.leftBefore + first;
        uint32_t* hole = d.hole + first;
        uint32_t* backLeft = d.backLeft + first;
        uint32_t running = 0;
        for (uint32_t base = 0; base < n; base += nT) {       // pass 1: lefts before each position
            const uint32_t k = base + tid;
            const bool isLeft = k < n && cen[perm[k]] < pos;
            uint32_t total;
            uint32_t ex = block_exclusive(s, isLeft, total);
            if (k < n) leftBefore[k] = running + ex;
            running += total;
        }
        const uint32_t nL = running;
        uint32_t mLocal = 0;
        for (uint32_t k = tid; k < n; k += nT) {              // pass 2: holes of the front region, lefts of the back region
            const bool isLeft = cen[perm[k]] < pos;
            const uint32_t L = leftBefore[k];
            if (k < nL) { if (!isLeft) { hole[k - L] = k; mLocal++; } }
            else if (isLeft) backLeft[nL - L - 1] = k;
        }
        if (mLocal) atomicAdd(&s.nLeft, mLocal);              // s.nLeft doubles as the hole counter m
        __syncthreads();
        const uint32_t m = s.nLeft;
        uint32_t* out = d.permTmp + first;
        for (uint32_t k = tid; k < n; k += nT) {              // pass 3: scatter
            const uint32_t t = perm[k];
            out[dest(k, n, cen[t] < pos, leftBefore[k], nL, m, hole, backLeft)] = t;
        }
        __syncthreads();
        const bool split = nL != 0 && nL != n;               // :1246-1248 (the triangles stay permuted either way)
        if (tid == 0 && split) s.kids = atomicAdd(&d.counters[0], 2u);
        {
            unsigned long long klo[3] = {~0ull, ~0ull, ~0ull}, khi[3] = {0ull, 0ull, 0ull};
            // children bounds (update_bvh_bounds :1196-1210) while copying the new order back; a thread's positions
            // are visited in increasing order, so per child it keeps running keys and flushes when the side changes
            int side = -1;
            auto flush = [&](int sd) {
                for (int c = 0; c < 3; ++c) { atomicMin(&s.kidLo[sd][c], klo[c]); atomicMax(&s.kidHi[sd][c], khi[c]); klo[c] = ~0ull; khi[c] = 0ull; }
            };
            for (uint32_t k = tid; k < n; k += nT) {
                const uint32_t t = out[k];
                perm[k] = t;
                if (!split) continue;
                const int sd = k < nL ? 0 : 1;
                if (sd != side) { if (side >= 0) flush(side); side = sd; }
                const uint32_t order = sd ? k - nL : k;
                for (int c = 0; c < 3; ++c) {
                    klo[c] = min(klo[c], bound_key(d.tlo[c][t], order));
                    khi[c] = max(khi[c], bound_key(d.thi[c][t], order));
                }
            }
            if (side >= 0) flush(side);
        }
        __syncthreads();
        if (split && tid < 2) {
            const uint32_t kid = s.kids + tid;
            TNode c;
            c.first = tid ? first + nL : first;
            c.count = tid ? n - nL : nL;
            c.left = -1;
            c.depth = node.depth + 1;
            for (int a = 0; a < 3; ++a) { c.lo[a] = bound_value(s.kidLo[tid][a]); c.hi[a] = bound_value(s.kidHi[tid][a]); }
            c.size = 1; c.finalIdx = 0; c.base = 0;
            d.nodes[kid] = c;
            if (c.count > 2 && c.depth < MAX_DEPTH) {        // :1214-1216
                const int cls = list_class(c.count);
                d.nextList[cls][atomicAdd(&d.counters[1 + cls], 1u)] = kid;
            }
            if (tid == 0) d.nodes[id].left = (int32_t)s.kids;
        }
        __syncthreads();
    }
}

// root node: bounds of the whole range in the original order (build_bvh :1186-1190), reduced by all SMs
__global__ void __launch_bounds__(HUGE_THREADS) k_root_bounds(Dev d, uint32_t n, unsigned long long* keys /* lo[3], hi[3] */) {
    __shared__ unsigned long long sLo[3], sHi[3];
    if (threadIdx.x < 3) { sLo[threadIdx.x] = ~0ull; sHi[threadIdx.x] = 0ull; }
    __syncthreads();
    unsigned long long klo[3] = {~0ull, ~0ull, ~0ull}, khi[3] = {0ull, 0ull, 0ull};
    for (uint32_t k = blockIdx.x * HUGE_THREADS + threadIdx.x; k < n; k += gridDim.x * HUGE_THREADS)
        for (int c = 0; c < 3; ++c) {
            klo[c] = min(klo[c], bound_key(d.tlo[c][k], k));
            khi[c] = max(khi[c], bound_key(d.thi[c][k], k));
        }
    for (int c = 0; c < 3; ++c) { atomicMin(&sLo[c], klo[c]); atomicMax(&sHi[c], khi[c]); }
    __syncthreads();
    if (threadIdx.x < 3) { atomicMin(&keys[threadIdx.x], sLo[threadIdx.x]); atomicMax(&keys[3 + threadIdx.x], sHi[threadIdx.x]); }
}
__global__ void k_root(Dev d, uint32_t n, const unsigned long long* keys) {
    if (blockIdx.x || threadIdx.x) return;
    TNode r;
    r.first = 0; r.count = n; r.left = -1; r.depth = 0; r.size = 1; r.finalIdx = 0; r.base = 0;
    for (int a = 0; a < 3; ++a) { r.lo[a] = bound_value(keys[a]); r.hi[a] = bound_value(keys[3 + a]); }
    d.nodes[0] = r;
    d.counters[0] = 1;
    d.counters[1] = d.counters[2] = d.counters[3] = d.counters[4] = 0;
    if (n > 2) { const int cls = list_class(n); d.nextList[cls][0] = 0; d.counters[1 + cls] = 1; }
}

// ---- nodes above HUGE_MIN triangles: the steps of k_split, each as one launch over chunks of HUGE_CHUNK positions ----
struct HugeNode {
    uint32_t id, first, count, depth, chunk0, nChunks;
    float nlo[3], nhi[3];                    // the node's own bounds (for the "stay" cost)
    uint32_t cLo[3], cHi[3];
    uint32_t binCount[3][BINS];
    uint32_t binLo[3][BINS][3], binHi[3][BINS][3];
    int axis, doSplit;
    float pos;
    uint32_t nL, m;
    unsigned long long kidLo[2][3], kidHi[2][3];
};
struct HugeChunk { uint32_t slot, begin, len; };

__global__ void k_huge_init(Dev d, HugeNode* hn, uint32_t nHuge, const uint32_t* __restrict__ list) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nHuge) return;
    HugeNode& h = hn[i];
    const TNode node = d.nodes[list[i]];
    h.id = list[i]; h.first = node.first; h.count = node.count; h.depth = node.depth;
    for (int a = 0; a < 3; ++a) { h.nlo[a] = node.lo[a]; h.nhi[a] = node.hi[a]; h.cLo[a] = f2o(1e30f); h.cHi[a] = f2o(-1e30f); }
    for (int k = 0; k < 3 * BINS; ++k) {
        (&h.binCount[0][0])[k] = 0u;
        for (int c = 0; c < 3; ++c) { (&h.binLo[0][0][0])[k * 3 + c] = f2o(1e30f); (&h.binHi[0][0][0])[k * 3 + c] = f2o(-1e30f); }
    }
    for (int k = 0; k < 6; ++k) { (&h.kidLo[0][0])[k] = ~0ull; (&h.kidHi[0][0])[k] = 0ull; }
    h.axis = 0; h.doSplit = 0; h.pos = 0.f; h.nL = 0; h.m = 0;
}

__global__ void __launch_bounds__(HUGE_THREADS) k_huge_minmax(Dev d, HugeNode* hn, const HugeChunk* __restrict__ chunks) {
    const HugeChunk ck = chunks[blockIdx.x];
    HugeNode& h = hn[ck.slot];
    const uint32_t* perm = d.perm + h.first + ck.begin;
    uint32_t lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
    for (uint32_t k = threadIdx.x; k < ck.len; k += HUGE_THREADS) {
        const uint32_t t = perm[k];
        for (int a = 0; a < 3; ++a) { uint32_t o = f2o(d.cen[a][t]); lo[a] = min(lo[a], o); hi[a] = max(hi[a], o); }
    }
    __shared__ uint32_t sLo[3], sHi[3];
    if (threadIdx.x < 3) { sLo[threadIdx.x] = 0xffffffffu; sHi[threadIdx.x] = 0u; }
    __syncthreads();
    for (int a = 0; a < 3; ++a) {
        uint32_t l = __reduce_min_sync(0xffffffffu, lo[a]), hh = __reduce_max_sync(0xffffffffu, hi[a]);
        if ((threadIdx.x & 31u) == 0) { atomicMin(&sLo[a], l); atomicMax(&sHi[a], hh); }
    }
    __syncthreads();
    if (threadIdx.x < 3) { atomicMin(&h.cLo[threadIdx.x], sLo[threadIdx.x]); atomicMax(&h.cHi[threadIdx.x], sHi[threadIdx.x]); }
}

__global__ void __launch_bounds__(HUGE_THREADS) k_huge_bins(Dev d, HugeNode* hn, const HugeChunk* __restrict__ chunks) {
    __shared__ uint32_t binCount[3][BINS], binLo[3][BINS][3], binHi[3][BINS][3];
    const HugeChunk ck = chunks[blockIdx.x];
    HugeNode& h = hn[ck.slot];
    const uint32_t* perm = d.perm + h.first + ck.begin;
    for (uint32_t i = threadIdx.x; i < 3 * BINS; i += HUGE_THREADS) {
        (&binCount[0][0])[i] = 0u;
        for (int c = 0; c < 3; ++c) { (&binLo[0][0][0])[i * 3 + c] = f2o(1e30f); (&binHi[0][0][0])[i * 3 + c] = f2o(-1e30f); }
    }
    float cLo[3], cHi[3], scale[3];
    for (int a = 0; a < 3; ++a) { cLo[a] = o2f(h.cLo[a]); cHi[a] = o2f(h.cHi[a]); scale[a] = (float)BINS / (cHi[a] - cLo[a]); }
    __syncthreads();
    for (uint32_t k = threadIdx.x; k < ck.len; k += HUGE_THREADS) {
        const uint32_t t = perm[k];
        uint32_t blo[3], bhi[3];
        for (int c = 0; c < 3; ++c) { blo[c] = f2o(d.tlo[c][t]); bhi[c] = f2o(d.thi[c][t]); }
        for (int a = 0; a < 3; ++a) {
            if (cLo[a] == cHi[a]) continue;
            float f = floorf((d.cen[a][t] - cLo[a]) * scale[a]);
            const float lim = (float)(BINS - 1);
            int b = (int)(lim < f ? lim : f);
            atomicAdd(&binCount[a][b], 1u);
            for (int c = 0; c < 3; ++c) { atomicMin(&binLo[a][b][c], blo[c]); atomicMax(&binHi[a][b][c], bhi[c]); }
        }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < 3 * BINS; i += HUGE_THREADS) {
        uint32_t cnt = (&binCount[0][0])[i];
        if (!cnt) continue;
        atomicAdd(&(&h.binCount[0][0])[i], cnt);
        for (int c = 0; c < 3; ++c) { atomicMin(&(&h.binLo[0][0][0])[i * 3 + c], (&binLo[0][0][0])[i * 3 + c]); atomicMax(&(&h.binHi[0][0][0])[i * 3 + c], (&binHi[0][0][0])[i * 3 + c]); }
    }
}

__global__ void k_huge_sah(HugeNode* hn, uint32_t nHuge) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nHuge) return;
    HugeNode& h = hn[i];
    float cLo[3], cHi[3];
    for (int a = 0; a < 3; ++a) { cLo[a] = o2f(h.cLo[a]); cHi[a] = o2f(h.cHi[a]); }
    float best = 1e30f, pos = 0.f;
    int axis = 0;
    sah_sweep(h.binCount, h.binLo, h.binHi, cLo, cHi, best, axis, pos);
    float ex = h.nhi[0] - h.nlo[0], ey = h.nhi[1] - h.nlo[1], ez = h.nhi[2] - h.nlo[2];
    float stay = (float)h.count * (ex * ey + ey * ez + ez * ex);
    h.doSplit = best >= stay ? 0 : 1;                         // :1226-1229
    h.axis = axis; h.pos = pos;
}

__global__ void __launch_bounds__(HUGE_THREADS) k_huge_count(Dev d, const HugeNode* hn, const HugeChunk* __restrict__ chunks, uint32_t* chunkLeft) {
    const HugeChunk ck = chunks[blockIdx.x];
    const HugeNode& h = hn[ck.slot];
    uint32_t nLeft = 0;
    if (h.doSplit) {
        const float* cen = d.cen[h.axis];
        const uint32_t* perm = d.perm + h.first + ck.begin;
        for (uint32_t k = threadIdx.x; k < ck.len; k += HUGE_THREADS) nLeft += cen[perm[k]] < h.pos ? 1u : 0u;
    }
    __shared__ uint32_t total;
    if (threadIdx.x == 0) total = 0;
    __syncthreads();
    nLeft = __reduce_add_sync(0xffffffffu, nLeft);
    if ((threadIdx.x & 31u) == 0 && nLeft) atomicAdd(&total, nLeft);
    __syncthreads();
    if (threadIdx.x == 0) chunkLeft[blockIdx.x] = total;
}

__global__ void k_huge_scan(HugeNode* hn, uint32_t nHuge, uint32_t* chunkLeft) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nHuge) return;
    HugeNode& h = hn[i];
    uint32_t run = 0;
    for (uint32_t c = h.chunk0; c < h.chunk0 + h.nChunks; ++c) { uint32_t v = chunkLeft[c]; chunkLeft[c] = run; run += v; }
    h.nL = run;
}

// pass 1 + 2 of k_split's partition: lefts before each position, holes of the front region, lefts of the back region
__global__ void __launch_bounds__(HUGE_THREADS) k_huge_marks(Dev d, HugeNode* hn, const HugeChunk* __restrict__ chunks, const uint32_t* __restrict__ chunkLeft) {
    __shared__ SplitShared s;
    const HugeChunk ck = chunks[blockIdx.x];
    HugeNode& h = hn[ck.slot];
    if (!h.doSplit) return;
    const float* cen = d.cen[h.axis];
    const float pos = h.pos;
    const uint32_t nL = h.nL;
    const uint32_t* perm = d.perm + h.first;
    uint32_t* leftBefore = d.leftBefore + h.first;
    uint32_t* hole = d.hole + h.first;
    uint32_t* backLeft = d.backLeft + h.first;
    uint32_t running = chunkLeft[blockIdx.x], mLocal = 0;
    for (uint32_t base = 0; base < ck.len; base += HUGE_THREADS) {
        const uint32_t k = base + threadIdx.x, kr = ck.begin + k;
        const bool isLeft = k < ck.len && cen[perm[kr]] < pos;
        uint32_t total;
        const uint32_t L = running + block_exclusive(s, isLeft, total);
        if (k < ck.len) {
            leftBefore[kr] = L;
            if (kr < nL) { if (!isLeft) { hole[kr - L] = kr; mLocal++; } }
            else if (isLeft) backLeft[nL - L - 1] = kr;
        }
        running += total;
    }
    mLocal = __reduce_add_sync(0xffffffffu, mLocal);
    if ((threadIdx.x & 31u) == 0 && mLocal) atomicAdd(&h.m, mLocal);
}

__global__ void __launch_bounds__(HUGE_THREADS) k_huge_scatter(Dev d, const HugeNode* hn, const HugeChunk* __restrict__ chunks) {
    const HugeChunk ck = chunks[blockIdx.x];
    const HugeNode& h = hn[ck.slot];
    if (!h.doSplit) return;
    const float* cen = d.cen[h.axis];
    const uint32_t* perm = d.perm + h.first;
    uint32_t* out = d.permTmp + h.first;
    for (uint32_t k = threadIdx.x; k < ck.len; k += HUGE_THREADS) {
        const uint32_t kr = ck.begin + k, t = perm[kr];
        out[dest(kr, h.count, cen[t] < h.pos, (d.leftBefore + h.first)[kr], h.nL, h.m, d.hole + h.first, d.backLeft + h.first)] = t;
    }
}

// copy the new order back and reduce the children's bounds (update_bvh_bounds :1196-1210)
__global__ void __launch_bounds__(HUGE_THREADS) k_huge_finish(Dev d, HugeNode* hn, const HugeChunk* __restrict__ chunks) {
    __shared__ unsigned long long sLo[2][3], sHi[2][3];
    const HugeChunk ck = chunks[blockIdx.x];
    HugeNode& h = hn[ck.slot];
    if (!h.doSplit) return;
    if (threadIdx.x < 6) { (&sLo[0][0])[threadIdx.x] = ~0ull; (&sHi[0][0])[threadIdx.x] = 0ull; }
    __syncthreads();
    const uint32_t nL = h.nL;
    const bool split = nL != 0 && nL != h.count;
    uint32_t* perm = d.perm + h.first;
    const uint32_t* out = d.permTmp + h.first;
    unsigned long long klo[3] = {~0ull, ~0ull, ~0ull}, khi[3] = {0ull, 0ull, 0ull};
    int side = -1;
    auto flush = [&](int sd) {
        for (int c = 0; c < 3; ++c) { atomicMin(&sLo[sd][c], klo[c]); atomicMax(&sHi[sd][c], khi[c]); klo[c] = ~0ull; khi[c] = 0ull; }
    };
    for (uint32_t k = threadIdx.x; k < ck.len; k += HUGE_THREADS) {
        const uint32_t kr = ck.begin + k, t = out[kr];
        perm[kr] = t;
        if (!split) continue;
        const int sd = kr < nL ? 0 : 1;
        if (sd != side) { if (side >= 0) flush(side); side = sd; }
        const uint32_t order = sd ? kr - nL : kr;
        for (int c = 0; c < 3; ++c) {
            klo[c] = min(klo[c], bound_key(d.tlo[c][t], order));
            khi[c] = max(khi[c], bound_key(d.thi[c][t], order));
        }
    }
    if (side >= 0) flush(side);
    __syncthreads();
    if (split && threadIdx.x < 6) {
        const unsigned long long lo = (&sLo[0][0])[threadIdx.x], hi = (&sHi[0][0])[threadIdx.x];
        if (lo != ~0ull) atomicMin(&(&h.kidLo[0][0])[threadIdx.x], lo);
        if (hi != 0ull) atomicMax(&(&h.kidHi[0][0])[threadIdx.x], hi);
    }
}

__global__ void k_huge_children(Dev d, const HugeNode* hn, uint32_t nHuge) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nHuge) return;
    const HugeNode& h = hn[i];
    if (!h.doSplit || h.nL == 0 || h.nL == h.count) return;   // leaf (:1226, :1246); the triangles stay permuted
    const uint32_t kids = atomicAdd(&d.counters[0], 2u);
    for (uint32_t side = 0; side < 2; ++side) {
        TNode c;
        c.first = side ? h.first + h.nL : h.first;
        c.count = side ? h.count - h.nL : h.nL;
        c.left = -1;
        c.depth = h.depth + 1;
        for (int a = 0; a < 3; ++a) { c.lo[a] = bound_value(h.kidLo[side][a]); c.hi[a] = bound_value(h.kidHi[side][a]); }
        c.size = 1; c.finalIdx = 0; c.base = 0;
        d.nodes[kids + side] = c;
        if (c.count > 2 && c.depth < MAX_DEPTH) {
            const int cls = list_class(c.count);
            d.nextList[cls][atomicAdd(&d.counters[1 + cls], 1u)] = kids + side;
        }
    }
    d.nodes[h.id].left = (int32_t)kids;
}

// Temp nodes of one level occupy a contiguous id range (children are allocated while their parents' level runs).
__global__ void k_sizes(TNode* nodes, uint32_t begin, uint32_t end) {           // bottom-up, one level per launch
    uint32_t i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= end) return;
    int32_t l = nodes[i].left;
    nodes[i].size = l < 0 ? 1u : 1u + nodes[l].size + nodes[l + 1].size;
}
// Depth-first numbering: subdivide_bvh takes the pair {nodesUsed, nodesUsed + 1} for its children, then builds the
// whole left subtree before the right one (:1250-1276).
__global__ void k_number(TNode* nodes, uint32_t begin, uint32_t end) {          // top-down, one level per launch
    uint32_t i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= end) return;
    const TNode n = nodes[i];
    if (n.left < 0) return;
    TNode& L = nodes[n.left];
    TNode& R = nodes[n.left + 1];
    L.finalIdx = n.base; R.finalIdx = n.base + 1;
    L.base = n.base + 2;
    R.base = n.base + 2 + (L.size - 1);
}
__global__ void k_emit(const TNode* __restrict__ nodes, uint32_t nTemp, uint32_t rootIndex, uint32_t firstTri, ptl::BVHNode* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nTemp) return;
    const TNode n = nodes[i];
    ptl::BVHNode o;
    o.minX = n.lo[0]; o.maxX = n.hi[0]; o.minY = n.lo[1]; o.maxY = n.hi[1]; o.minZ = n.lo[2]; o.maxZ = n.hi[2];
    o.index = n.left < 0 ? firstTri + n.first : n.base;
    o.triCount = n.left < 0 ? n.count : 0u;
    out[n.finalIdx - rootIndex] = o;
}
__global__ void k_gather_triangles(const ptl::Triangle* __restrict__ in, const uint32_t* __restrict__ perm, uint32_t n, ptl::Triangle* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[perm[i]];
}

struct Buffers {                             // freed on every exit path
    std::vector<void*> ptrs;
    ~Buffers() { for (void* p : ptrs) cudaFree(p); }
    template <class T> cudaError_t get(T** p, size_t count) {
        cudaError_t e = cudaMalloc((void**)p, std::max<size_t>(count, 1) * sizeof(T));
        if (e == cudaSuccess) ptrs.push_back(*p);
        return e;
    }
};

struct Events {
    cudaEvent_t e[2] = {nullptr, nullptr};
    cudaError_t create() { cudaError_t r = cudaEventCreate(&e[0]); return r != cudaSuccess ? r : cudaEventCreate(&e[1]); }
    ~Events() { for (cudaEvent_t x : e) if (x) cudaEventDestroy(x); }
};

}  // namespace bvh2

#define BCU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { char m_[256]; snprintf(m_, sizeof m_, "b200pt_build_bvh: %s: %s", #x, cudaGetErrorString(e_)); pt::set_global_error(m_); return 2; } } while (0)

extern "C" int b200pt_build_bvh(int device, const void* triPointsV, uint32_t nPoints, void* trianglesV, uint32_t firstTri, uint32_t count,
                                uint32_t rootIndex, void* nodesOutV, uint32_t* nNodesOut, double* deviceMs) {
    using namespace bvh2;
    if (!triPointsV || !trianglesV || !nodesOutV || !nNodesOut) { pt::set_global_error("b200pt_build_bvh: null argument"); return 1; }
    if (count == 0) { pt::set_global_error("b200pt_build_bvh: empty triangle range (undefined in the reference, :1186)"); return 1; }
    const ptl::TrianglePoint* points = (const ptl::TrianglePoint*)triPointsV;
    ptl::Triangle* tris = (ptl::Triangle*)trianglesV + firstTri;
    uint32_t vLo = 0xffffffffu, vHi = 0;
    for (uint32_t i = 0; i < count; ++i) {
        const ptl::Triangle& t = tris[i];
        vLo = std::min(vLo, std::min(t.v0, std::min(t.v1, t.v2)));
        vHi = std::max(vHi, std::max(t.v0, std::max(t.v1, t.v2)));
    }
    if (vHi >= nPoints) { pt::set_global_error("b200pt_build_bvh: vertex index out of range"); return 1; }
    BCU(cudaSetDevice(device));
    Buffers mem;
    const size_t nP = (size_t)vHi - vLo + 1, n = count, maxNodes = 2 * n - 1;
    ptl::TrianglePoint* dPts; ptl::Triangle *dTris, *dTrisOut; ptl::BVHNode* dOut;
    float* f[9]; uint32_t *perm, *permTmp, *leftBefore, *hole, *backLeft, *counters, *lists[8], *chunkLeft;
    TNode* nodes; HugeNode* hugeNodes; HugeChunk* hugeChunks; unsigned long long* rootKeys;
    const size_t maxHuge = n / HUGE_MIN + 1, maxChunks = n / HUGE_CHUNK + maxHuge + 1;
    BCU(mem.get(&dPts, nP)); BCU(mem.get(&dTris, n)); BCU(mem.get(&dTrisOut, n)); BCU(mem.get(&dOut, maxNodes));
    for (auto& p : f) BCU(mem.get(&p, n));
    BCU(mem.get(&perm, n)); BCU(mem.get(&permTmp, n)); BCU(mem.get(&leftBefore, n)); BCU(mem.get(&hole, n)); BCU(mem.get(&backLeft, n));
    BCU(mem.get(&counters, 8)); BCU(mem.get(&nodes, maxNodes));
    BCU(mem.get(&hugeNodes, maxHuge)); BCU(mem.get(&hugeChunks, maxChunks)); BCU(mem.get(&chunkLeft, maxChunks)); BCU(mem.get(&rootKeys, 6));
    for (auto& p : lists) BCU(mem.get(&p, n / 3 + 1));       // a listed node holds at least 3 triangles
    cudaStream_t st = 0;
    Events ev;                                               // destroyed on every exit path
    BCU(ev.create());
    cudaEvent_t e0 = ev.e[0], e1 = ev.e[1];
    BCU(cudaMemcpyAsync(dPts, points + vLo, nP * sizeof(ptl::TrianglePoint), cudaMemcpyHostToDevice, st));
    BCU(cudaMemcpyAsync(dTris, tris, n * sizeof(ptl::Triangle), cudaMemcpyHostToDevice, st));
    BCU(cudaEventRecord(e0, st));
    k_prepare<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(dPts, vLo, dTris, count, f[0], f[1], f[2], f[3], f[4], f[5], f[6], f[7], f[8], perm);
    Dev d;
    for (int a = 0; a < 3; ++a) { d.cen[a] = f[a]; d.tlo[a] = f[3 + a]; d.thi[a] = f[6 + a]; }
    d.perm = perm; d.permTmp = permTmp; d.leftBefore = leftBefore; d.hole = hole; d.backLeft = backLeft;
    d.nodes = nodes; d.counters = counters;
    int cur = 0;
    for (int c = 0; c < 4; ++c) d.nextList[c] = lists[c];
    {
        const unsigned long long init[6] = {~0ull, ~0ull, ~0ull, 0ull, 0ull, 0ull};
        BCU(cudaMemcpyAsync(rootKeys, init, sizeof init, cudaMemcpyHostToDevice, st));
        k_root_bounds<<<(unsigned)std::min<size_t>((n + HUGE_CHUNK - 1) / HUGE_CHUNK, 4096), HUGE_THREADS, 0, st>>>(d, count, rootKeys);
        k_root<<<1, 32, 0, st>>>(d, count, rootKeys);
    }
    std::vector<uint32_t> levelStart{0};                     // temp-id range of each level
    std::vector<HugeNode> hHuge;
    std::vector<HugeChunk> hChunks;
    const unsigned threadsOf[3] = {32, 256, 1024};
    for (;;) {
        uint32_t h[8];
        BCU(cudaMemcpyAsync(h, counters, sizeof h, cudaMemcpyDeviceToHost, st));
        BCU(cudaStreamSynchronize(st));
        levelStart.push_back(h[0]);
        if (h[1] + h[2] + h[3] + h[4] == 0) break;
        const uint32_t* curList[4] = {lists[cur * 4], lists[cur * 4 + 1], lists[cur * 4 + 2], lists[cur * 4 + 3]};
        cur ^= 1;
        for (int c = 0; c < 4; ++c) d.nextList[c] = lists[cur * 4 + c];
        BCU(cudaMemsetAsync(counters + 1, 0, 16, st));
        for (int c = 0; c < 3; ++c)
            if (h[1 + c]) k_split<<<std::min<uint32_t>(h[1 + c], 1u << 20), threadsOf[c], 0, st>>>(d, curList[c], h[1 + c]);
        if (const uint32_t nHuge = h[4]) {                   // the few nodes that would otherwise each sit on one SM
            k_huge_init<<<(nHuge + 63) / 64, 64, 0, st>>>(d, hugeNodes, nHuge, curList[3]);
            hHuge.resize(nHuge);
            BCU(cudaMemcpyAsync(hHuge.data(), hugeNodes, nHuge * sizeof(HugeNode), cudaMemcpyDeviceToHost, st));
            BCU(cudaStreamSynchronize(st));
            hChunks.clear();
            for (uint32_t i = 0; i < nHuge; ++i) {
                hHuge[i].chunk0 = (uint32_t)hChunks.size();
                hHuge[i].nChunks = (hHuge[i].count + HUGE_CHUNK - 1) / HUGE_CHUNK;
                for (uint32_t b = 0; b < hHuge[i].count; b += HUGE_CHUNK) hChunks.push_back({i, b, std::min(HUGE_CHUNK, hHuge[i].count - b)});
            }
            const uint32_t nChunks = (uint32_t)hChunks.size();
            if (nHuge > maxHuge || nChunks > maxChunks) { pt::set_global_error("b200pt_build_bvh: chunk table overflow"); return 2; }
            BCU(cudaMemcpyAsync(hugeNodes, hHuge.data(), nHuge * sizeof(HugeNode), cudaMemcpyHostToDevice, st));
            BCU(cudaMemcpyAsync(hugeChunks, hChunks.data(), nChunks * sizeof(HugeChunk), cudaMemcpyHostToDevice, st));
            const unsigned gN = (nHuge + 63) / 64;
            k_huge_minmax<<<nChunks, HUGE_THREADS, 0, st>>>(d, hugeNodes, hugeChunks);
            k_huge_bins<<<nChunks, HUGE_THREADS, 0, st>>>(d, hugeNodes, hugeChunks);
            k_huge_sah<<<gN, 64, 0, st>>>(hugeNodes, nHuge);
            k_huge_count<<<nChunks, HUGE_THREADS, 0, st>>>(d, hugeNodes, hugeChunks, chunkLeft);
            k_huge_scan<<<gN, 64, 0, st>>>(hugeNodes, nHuge, chunkLeft);
            k_huge_marks<<<nChunks, HUGE_THREADS, 0, st>>>(d, hugeNodes, hugeChunks, chunkLeft);
            k_huge_scatter<<<nChunks, HUGE_THREADS, 0, st>>>(d, hugeNodes, hugeChunks);
            k_huge_finish<<<nChunks, HUGE_THREADS, 0, st>>>(d, hugeNodes, hugeChunks);
            k_huge_children<<<gN, 64, 0, st>>>(d, hugeNodes, nHuge);
        }
        BCU(cudaGetLastError());
    }
    // levelStart = {0, 1, end of level 1, ...}: level L holds temp ids [levelStart[L], levelStart[L + 1])
    const uint32_t nTemp = levelStart.back();
    const int nLevels = (int)levelStart.size() - 1;
    for (int L = nLevels - 1; L >= 0; --L) {
        uint32_t b = levelStart[L], e = levelStart[L + 1];
        if (e > b) k_sizes<<<(e - b + 255) / 256, 256, 0, st>>>(nodes, b, e);
    }
    {   // the root is bvhNodes[rootIndex]; its children take the next pair
        TNode patch;
        BCU(cudaMemcpyAsync(&patch, nodes, sizeof patch, cudaMemcpyDeviceToHost, st));
        BCU(cudaStreamSynchronize(st));
        patch.finalIdx = rootIndex; patch.base = rootIndex + 1;
        BCU(cudaMemcpyAsync(nodes, &patch, sizeof patch, cudaMemcpyHostToDevice, st));
        *nNodesOut = patch.size;
        if (patch.size != nTemp) { pt::set_global_error("b200pt_build_bvh: node count mismatch"); return 2; }
    }
    for (int L = 0; L < nLevels; ++L) {
        uint32_t b = levelStart[L], e = levelStart[L + 1];
        if (e > b) k_number<<<(e - b + 255) / 256, 256, 0, st>>>(nodes, b, e);
    }
    k_emit<<<(nTemp + 255) / 256, 256, 0, st>>>(nodes, nTemp, rootIndex, firstTri, dOut);
    k_gather_triangles<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(dTris, perm, count, dTrisOut);
    BCU(cudaEventRecord(e1, st));
    BCU(cudaGetLastError());
    BCU(cudaMemcpyAsync(nodesOutV, dOut, (size_t)nTemp * sizeof(ptl::BVHNode), cudaMemcpyDeviceToHost, st));
    BCU(cudaMemcpyAsync(tris, dTrisOut, n * sizeof(ptl::Triangle), cudaMemcpyDeviceToHost, st));
    BCU(cudaStreamSynchronize(st));
    if (deviceMs) { float ms = 0.f; cudaEventElapsedTime(&ms, e0, e1); *deviceMs = ms; }
    return 0;
}
