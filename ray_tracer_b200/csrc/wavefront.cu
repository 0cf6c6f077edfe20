// Wavefront path-tracing pipeline for sm_100a: thin kernels around the stage bodies of
// pt_wavefront.h, device-side queues with warp-aggregated compaction, no host round trip inside
// a batch. Compiled with -fmad=false (pt_math.h); box tests use explicit FMAs.
//
// One batch = `slots` progressive frames x the row band x W pixels, all in flight at once
// (frames have independent seeds, raytrace.comp:562-564; samples inside one dispatch do not, A-Q5).
// Launch sequence per batch:  generate, { extend_j, shade_j, connect_j } for j = 0..bounceLimit,
// accumulate, fold.  Queue sizes live in device memory (qCount[j], sCount[j]); every kernel is a
// grid-stride loop over its queue sized to the SM count, so an empty queue costs one idle launch.
#include <algorithm>
#include <cstdio>
#include <mutex>

#include "../../include/b200pt.h"

#include "bvh8_build.h"
#include "pt_wavefront.h"
#include "pt_reftrace.h"
#include "wavefront.h"
#include "wide_build.h"

namespace pt {

#define PT_N_STAGES 8
struct WavefrontState {
    // re-laid-out scene
    DeviceBuffer dNodes, dPrims, dShade, dUv, dNodesE, dPrimsE, dShadeE, dXforms, dObjXform, dUp;
    WideScene scene{};
    uint32_t maxDepth = 0;
    // path state
    DeviceBuffer dPaths[9];
    DeviceBuffer dHit, dQueue[2], dShadowQ, dJobQ, dCounts, dRayCounters, dSampleSum;
    DeviceBuffer dWalkQ;               // rays whose final hit needs the leaf -> root reachability walk (k_verify_walk)
    DeviceBuffer dExactQ, dExactNeeQ;  // contested rays / light-sample queries: re-traced in the reference's order (pt_traverse.h)
    DeviceBuffer dPrimOf, dInstBase, dTriFirst;
    PathArrays arrays{};
    WideDeviceResult wideRes;          // device-built tree (owned here; dNodes / dPrims / dShade / dUv alias it while wideOnDevice)
    WideBuildCache wideCache;
    bool wideOnDevice = false;
    std::vector<uint8_t> objEmissive;  // per object: material emits (decides which tree a triangle is in)
    std::vector<uint32_t> hLeafOf;     // triangle -> reference BVH2 leaf, parent links: kept for the refit after a transform edit
    std::vector<UpNode> hUp;
    bool haveScene = false, anyEmissiveTri = false;
    size_t capacity = 0;
    uint32_t countSlots = 0;
    int smCount = 148;
    int gridOcclude = 0;
    int gridShade = 0, gridPrep = 0, gridFinish = 0;
    int gridExtend = 0;               // resident blocks of the persistent kernels (occupancy x SMs)
    int gridVerify = 0;               // one wave of k_verify_extend
    // per-stage device time: an event between consecutive launches (generate, extend, shade, connect, accumulate)
    std::vector<cudaEvent_t> events;
    std::vector<int> eventStage;      // stage of the launch that FOLLOWS event i (-1: batch end)
    size_t eventsUsed = 0;
    double stageMs[PT_N_STAGES] = {0};
    uint64_t stageLaunches[PT_N_STAGES] = {0};
    // second stream for the connect chain of a bounce (option overlap_connect)
    cudaStream_t sideStream = nullptr;
    cudaEvent_t evFork = nullptr, evJoin = nullptr;
    // third stream: the reference-order re-trace of bounce j+1's contested rays runs beside the connect chain of bounce j
    cudaStream_t exactStream = nullptr;
    cudaEvent_t evX0 = nullptr, evX1 = nullptr;
};

enum { ST_GENERATE = 0, ST_EXTEND = 1, ST_SHADE = 2, ST_CONNECT = 3, ST_ACCUMULATE = 4, ST_OCCLUDE = 5, ST_FINISH = 6, ST_EXACT = 7 };

static void wf_mark(b200pt_ctx* c, int stage) {
    WavefrontState* w = c->wf;
    if (!c->profileStages) return;
    if (w->eventsUsed == w->events.size()) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        w->events.push_back(e);
        w->eventStage.push_back(-1);
    }
    w->eventStage[w->eventsUsed] = stage;
    cudaEventRecord(w->events[w->eventsUsed++], c->stream);
}

// stream must be idle: turns the recorded events into per-stage totals
void wavefront_resolve_stage_times(b200pt_ctx* c) {
    WavefrontState* w = c->wf;
    if (!w) return;
    for (size_t i = 0; i + 1 < w->eventsUsed; ++i) {
        int st = w->eventStage[i];
        if (st < 0) continue;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, w->events[i], w->events[i + 1]) == cudaSuccess) { w->stageMs[st] += ms; w->stageLaunches[st]++; }
    }
    w->eventsUsed = 0;
}
void wavefront_stage_times(b200pt_ctx* c, double* ms8, uint64_t* launches8, bool reset) {
    WavefrontState* w = c->wf;
    for (int i = 0; i < PT_N_STAGES; ++i) {
        ms8[i] = w ? w->stageMs[i] : 0.0; launches8[i] = w ? w->stageLaunches[i] : 0;
        if (w && reset) { w->stageMs[i] = 0.0; w->stageLaunches[i] = 0; }
    }
}

#define WF_THREADS 128
#define WF_COUNT_ARRAYS 10       /* per-bounce device counters: qCount sCount workE workO oCount primaryCount mCount xCount xnCount wCount */
#define WF_RAY_COUNTERS 16      /* device-side u64 counters: k_fold's five, then work counters [8..11] (option count_work) */

static __device__ __forceinline__ uint32_t warp_push(uint32_t* counter, bool pred) {
    // warp-aggregated queue append: one atomic per warp; returns this lane's slot (valid when pred)
    unsigned mask = __ballot_sync(0xffffffffu, pred);
    if (mask == 0) return 0;
    int lane = threadIdx.x & 31;
    int leader = __ffs(mask) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(counter, (uint32_t)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + (uint32_t)__popc(mask & ((1u << lane) - 1u));
}

__global__ void __launch_bounds__(WF_THREADS) k_generate(const __grid_constant__ WfParams P, PathArrays A, uint32_t* queue, uint32_t* qCount,
                                                         uint32_t* primaryCount) {
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < P.nPaths; p += gridDim.x * blockDim.x) {
        stage_generate(P, A, p);
        queue[p] = p;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) { qCount[0] = P.nPaths; *primaryCount = P.rows * (uint32_t)P.W; }
}

// ---- persistent traversal kernels ------------------------------------------------------------
// One warp = 32 ray slots. Every iteration the warp votes: lanes whose ray is finished are refilled
// from the queue (chunks of PT_CHUNK indices per global atomic, handed out by ballot rank) once
// enough of them are idle; then the whole warp runs ONE kind of step — a triangle test for the lanes
// that hold a triangle group, or a node visit for the lanes that hold a node group — so that the
// divergent halves of the classic while-while loop each run with as many lanes as possible.
// Triangle groups met while the warp is visiting nodes are postponed on the lane's stack
// (trav_node_step). Grid = exactly the resident blocks of the device: no second wave.
#ifndef PT_CHUNK
#define PT_CHUNK 128u
#endif
#define PT_FETCH_MIN_IDLE 8        /* refill once this many lanes are idle (or all are) */

struct WarpFeeder {                // warp-uniform
    uint32_t chunk = PT_CHUNK;     // indices per global atomic (feeder_chunk)
    uint32_t next = 0, end = 0;
    bool exhausted = false;
};
// Hands queue indices to idle lanes (idleMask = ballot of idle lanes, warp-uniform). Returns whether THIS
// lane got one.
static __device__ __forceinline__ bool feeder_take(WarpFeeder& f, uint32_t* workCounter, uint32_t n, unsigned idleMask, uint32_t& index) {
    const unsigned lane = threadIdx.x & 31u;
    if (f.next == f.end) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(workCounter, f.chunk);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) { f.exhausted = true; return false; }
        f.next = base; f.end = min(base + f.chunk, n);
    }
    uint32_t give = min((uint32_t)__popc(idleMask), f.end - f.next);
    uint32_t rank = (uint32_t)__popc(idleMask & ((1u << lane) - 1u));
    bool got = ((idleMask >> lane) & 1u) && rank < give;
    index = f.next + rank;
    f.next += give;
    return got;
}

// A queue too short to give every resident warp a full chunk is handed out a warp's width at a time: with 128 a launch
// of 259 k rays (one eighth of a 1080p frame) kept 2 000 of the 4 700 resident warps busy for four generations of rays each,
// with 32 all of them for two — the length of a small launch is the time ONE warp needs.
#ifndef PT_CHUNK_SMALL
#define PT_CHUNK_SMALL 32u
#endif
static __device__ __forceinline__ uint32_t feeder_chunk(uint32_t n) {
    const uint32_t warps = gridDim.x * (blockDim.x >> 5);
    return n < PT_CHUNK * warps ? PT_CHUNK_SMALL : PT_CHUNK;
}

#ifndef PT_STREAM_PREFETCH
#define PT_STREAM_PREFETCH 0  /* grid-stride kernels request the next round's path state into L2 */
#endif
static __device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
#ifndef PT_STACK_TOS
#define PT_STACK_TOS 0        /* top of the per-lane stack in registers (TopCachedStack, pt_traverse.h) */
#endif
#ifndef PT_SPLIT_STACK
#define PT_SPLIT_STACK 0      /* A/B on B200: shared-memory stack 250.2, local-memory stack 252.3 Mpaths/s: local (L1-resident) wins */
#endif
#if PT_SPLIT_STACK
struct SplitStack {
    uint2* sh;                                            // this thread's column of the block's shared array
    uint32_t base[PT_BVH8_STACK - PT_BVH8_SHARED_STACK], bits[PT_BVH8_STACK - PT_BVH8_SHARED_STACK];
    __device__ __forceinline__ void put(int i, uint32_t b, uint32_t v) {
        if (i < PT_BVH8_SHARED_STACK) sh[i * WF_THREADS] = make_uint2(b, v);
        else { base[i - PT_BVH8_SHARED_STACK] = b; bits[i - PT_BVH8_SHARED_STACK] = v; }
    }
    __device__ __forceinline__ void get(int i, uint32_t& b, uint32_t& v) {
        if (i < PT_BVH8_SHARED_STACK) { uint2 e = sh[i * WF_THREADS]; b = e.x; v = e.y; }
        else { b = base[i - PT_BVH8_SHARED_STACK]; v = bits[i - PT_BVH8_SHARED_STACK]; }
    }
};
#define PT_DECLARE_STACK(name)                                              \
    __shared__ uint2 name##_shared[PT_BVH8_SHARED_STACK * WF_THREADS];      \
    SplitStack name;                                                        \
    name.sh = name##_shared + threadIdx.x
#elif PT_STACK_TOS
#define PT_DECLARE_STACK(name) TopCachedStack name
#else
#define PT_DECLARE_STACK(name) PrivateStack name
#endif

// One vote per iteration: a single warp reduction of (busy | holds triangles << 8 | can visit a node << 16)
// gives the three lane counts the schedule needs.
static __device__ __forceinline__ unsigned warp_votes(bool busy, bool hasTri, bool canNode) {
    return __reduce_add_sync(0xffffffffu, (busy ? 1u : 0u) | (hasTri ? 0x100u : 0u) | (canNode ? 0x10000u : 0u));
}

#ifndef PT_TRI_PER_STEP
#define PT_TRI_PER_STEP 2          /* triangles a lane tests per triangle step of the warp: 1 / 2 / 3 -> 310.1 / 318.9 / 311.5 Mpaths/s (fewer votes per triangle) */
#endif
#ifndef PT_NODE_PER_STEP
#define PT_NODE_PER_STEP 1         /* node visits per node step for lanes that came out of the first one without triangles (A/B knob) */
#endif
#ifndef PT_TRI_PER_STEP_OCCL
#define PT_TRI_PER_STEP_OCCL 2     /* the same for k_occlude: 1 / 2 / 3 -> 321.2 / 326.4 / 324.8 */
#endif
#ifndef PT_EXTEND_MIN_BLOCKS
#define PT_EXTEND_MIN_BLOCKS 7     /* 72 registers: 7 blocks x 4 warps per SM (A/B: 6 -> 80 regs, 8 -> 64 regs + spills) */
#endif
#ifndef PT_EXTEND_MIN_BLOCKS_OPT
#define PT_EXTEND_MIN_BLOCKS_OPT 8 /* the kernels without the in-loop reachability confirmation fit 64 registers (k_extend: no spill, k_occlude: 20 bytes) */
#endif
// COUNT (option "count_work"): node visits and triangle tests of the launch are added to work[0], work[1] — the
// kernel's own algorithmic work in the WIDE layout (80 B per node visit, 64 B per triangle test), which the bench's
// roofline divides by the measured L2 gather bandwidth.
static __device__ __forceinline__ void flush_work(unsigned long long* work, uint32_t nodes, uint32_t tris) {
    nodes = __reduce_add_sync(0xffffffffu, nodes);
    tris = __reduce_add_sync(0xffffffffu, tris);
    if ((threadIdx.x & 31u) == 0u) { atomicAdd(work, (unsigned long long)nodes); atomicAdd(work + 1, (unsigned long long)tris); }
}
// OPT (option exact_order = 1, the default): candidate hits are accepted without the reachability confirmation
// (pt_traverse.h: trav_tri_step_closest<false>), a contested ray is only MARKED (PT_HIT_CONTESTED in the stored id), and
// k_verify_extend confirms every ray's final hit at full lane count and queues the rays for k_exact_extend. Without OPT
// (exact_order = 0: no re-trace to fall back on) the confirmation stays inside the loop.
template <bool COUNT, bool OPT>
__global__ void __launch_bounds__(WF_THREADS, OPT ? PT_EXTEND_MIN_BLOCKS_OPT : PT_EXTEND_MIN_BLOCKS) k_extend(const __grid_constant__ WfParams P, const __grid_constant__ WideScene sc, PathArrays A,
                                                       const uint32_t* __restrict__ queue, const uint32_t* __restrict__ qCount,
                                                       uint32_t* workCounter, Hit* __restrict__ hitOut, uint32_t* exactQueue, uint32_t* exactCount,
                                                       unsigned long long* work) {
    uint32_t cntNode = 0, cntTri = 0;
    const uint32_t n = *qCount;
    const uint32_t sphereCount = P.pc.params.sphereCount, objectCount = P.pc.params.objectCount;
    PT_DECLARE_STACK(stack);
    WarpFeeder feed;
    feed.chunk = feeder_chunk(n);
    Trav t;
    Hit hit;
    uint32_t key = 0xffffffffu, p = 0;
    bool busy = false;
    hit.dst = PT_FAR; hit.id = PT_HIT_MISS; hit.u = 0.f; hit.v = 0.f;
    t.gBits = 0; t.tBits = 0; t.sp = 0;
    for (;;) {
        // lanes with nothing in hand take the next group from their stack, or finish
        if (busy && !trav_has_node(t) && !trav_has_tri(t) && !trav_pop(t, stack)) {
            if (OPT) {
                if (trav_contested(t)) hit.id |= PT_HIT_CONTESTED;                  // contested implies a hit (pt_traverse.h)
                hitOut[p] = hit;
            } else {
                hitOut[p] = hit;
                if (trav_contested(t)) exactQueue[atomicAdd(exactCount, 1u)] = p;  // rare (coincident surfaces): k_exact_extend overwrites the hit
            }
            busy = false;
        }
        const bool hasTri = busy && trav_has_tri(t);
        const bool canNode = busy && trav_has_node(t) && (!hasTri || t.sp < PT_BVH8_POSTPONE_LIMIT);
        const unsigned votes = warp_votes(busy, hasTri, canNode);
        const unsigned nBusy = votes & 0xffu, nTri = (votes >> 8) & 0xffu, nNode = votes >> 16;
        if (!feed.exhausted && 32u - nBusy >= P.fetchMinIdle) {
            uint32_t qi;
            if (feeder_take(feed, workCounter, n, __ballot_sync(0xffffffffu, !busy), qi)) {
                p = queue[qi];
                Ray r; r.o = xyz(A.rayO[p]); r.d = xyz(A.rayD[p]);
                spheres_closest(sc, r, sphereCount, hit, key);
                trav_begin(t, r);
                busy = true;
            }
            continue;
        }
        if (nBusy == 0u) break;                    // queue exhausted and every lane done
        if (nNode == 0u || 32u * nTri >= P.triVote32 * nBusy) {
            if (hasTri) {
#pragma unroll 1
                for (int r = 0; r < PT_TRI_PER_STEP && trav_has_tri(t); ++r) {
                    trav_tri_step_closest<!OPT>(sc.all, sc.xforms, sc.up, sc.nSpheres, objectCount, sc.extent, t, hit, key);
                    if (COUNT) cntTri++;
                }
            }
        } else {
            if (canNode) {
                trav_node_step(sc.all, t, stack, cull_dst(hit.dst, sc.extent)); if (COUNT) cntNode++;
#pragma unroll 1
                for (int r = 1; r < PT_NODE_PER_STEP && trav_has_node(t) && !trav_has_tri(t); ++r) { trav_node_step(sc.all, t, stack, cull_dst(hit.dst, sc.extent)); if (COUNT) cntNode++; }
            }
        }
    }
    if (COUNT) flush_work(work, cntNode, cntTri);
}

// Block-aggregated queue append: ONE global atomic per block per round instead of one per warp. With 16 M paths per
// launch the per-warp form issued 0.5 M atomics on each of two addresses and shade's stall samples pointed at them
// (54 %, profiles/r01/k_shade_v12_by_line.txt); measured gain 0.4 ms of 9.5 per step — the wait was mostly overlapped.
// Every thread of the block must call it (uniform trip counts).
#ifndef SHADE_THREADS
#define SHADE_THREADS 128       // 128 x 8 blocks at 64 registers: 6.3 ms per step; 256 x 3 at 75 registers: 7.0 ms
#endif
template <int NQ>
static __device__ __forceinline__ void block_push(uint32_t* const (&counters)[NQ], const bool (&pred)[NQ], uint32_t (&slot)[NQ]) {
    __shared__ uint32_t warpCount[NQ][SHADE_THREADS / 32];
    __shared__ uint32_t blockBase[NQ];
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    unsigned mask[NQ];
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        mask[q] = __ballot_sync(0xffffffffu, pred[q]);
        if (lane == 0) warpCount[q][warp] = (uint32_t)__popc(mask[q]);
    }
    __syncthreads();
    if (threadIdx.x < NQ) {
        uint32_t total = 0;
        for (unsigned k = 0; k < blockDim.x / 32; ++k) total += warpCount[threadIdx.x][k];
        blockBase[threadIdx.x] = total ? atomicAdd(counters[threadIdx.x], total) : 0u;
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        uint32_t before = 0;
        for (unsigned k = 0; k < warp; ++k) before += warpCount[q][k];
        slot[q] = blockBase[q] + before + (uint32_t)__popc(mask[q] & ((1u << lane) - 1u));
    }
    __syncthreads();            // warpCount is reused by the next round
}

// After k_extend<*, true>: every ray's final hit, at full lane count. A triangle hit that the reference would not have
// tested (pt_traverse.h: ref_reachable) and every ray marked contested go to k_exact_extend; for the former the stored
// distance becomes PT_FAR, which tells the re-trace not to use it as its starting bound. Two kernels: the quick accept
// for every ray here, the leaf -> root walk (1 % of the hits) from a compacted queue in k_verify_walk — inside one
// kernel a warp with one walking lane waited for it (10 of 32 lanes per instruction, 7.2 ms per step instead of ~3).
__global__ void __launch_bounds__(WF_THREADS) k_verify_extend(const __grid_constant__ WideScene sc, PathArrays A, const uint32_t* __restrict__ queue,
                                                              const uint32_t* __restrict__ qCount, const Hit* __restrict__ hit, uint32_t* exactQueue,
                                                              uint32_t* exactCount, uint32_t* walkQueue, uint32_t* walkCount) {
    const uint32_t n = *qCount;
    uint32_t* const counters[2] = {walkCount, exactCount};
    // uniform trip count: the two appends are aggregated per block (one atomic per block and queue: the per-ray form put
    // 0.6 M atomics per launch on two addresses)
    for (uint32_t i0 = blockIdx.x * blockDim.x; i0 < n; i0 += gridDim.x * blockDim.x) {
        const uint32_t i = i0 + threadIdx.x;
        bool pred[2] = {false, false};
        uint32_t p = 0;
        if (i < n) {
            p = queue[i];
            const Hit h = hit[p];
            if (h.id != PT_HIT_MISS) {
                Ray r; r.o = xyz(A.rayO[p]); r.d = xyz(A.rayD[p]);
                if (!hit_reachable_quick(sc, r, h)) pred[0] = true;
                else if (h.id & PT_HIT_CONTESTED) pred[1] = true;
            }
        }
        uint32_t slot[2];
        block_push<2>(counters, pred, slot);
        if (pred[0]) walkQueue[slot[0]] = p;
        if (pred[1]) exactQueue[slot[1]] = p;
    }
}
__global__ void __launch_bounds__(WF_THREADS) k_verify_walk(const __grid_constant__ WideScene sc, PathArrays A, const uint32_t* __restrict__ walkQueue,
                                                            const uint32_t* __restrict__ walkCount, Hit* hit, uint32_t* exactQueue, uint32_t* exactCount) {
    const uint32_t n = *walkCount;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t p = walkQueue[i];
        const Hit h = hit[p];
        Ray r; r.o = xyz(A.rayO[p]); r.d = xyz(A.rayD[p]);
        const bool ok = hit_reachable_walk(sc, r, h);
        if (!ok) hit[p].dst = PT_FAR;
        if (!ok || (h.id & PT_HIT_CONTESTED)) exactQueue[atomicAdd(exactCount, 1u)] = p;
    }
}
static void launch_verify(b200pt_ctx* c, int grid, cudaStream_t st, const WideScene& sc, const PathArrays& A, const uint32_t* queue,
                          const uint32_t* count, uint32_t* exactQ, uint32_t* xCount, uint32_t* walkQ, uint32_t* wCount) {
    k_verify_extend<<<grid, WF_THREADS, 0, st>>>(sc, A, queue, count, A.hit, exactQ, xCount, walkQ, wCount);
    k_verify_walk<<<grid, WF_THREADS, 0, st>>>(sc, A, walkQ, wCount, A.hit, exactQ, xCount);      // a chain of dependent parent loads per ray: as many resident warps as fit
    c->counters[0] += 2;
}
static void launch_extend(b200pt_ctx* c, int grid, const WfParams& P, const WideScene& sc, const PathArrays& A, const uint32_t* queue, const uint32_t* count,
                          uint32_t* workCounter, uint32_t* exactQ, uint32_t* xCount, unsigned long long* workCnt) {
    const bool countWork = c->countWork && workCnt;
    if (c->exactOrder) {
        if (countWork) k_extend<true, true><<<grid, WF_THREADS, 0, c->stream>>>(P, sc, A, queue, count, workCounter, A.hit, exactQ, xCount, workCnt);
        else k_extend<false, true><<<grid, WF_THREADS, 0, c->stream>>>(P, sc, A, queue, count, workCounter, A.hit, exactQ, xCount, nullptr);
    } else {
        if (countWork) k_extend<true, false><<<grid, WF_THREADS, 0, c->stream>>>(P, sc, A, queue, count, workCounter, A.hit, exactQ, xCount, workCnt);
        else k_extend<false, false><<<grid, WF_THREADS, 0, c->stream>>>(P, sc, A, queue, count, workCounter, A.hit, exactQ, xCount, nullptr);
    }
}

#ifndef PT_SHADE_MIN_BLOCKS
#define PT_SHADE_MIN_BLOCKS 8
#endif
__global__ void __launch_bounds__(SHADE_THREADS, PT_SHADE_MIN_BLOCKS) k_shade(const __grid_constant__ WfParams P, const __grid_constant__ WideScene sc, PathArrays A,
                                                         const uint32_t* __restrict__ queue, const uint32_t* __restrict__ qCount,
                                                         uint32_t* nextQueue, uint32_t* nextCount, uint32_t* shadowQueue, uint32_t* shadowCount,
                                                         uint32_t j) {
    const uint32_t n = *qCount;
    uint32_t* const counters[2] = {nextCount, shadowCount};
#if PT_STREAM_PREFETCH
    // The path state of the NEXT round of this thread is requested into L2 while this round is shaded (the index of
    // the round after that is fetched at the same time): the kernel is bound by the latency of its dependent loads
    // (path state -> shading record -> material), not by bandwidth (ncu: 55 % of DRAM peak, long-scoreboard stalls).
    const uint32_t stride = gridDim.x * blockDim.x;
    uint32_t pNext = 0;
    { const uint32_t i1 = blockIdx.x * blockDim.x + threadIdx.x + stride; if (i1 < n) pNext = queue[i1]; }
#endif
    for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x) {      // block-uniform trip count
        const uint32_t i = base + threadIdx.x;
        uint32_t flags = 0, p = 0;
#if PT_STREAM_PREFETCH
        if (i + stride < n) {
            const uint32_t q = pNext;
            prefetch_l2(&A.rayO[q]); prefetch_l2(&A.rayD[q]); prefetch_l2(&A.total[q]); prefetch_l2(&A.atten[q]); prefetch_l2(&A.direct[q]);
            if (!(j == 0 && P.sharePrimary)) prefetch_l2(&A.hit[q]);
        }
        if (i + 2u * stride < n) pNext = queue[i + 2u * stride];
#endif
        if (i < n) { p = queue[i]; flags = stage_shade(P, sc, A, p, j); }
        const bool go = (flags & PT_SHADE_CONTINUE) != 0;
        const bool pred[2] = {go, go && (flags & PT_SHADE_CONNECT) != 0};
        uint32_t slot[2];
        block_push<2>(counters, pred, slot);
        if (pred[0]) nextQueue[slot[0]] = p;
        if (pred[1]) shadowQueue[slot[1]] = p;
    }
}

// connect, part 1: emissive candidate of the light-sample ray; emits the occlusion jobs (path indices)
#ifndef PT_PREP_MIN_BLOCKS
#define PT_PREP_MIN_BLOCKS 8        // 64 registers: 2.98 ms per step; unconstrained (80 registers, 6 blocks) 3.12; 10 blocks 3.26
#endif
__global__ void __launch_bounds__(WF_THREADS, PT_PREP_MIN_BLOCKS) k_connect_prep(const __grid_constant__ WfParams P, const __grid_constant__ WideScene sc, PathArrays A,
                                                             const uint32_t* __restrict__ queue, const uint32_t* __restrict__ sCount,
                                                             uint32_t* jobQueue, uint32_t* jobCount, uint32_t* exactQueue, uint32_t* exactCount) {
    uint32_t n = *sCount;
    uint32_t nRound = (n + 31u) & ~31u;           // whole warps stay together for the ballot
#if PT_STREAM_PREFETCH
    const uint32_t stride = gridDim.x * blockDim.x;
    uint32_t pNext = 0;
    { const uint32_t i1 = blockIdx.x * blockDim.x + threadIdx.x + stride; if (i1 < n) pNext = queue[i1]; }
#endif
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nRound; i += gridDim.x * blockDim.x) {
        uint32_t jobs = 0, p = 0;
#if PT_STREAM_PREFETCH
        if (i + stride < n) { const uint32_t q = pNext; prefetch_l2(&A.qA[q]); prefetch_l2(&A.qL[q]); prefetch_l2(&A.qO[q]); prefetch_l2(&A.hit[q]); }
        if (i + 2u * stride < n) pNext = queue[i + 2u * stride];
#endif
        if (i < n) { p = queue[i]; jobs = stage_connect_prep(P, sc, A, p); }
        uint32_t s0 = warp_push(jobCount, jobs == PT_CONNECT_JOB_LIGHT);
        if (jobs == PT_CONNECT_JOB_LIGHT) jobQueue[s0] = p;
        if (jobs == PT_CONNECT_JOB_EXACT) exactQueue[atomicAdd(exactCount, 1u)] = p;      // rare
    }
}

// connect, part 2: bounded any-hit walks of the full BVH, same warp scheduling as k_extend
template <bool COUNT, bool OPT>
__global__ void __launch_bounds__(WF_THREADS, OPT ? PT_EXTEND_MIN_BLOCKS_OPT : PT_EXTEND_MIN_BLOCKS) k_occlude(const __grid_constant__ WfParams P, const __grid_constant__ WideScene sc, PathArrays A,
                                                        const uint32_t* __restrict__ jobQueue, const uint32_t* __restrict__ jobCount,
                                                        uint32_t* workCounter, uint32_t* exactQueue, uint32_t* exactCount, unsigned long long* work) {
    uint32_t cntNode = 0, cntTri = 0;
    const uint32_t n = *jobCount;
    const uint32_t objectCount = P.pc.params.objectCount;
    PT_DECLARE_STACK(stack);
    WarpFeeder feed;
    feed.chunk = feeder_chunk(n);
    Trav t;
    float tE = 0.f, ws = 1.f;
    uint32_t keyE = 0, job = 0;
    bool busy = false;
    t.gBits = 0; t.tBits = 0; t.sp = 0;
    for (;;) {
        if (busy && !trav_has_node(t) && !trav_has_tri(t) && !trav_pop(t, stack)) busy = false;   // not occluded
        const bool hasTri = busy && trav_has_tri(t);
        const bool canNode = busy && trav_has_node(t) && (!hasTri || t.sp < PT_BVH8_POSTPONE_LIMIT);
        const unsigned votes = warp_votes(busy, hasTri, canNode);
        const unsigned nBusy = votes & 0xffu, nTri = (votes >> 8) & 0xffu, nNode = votes >> 16;
        if (!feed.exhausted && 32u - nBusy >= P.fetchMinIdleOccl) {
            uint32_t qi;
            if (feeder_take(feed, workCounter, n, __ballot_sync(0xffffffffu, !busy), qi)) {
                job = jobQueue[qi];
                Ray r;
                connect_job_ray(A, job, r, tE, keyE, ws);
                trav_begin(t, r);
                busy = true;
            }
            continue;
        }
        if (nBusy == 0u) break;
        if (nNode == 0u || 32u * nTri >= P.triVoteOccl32 * nBusy) {
            if (hasTri) {
#pragma unroll 1
                for (int r = 0; r < PT_TRI_PER_STEP_OCCL && trav_has_tri(t); ++r) {
                    if (COUNT) cntTri++;
                    uint32_t occluder = 0; float occDst = 0.f;
                    const int o = trav_tri_step_occluded<!OPT>(sc.all, sc.xforms, sc.up, sc.nSpheres, objectCount, sc.extent, ws, t, tE, keyE, &occluder, &occDst);
                    if (o != PT_OCCL_NO) {
                        if (o == PT_OCCL_YES) { if (OPT) connect_job_occluded_by(A, job, occluder, occDst); else connect_job_occluded(A, job); }
                        else exactQueue[atomicAdd(exactCount, 1u)] = job;      // contested: k_exact_nee answers in the reference's order
                        busy = false;
                        break;
                    }
                }
            }
        } else {
            if (canNode) {
                trav_node_step(sc.all, t, stack, tE + contest_window(tE, sc.extent) * ws); if (COUNT) cntNode++;
#pragma unroll 1
                for (int r = 1; r < PT_NODE_PER_STEP && trav_has_node(t) && !trav_has_tri(t); ++r) { trav_node_step(sc.all, t, stack, tE + contest_window(tE, sc.extent) * ws); if (COUNT) cntNode++; }
            }
        }
    }
    if (COUNT) flush_work(work, cntNode, cntTri);
}

// After k_occlude<*, true>: the occluders found, at full lane count; a query whose occluder the reference would not have tested
// is answered by k_exact_nee instead (qE.x < 0 there means: no bound for the search).
__global__ void __launch_bounds__(WF_THREADS) k_verify_occluded(const __grid_constant__ WideScene sc, PathArrays A, const uint32_t* __restrict__ jobQueue,
                                                                const uint32_t* __restrict__ jobCount, uint32_t* exactQueue, uint32_t* exactCount) {
    const uint32_t n = *jobCount;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t job = jobQueue[i];
        if (!connect_job_occluder_reachable(sc, A, job)) exactQueue[atomicAdd(exactCount, 1u)] = job;
    }
}

// connect, part 3: pdf, power-heuristic weight, direct light of the paths whose light-sample ray reached the light (the
// others keep the zero shade stored: one 16-byte read each); queues the (practically never occurring) paths whose
// misWeight is live
__global__ void __launch_bounds__(WF_THREADS) k_connect_finish(const __grid_constant__ WideScene sc, PathArrays A,
                                                               const uint32_t* __restrict__ queue, const uint32_t* __restrict__ sCount,
                                                               uint32_t* misQueue, uint32_t* misCount) {
    uint32_t n = *sCount;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        uint32_t p = queue[i];
        if (connect_job_finish(sc, A, p)) misQueue[atomicAdd(misCount, 1u)] = p;
    }
}
__global__ void __launch_bounds__(WF_THREADS) k_connect_mis(const __grid_constant__ WfParams P, const __grid_constant__ RefScene ref, PathArrays A,
                                                            const uint32_t* __restrict__ misQueue, const uint32_t* __restrict__ misCount) {
    uint32_t n = *misCount;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        connect_mis_weight_exact(ref, P.pc.params, A, misQueue[i]);
}

// Contested rays (pt_traverse.h: two reachable candidates closer together than the rounding slack of their distances,
// coincident surfaces mostly): calculateIntersections in the reference's own order over the reference's own buffers.
// One thread per ray; a few per thousand of the rays of a launch.
__global__ void __launch_bounds__(WF_THREADS) k_exact_extend(const __grid_constant__ WfParams P, const __grid_constant__ WideScene sc, const __grid_constant__ RefScene ref,
                                                             PathArrays A, const uint32_t* __restrict__ queue, const uint32_t* __restrict__ count, Hit* hitOut,
                                                             unsigned long long* total) {
    const uint32_t n = *count;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t p = queue[i];
        Ray r; r.o = xyz(A.rayO[p]); r.d = xyz(A.rayD[p]);
        hitOut[p] = exact_closest_hit(ref, sc, P.pc.params, r, hitOut[p].dst);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && total) atomicAdd(total, (unsigned long long)n);
}
__global__ void __launch_bounds__(WF_THREADS) k_exact_nee(const __grid_constant__ WfParams P, const __grid_constant__ WideScene sc, const __grid_constant__ RefScene ref, PathArrays A,
                                                          const uint32_t* __restrict__ queue, const uint32_t* __restrict__ count, unsigned long long* total) {
    const uint32_t n = *count;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) connect_exact(ref, sc, P.pc.params, A, queue[i]);
    if (blockIdx.x == 0 && threadIdx.x == 0 && total) atomicAdd(total, (unsigned long long)n);
}

// (object, reference triangle) -> record index in the wide layout, from the finished tree
__global__ void k_wide_prim_of(const PrimRecord* __restrict__ prims, uint32_t n, const uint32_t* __restrict__ instBase, const uint32_t* __restrict__ triFirst,
                               uint32_t* primOf) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t o = prims[i].objectIndex, t = prims[i].triIndexFlags & PT_PRIM_TRIMASK;
    primOf[instBase[o] + t - triFirst[o]] = i;
}

__global__ void __launch_bounds__(WF_THREADS) k_accumulate(const __grid_constant__ WfParams P, PathArrays A, float4* accum) {
    uint32_t n = P.rows * (uint32_t)P.W;
    for (uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x; pix < n; pix += gridDim.x * blockDim.x)
        stage_accumulate(P, A, accum, pix);
}

// Multi-sample dispatches (raysPerPixel / sampleLimit > 1, raytrace.comp:568-574): outColor = (sum of the samples in
// order) / samples. After the last sample the mean replaces A.total, which accumulate / dispatch_resolve read.
__global__ void __launch_bounds__(WF_THREADS) k_sample_sum(PathArrays A, float4* sum, uint32_t nPaths, uint32_t sample, uint32_t samples) {
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < nPaths; p += gridDim.x * blockDim.x) {
        v3 acc = sample ? xyz(sum[p]) : splat3(0.f);
        acc = acc + xyz(A.total[p]);
        if (sample + 1 == samples) A.total[p] = f4(acc / float(samples), 0.f);
        else sum[p] = f4(acc, 0.f);
    }
}

// One literal dispatch served by the wavefront: main() after trace() (raytrace.comp:568-593) for 1-sample dispatches.
// outColor of the single frame slot goes through the same ref_resolve as the per-pixel kernel (progressive blend with
// the stored image, NaN/Inf -> magenta) and, with option storage_rgba8, the same unorm8 rounding.
__global__ void __launch_bounds__(WF_THREADS) k_dispatch_resolve(const __grid_constant__ WfParams P, PathArrays A, float4* image, int storeUnorm8) {
    uint32_t n = P.rows * (uint32_t)P.W;
    for (uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x; pix < n; pix += gridDim.x * blockDim.x) {
        RefPixel px;
        px.outColor = xyz(A.total[pix]);
        px.stats0 = 0.f; px.stats1 = 0.f;
        float4 old = image[pix];
        v3 fin = ref_resolve(P.pc, px, mk3(old.x, old.y, old.z));
        if (storeUnorm8) fin = mk3(unorm8_round(fin.x), unorm8_round(fin.y), unorm8_round(fin.z));
        image[pix] = make_float4(fin.x, fin.y, fin.z, 1.f);
    }
}

// ray counters of one batch. Algorithmic (what the reference's trace() does for these paths): [0] += extension rays
// (sum of qCount: one calculateIntersections per path per bounce), [1] += connect items (diffuse bounces whose NEE
// query reached the connect stage; the reference traces three rays for each, raytrace.comp:443,447,453).
// Traced (what the kernels really walked): [2] += rays of the k_extend launches (bounce 0 is ONE ray per pixel when
// the batch shares the primary hit between its frame slots), [3] += occlusion walks (k_occlude jobs), [4] += exact
// pdf rays (k_connect_mis, lazily evaluated).
__global__ void k_fold(const uint32_t* qCount, const uint32_t* sCount, const uint32_t* oCount, const uint32_t* mCount,
                       const uint32_t* primaryCount, uint32_t sharePrimary, uint32_t nBounce, unsigned long long* rayCounters) {
    if (blockIdx.x || threadIdx.x) return;
    unsigned long long e = 0, s = 0, o = 0, m = 0;
    for (uint32_t j = 0; j < nBounce; ++j) { e += qCount[j]; s += sCount[j]; o += oCount[j]; m += mCount[j]; }
    rayCounters[0] += e;
    rayCounters[1] += s;
    rayCounters[2] += sharePrimary ? e - qCount[0] + *primaryCount : e;
    rayCounters[3] += o;
    rayCounters[4] += m;
}

// b200pt_trace_batch: caller rays go through the SAME persistent closest-hit kernel as the renderer (k_extend) —
// k_load_rays fills the path slots and an identity queue, k_hit_records turns the Hit of every ray into the 48-byte
// record (surface data as calculateIntersections returns it, raytrace.comp:316-322).
__global__ void __launch_bounds__(WF_THREADS) k_load_rays(const float4* __restrict__ rays, uint32_t n, PathArrays A, uint32_t* queue, uint32_t* qCount) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        A.rayO[i] = rays[2 * (size_t)i];
        A.rayD[i] = rays[2 * (size_t)i + 1];
        queue[i] = i;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) *qCount = n;
}
__global__ void __launch_bounds__(WF_THREADS) k_hit_records(const __grid_constant__ WideScene sc, PathArrays A, uint32_t n, HitRecord48* out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        Ray r; r.o = xyz(A.rayO[i]); r.d = xyz(A.rayD[i]);
        const Hit h = A.hit[i];
        HitRecord48 rec;
        memset(&rec, 0, sizeof(rec));
        rec.dst = h.dst; rec.objectHitIndex = 0xffffffffu;
        if (h.id != PT_HIT_MISS) {
            v3 hp, nn; uint32_t mat, obj, prim; bool front;
            hit_surface(sc, r, h, hp, nn, mat, front, obj, prim);
            rec.kind = (h.id & PT_HIT_SPHERE) ? 1u : 2u;
            rec.objectHitIndex = obj; rec.triHitIndex = prim; rec.materialIndex = mat; rec.frontFace = front ? 1u : 0u;
            rec.normal[0] = nn.x; rec.normal[1] = nn.y; rec.normal[2] = nn.z;
            rec.hitPoint[0] = hp.x; rec.hitPoint[1] = hp.y; rec.hitPoint[2] = hp.z;
        }
        out[i] = rec;
    }
}

// Vulkan unorm8 store of the float image: clamp, scale, round to nearest even (f4 of SURVEY.md §8)
__global__ void k_image_to_rgba8(const float4* __restrict__ img, uchar4* out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 c = img[i];
    auto q = [](float f) -> unsigned char {
        if (!(f > 0.f)) return 0;
        if (f >= 1.f) return 255;
        return (unsigned char)__float2int_rn(f * 255.f);
    };
    out[i] = make_uchar4(q(c.x), q(c.y), q(c.z), 255);
}

// ---------------------------------------------------------------------------------------------
static int wf_fail(b200pt_ctx* c, int code, const char* what, cudaError_t e) {
    c->error = std::string(what) + ": " + (e == cudaSuccess ? "" : cudaGetErrorString(e));
    return code;
}
#define WCU(call)                                                        \
    do {                                                                 \
        cudaError_t e_ = (call);                                         \
        if (e_ != cudaSuccess) return wf_fail(c, 2, #call, e_);          \
    } while (0)

static cudaError_t wf_resize(DeviceBuffer& b, size_t bytes) {
    if (b.p && bytes <= b.bytes) return cudaSuccess;
    if (b.p) cudaFree(b.p);
    b.p = nullptr; b.bytes = 0;
    if (!bytes) return cudaSuccess;
    cudaError_t e = cudaMalloc(&b.p, bytes);
    if (e == cudaSuccess) b.bytes = bytes;
    return e;
}
static void wf_free(DeviceBuffer& b) { if (b.p) cudaFree(b.p); b.p = nullptr; b.bytes = 0; }
// frees a device-built tree (the four DeviceBuffers only alias it); host-uploaded trees stay until overwritten
static void wf_release_tree(WavefrontState* w) {
    if (!w->wideOnDevice) return;
    w->dNodes = DeviceBuffer{}; w->dPrims = DeviceBuffer{}; w->dShade = DeviceBuffer{}; w->dUv = DeviceBuffer{};
    WideDeviceResult& r = w->wideRes;
    if (r.nodes) cudaFree(r.nodes);
    if (r.prims) cudaFree(r.prims);
    if (r.shade) cudaFree(r.shade);
    if (r.uv) cudaFree(r.uv);
    if (r.nodeBox) cudaFree(r.nodeBox);
    r = WideDeviceResult{};
    w->wideOnDevice = false;
}

template <class T>
static cudaError_t wf_upload(b200pt_ctx* c, DeviceBuffer& d, const std::vector<T>& h) {
    cudaError_t e = wf_resize(d, std::max<size_t>(h.size() * sizeof(T), 16));
    if (e == cudaSuccess && !h.empty()) e = cudaMemcpyAsync(d.p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, c->stream);
    return e;
}

int wavefront_build_scene(b200pt_ctx* c, int changed) {
    if (!c->wf) {
        c->wf = new WavefrontState();
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, c->device) == cudaSuccess) c->wf->smCount = prop.multiProcessorCount;
    }
    WavefrontState* w = c->wf;
    WideSceneHost host;
    const auto tBuild = std::chrono::steady_clock::now();
    // Only spheres or materials changed and no object moved between the emissive and the non-emissive set: triangles,
    // trees and records stay; the sphere slot masks and the scene-wide flags are recomputed.
    std::vector<uint8_t> objEmissive(c->hObjects.size());
    for (size_t o = 0; o < c->hObjects.size(); ++o)
        objEmissive[o] = !(c->hMaterials[c->hObjects[o].materialIndex].emissionStrength == 0.f);
    if ((changed == WF_CHANGED_MATERIALS || changed == WF_CHANGED_SPHERES) && w->haveScene && objEmissive == w->objEmissive) {
        wide_sphere_masks(c->hSpheres, c->hMaterials, host);
        WideScene& s = w->scene;
        s.spheres = c->ref.spheres; s.materials = c->ref.materials;
        s.anyEmissive = (host.anyEmissive || w->anyEmissiveTri) ? 1u : 0u;
        s.emissionFinite = host.emissionFinite ? 1u : 0u;
        s.uniqSphereMask = host.uniqSphereMask; s.emisSphereMask = host.emisSphereMask; s.plainSphereMask = host.plainSphereMask;
        s.extent = wide_extent(c->hSpheres, c->hObjects, c->hNodes);
        c->lastWideDeviceMs = 0.0;
        c->lastSceneBuildMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tBuild).count();
        return 0;
    }
    // Only object matrices changed (the UI's "update objects", src/vk_engine.cpp:1594-1603): refit instead of rebuilding.
    if (changed == WF_CHANGED_TRANSFORMS && c->refitOnUpdate && w->haveScene && w->wideOnDevice && !w->wideRes.nDropped && !w->wideRes.levelCount.empty() &&
        objEmissive == w->objEmissive && w->hLeafOf.size() == c->hTris.size() && w->hUp.size() == c->hNodes.size()) {
        bool finite = true;
        for (const auto& o : c->hObjects) for (int k = 0; k < 16; ++k) finite = finite && std::fabs(o.m[k]) < 1e30f;
        const char* err = finite ? build_emissive_scene(c->hMaterials, c->hPoints, c->hTris, c->hObjects, c->hNodes, w->hLeafOf, w->hUp, host) : "non-finite matrix";
        if (!err) {
            WCU(wf_upload(c, w->dNodesE, host.emis.nodes));
            WCU(wf_upload(c, w->dPrimsE, host.emis.prims));
            WCU(wf_upload(c, w->dShadeE, host.emis.shade));
            WCU(wf_upload(c, w->dXforms, host.xforms));
            WCU(wf_upload(c, w->dObjXform, host.objXform));
            WideDeviceInputs in{(const ptl::TrianglePoint*)c->dPoints.p, (const ptl::Triangle*)c->dTris.p, (const ptl::RenderObject*)c->dObjects.p,
                                (const ptl::Material*)c->dMaterials.p};
            float ms = 0.f;
            err = wide_refit_device(c->stream, in, (const uint32_t*)w->dObjXform.p, w->wideRes, &ms);      // synchronises the stream
            if (!err) {
                WideScene& s = w->scene;
                s.emis = {(const Bvh8Node*)w->dNodesE.p, (const PrimRecord*)w->dPrimsE.p, (uint32_t)host.emis.nodes.size(), (uint32_t)host.emis.prims.size()};
                s.shadeEmis = (const ShadeRecord*)w->dShadeE.p;
                s.xforms = (const Xform*)w->dXforms.p; s.objXform = (const uint32_t*)w->dObjXform.p;
                s.nXforms = (uint32_t)host.xforms.size();
                s.extent = wide_extent(c->hSpheres, c->hObjects, c->hNodes);
                for (int a = 0; a < 3; ++a) { s.emisLo[a] = host.emisLo[a]; s.emisHi[a] = host.emisHi[a]; }
                c->lastWideDeviceMs = ms;
                c->lastRefit = 1;
                c->lastSceneBuildMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tBuild).count();
                return 0;
            }
            WCU(cudaStreamSynchronize(c->stream));
        }
        host = WideSceneHost();      // fall through to the full re-layout
    }
    c->lastRefit = 0;
    w->objEmissive = objEmissive; w->haveScene = false;
    const bool onDevice = c->wideBuild != 0;
    if (const char* err = build_wide_scene(c->hSpheres, c->hMaterials, c->hPoints, c->hTris, c->hObjects, c->hNodes, host, !onDevice)) {
        c->error = std::string("scene re-layout: ") + err;
        return 1;
    }
    uint32_t nNodesAll, nPrimsAll, depthAll;
    c->lastWideDeviceMs = 0.0;
    const bool timing = std::getenv("B200PT_BUILD_TIMING") != nullptr;
    if (timing) fprintf(stderr, "[relayout] host part %.3f ms\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tBuild).count());
    if (onDevice) {
        WideDeviceInputs in{(const ptl::TrianglePoint*)c->dPoints.p, (const ptl::Triangle*)c->dTris.p, (const ptl::RenderObject*)c->dObjects.p,
                            (const ptl::Material*)c->dMaterials.p};
        if (!w->wideOnDevice) { wf_free(w->dNodes); wf_free(w->dPrims); wf_free(w->dShade); wf_free(w->dUv); }
        WideDeviceResult& res = w->wideRes;
        if (const char* err = wide_build_device(c->stream, in, host, res, &w->wideCache)) { c->error = err; wf_release_tree(w); return 2; }
        w->wideOnDevice = true;
        w->dNodes = {res.nodes, (size_t)res.nNodes * sizeof(Bvh8Node)}; w->dPrims = {res.prims, (size_t)res.nPrims * sizeof(PrimRecord)};
        w->dShade = {res.shade, (size_t)res.nPrims * sizeof(ShadeRecord)}; w->dUv = {res.uv, (size_t)res.nPrims * sizeof(UvRecord)};
        nNodesAll = res.nNodes; nPrimsAll = res.nPrims; depthAll = res.maxDepth;
        c->lastWideDeviceMs = res.deviceMs;
        if (timing) fprintf(stderr, "[relayout] + device build (alloc, copies, kernels %.3f ms): %.3f ms\n", res.deviceMs,
                            std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tBuild).count());
        if (depthAll > PT_BVH8_MAX_DEPTH) { c->error = "scene re-layout: wide BVH deeper than the traversal stack"; return 1; }
    } else {
        wf_release_tree(w);
        WCU(wf_upload(c, w->dNodes, host.all.nodes));
        WCU(wf_upload(c, w->dPrims, host.all.prims));
        WCU(wf_upload(c, w->dShade, host.all.shade));
        WCU(wf_upload(c, w->dUv, host.all.uvs));
        nNodesAll = (uint32_t)host.all.nodes.size(); nPrimsAll = (uint32_t)host.all.prims.size(); depthAll = host.all.maxDepth;
    }
    WCU(wf_upload(c, w->dNodesE, host.emis.nodes));
    WCU(wf_upload(c, w->dPrimsE, host.emis.prims));
    WCU(wf_upload(c, w->dShadeE, host.emis.shade));
    WCU(wf_upload(c, w->dXforms, host.xforms));
    WCU(wf_upload(c, w->dObjXform, host.objXform));
    WCU(wf_upload(c, w->dUp, host.up));
    if (nPrimsAll >= PT_HIT_CONTESTED) { c->error = "wavefront: more than 2^29 instance triangles"; return 1; }      // hit ids keep bit 29 for PT_HIT_CONTESTED
    // (object, triangle) -> record index, for hits found by the reference-order re-trace of contested rays
    WCU(wf_upload(c, w->dInstBase, host.instBase));
    WCU(wf_upload(c, w->dTriFirst, host.triFirst));
    WCU(wf_resize(w->dPrimOf, std::max<size_t>((size_t)host.nInstSlots * 4, 16)));
    WCU(cudaMemsetAsync(w->dPrimOf.p, 0xff, std::max<size_t>((size_t)host.nInstSlots * 4, 16), c->stream));
    if (nPrimsAll) k_wide_prim_of<<<(nPrimsAll + 255) / 256, 256, 0, c->stream>>>((const PrimRecord*)w->dPrims.p, nPrimsAll, (const uint32_t*)w->dInstBase.p,
                                                                                (const uint32_t*)w->dTriFirst.p, (uint32_t*)w->dPrimOf.p);
    WCU(cudaStreamSynchronize(c->stream));        // `host` is a local
    WideScene& s = w->scene;
    s.all = {(const Bvh8Node*)w->dNodes.p, (const PrimRecord*)w->dPrims.p, nNodesAll, nPrimsAll};
    s.emis = {(const Bvh8Node*)w->dNodesE.p, (const PrimRecord*)w->dPrimsE.p, (uint32_t)host.emis.nodes.size(), (uint32_t)host.emis.prims.size()};
    s.shade = (const ShadeRecord*)w->dShade.p;
    s.shadeEmis = (const ShadeRecord*)w->dShadeE.p;
    s.uv = (const UvRecord*)w->dUv.p;
    s.tex = (const TexEntry*)c->dTexTable.p; s.srgbLut = (const float*)c->dSrgbLut.p;
    s.xforms = (const Xform*)w->dXforms.p;
    s.objXform = (const uint32_t*)w->dObjXform.p;
    s.up = (const UpNode*)w->dUp.p;
    s.primOf = (const uint32_t*)w->dPrimOf.p; s.instBase = (const uint32_t*)w->dInstBase.p; s.triFirst = (const uint32_t*)w->dTriFirst.p;
    s.spheres = c->ref.spheres; s.materials = c->ref.materials;
    s.nXforms = (uint32_t)host.xforms.size(); s.nSpheres = (uint32_t)c->hSpheres.size();
    s.nMaterials = (uint32_t)c->hMaterials.size(); s.nObjects = (uint32_t)c->hObjects.size();
    s.anyEmissive = host.anyEmissive ? 1u : 0u;
    s.emissionFinite = host.emissionFinite ? 1u : 0u;
    s.extent = host.extent;
    for (int a = 0; a < 3; ++a) {     s.emisLo[a] = host.emisLo[a];     s.emisHi[a] = host.emisHi[a]; }
    s.uniqSphereMask = host.uniqSphereMask; s.emisSphereMask = host.emisSphereMask; s.plainSphereMask = host.plainSphereMask;
    w->maxDepth = depthAll;
    w->hLeafOf = onDevice ? host.leafOf : std::vector<uint32_t>();
    w->hUp = host.up;
    w->haveScene = true;
    w->anyEmissiveTri = !host.emis.prims.empty();
    c->lastSceneBuildMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tBuild).count();
    return 0;
}

void wavefront_invalidate_frame_state(b200pt_ctx*) {}

void wavefront_set_textures(b200pt_ctx* c) {
    if (!c->wf) return;
    c->wf->scene.tex = (const TexEntry*)c->dTexTable.p;
    c->wf->scene.srgbLut = (const float*)c->dSrgbLut.p;
}

void wavefront_destroy(b200pt_ctx* c) {
    WavefrontState* w = c->wf;
    if (!w) return;
    wf_release_tree(w);
    DeviceBuffer* all[] = {&w->dNodes, &w->dPrims, &w->dShade, &w->dUv, &w->dNodesE, &w->dPrimsE, &w->dShadeE, &w->dXforms, &w->dObjXform, &w->dUp,
                           &w->dHit, &w->dQueue[0], &w->dQueue[1], &w->dShadowQ, &w->dJobQ, &w->dCounts, &w->dRayCounters, &w->dSampleSum,
                           &w->dExactQ, &w->dExactNeeQ, &w->dWalkQ, &w->dPrimOf, &w->dInstBase, &w->dTriFirst};
    for (DeviceBuffer* b : all) wf_free(*b);
    for (auto& b : w->dPaths) wf_free(b);
    for (cudaEvent_t e : w->events) cudaEventDestroy(e);
    if (w->evFork) cudaEventDestroy(w->evFork);
    if (w->evJoin) cudaEventDestroy(w->evJoin);
    if (w->sideStream) cudaStreamDestroy(w->sideStream);
    if (w->evX0) cudaEventDestroy(w->evX0);
    if (w->evX1) cudaEventDestroy(w->evX1);
    if (w->exactStream) cudaStreamDestroy(w->exactStream);
    delete w;
    c->wf = nullptr;
}

static int wf_reserve(b200pt_ctx* c, size_t paths, uint32_t bounces) {
    WavefrontState* w = c->wf;
    if (paths > w->capacity) {
        for (auto& b : w->dPaths) WCU(wf_resize(b, paths * 16));
        WCU(wf_resize(w->dHit, paths * sizeof(Hit)));
        WCU(wf_resize(w->dQueue[0], paths * 4));
        WCU(wf_resize(w->dQueue[1], paths * 4));
        WCU(wf_resize(w->dShadowQ, paths * 4));
        WCU(wf_resize(w->dJobQ, paths * 8));
        WCU(wf_resize(w->dExactQ, paths * 4));
        WCU(wf_resize(w->dExactNeeQ, paths * 4));
        WCU(wf_resize(w->dWalkQ, paths * 4));
        w->capacity = paths;
        PathArrays& A = w->arrays;
        A.rayO = (float4*)w->dPaths[0].p; A.rayD = (float4*)w->dPaths[1].p; A.atten = (float4*)w->dPaths[2].p;
        A.total = (float4*)w->dPaths[3].p; A.direct = (float4*)w->dPaths[4].p; A.qO = (float4*)w->dPaths[5].p;
        A.qL = (float4*)w->dPaths[6].p; A.qA = (float4*)w->dPaths[7].p; A.qE = (float4*)w->dPaths[8].p; A.hit = (Hit*)w->dHit.p;
    }
    uint32_t need = bounces + 3;
    if (need > w->countSlots) {
        WCU(wf_resize(w->dCounts, (size_t)need * WF_COUNT_ARRAYS * 4));      // qCount | sCount | extend work counters | occlusion work counters | job counts
        w->countSlots = need;
    }
    if (!w->dRayCounters.p) {
        WCU(wf_resize(w->dRayCounters, WF_RAY_COUNTERS * 8));
        WCU(cudaMemsetAsync(w->dRayCounters.p, 0, WF_RAY_COUNTERS * 8, c->stream));
    }
    return 0;
}

int wavefront_render_frames(b200pt_ctx* c, const ptl::PushConstants& pc, uint32_t first, uint32_t n, uint32_t stride,
                            uint32_t row0, uint32_t row1, float4* dispatchImage, uint32_t rowStride) {
    WavefrontState* w = c->wf;
    if (!w || !w->haveScene) { c->error = "wavefront: no scene (the last upload or update failed)"; return 3; }
    if (n == 0 || row1 <= row0) return 0;
    const uint32_t samples = pc.params.singleRender ? pc.params.sampleLimit : pc.params.raysPerPixel;
    if (samples == 0) {
        // no sample: outColor = 0/0 (raytrace.comp:574) — the per-pixel kernel produces exactly that
        if (rowStride != 1) { c->error = "render: a dispatch without samples cannot be split by interleaved rows"; return 1; }
        cudaError_t e = launch_ref_frames(c, pc, first, n, stride, row0, row1);
        if (e != cudaSuccess) return wf_fail(c, 2, "launch_ref_frames", e);
        return 0;
    }
    if (rowStride == 0) rowStride = 1;
    uint32_t rows = (row1 - row0 + rowStride - 1) / rowStride;
    size_t perSlot = (size_t)rows * c->W;
    uint32_t slots = (uint32_t)std::min<uint64_t>((uint64_t)c->framesInFlight, n);
    // keep a batch below 2^31 paths and ~64M paths (10 GB of path state)
    while (slots > 1 && perSlot * slots > ((size_t)1 << 26)) slots--;
    if (int rc = wf_reserve(c, perSlot * slots, pc.params.bounceLimit)) return rc;
    // grid-stride kernels: exactly one wave of resident blocks. The former fixed grids (4 or 8 blocks per SM) were 1.33
    // waves for k_shade and k_connect_prep: the last third ran at a fraction of the occupancy (shade 9.05 -> 7.0 ms,
    // prep 3.8 -> 3.1 ms per step; 2, 4 or 8 whole waves measured the same or slightly slower).
    if (!w->gridShade) {
        int perSm = 0;
        WCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_shade, SHADE_THREADS, 0));
        w->gridShade = std::max(1, perSm) * w->smCount;
        WCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_connect_prep, WF_THREADS, 0));
        w->gridPrep = std::max(1, perSm) * w->smCount;
        WCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_connect_finish, WF_THREADS, 0));
        w->gridFinish = std::max(1, perSm) * w->smCount;
    }
    const int grid = w->smCount * 8;
    const int gridShade = w->gridShade;
    WfParams P;
    P.pc = pc; P.cp = camera_plane(pc.cam); P.W = c->W; P.H = c->H; P.row0 = row0; P.rows = rows; P.rowStride = rowStride; P.stride = stride;
    P.useTextures = (c->useTextures && c->dTexTable.p) ? 1u : 0u;
    P.triVote32 = (uint32_t)c->triVote32; P.fetchMinIdle = (uint32_t)c->fetchMinIdle;
    P.triVoteOccl32 = (uint32_t)c->triVoteOccl32; P.fetchMinIdleOccl = (uint32_t)c->fetchMinIdleOccl;
    uint32_t nB = pc.params.bounceLimit + 1;
    uint32_t* qCount = (uint32_t*)w->dCounts.p;
    uint32_t* sCount = qCount + w->countSlots;
    uint32_t* workE = sCount + w->countSlots;
    uint32_t* workO = workE + w->countSlots;
    uint32_t* oCount = workO + w->countSlots;
    uint32_t* primaryCount = oCount + w->countSlots;
    uint32_t* mCount = primaryCount + w->countSlots;
    uint32_t* xCount = mCount + w->countSlots;          // contested extension rays per bounce
    uint32_t* xnCount = xCount + w->countSlots;         // contested light-sample queries per bounce
    uint32_t* wCount = xnCount + w->countSlots;         // hits that need the reachability walk, per bounce
    uint32_t* walkQ = (uint32_t*)w->dWalkQ.p;
    unsigned long long* exactTotals = (unsigned long long*)w->dRayCounters.p + 12;      // [12] extension rays, [13] light-sample queries re-traced exactly
    const int gridExact = w->smCount * 7;      // 72 registers: 7 resident blocks per SM; the re-trace is bound by the latency of its dependent node loads (4 blocks: 4.9 -> ? ms per step)
    if (!w->gridOcclude) {
        int perSm = 0;
        WCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_occlude<false, true>, WF_THREADS, 0));
        w->gridOcclude = std::max(1, perSm) * w->smCount;
    }
    if (!w->gridExtend) {
        int perSm = 0;
        WCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_extend<false, true>, WF_THREADS, 0));
        w->gridExtend = std::max(1, perSm) * w->smCount;
        WCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_verify_extend, WF_THREADS, 0));
        w->gridVerify = std::max(1, perSm) * w->smCount;
    }
    for (uint32_t done = 0; done < n; done += slots) {
        P.nSlots = std::min(slots, n - done);
        P.firstFrame = first + done * stride;
        P.nPaths = (uint32_t)(perSlot * P.nSlots);
        if (samples > 1) WCU(wf_resize(w->dSampleSum, (size_t)P.nPaths * 16));
        for (uint32_t smp = 0; smp < samples; ++smp) {            // samples of one dispatch run one after the other: they share the RNG stream
            P.continueSample = smp ? 1u : 0u;
            WCU(cudaMemsetAsync(w->dCounts.p, 0, (size_t)w->countSlots * WF_COUNT_ARRAYS * 4, c->stream));
            uint32_t* qa = (uint32_t*)w->dQueue[0].p;
            uint32_t* qb = (uint32_t*)w->dQueue[1].p;
            if (w->eventsUsed > 4096) { WCU(cudaStreamSynchronize(c->stream)); wavefront_resolve_stage_times(c); }
            wf_mark(c, ST_GENERATE);
            // the primary ray of a pixel is the same in every frame (no jitter: raytrace.comp:541, A-Q4): with several
            // frame slots in a batch it is traced once (slot 0) and the hit is shared; shade reads hit[p % perSlot] at j = 0
            P.sharePrimary = (c->sharePrimary && P.nSlots > 1) ? 1u : 0u;
            k_generate<<<grid, WF_THREADS, 0, c->stream>>>(P, w->arrays, qa, qCount, primaryCount);
            // Order per bounce: shade(j), extend(j+1), connect(j). Connect comes AFTER the next extension so that NEE rays of
            // paths whose next ray escapes (their direct light is dropped, raytrace.comp:531-534) are not traced at all.
            // With FEW frames in the batch (<= 8; measured gain 8 % at one frame, 0.5 % at eight, loss 0.7 % at 24) the launches are small and the drain phase of every persistent launch (its last
            // rays run in partly filled warps) is a large share of it: the connect chain of bounce j then runs on a second
            // stream beside extend(j+1) — one's drain overlaps the other's work — and gives up dropping the NEE rays of
            // escaping paths, which needs that extension's result (same image: trace() discards their direct light).
            const bool overlap = !c->profileStages && (c->overlapConnect == 2 || (c->overlapConnect == 1 && P.nSlots <= 8));
            if (overlap && !w->sideStream) {
                WCU(cudaStreamCreateWithFlags(&w->sideStream, cudaStreamNonBlocking));
                WCU(cudaEventCreateWithFlags(&w->evFork, cudaEventDisableTiming));
                WCU(cudaEventCreateWithFlags(&w->evJoin, cudaEventDisableTiming));
            }
            P.connectAfterExtend = overlap ? 0u : 1u;
            // k_exact_extend of bounce j+1 (a few rays per thousand, one thread each: latency bound) is only needed by
            // shade(j+1): it runs on its own stream beside the connect chain of bounce j. (prep reads hit.id of those paths
            // meanwhile, only to see whether the ray escaped: a contested ray has candidates, so both values say "hit".)
            const bool exactAside = !c->profileStages && c->exactOrder;
            if (exactAside && !w->exactStream) {
                WCU(cudaStreamCreateWithFlags(&w->exactStream, cudaStreamNonBlocking));
                WCU(cudaEventCreateWithFlags(&w->evX0, cudaEventDisableTiming));
                WCU(cudaEventCreateWithFlags(&w->evX1, cudaEventDisableTiming));
            }
            wf_mark(c, ST_EXTEND);
            unsigned long long* workCnt = (unsigned long long*)w->dRayCounters.p + 8;     // [8,9] extend, [10,11] occlude
            uint32_t* exactQ = (uint32_t*)w->dExactQ.p;
            uint32_t* exactNeeQ = (uint32_t*)w->dExactNeeQ.p;
            launch_extend(c, w->gridExtend, P, w->scene, w->arrays, qa, P.sharePrimary ? primaryCount : qCount, workE, exactQ, xCount, workCnt);
            wf_mark(c, ST_EXACT);
            if (c->exactOrder) {
                launch_verify(c, w->gridVerify, c->stream, w->scene, w->arrays, qa, P.sharePrimary ? primaryCount : qCount, exactQ, xCount, walkQ, wCount);
                k_exact_extend<<<gridExact, WF_THREADS, 0, c->stream>>>(P, w->scene, c->ref, w->arrays, exactQ, xCount, w->arrays.hit, exactTotals);
            }
            for (uint32_t j = 0; j < nB; ++j) {
                wf_mark(c, ST_SHADE);
                k_shade<<<gridShade, SHADE_THREADS, 0, c->stream>>>(P, w->scene, w->arrays, qa, qCount + j, qb, qCount + j + 1,
                                                            (uint32_t*)w->dShadowQ.p, sCount + j, j);
                if (j + 1 < nB) {
                    cudaStream_t cs = c->stream;
                    if (overlap) {
                        WCU(cudaEventRecord(w->evFork, c->stream));
                        WCU(cudaStreamWaitEvent(w->sideStream, w->evFork, 0));
                        cs = w->sideStream;
                    }
                    wf_mark(c, ST_EXTEND);
                    launch_extend(c, w->gridExtend, P, w->scene, w->arrays, qb, qCount + j + 1, workE + j + 1, exactQ, xCount + j + 1, workCnt);
                    wf_mark(c, ST_EXACT);
                    if (exactAside) {
                        WCU(cudaEventRecord(w->evX0, c->stream));
                        WCU(cudaStreamWaitEvent(w->exactStream, w->evX0, 0));
                    }
                    if (c->exactOrder) {
                        cudaStream_t xs = exactAside ? w->exactStream : c->stream;
                        launch_verify(c, w->gridVerify, xs, w->scene, w->arrays, qb, qCount + j + 1, exactQ, xCount + j + 1, walkQ, wCount + j + 1);
                        k_exact_extend<<<gridExact, WF_THREADS, 0, xs>>>(P, w->scene, c->ref, w->arrays, exactQ, xCount + j + 1, w->arrays.hit, exactTotals);
                    }
                    if (exactAside) WCU(cudaEventRecord(w->evX1, w->exactStream));
                    wf_mark(c, ST_CONNECT);
                    k_connect_prep<<<w->gridPrep, WF_THREADS, 0, cs>>>(P, w->scene, w->arrays, (const uint32_t*)w->dShadowQ.p, sCount + j,
                                                                       (uint32_t*)w->dJobQ.p, oCount + j, exactNeeQ, xnCount + j);
                    wf_mark(c, ST_OCCLUDE);
                    uint32_t* misQ = (uint32_t*)w->dJobQ.p + w->capacity;      // second half of the job buffer
                    if (c->exactOrder) {
                        if (c->countWork) k_occlude<true, true><<<w->gridOcclude, WF_THREADS, 0, cs>>>(P, w->scene, w->arrays, (const uint32_t*)w->dJobQ.p, oCount + j, workO + j, exactNeeQ, xnCount + j, workCnt + 2);
                        else k_occlude<false, true><<<w->gridOcclude, WF_THREADS, 0, cs>>>(P, w->scene, w->arrays, (const uint32_t*)w->dJobQ.p, oCount + j, workO + j, exactNeeQ, xnCount + j, nullptr);
                    } else {
                        if (c->countWork) k_occlude<true, false><<<w->gridOcclude, WF_THREADS, 0, cs>>>(P, w->scene, w->arrays, (const uint32_t*)w->dJobQ.p, oCount + j, workO + j, exactNeeQ, xnCount + j, workCnt + 2);
                        else k_occlude<false, false><<<w->gridOcclude, WF_THREADS, 0, cs>>>(P, w->scene, w->arrays, (const uint32_t*)w->dJobQ.p, oCount + j, workO + j, exactNeeQ, xnCount + j, nullptr);
                    }
                    wf_mark(c, ST_EXACT);
                    if (c->exactOrder) {
                        k_verify_occluded<<<w->gridVerify, WF_THREADS, 0, cs>>>(w->scene, w->arrays, (const uint32_t*)w->dJobQ.p, oCount + j, exactNeeQ, xnCount + j);
                        k_exact_nee<<<gridExact, WF_THREADS, 0, cs>>>(P, w->scene, c->ref, w->arrays, exactNeeQ, xnCount + j, exactTotals + 1);
                        c->counters[0] += 1;
                    }
                    wf_mark(c, ST_FINISH);
                    k_connect_finish<<<w->gridFinish, WF_THREADS, 0, cs>>>(w->scene, w->arrays, (const uint32_t*)w->dShadowQ.p, sCount + j, misQ, mCount + j);
                    k_connect_mis<<<w->smCount, WF_THREADS, 0, cs>>>(P, c->ref, w->arrays, misQ, mCount + j);
                    if (overlap) {
                        WCU(cudaEventRecord(w->evJoin, w->sideStream));
                        WCU(cudaStreamWaitEvent(c->stream, w->evJoin, 0));
                    }
                    if (exactAside) WCU(cudaStreamWaitEvent(c->stream, w->evX1, 0));
                }
                std::swap(qa, qb);
                c->counters[0] += (j + 1 < nB) ? 8 : 1;
            }
            k_fold<<<1, 32, 0, c->stream>>>(qCount, sCount, oCount, mCount, primaryCount, P.sharePrimary, nB, (unsigned long long*)w->dRayCounters.p);
            c->counters[0] += 4;                                      // generate, extend(0) + its exact re-trace, fold (the verify pair counts itself)
            if (samples > 1) {
                k_sample_sum<<<grid, WF_THREADS, 0, c->stream>>>(w->arrays, (float4*)w->dSampleSum.p, P.nPaths, smp, samples);
                c->counters[0] += 1;
            }
        }
        wf_mark(c, ST_ACCUMULATE);
        if (dispatchImage) k_dispatch_resolve<<<grid, WF_THREADS, 0, c->stream>>>(P, w->arrays, dispatchImage, c->storageRgba8);
        else k_accumulate<<<grid, WF_THREADS, 0, c->stream>>>(P, w->arrays, c->accum);
        wf_mark(c, -1);
        c->counters[0] += 1;
        c->counters[1] += (uint64_t)P.nPaths * samples;
        WCU(cudaGetLastError());
    }
    return 0;
}

int wavefront_trace_batch(b200pt_ctx* c, const ptl::PushConstants& pc, const float* dRays, uint64_t n, HitRecord48* dHits) {
    WavefrontState* w = c->wf;
    if (!w || !w->haveScene) { c->error = "wavefront: no scene (the last upload or update failed)"; return 3; }
    if (n == 0) return 0;
    const uint64_t chunkMax = (uint64_t)1 << 25;
    if (int rc = wf_reserve(c, (size_t)std::min<uint64_t>(n, chunkMax), pc.params.bounceLimit)) return rc;
    if (!w->gridExtend) {
        int perSm = 0;
        WCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_extend<false, true>, WF_THREADS, 0));
        w->gridExtend = std::max(1, perSm) * w->smCount;
        WCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_verify_extend, WF_THREADS, 0));
        w->gridVerify = std::max(1, perSm) * w->smCount;
    }
    WfParams P;
    memset(&P, 0, sizeof(P));
    P.pc = pc; P.W = c->W; P.H = c->H; P.rowStride = 1;
    P.triVote32 = (uint32_t)c->triVote32; P.fetchMinIdle = (uint32_t)c->fetchMinIdle;
    uint32_t* qCount = (uint32_t*)w->dCounts.p;
    uint32_t* workE = qCount + 2 * w->countSlots;
    unsigned long long* workCnt = (unsigned long long*)w->dRayCounters.p + 8;
    const int grid = w->smCount * 8;
    for (uint64_t done = 0; done < n; done += chunkMax) {
        const uint32_t m = (uint32_t)std::min<uint64_t>(chunkMax, n - done);
        WCU(cudaMemsetAsync(w->dCounts.p, 0, (size_t)w->countSlots * WF_COUNT_ARRAYS * 4, c->stream));
        k_load_rays<<<grid, WF_THREADS, 0, c->stream>>>((const float4*)dRays + 2 * done, m, w->arrays, (uint32_t*)w->dQueue[0].p, qCount);
        wf_mark(c, ST_EXTEND);
        uint32_t* xCount = qCount + 7 * w->countSlots;
        launch_extend(c, w->gridExtend, P, w->scene, w->arrays, (const uint32_t*)w->dQueue[0].p, qCount, workE, (uint32_t*)w->dExactQ.p, xCount, workCnt);
        wf_mark(c, -1);
        if (c->exactOrder) launch_verify(c, w->gridVerify, c->stream, w->scene, w->arrays, (const uint32_t*)w->dQueue[0].p, qCount, (uint32_t*)w->dExactQ.p, xCount,
                                         (uint32_t*)w->dWalkQ.p, qCount + 9 * w->countSlots);
        k_exact_extend<<<w->smCount * 4, WF_THREADS, 0, c->stream>>>(P, w->scene, c->ref, w->arrays, (const uint32_t*)w->dExactQ.p, xCount, w->arrays.hit,
                                                                   (unsigned long long*)w->dRayCounters.p + 12);
        k_hit_records<<<grid, WF_THREADS, 0, c->stream>>>(w->scene, w->arrays, m, dHits + done);
        c->counters[0] += 4;
    }
    c->counters[2] += n;
    c->counters[4] += n;
    WCU(cudaGetLastError());
    return 0;
}

// ---- b200pt_trace_batch_device_static: the same launches with NO allocation ------------------------------------
// wavefront_trace_batch keeps its scratch (ray slots, hits, queues, counters) in buffers that wf_reserve grows with
// cudaMalloc. Callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid it) get a variant
// whose scratch is this file-scope device array: a fixed size, held by every GPU that loads the library until exit.
// Layout: 256-byte header {qCount, k_extend's work counter, contested count, walk count, u64 re-trace total}, then per-ray
// arrays origin (16 B) | direction (16 B) | Hit (16 B) | queue (4 B) | contested queue (4 B) | walk queue (4 B).
// The NAME below is a host-side stand-in in host code, not a device address: its device address is looked up once per
// device (cudaGetSymbolAddress with that device current) and only that pointer is handed to kernels.
// Every call shares the array and finds the previous call's contents in it: the header is cleared at the start of each
// call, the per-ray arrays are written before they are read. Calls that overlap on one GPU (two streams, two contexts,
// two threads) race on it.
#define WF_STATIC_HEADER 256
#define WF_STATIC_PER_RAY 60
static __device__ __align__(256) unsigned char g_staticScratch[WF_STATIC_HEADER + (size_t)B200PT_STATIC_MAX_RAYS * WF_STATIC_PER_RAY];
static_assert(sizeof(Hit) == 16, "scratch layout");
static unsigned char* g_staticScratchAddr[64] = {nullptr};      // per device ordinal
static std::mutex g_staticScratchMutex;

int wavefront_trace_batch_static(b200pt_ctx* c, const ptl::PushConstants& pc, const float* dRays, uint64_t n, HitRecord48* dHits) {
    WavefrontState* w = c->wf;
    if (!w || !w->haveScene) { c->error = "wavefront: no scene (the last upload or update failed)"; return 3; }
    if (n > (uint64_t)B200PT_STATIC_MAX_RAYS) {
        char msg[160];
        snprintf(msg, sizeof(msg), "trace_batch_device_static: %llu rays exceed the static scratch (B200PT_STATIC_MAX_RAYS = %u); split the batch or use b200pt_trace_batch_device",
                 (unsigned long long)n, (unsigned)B200PT_STATIC_MAX_RAYS);
        c->error = msg;
        return 1;
    }
    if (n == 0) return 0;
    if (c->device < 0 || c->device >= 64) { c->error = "trace_batch_device_static: device ordinal above 63"; return 1; }
    unsigned char* base;
    {
        std::lock_guard<std::mutex> lock(g_staticScratchMutex);
        if (!g_staticScratchAddr[c->device]) {                   // once per device; the caller made c->device current
            void* q = nullptr;
            cudaError_t e = cudaGetSymbolAddress(&q, g_staticScratch);
            if (e != cudaSuccess || !q) return wf_fail(c, 2, "cudaGetSymbolAddress(g_staticScratch)", e);
            g_staticScratchAddr[c->device] = (unsigned char*)q;
        }
        base = g_staticScratchAddr[c->device];
    }
    if (!w->gridExtend) {
        int perSm = 0;
        WCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_extend<false, true>, WF_THREADS, 0));
        w->gridExtend = std::max(1, perSm) * w->smCount;
        WCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_verify_extend, WF_THREADS, 0));
        w->gridVerify = std::max(1, perSm) * w->smCount;
    }
    const size_t cap = B200PT_STATIC_MAX_RAYS;
    uint32_t* header = (uint32_t*)base;
    PathArrays A;
    memset(&A, 0, sizeof(A));
    A.rayO = (float4*)(base + WF_STATIC_HEADER);
    A.rayD = A.rayO + cap;
    A.hit = (Hit*)(A.rayD + cap);
    uint32_t* queue = (uint32_t*)(A.hit + cap);
    uint32_t* exactQ = queue + cap;
    uint32_t* walkQ = exactQ + cap;
    uint32_t* qCount = header, *workE = header + 1, *xCount = header + 2;
    unsigned long long* total = (unsigned long long*)(header + 4);
    WfParams P;
    memset(&P, 0, sizeof(P));
    P.pc = pc; P.W = c->W; P.H = c->H; P.rowStride = 1;
    P.triVote32 = (uint32_t)c->triVote32; P.fetchMinIdle = (uint32_t)c->fetchMinIdle;
    const uint32_t m = (uint32_t)n;
    const int grid = std::min<int>(w->smCount * 8, (int)((m + WF_THREADS - 1) / WF_THREADS));
    WCU(cudaMemsetAsync(header, 0, WF_STATIC_HEADER, c->stream));
    k_load_rays<<<grid, WF_THREADS, 0, c->stream>>>((const float4*)dRays, m, A, queue, qCount);
    launch_extend(c, w->gridExtend, P, w->scene, A, queue, qCount, workE, exactQ, xCount, nullptr);      // no work counters here
    if (c->exactOrder) {
        launch_verify(c, w->gridVerify, c->stream, w->scene, A, queue, qCount, exactQ, xCount, walkQ, header + 3);
        k_exact_extend<<<w->smCount * 4, WF_THREADS, 0, c->stream>>>(P, w->scene, c->ref, A, exactQ, xCount, A.hit, total);
    }
    k_hit_records<<<grid, WF_THREADS, 0, c->stream>>>(w->scene, A, m, dHits);
    c->counters[0] += 4;
    WCU(cudaGetLastError());
    return 0;
}

// called with the stream idle: moves the device ray counters into the context's counters
void wavefront_fold_counters(b200pt_ctx* c) {
    WavefrontState* w = c->wf;
    if (!w || !w->dRayCounters.p) return;
    unsigned long long h[WF_RAY_COUNTERS] = {0};
    if (cudaMemcpy(h, w->dRayCounters.p, sizeof(h), cudaMemcpyDeviceToHost) != cudaSuccess) return;
    for (int i = 0; i < 5; ++i) c->counters[2 + i] += h[i];           // [2..6]: see k_fold
    for (int i = 0; i < 4; ++i) c->counters[7 + i] += h[8 + i];       // [7..10]: node visits / triangle tests of k_extend, k_occlude
    for (int i = 0; i < 2; ++i) c->counters[11 + i] += h[12 + i];     // [11, 12]: extension rays / light-sample queries re-traced in reference order
    cudaMemset(w->dRayCounters.p, 0, sizeof(h));
}

cudaError_t launch_image_to_rgba8(b200pt_ctx* c, const float4* image, uchar4* out, size_t n) {
    k_image_to_rgba8<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(image, out, n);
    return cudaGetLastError();
}

}  // namespace pt
