// Multi-GPU entries of the C ABI (include/b200pt.h, "multi-GPU"): one context per GPU, the contexts of a job joined by
// an NCCL communicator. The reference has a single queue on a single GPU (src/vk_engine.cpp:115-117); what is split
// here is its one dispatch point (run_compute, :1623-1676):
//   * sample split (throughput): rank r renders the progressive frames f = first + r, first + r + N, ... — frames have
//     independent seeds (shaders/raytrace.comp:562-564) — and one ncclReduce(sum, f32) puts the W*H*4 accumulators on the
//     root;
//   * interleaved rows (one 1-spp frame, latency): rank r renders rows r, r + N, ... — a pixel's seed depends only on
//     (x, y, W, frame), so any subset of pixels is exact, and interleaving balances rows of different cost — the root
//     receives every rank's rows with grouped ncclSend / ncclRecv and scatters them into its accumulator.
// NCCL is bound at run time (dlopen "libnccl.so.2"): the library neither links against it nor needs it on one GPU;
// with torch in the process the already loaded library of that name is the one that answers.
#include <dlfcn.h>

#include <cstdio>
#include <cstring>
#include <mutex>

#include "../../include/b200pt.h"
#include "pt_context.h"
#include "wavefront.h"

namespace {

// the part of nccl.h (2.x, stable since 2.2) that is used; declared here so that the build needs no NCCL headers
typedef struct ncclComm* ncclComm_t;
struct ncclUniqueId { char internal[128]; };
enum { ncclSuccess = 0 };
enum { ncclFloat32 = 7 };
enum { ncclSum = 0 };
struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Reduce)(const void*, void*, size_t, int, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    std::string error;
};
NcclApi g_nccl;
std::once_flag g_ncclOnce;

const NcclApi* nccl_api() {
    std::call_once(g_ncclOnce, [] {
        NcclApi& a = g_nccl;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) { a.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (a.handle) break; }
        if (!a.handle) { a.error = std::string("cannot load libnccl.so.2: ") + dlerror(); return; }
#define SYM(field, name) a.field = (decltype(a.field))dlsym(a.handle, name); if (!a.field) { a.error = "libnccl lacks " name; return; }
        SYM(GetUniqueId, "ncclGetUniqueId") SYM(CommInitRank, "ncclCommInitRank") SYM(CommDestroy, "ncclCommDestroy")
        SYM(Reduce, "ncclReduce") SYM(Send, "ncclSend") SYM(Recv, "ncclRecv") SYM(GroupStart, "ncclGroupStart")
        SYM(GroupEnd, "ncclGroupEnd") SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
    });
    return g_nccl.error.empty() ? &g_nccl : nullptr;
}

int cfail(b200pt_ctx* c, int code, const std::string& what) { if (c) c->error = what; return code; }
#define NC(c, call)                                                                                             \
    do {                                                                                                        \
        int r_ = (call);                                                                                        \
        if (r_ != ncclSuccess) return cfail(c, B200PT_ERR_CUDA, std::string(#call ": ") + api->GetErrorString(r_)); \
    } while (0)
#define CC(c, call)                                                                                             \
    do {                                                                                                        \
        cudaError_t e_ = (call);                                                                                \
        if (e_ != cudaSuccess) return cfail(c, B200PT_ERR_CUDA, std::string(#call ": ") + cudaGetErrorString(e_)); \
    } while (0)

// rows r0, r0 + stride, ... of a W-wide float4 image <-> a compact buffer of those rows
__global__ void k_pack_rows(const float4* __restrict__ image, float4* __restrict__ packed, uint32_t W, uint32_t nRows, uint32_t r0, uint32_t stride) {
    const uint32_t n = nRows * W;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        packed[i] = image[(size_t)(r0 + (i / W) * stride) * W + i % W];
}
// every other rank's rows at once: packed = rank 0's rows (unused), rank 1's rows, ... each `slot` float4 apart
__global__ void k_unpack_rows(float4* __restrict__ image, const float4* __restrict__ packed, uint32_t W, uint32_t H, uint32_t nranks, uint32_t root,
                              size_t slot) {
    const uint32_t n = H * W;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t y = i / W, x = i % W, r = y % nranks;
        if (r == root) continue;
        image[i] = packed[(size_t)r * slot + (size_t)(y / nranks) * W + x];
    }
}

uint32_t rows_of(uint32_t H, uint32_t rank, uint32_t nranks) { return rank < H ? (H - rank + nranks - 1) / nranks : 0; }

}  // namespace

struct b200pt_comm {
    ncclComm_t comm = nullptr;
    int nranks = 1, rank = 0;
    pt::DeviceBuffer staging;     // packed rows: own rows on a sender, all ranks' rows on the root
};

extern "C" {

int b200pt_comm_unique_id(void* out128) {
    const NcclApi* api = nccl_api();
    if (!api) { pt::set_global_error(g_nccl.error.c_str()); return B200PT_ERR_STATE; }
    if (!out128) return B200PT_ERR_ARG;
    ncclUniqueId id;
    int r = api->GetUniqueId(&id);
    if (r != ncclSuccess) { pt::set_global_error(api->GetErrorString(r)); return B200PT_ERR_CUDA; }
    std::memcpy(out128, &id, sizeof(id));
    return 0;
}

int b200pt_comm_init_rank(b200pt_ctx* c, int nranks, int rank, const void* id128) {
    if (!c || !id128 || nranks < 1 || rank < 0 || rank >= nranks) return cfail(c, B200PT_ERR_ARG, "comm_init_rank: bad argument");
    if (c->comm) return cfail(c, B200PT_ERR_STATE, "comm_init_rank: the context already has a communicator");
    const NcclApi* api = nccl_api();
    if (!api) return cfail(c, B200PT_ERR_STATE, g_nccl.error);
    CC(c, cudaSetDevice(c->device));
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof(id));
    b200pt_comm* k = new b200pt_comm();
    k->nranks = nranks; k->rank = rank;
    int r = api->CommInitRank(&k->comm, nranks, id, rank);
    if (r != ncclSuccess) { delete k; return cfail(c, B200PT_ERR_CUDA, std::string("ncclCommInitRank: ") + api->GetErrorString(r)); }
    c->comm = k;
    return 0;
}

int b200pt_comm_destroy(b200pt_ctx* c) {
    if (!c) return B200PT_ERR_ARG;
    if (!c->comm) return 0;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    if (const NcclApi* api = nccl_api()) api->CommDestroy(c->comm->comm);
    if (c->comm->staging.p) cudaFree(c->comm->staging.p);
    delete c->comm;
    c->comm = nullptr;
    return 0;
}

int b200pt_comm_info(b200pt_ctx* c, int* nranks, int* rank) {
    if (!c) return B200PT_ERR_ARG;
    if (nranks) *nranks = c->comm ? c->comm->nranks : 1;
    if (rank) *rank = c->comm ? c->comm->rank : 0;
    return 0;
}

// frames of [first, first + n) that `rank` of `nranks` renders: first + rank, first + rank + nranks, ...
static uint32_t share_count(uint32_t n, uint32_t rank, uint32_t nranks) { return n > rank ? (n - rank + nranks - 1) / nranks : 0; }

int b200pt_render_frames_split(b200pt_ctx* c, const void* pc208, uint32_t first, uint32_t n) {
    if (!c) return B200PT_ERR_ARG;
    const uint32_t N = c->comm ? (uint32_t)c->comm->nranks : 1u, r = c->comm ? (uint32_t)c->comm->rank : 0u;
    const uint32_t mine = share_count(n, r, N);
    if (!mine) return 0;
    return b200pt_render_frames(c, pc208, first + r, mine, N);
}

int b200pt_reduce_accum(b200pt_ctx* c, int root, int zero_others) {
    if (!c) return B200PT_ERR_ARG;
    if (!c->comm) return 0;
    const NcclApi* api = nccl_api();
    if (!api) return cfail(c, B200PT_ERR_STATE, g_nccl.error);
    if (root < 0 || root >= c->comm->nranks) return cfail(c, B200PT_ERR_ARG, "reduce_accum: bad root");
    CC(c, cudaSetDevice(c->device));
    const size_t count = (size_t)c->W * c->H * 4;
    NC(c, api->Reduce(c->accum, c->accum, count, ncclFloat32, ncclSum, root, c->comm->comm, c->stream));
    if (zero_others && c->comm->rank != root) CC(c, cudaMemsetAsync(c->accum, 0, count * 4, c->stream));
    c->counters[0] += 1;
    return 0;
}

int b200pt_render_frame_interleaved(b200pt_ctx* c, const void* pc208, uint32_t frame, int root) {
    if (!c || !pc208) return cfail(c, B200PT_ERR_ARG, "render_frame_interleaved: NULL argument");
    if (!c->comm || c->comm->nranks == 1) return b200pt_render_frames(c, pc208, frame, 1, 1);
    const NcclApi* api = nccl_api();
    if (!api) return cfail(c, B200PT_ERR_STATE, g_nccl.error);
    b200pt_comm* k = c->comm;
    const uint32_t N = (uint32_t)k->nranks, r = (uint32_t)k->rank, W = (uint32_t)c->W, H = (uint32_t)c->H;
    if (root < 0 || root >= k->nranks) return cfail(c, B200PT_ERR_ARG, "render_frame_interleaved: bad root");
    if (c->pipeline != 1) return cfail(c, B200PT_ERR_STATE, "render_frame_interleaved: needs the wavefront pipeline");
    ptl::PushConstants pc;
    std::memcpy(&pc, pc208, sizeof(pc));
    if (!c->sceneLoaded) return cfail(c, B200PT_ERR_STATE, "render: no scene uploaded");
    if (pc.params.sphereCount > c->hSpheres.size() || pc.params.objectCount > c->hObjects.size())
        return cfail(c, B200PT_ERR_ARG, "render: sphereCount / objectCount above the uploaded counts");
    CC(c, cudaSetDevice(c->device));
    const uint32_t mine = rows_of(H, r, N);
    if (mine) if (int rc = pt::wavefront_render_frames(c, pc, frame, 1, 1, r, H, nullptr, N)) return rc;
    // the root's accumulator receives every other rank's rows (their accumulated values: a snapshot of the split image)
    const size_t slot = (size_t)rows_of(H, 0, N) * W;               // float4 per rank in the staging buffer (rank 0 has the most rows)
    const size_t need = ((int)r == root ? slot * N : slot) * sizeof(float4);
    if (k->staging.bytes < need) {
        if (k->staging.p) cudaFree(k->staging.p);
        k->staging = pt::DeviceBuffer{};
        CC(c, cudaMalloc(&k->staging.p, need));
        k->staging.bytes = need;
    }
    float4* st = (float4*)k->staging.p;
    const int grid = 148 * 4;
    if ((int)r != root) {
        if (mine) {
            k_pack_rows<<<grid, 256, 0, c->stream>>>(c->accum, st, W, mine, r, N);
            NC(c, api->Send(st, (size_t)mine * W * 4, ncclFloat32, root, k->comm, c->stream));
            c->counters[0] += 2;
        }
    } else {
        NC(c, api->GroupStart());
        for (uint32_t q = 0; q < N; ++q) {
            const uint32_t rq = rows_of(H, q, N);
            if ((int)q == root || !rq) continue;
            int rr = api->Recv(st + slot * q, (size_t)rq * W * 4, ncclFloat32, (int)q, k->comm, c->stream);
            if (rr != ncclSuccess) { api->GroupEnd(); return cfail(c, B200PT_ERR_CUDA, std::string("ncclRecv: ") + api->GetErrorString(rr)); }
        }
        NC(c, api->GroupEnd());
        k_unpack_rows<<<grid, 256, 0, c->stream>>>(c->accum, st, W, H, N, (uint32_t)root, slot);
        c->counters[0] += 2;
    }
    CC(c, cudaGetLastError());
    return 0;
}

}  // extern "C"
