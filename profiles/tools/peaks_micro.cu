// peaks_micro — measured ceilings for the traversal roofline: the denominators the
// bench's `roofline` block uses besides the driver's HBM copy figure. Built by profiles/tools/build_peaks_micro.sh,
// run on the B200 under gpurun; prints one JSON object (-> profiles/peaks_micro.json).
//
//   hbm_read_gbs           stream read of a 4 GiB buffer (16-byte loads)
//   l2_read_gbs[MB]        the same loop over a buffer that stays in L2 (21 / 33 / 64 MB)
//   l2_gather_gbs[...]     every lane reads one RANDOM record (80-byte wide-BVH node = 5 x 16 B, 64-byte triangle
//                          record = 4 x 16 B) of an L2-resident array through the read-only path (__ldg, as the
//                          traversal kernels do): the incoherent-access ceiling, in GB/s and G records/s
//   ffma_tflops            FP32 FMA rate, 8 independent chains per thread
//   issue_ginst            warp instructions per second with FMA- and ALU-pipe instructions interleaved (the
//                          schedulers' issue ceiling: 4 per SM per clock)
//   alu_ginst / fmnmx / prmt / i2f / imad   single-pipe rates (warp instructions per second)
//   ffma2_ginst / issue_ffma2_lop3_ginst   packed FP32 FMA (FFMA2, sm_100): its own rate, and interleaved with LOP3
// Every figure is the best of 5 timed repetitions after a warm-up, CUDA events on the launching stream.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

template <class F>
static float best_ms(F&& launch, int reps = 5) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    launch();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        CK(cudaEventRecord(a));
        launch();
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        float ms; CK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    cudaEventDestroy(a); cudaEventDestroy(b);
    return best;
}

// ---- streaming reads -------------------------------------------------------------------------------------
// L2 = true: ld.global.cg (L2 only). Without it the 21 / 33 MB loops are served by the L1s: every SM re-reads the same
// 1/148th of the buffer (223 KB at 33 MB, the size of its L1) and the figure comes out above the L2's own ceiling.
template <bool L2>
__global__ void k_stream_read(const uint4* __restrict__ p, size_t n16, int passes, uint32_t* sink) {
    uint32_t acc = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int pass = 0; pass < passes; ++pass)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
            uint4 v = L2 ? __ldcg(p + i) : p[i];
            acc += v.x ^ v.y ^ v.z ^ v.w;
        }
    if (acc == 0x12345678u) *sink = acc;
}

// ---- random record gathers (one record per lane per round) ---------------------------------------------
template <int VEC>   // VEC 16-byte loads per record
__global__ void k_gather(const uint4* __restrict__ p, uint32_t nRecords, int rounds, uint32_t* sink) {
    uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    uint32_t acc = 0;
    for (int r = 0; r < rounds; ++r) {
        s = s * 747796405u + 2891336453u;
        uint32_t w = ((s >> ((s >> 28) + 4)) ^ s) * 277803737u;
        uint32_t idx = (uint32_t)(((uint64_t)(w ^ (w >> 22)) * nRecords) >> 32);
        const uint4* q = p + (size_t)idx * VEC;
#pragma unroll
        for (int k = 0; k < VEC; ++k) { uint4 v = __ldg(q + k); acc += v.x ^ v.y ^ v.z ^ v.w; }
    }
    if (acc == 0x12345678u) *sink = acc;
}

// ---- instruction-rate kernels: CH independent chains, UNROLL instructions per chain per iteration ---------------
#define CH 8
#define UNROLL 32
__global__ void k_ffma(float* out, int iters, float x, float y) {
    float a[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) a[c] = threadIdx.x * 1e-3f + c;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < UNROLL; ++u)
#pragma unroll
            for (int c = 0; c < CH; ++c) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[c]) : "f"(x), "f"(y));
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < CH; ++c) s += a[c];
    if (s == 1.2345f) *out = s;
}
// OP: 0 lop3 (ALU), 1 max.f32 (FMNMX, ALU), 2 prmt (ALU), 3 cvt.rn.f32.u32 of a byte (I2F, XU), 4 mad.lo.u32 (IMAD, FMA pipe),
//     5 FFMA + LOP3 interleaved (two pipes: issue ceiling)
// (ptxas fuses two chained max.f32 into one 3-input FMNMX3 and folds chained integer adds; counts below follow the SASS)
template <int OP>
__global__ void k_rate(uint32_t* out, int iters, uint32_t x, uint32_t y) {
    uint32_t a[CH];
    float f[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) { a[c] = threadIdx.x + c * 977u; f[c] = threadIdx.x * 1e-3f + c; }
    const float fx = __uint_as_float(x | 0x3f800000u), fy = __uint_as_float(y | 0x3f000000u);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < UNROLL; ++u)
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                if (OP == 0) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[c]) : "r"(x), "r"(y));
                if (OP == 1) asm volatile("max.f32 %0, %0, %1;" : "+f"(f[c]) : "f"(fx));
                if (OP == 2) asm volatile("prmt.b32 %0, %0, %1, 0x7604;" : "+r"(a[c]) : "r"(x));
                if (OP == 3) asm volatile("{ .reg .u32 b; .reg .f32 t; xor.b32 %0, %0, %2; and.b32 b, %0, 0xff; cvt.rn.f32.u32 t, b; add.f32 %1, %1, t; }"
                                          : "+r"(a[c]), "+f"(f[c]) : "r"(x));       // LOP3 + I2F (byte, XU pipe) + FADD: bound by the XU pipe
                if (OP == 4) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[c]) : "r"(x), "r"(y));
                if (OP == 5) {
                    if (c & 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[c]) : "r"(x), "r"(y));
                    else asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[c]) : "f"(fx), "f"(fy));
                }
            }
    }
    uint32_t s = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c) s += a[c] + __float_as_uint(f[c]);
    if (s == 0x12345678u) *out = s;
}

// packed FP32 FMA of sm_100 (fma.rn.f32x2 -> FFMA2: two FMAs per issued instruction). PAIRED = 0: 8 chains of FFMA2;
// PAIRED = 1: FFMA2 and LOP3 interleaved (does an FFMA2 take one issue slot?)
template <int WITH_LOP3>
__global__ void k_ffma2(uint32_t* out, int iters, float x, float y, uint32_t ux, uint32_t uy) {
    unsigned long long a[CH];
    uint32_t b[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) {
        float lo = threadIdx.x * 1e-3f + c, hi = lo + 0.5f;
        asm("mov.b64 %0, {%1, %2};" : "=l"(a[c]) : "f"(lo), "f"(hi));
        b[c] = threadIdx.x + c * 977u;
    }
    unsigned long long xs, ys;
    asm("mov.b64 %0, {%1, %1};" : "=l"(xs) : "f"(x));
    asm("mov.b64 %0, {%1, %1};" : "=l"(ys) : "f"(y));
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < UNROLL; ++u)
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                if (!WITH_LOP3 || !(c & 1)) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a[c]) : "l"(xs), "l"(ys));
                else asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[c]) : "r"(ux), "r"(uy));
            }
    }
    uint32_t s = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c) s += (uint32_t)a[c] + (uint32_t)(a[c] >> 32) + b[c];
    if (s == 0x12345678u) *out = s;
}

int main(int argc, char** argv) {
    int dev = 0;
    CK(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, dev));
    const int sms = prop.multiProcessorCount;
    int clockKHz = 0;
    cudaDeviceGetAttribute(&clockKHz, cudaDevAttrClockRate, dev);
    uint32_t* sink; CK(cudaMalloc(&sink, 64));
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"sm_clock_mhz_attr\": %.0f, \"l2_bytes\": %d", prop.name, sms, clockKHz / 1e3, prop.l2CacheSize);

    // ---- HBM and L2 streaming reads ----
    {
        const size_t big = (size_t)4 << 30;
        uint4* buf; CK(cudaMalloc(&buf, big));
        CK(cudaMemset(buf, 1, big));
        const int grid = sms * 8, block = 256;
        float ms = best_ms([&] { k_stream_read<false><<<grid, block>>>(buf, big / 16, 1, sink); });
        printf(", \"hbm_read_gbs\": %.1f", big / (ms * 1e-3) / 1e9);
        printf(", \"l2_read_gbs\": {");
        const int mbs[] = {21, 33, 64};
        for (int k = 0; k < 3; ++k) {
            size_t bytes = (size_t)mbs[k] << 20;
            int passes = 400;
            float m = best_ms([&] { k_stream_read<true><<<grid, block>>>(buf, bytes / 16, passes, sink); });
            printf("%s\"%d\": %.1f", k ? ", " : "", mbs[k], (double)bytes * passes / (m * 1e-3) / 1e9);
        }
        printf("}");
        // ---- random gathers ----
        printf(", \"l2_gather\": {");
        struct G { const char* name; int vec; int mb; } gs[] = {{"node80_21MB", 5, 21}, {"node80_4MB", 5, 4}, {"prim64_17MB", 4, 17}, {"prim64_33MB", 4, 33},
                                                                {"node80_2GB", 5, 2048}, {"prim64_2GB", 4, 2048}};
        bool first = true;
        for (const G& g : gs) {
            uint32_t nRec = (uint32_t)(((size_t)g.mb << 20) / (16 * g.vec));
            const int rounds = 256, gthreads = 128, ggrid = sms * 16;
            float m;
            if (g.vec == 5) m = best_ms([&] { k_gather<5><<<ggrid, gthreads>>>(buf, nRec, rounds, sink); });
            else m = best_ms([&] { k_gather<4><<<ggrid, gthreads>>>(buf, nRec, rounds, sink); });
            double recs = (double)ggrid * gthreads * rounds;
            printf("%s\"%s\": {\"gbs\": %.1f, \"grecords_s\": %.2f}", first ? "" : ", ", g.name, recs * 16 * g.vec / (m * 1e-3) / 1e9, recs / (m * 1e-3) / 1e9);
            first = false;
        }
        printf("}");
        cudaFree(buf);
    }
    // ---- instruction rates ----
    {
        const int block = 256, grid = sms * 8, iters = 400;
        const double instPerThread = (double)iters * UNROLL * CH;
        const double warpInst = instPerThread * grid * block / 32.0;
        float* fout; CK(cudaMalloc(&fout, 64));
        float ms = best_ms([&] { k_ffma<<<grid, block>>>(fout, iters, 1.0001f, 0.5f); });
        printf(", \"ffma_tflops\": %.2f, \"ffma_ginst\": %.1f", warpInst * 32 * 2 / (ms * 1e-3) / 1e12, warpInst / (ms * 1e-3) / 1e9);
        const char* names[] = {"lop3_ginst", "fmnmx3_ginst", "prmt_ginst", "i2f_u8_ginst", "imad_ginst", "issue_ffma_lop3_ginst"};
        float t[6];
        t[0] = best_ms([&] { k_rate<0><<<grid, block>>>(sink, iters, 0x55u, 0x33u); });
        t[1] = best_ms([&] { k_rate<1><<<grid, block>>>(sink, iters, 0x55u, 0x33u); });
        t[2] = best_ms([&] { k_rate<2><<<grid, block>>>(sink, iters, 0x55u, 0x33u); });
        t[3] = best_ms([&] { k_rate<3><<<grid, block>>>(sink, iters, 0x55u, 0x33u); });
        t[4] = best_ms([&] { k_rate<4><<<grid, block>>>(sink, iters, 0x55u, 0x33u); });
        t[5] = best_ms([&] { k_rate<5><<<grid, block>>>(sink, iters, 0x55u, 0x33u); });
        for (int k = 0; k < 6; ++k) {
            double n = warpInst * (k == 1 ? 0.5 : 1.0);     // two max.f32 = one FMNMX3; the I2F figure counts conversions only
            printf(", \"%s\": %.1f", names[k], n / (t[k] * 1e-3) / 1e9);
        }
        float t2 = best_ms([&] { k_ffma2<0><<<grid, block>>>(sink, iters, 1.0001f, 0.5f, 0x55u, 0x33u); });
        float t3 = best_ms([&] { k_ffma2<1><<<grid, block>>>(sink, iters, 1.0001f, 0.5f, 0x55u, 0x33u); });
        printf(", \"ffma2_ginst\": %.1f, \"ffma2_tflops\": %.2f, \"issue_ffma2_lop3_ginst\": %.1f", warpInst / (t2 * 1e-3) / 1e9,
               warpInst * 32 * 4 / (t2 * 1e-3) / 1e12, warpInst / (t3 * 1e-3) / 1e9);
        printf(", \"issue_nominal_ginst\": %.1f", sms * 4.0 * clockKHz / 1e6);
        cudaFree(fout);
    }
    printf("}\n");
    return 0;
}
