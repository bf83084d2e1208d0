// tools/peaks.cu — hand-written microbenchmarks for the roofline denominators the traversal kernels are judged against
// (VERDICT r01 "missing" #2 / BASELINE.md §2): none of these is in MEASURED_PEAKS.json (which has HBM copy bandwidth and bf16 GEMM only).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/peaks tools/peaks.cu && /tmp/peaks > profiles/peaks_r02.json
//
// Measured, each with CUDA events after a warm-up, best of 5:
//   l2_ld128_gbs[footprint]  16-byte loads (LDG.128, the width of every node / slot fetch) by all SMs over a buffer that stays L2-resident
//                            (32 MB and 100 MB footprints; 126 MB L2) and one that does not (4 GB -> HBM), grid-stride, fully coalesced
//   l2_gather128_gbs         the same bytes fetched as INDEPENDENT random 128-byte node reads (one BVH4 node = 8 x 16 B; every lane its own node,
//                            as an incoherent traversal does): sector-efficiency-limited L2 bandwidth
//   fp32_ffma_tflops / warp_inst_per_s   dependent-chain-free FFMA issue rate, 8 independent accumulators per thread
//   fp64_dfma_tflops         the same with DFMA
//   l2_latency_cycles / l1_latency_cycles / hbm_latency_cycles   pointer chase, one thread, stride larger than a line, footprints 48 MB / 16 KB / 2 GB
//   issue_peak_warp_inst_per_s = SMs x 4 schedulers x SM clock (one warp instruction per scheduler per cycle) — stated next to the measured FFMA rate
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void __launch_bounds__(256) k_ld128(const uint4* __restrict__ p, size_t n16, int reps, uint4* sink) {
    uint4 acc = make_uint4(0, 0, 0, 0);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < reps; r++)
        for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += stride) {
            const uint4 v = __ldg(p + i);
            acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
        }
    if (acc.x == 0x12345678u && acc.y == 0x9abcdef0u) *sink = acc;     // never true in practice; keeps the loads alive
}
// every lane reads whole 128-byte records at pseudo-random record indices (LCG per thread), 8 x LDG.128 each — the access pattern of one BVH4 node step
__global__ void __launch_bounds__(256) k_gather128(const uint4* __restrict__ p, uint32_t n_rec_mask, int steps, uint4* sink) {
    uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    uint4 acc = make_uint4(0, 0, 0, 0);
    for (int k = 0; k < steps; k++) {
        s = s * 1664525u + 1013904223u;
        const uint4* q = p + (size_t)((s >> 4) & n_rec_mask) * 8;
#pragma unroll
        for (int w = 0; w < 8; w++) { const uint4 v = __ldg(q + w); acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w; }
    }
    if (acc.x == 0x12345678u && acc.y == 0x9abcdef0u) *sink = acc;
}
template <class T> __global__ void __launch_bounds__(256) k_fma(T a, T b, int iters, T* sink) {
    T x0 = a, x1 = a + T(1), x2 = a + T(2), x3 = a + T(3), x4 = a + T(4), x5 = a + T(5), x6 = a + T(6), x7 = a + T(7);
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
            x0 = fma(x0, b, a); x1 = fma(x1, b, a); x2 = fma(x2, b, a); x3 = fma(x3, b, a);
            x4 = fma(x4, b, a); x5 = fma(x5, b, a); x6 = fma(x6, b, a); x7 = fma(x7, b, a);
        }
    }
    const T s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
    if (s == T(123456789)) *sink = s;
}
__global__ void k_chase(const uint32_t* __restrict__ next, uint32_t start, int steps, uint32_t* out, long long* cycles) {
    uint32_t i = start;
    for (int k = 0; k < 256; k++) i = next[i];                      // warm the TLB / first touches
    const long long t0 = clock64();
    for (int k = 0; k < steps; k++) i = next[i];
    const long long t1 = clock64();
    *out = i; *cycles = t1 - t0;
}

static float best_ms(int reps, void (*fn)(void*), void* arg) {
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    fn(arg); CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; r++) {
        CK(cudaEventRecord(a)); fn(arg); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        float ms; CK(cudaEventElapsedTime(&ms, a, b)); best = std::min(best, ms);
    }
    CK(cudaGetLastError());
    return best;
}

struct LdArg { const uint4* p; size_t n16; int reps; uint4* sink; int grid; };
struct GaArg { const uint4* p; uint32_t mask; int steps; uint4* sink; int grid; };
struct FmaArg { int iters; void* sink; int grid; };

int main() {
    int dev = 0; CK(cudaSetDevice(dev));
    cudaDeviceProp pr; CK(cudaGetDeviceProperties(&pr, dev));
    int sms = pr.multiProcessorCount, clock_khz = 0;
    CK(cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, dev));
    uint4* sink; CK(cudaMalloc(&sink, 64));
    const size_t big = (size_t)4 << 30;
    uint4* buf; CK(cudaMalloc(&buf, big)); CK(cudaMemset(buf, 1, big));
    printf("{\n \"gpu\": \"%s\", \"sms\": %d, \"sm_clock_mhz_nominal\": %.0f, \"l2_bytes\": %d,\n", pr.name, sms, clock_khz / 1e3, pr.l2CacheSize);

    // ---- streaming 16-byte loads
    const size_t foot[3] = {(size_t)32 << 20, (size_t)100 << 20, big};
    const char* fname[3] = {"32MB", "100MB", "4GB_hbm"};
    printf(" \"ld128_gbs\": {");
    for (int f = 0; f < 3; f++) {
        LdArg a{buf, foot[f] / 16, f == 2 ? 1 : (int)(((size_t)8 << 30) / foot[f]), sink, sms * 8};
        float ms = best_ms(5, [](void* v) { LdArg* a = (LdArg*)v; k_ld128<<<a->grid, 256>>>(a->p, a->n16, a->reps, a->sink); }, &a);
        printf("%s\"%s\": %.1f", f ? ", " : "", fname[f], (double)foot[f] * a.reps / (ms * 1e-3) / 1e9);
    }
    printf("},\n");
    // ---- random 128-byte record gathers (independent per lane)
    printf(" \"gather128_gbs\": {");
    const size_t gfoot[3] = {(size_t)32 << 20, (size_t)64 << 20, (size_t)2 << 30};
    const char* gname[3] = {"32MB", "64MB", "2GB_hbm"};
    for (int f = 0; f < 3; f++) {
        GaArg a{buf, (uint32_t)(gfoot[f] / 128 - 1), 256, sink, sms * 8};
        float ms = best_ms(5, [](void* v) { GaArg* a = (GaArg*)v; k_gather128<<<a->grid, 256>>>(a->p, a->mask, a->steps, a->sink); }, &a);
        printf("%s\"%s\": %.1f", f ? ", " : "", gname[f], (double)a.grid * 256 * a.steps * 128 / (ms * 1e-3) / 1e9);
    }
    printf("},\n");
    // ---- FMA issue rates
    {
        FmaArg a{4096, sink, sms * 8};
        float ms32 = best_ms(5, [](void* v) { FmaArg* a = (FmaArg*)v; k_fma<float><<<a->grid, 256>>>(1.0001f, 0.9999f, a->iters, (float*)a->sink); }, &a);
        FmaArg b{1024, sink, sms * 8};
        float ms64 = best_ms(5, [](void* v) { FmaArg* a = (FmaArg*)v; k_fma<double><<<a->grid, 256>>>(1.0001, 0.9999, a->iters, (double*)a->sink); }, &b);
        const double n32 = (double)a.grid * 256 * a.iters * 64, n64 = (double)b.grid * 256 * b.iters * 64;
        printf(" \"fp32_ffma_tflops\": %.2f, \"fp32_warp_inst_per_s\": %.4e,\n", 2 * n32 / (ms32 * 1e-3) / 1e12, n32 / 32 / (ms32 * 1e-3));
        printf(" \"fp64_dfma_tflops\": %.2f, \"fp64_warp_inst_per_s\": %.4e,\n", 2 * n64 / (ms64 * 1e-3) / 1e12, n64 / 32 / (ms64 * 1e-3));
    }
    // ---- dependent-load latency
    {
        const size_t feet[3] = {(size_t)16 << 10, (size_t)48 << 20, (size_t)2 << 30};
        const char* names[3] = {"l1_16KB", "l2_48MB", "hbm_2GB"};
        printf(" \"load_latency_cycles\": {");
        for (int f = 0; f < 3; f++) {
            const size_t n = feet[f] / 4, step = 64;                // one element every 256 bytes
            const size_t m = n / step;
            std::vector<uint32_t> perm(m);
            for (size_t i = 0; i < m; i++) perm[i] = (uint32_t)i;
            uint64_t s = 88172645463325252ull;
            for (size_t i = m - 1; i > 0; i--) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; std::swap(perm[i], perm[s % (i + 1)]); }
            std::vector<uint32_t> h(n, 0);
            for (size_t i = 0; i < m; i++) h[(size_t)perm[i] * step] = (uint32_t)((size_t)perm[(i + 1) % m] * step);
            uint32_t* d = (uint32_t*)buf;
            CK(cudaMemcpy(d, h.data(), n * 4, cudaMemcpyHostToDevice));
            uint32_t* out; long long* cyc; CK(cudaMalloc(&out, 4)); CK(cudaMalloc(&cyc, 8));
            const int steps = f == 0 ? 2048 : 20000;
            long long best = 1ll << 62;
            for (int r = 0; r < 3; r++) {
                k_chase<<<1, 1>>>(d, (uint32_t)((size_t)perm[0] * step), steps, out, cyc);
                long long c; CK(cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost)); best = std::min(best, c);
            }
            printf("%s\"%s\": %.1f", f ? ", " : "", names[f], (double)best / steps);
            cudaFree(out); cudaFree(cyc);
        }
        printf("},\n");
    }
    printf(" \"issue_peak_warp_inst_per_s_at_nominal_clock\": %.4e,\n", (double)sms * 4 * clock_khz * 1e3);
    printf(" \"how\": \"tools/peaks.cu: CUDA events, warm-up + best of 5; ld128 = grid-stride LDG.128 over the footprint; gather128 = independent random 128-byte records per lane (one BVH4 node); fma = 8 independent accumulators per thread, 8 blocks x 256 threads per SM; latency = single-thread pointer chase, 256-byte stride, clock64\"\n}\n");
    return 0;
}
