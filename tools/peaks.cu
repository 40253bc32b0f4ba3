// tools/peaks.cu -- microbenchmarks for the roofline denominators of the render kernels:
// the rates MEASURED_PEAKS.json does not carry.  Measurement infrastructure, not product.
//
//   fp32 FFMA issue rate            one FFMA per lane per issue slot         -> TFLOP/s (2 flop per FFMA)
//   fp32 FFMA2 (fma.rn.f32x2)       two FFMA per lane per issue slot          -> TFLOP/s
//   MUFU.EX2 (ex2.approx.ftz.f32)                                              -> G results / s
//   MUFU.RCP (rcp.approx.ftz.f32)                                              -> G results / s
//   red.global.add.f32 / .v4.f32    into an L2-resident footprint (8 MB) and an HBM-resident one (2 GB),
//                                   a warp to consecutive addresses (the access pattern of the render backward's
//                                   per-(strip, Gaussian) reductions and of accumulate_packed)  -> G atomics / s, GB/s
//
// Every figure: best of 5 launches after 2 warm-ups, CUDA events on the launching stream.  One JSON object on stdout.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/b200_peaks tools/peaks.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int ITERS = 4096;
constexpr int CHAINS = 8;

__global__ void __launch_bounds__(256) ffma_kernel(float* out, float a, float b) {
    float x[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; k++) x[k] = (float)(threadIdx.x + k);
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int k = 0; k < CHAINS; k++) x[k] = fmaf(x[k], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < CHAINS; k++) s += x[k];
    if (s == 12345.678f) out[0] = s;
}

__global__ void __launch_bounds__(256) ffma2_kernel(float* out, float a, float b) {
    unsigned long long x[CHAINS], aa, bb;
    asm("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a));
    asm("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(b));
#pragma unroll
    for (int k = 0; k < CHAINS; k++) { float v = (float)(threadIdx.x + k); asm("mov.b64 %0, {%1, %1};" : "=l"(x[k]) : "f"(v)); }
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int k = 0; k < CHAINS; k++) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(x[k]) : "l"(aa), "l"(bb));
    }
    unsigned long long s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; k++) s ^= x[k];
    if (s == 0x123456789abcdefull) out[0] = 1.f;
}

template <int OP>   // 0: ex2.approx.ftz, 1: rcp.approx.ftz
__global__ void __launch_bounds__(256) mufu_kernel(float* out) {
    float x[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; k++) x[k] = -0.001f * (float)(threadIdx.x + k + 1);
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int k = 0; k < CHAINS; k++) {
                if (OP == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[k]));
                else {   // 1/x + 1: without the add ptxas folds rcp(rcp(x)) back into x; the FADD pipe is 8x wider than MUFU
                    asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x[k]));
                    x[k] += 1.0f;
                }
            }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < CHAINS; k++) s += x[k];
    if (s == 12345.678f) out[0] = s;
}

// Every warp walks `rounds` pseudo-random 128-byte (scalar: one red.f32 per lane) or 512-byte (v4) lines of the
// footprint; lanes hit consecutive addresses.
template <bool V4>
__global__ void __launch_bounds__(256) red_kernel(float* buf, size_t n_lines, int rounds) {
    const unsigned warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    unsigned long long h = 0x9e3779b97f4a7c15ull * (warp + 1);
    for (int r = 0; r < rounds; r++) {
        h = h * 6364136223846793005ull + 1442695040888963407ull;
        const size_t line = (size_t)(h >> 24) % n_lines;
        if (V4) {
            float4* p = reinterpret_cast<float4*>(buf) + line * 32 + lane;
            atomicAdd(p, make_float4(1.f, 1.f, 1.f, 1.f));
        } else {
            atomicAdd(buf + line * 32 + lane, 1.0f);
        }
    }
}

template <typename F>
static float best_ms(F launch) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int i = 0; i < 7; i++) {
        CK(cudaEventRecord(e0));
        launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (i >= 2 && ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

int main() {
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    float* out; CK(cudaMalloc(&out, 256));
    const int blocks = sms * 8, threads = 256;
    const double lanes = (double)blocks * threads;
    const double n_op = lanes * ITERS * 4.0 * CHAINS;

    const float t_ffma = best_ms([&] { ffma_kernel<<<blocks, threads>>>(out, 0.999f, 0.001f); });
    const float t_ffma2 = best_ms([&] { ffma2_kernel<<<blocks, threads>>>(out, 0.999f, 0.001f); });
    const float t_ex2 = best_ms([&] { mufu_kernel<0><<<blocks, threads>>>(out); });
    const float t_rcp = best_ms([&] { mufu_kernel<1><<<blocks, threads>>>(out); });

    int clock_khz = 0; CK(cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, 0));
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"sm_clock_attr_mhz\": %.0f,\n", prop.name, sms, clock_khz / 1000.0);
    printf(" \"ffma_tflops\": %.2f, \"ffma_ginst_s\": %.1f,\n", 2.0 * n_op / (t_ffma * 1e-3) / 1e12, n_op / 32.0 / (t_ffma * 1e-3) / 1e9);
    printf(" \"ffma2_tflops\": %.2f, \"ffma2_ginst_s\": %.1f,\n", 4.0 * n_op / (t_ffma2 * 1e-3) / 1e12, n_op / 32.0 / (t_ffma2 * 1e-3) / 1e9);
    printf(" \"mufu_ex2_gops\": %.1f, \"mufu_rcp_gops\": %.1f,\n", n_op / (t_ex2 * 1e-3) / 1e9, n_op / (t_rcp * 1e-3) / 1e9);

    // red.global.add
    const size_t small_bytes = (size_t)8 << 20, big_bytes = (size_t)2 << 30;
    float* buf; CK(cudaMalloc(&buf, big_bytes));
    CK(cudaMemset(buf, 0, big_bytes));
    const int rblocks = sms * 16, rounds = 512;
    const double warps = (double)rblocks * threads / 32.0;
    struct { const char* name; bool v4; size_t bytes; } cfg[4] = {
        {"red_f32_l2", false, small_bytes}, {"red_f32_hbm", false, big_bytes},
        {"red_v4_l2", true, small_bytes}, {"red_v4_hbm", true, big_bytes}};
    for (int c = 0; c < 4; c++) {
        const size_t line_bytes = cfg[c].v4 ? 512 : 128;
        const size_t n_lines = cfg[c].bytes / line_bytes;
        float t;
        if (cfg[c].v4) t = best_ms([&] { red_kernel<true><<<rblocks, threads>>>(buf, n_lines, rounds); });
        else t = best_ms([&] { red_kernel<false><<<rblocks, threads>>>(buf, n_lines, rounds); });
        const double ops = warps * rounds * 32.0;                 // lane-level atomic instructions' worth of elements
        const double bytes = warps * rounds * (double)line_bytes;
        printf(" \"%s\": {\"g_lane_ops_s\": %.2f, \"g_f32_adds_s\": %.2f, \"gbs\": %.1f, \"g_sectors_s\": %.2f}%s\n", cfg[c].name,
               ops / (t * 1e-3) / 1e9, ops * (cfg[c].v4 ? 4.0 : 1.0) / (t * 1e-3) / 1e9, bytes / (t * 1e-3) / 1e9,
               bytes / 32.0 / (t * 1e-3) / 1e9, c == 3 ? "" : ",");
    }
    printf("}\n");
    return 0;
}
