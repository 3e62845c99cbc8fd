// Microbenchmark: latency and low-occupancy throughput of the packed FP32 instructions (FFMA2/FMUL2) on
// sm_100a. NCHAIN independent dependent-chains per thread, WPS warps per SM sub-partition.
//   cycles per packed instruction per SMSP = the number the PairHMM kernel's schedule has to live with.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm volatile("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
struct P { float c1, c2; };
// MODE 0: a = fma2(a, UR, c)   (chain through a; c loop-invariant vector pair)
// MODE 1: a = mul2(a, b)
// MODE 2: alternating mul2 / fma2(UR) on the same chain (like the M chain)
// MODE 3: scalar FFMA chain (reference)
template <int NCHAIN, int MODE>
__global__ void k(float *out, const P p, int n, long long *clk) {
    u64 a[NCHAIN], b[NCHAIN], c[NCHAIN]; float s[NCHAIN];
    for (int i = 0; i < NCHAIN; i++) {
        a[i] = pk(threadIdx.x * 0.001f + i, threadIdx.x * 0.002f + i);
        b[i] = pk(1.0f + threadIdx.x * 1e-7f, 1.0f - threadIdx.x * 1e-7f);
        c[i] = pk(i * 1e-3f + threadIdx.x * 1e-5f, threadIdx.x * 1e-5f);
        s[i] = threadIdx.x + i;
    }
    const u64 k1 = pk(p.c1, p.c1);
    long long t0 = clock64();
    for (int it = 0; it < n; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
#pragma unroll
            for (int i = 0; i < NCHAIN; i++) {
                if (MODE == 0) a[i] = fma2(a[i], k1, c[i]);
                if (MODE == 1) a[i] = mul2(a[i], b[i]);
                if (MODE == 2) a[i] = (r & 1) ? fma2(a[i], k1, c[i]) : mul2(a[i], b[i]);
                if (MODE == 3) s[i] = fmaf(s[i], p.c1, p.c2);
            }
        }
    }
    long long t1 = clock64();
    float acc = 0;
    for (int i = 0; i < NCHAIN; i++) acc += __uint_as_float((uint32_t)a[i]) + __uint_as_float((uint32_t)(a[i] >> 32)) + s[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (blockIdx.x == 0 && threadIdx.x == 0) *clk = t1 - t0;
}
template <int NCHAIN, int MODE> void run(const char *name, int warps_per_smsp) {
    float *d; long long *dc, hc; cudaMalloc(&d, 148 * 1024 * 4 * 4); cudaMalloc(&dc, 8);
    const int n = 2048, threads = 32 * 4 * warps_per_smsp;   // one CTA per SM
    k<NCHAIN, MODE><<<148, threads>>>(d, P{0.999f, 1e-3f}, 16, dc);
    k<NCHAIN, MODE><<<148, threads>>>(d, P{0.999f, 1e-3f}, n, dc);
    cudaDeviceSynchronize();
    cudaMemcpy(&hc, dc, 8, cudaMemcpyDeviceToHost);
    const double inst_per_warp = (double)n * 8 * NCHAIN;
    printf("%-28s chains=%2d warps/SMSP=%d : %.2f clk per instr per warp, %.3f instr/clk/SMSP\n", name, NCHAIN, warps_per_smsp,
           hc / inst_per_warp, inst_per_warp * warps_per_smsp / hc);
    cudaFree(d); cudaFree(dc);
}
int main() {
    run<1, 0>("FFMA2(UR) dependent", 1);  run<1, 1>("FMUL2 dependent", 1);  run<1, 3>("FFMA scalar dependent", 1);
    run<2, 0>("FFMA2(UR)", 1);  run<4, 0>("FFMA2(UR)", 1);  run<8, 0>("FFMA2(UR)", 1);
    run<4, 0>("FFMA2(UR)", 2);  run<8, 0>("FFMA2(UR)", 2);  run<8, 0>("FFMA2(UR)", 4);
    run<4, 2>("FMUL2/FFMA2 alternating", 1);  run<8, 2>("FMUL2/FFMA2 alternating", 1);
    run<4, 2>("FMUL2/FFMA2 alternating", 2);  run<8, 2>("FMUL2/FFMA2 alternating", 2);  run<8, 2>("FMUL2/FFMA2 alternating", 4);
    run<8, 1>("FMUL2", 2);  run<8, 3>("FFMA scalar", 2);  run<8, 3>("FFMA scalar", 4);
    return 0;
}
