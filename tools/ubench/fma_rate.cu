// Microbenchmark: issue rate of FP32 FMA-pipe instructions by operand form on sm_100a.
// Each variant runs ITER iterations of 16 independent chains per thread; reports warp-instructions
// per clock per SM sub-partition (1.0 = FP32 peak).
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 4096
template <int V>
__global__ void k(float *out, float cst, float cst2, int n) {
    float a[16], b[16], c[16];
    for (int i = 0; i < 16; i++) { a[i] = threadIdx.x * 0.001f + i; b[i] = 1.0f + i * 1e-6f + threadIdx.x * 1e-7f; c[i] = i * 1e-3f + threadIdx.x * 1e-5f; }
    for (int it = 0; it < n; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) {
            if (V == 0) a[i] = fmaf(a[i], b[i], c[i]);             // 3 distinct registers
            if (V == 1) a[i] = fmaf(a[i], cst, c[i]);              // 2 registers + constant bank
            if (V == 2) a[i] = a[i] * b[i];                        // FMUL 2 registers
            if (V == 3) a[i] = fmaf(a[i], cst, cst2);              // 1 register + 2 constants
            if (V == 4) a[i] = fmaf(a[i], b[i], a[i]);             // 2 distinct registers (a twice)
            if (V == 5) a[i] = fmaf(b[i], c[(i + 1) & 15], a[i]);  // accumulate form, 3 distinct
            if (V == 6) { a[i] = fmaf(a[i], b[i], c[i]); c[i] = c[i] * cst; }   // mixed FFMA3 + FMUL-const
        }
    }
    float s = 0;
    for (int i = 0; i < 16; i++) s += a[i] + c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int V> void run(const char *name, int per_iter) {
    float *d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<V><<<148 * 4, 256>>>(d, 0.999f, 1e-3f, 16);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<V><<<148 * 4, 256>>>(d, 0.999f, 1e-3f, ITER);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    // warps per SMSP = 4 blocks * 8 warps / 4 = 8
    double inst = (double)ITER * per_iter * 8;   // warp-instr per SMSP
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("%-40s %.3f ms  -> %.3f warp-inst/clk/SMSP at %d MHz (assumed)\n", name, ms, inst / (ms * 1e-3 * clk * 1e3), clk / 1000);
    cudaFree(d);
}
int main() {
    run<0>("FFMA a=a*b+c (3 regs)", 16);
    run<1>("FFMA a=a*const+c (2 regs + cbank)", 16);
    run<2>("FMUL a=a*b (2 regs)", 16);
    run<3>("FFMA a=a*const+const (1 reg)", 16);
    run<4>("FFMA a=a*b+a (2 distinct regs)", 16);
    run<5>("FFMA a=b*c'+a (3 regs, acc form)", 16);
    run<6>("FFMA3 + FMUL-const mix", 32);
    return 0;
}
