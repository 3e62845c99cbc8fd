// Microbenchmark: latency and throughput of DFMA / DMUL on sm_100a (B200), to put a measured number under the
// FP64 rescue kernel's roofline. NCHAIN independent dependent-chains per thread, WPS warps per SM sub-partition.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
struct P { double c1, c2; };
// MODE 0: a = fma(a, const, c)   MODE 1: a = a * b   MODE 2: alternating
template <int NCHAIN, int MODE>
__global__ void k(double *out, const P p, int n, long long *clk) {
    double a[NCHAIN], b[NCHAIN], c[NCHAIN];
    for (int i = 0; i < NCHAIN; i++) {
        a[i] = threadIdx.x * 0.001 + i;
        b[i] = 1.0 + threadIdx.x * 1e-12;
        c[i] = i * 1e-3 + threadIdx.x * 1e-5;
    }
    long long t0 = clock64();
    for (int it = 0; it < n; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
#pragma unroll
            for (int i = 0; i < NCHAIN; i++) {
                if (MODE == 0) a[i] = fma(a[i], p.c1, c[i]);
                if (MODE == 1) a[i] = a[i] * b[i];
                if (MODE == 2) a[i] = (r & 1) ? fma(a[i], p.c1, c[i]) : a[i] * b[i];
            }
        }
    }
    long long t1 = clock64();
    double acc = 0;
    for (int i = 0; i < NCHAIN; i++) acc += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (blockIdx.x == 0 && threadIdx.x == 0) *clk = t1 - t0;
}
template <int NCHAIN, int MODE> void run(const char *name, int warps_per_smsp) {
    double *d; long long *dc, hc; cudaMalloc(&d, 148 * 1024 * 8); cudaMalloc(&dc, 8);
    const int n = 1024, threads = 32 * 4 * warps_per_smsp;   // one CTA per SM
    k<NCHAIN, MODE><<<148, threads>>>(d, P{0.999, 1e-3}, 16, dc);
    k<NCHAIN, MODE><<<148, threads>>>(d, P{0.999, 1e-3}, n, dc);
    cudaDeviceSynchronize();
    cudaMemcpy(&hc, dc, 8, cudaMemcpyDeviceToHost);
    const double inst_per_warp = (double)n * 8 * NCHAIN;
    printf("%-24s chains=%2d warps/SMSP=%d : %.2f clk per instr per warp, %.3f instr/clk/SMSP = %.1f FP64 lanes/clk/SM\n", name, NCHAIN,
           warps_per_smsp, hc / inst_per_warp, inst_per_warp * warps_per_smsp / hc, inst_per_warp * warps_per_smsp / hc * 32 * 4);
    cudaFree(d); cudaFree(dc);
}
int main() {
    run<1, 0>("DFMA dependent", 1);  run<1, 1>("DMUL dependent", 1);
    run<4, 0>("DFMA", 1);  run<8, 0>("DFMA", 1);  run<8, 0>("DFMA", 2);  run<8, 0>("DFMA", 4);
    run<8, 1>("DMUL", 2);  run<8, 2>("DMUL/DFMA alternating", 2);  run<8, 2>("DMUL/DFMA alternating", 3);
    return 0;
}
