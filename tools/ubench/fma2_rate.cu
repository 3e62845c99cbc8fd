// Microbenchmark: issue rate of the packed FP32 instructions of sm_100a (FFMA2 / FMUL2 / FADD2, PTX
// fma.rn.f32x2 ...) by operand form, alone and mixed with ALU-pipe work. Each variant runs ITER iterations
// of NCH independent chains per thread; reports warp-instructions per clock per SM sub-partition and the
// FP32 lane-ops per clock per SMSP they amount to (32 = FP32 FMA-pipe peak).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define ITER 4096
#define NCH 12
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ float lo(u64 v) { return __uint_as_float((uint32_t)v); }
__device__ __forceinline__ float hi(u64 v) { return __uint_as_float((uint32_t)(v >> 32)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

struct P { float2 c1, c2; float s1, s2; uint32_t bits; };

template <int V>
__global__ void k(float *out, const P p, int n) {
    u64 a[NCH], b[NCH], c[NCH];
    uint32_t m = p.bits + threadIdx.x;
    for (int i = 0; i < NCH; i++) {
        a[i] = pk(threadIdx.x * 0.001f + i, threadIdx.x * 0.002f + i);
        b[i] = pk(1.0f + i * 1e-6f + threadIdx.x * 1e-7f, 1.0f - i * 1e-6f - threadIdx.x * 1e-7f);
        c[i] = pk(i * 1e-3f + threadIdx.x * 1e-5f, i * 2e-3f + threadIdx.x * 1e-5f);
    }
    const u64 k1 = pk(p.c1.x, p.c1.y), k2 = pk(p.c2.x, p.c2.y);
    const u64 kb = pk(p.s1, p.s1);   // broadcast of one constant-bank scalar
    for (int it = 0; it < n; it++) {
#pragma unroll
        for (int i = 0; i < NCH; i++) {
            if (V == 0) a[i] = fma2(a[i], b[i], c[i]);                 // 3 distinct register pairs
            if (V == 1) a[i] = fma2(a[i], k1, c[i]);                   // 2 pairs + constant pair
            if (V == 2) a[i] = mul2(a[i], b[i]);                       // FMUL2, 2 pairs
            if (V == 3) a[i] = fma2(a[i], k1, k2);                     // 1 pair + 2 constant pairs
            if (V == 4) a[i] = fma2(a[i], kb, c[i]);                   // 2 pairs + broadcast scalar constant
            if (V == 5) a[i] = add2(a[i], b[i]);                       // FADD2
            if (V == 6) {                                              // FFMA2 (const) + 1 SEL per packed op
                a[i] = fma2(a[i], k1, c[i]);
                float x = lo(c[i]); x = (m >> i) & 1u ? x : p.s2; c[i] = pk(x, hi(c[i]));
            }
            if (V == 7) {                                              // kernel-like: FMUL2 + FFMA2c + FFMA2c + FMUL2(sel,sel)
                u64 s = mul2(a[i], b[i]);
                s = fma2(c[i], k1, s);
                s = fma2(b[i], k1, s);
                float e0 = (m >> i) & 1u ? p.s1 : p.s2, e1 = (m >> (i + 16)) & 1u ? p.s1 : p.s2;
                a[i] = mul2(s, pk(e0, e1));
            }
            if (V == 8) { a[i] = fma2(a[i], k1, c[i]); c[i] = mul2(c[i], b[i]); }   // FFMA2c + FMUL2 mix
            if (V == 9) { const float sv = lo(b[i]); a[i] = fma2(a[i], pk(sv, sv), c[i]); }   // 2 pairs + per-lane scalar broadcast (R.F32): 5 register reads
            if (V == 10) { const float sv = lo(b[i]); a[i] = fma2(a[i], pk(sv, sv), c[i]); c[i] = mul2(c[i], b[(i + 1) % NCH]); }   // the same mixed 1:1 with FMUL2
        }
        if (V == 6 || V == 7) m = m * 1664525u + 1013904223u;
    }
    float s = 0;
    for (int i = 0; i < NCH; i++) s += lo(a[i]) + hi(a[i]) + lo(c[i]) + hi(c[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int V> void run(const char *name, double packed_per_chain, double other_per_chain) {
    float *d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    P p{{0.999f, 0.998f}, {1e-3f, 2e-3f}, 0.997f, 0.5f, 0x5a5a1234u};
    k<V><<<148 * 4, 256>>>(d, p, 16);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<V><<<148 * 4, 256>>>(d, p, ITER);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double clks = ms * 1e-3 * clk * 1e3;
    const double packed = (double)ITER * NCH * packed_per_chain * 8;   // warp-instr per SMSP (8 warps per SMSP)
    const double other = (double)ITER * NCH * other_per_chain * 8;
    printf("%-52s %.3f ms  packed %.3f + other %.3f warp-inst/clk/SMSP = %.1f FP32 lane-ops/clk/SMSP (peak 32) @%d MHz\n",
           name, ms, packed / clks, other / clks, packed / clks * 64, clk / 1000);
    cudaFree(d);
}
int main() {
    run<0>("FFMA2 a=a*b+c (3 reg pairs)", 1, 0);
    run<1>("FFMA2 a=a*K+c (2 reg pairs + const pair)", 1, 0);
    run<2>("FMUL2 a=a*b (2 reg pairs)", 1, 0);
    run<3>("FFMA2 a=a*K+K2 (1 reg pair)", 1, 0);
    run<4>("FFMA2 a=a*bcast(const)+c", 1, 0);
    run<5>("FADD2 a=a+b", 1, 0);
    run<6>("FFMA2c + (SHF,LOP,SEL..) per packed op", 1, 3);
    run<7>("kernel-like M: FMUL2,FFMA2c,FFMA2c,FMUL2 + 2 SEL", 4, 6);
    run<8>("FFMA2c + FMUL2 mix", 2, 0);
    run<9>("FFMA2 a=a*bcast(per-lane scalar)+c (R.F32 form)", 1, 0);
    run<10>("FFMA2 R.F32 form + FMUL2 mix", 2, 0);
    return 0;
}
