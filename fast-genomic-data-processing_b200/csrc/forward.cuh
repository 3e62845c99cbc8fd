// PairHMM forward kernel template for sm_100a (no tensor cores: the M/X/Y recurrence is a wavefront).
//
// Work decomposition: one WARP sweeps one work unit = one read against a run of haplotypes.
//   * a group of W lanes (W = 32, 16, 8; 32/W reads per warp, all against the same haplotype run)
//     sweeps one read: lane l of the group owns K consecutive rows of the DP matrix (read rows l*K ..
//     l*K+K-1, top-aligned); rows past the end of the read are dead slots (all constants 0). The read's
//     last row is slot KL of lane (R-1)/K; KL is a template parameter so that the two running sums of
//     the last row stay in fixed registers. Narrower groups = more rows per lane = the per-column
//     overhead (2 shuffles, boundary selects, loop, loads) is amortised over more cells.
//   * lanes are skewed by one tile (kTile columns): at step t lane l works on tile t-l of the unit's
//     column stream. The only cross-lane traffic per column is two shuffles: the bottom row of lane l,
//     pre-combined with the transition constants of the row below it (D = diagonal term of M, U = the
//     X of that row). Lane 0 substitutes the matrix's row-0 boundary (D = 2^s/H * pGAPM[row 1], U = 0).
//   * the haplotypes of a unit are streamed back to back through the same pipeline, separated by one
//     SEP word (= column 0 of the next haplotype), so pipeline fill/drain is paid once per unit.
//   * UG ("uniform gap continuation"): when every gcp byte of the launch is the same value -- always
//     the case for Mutect2Cpp, which fills a constant array (PairHMMLikelihoodCalculationEngine.cpp:
//     284-292) -- pXX=pYY and pGAPM come from the kernel's constant bank instead of registers. This
//     matters on sm_100: an FFMA with three distinct REGISTER sources issues at ~0.6/clk/SMSP, with a
//     constant-bank operand at ~0.9 (tools/ubench/fma_rate.cu, measured on B200), and it is what keeps
//     every FMA-pipe instruction of the inner loop at <= 2 register reads.
// Numerics follow avx-pairhmm-template.h:177-192 / SURVEY.md Appendix A (association order kept).
#pragma once
#include "kernels.cuh"

namespace phmm {

constexpr unsigned kFull = 0xffffffffu;

// initNative sets MXCSR flush-to-zero on the reference's threads (IntelPairHmm.cc:230), which also flushes DOUBLE denormals;
// nvcc's -ftz covers FP32 only. A 2^1020-scaled FP64 sum that ends below DBL_MIN is zero in the reference (log10 = -inf):
// flushed here, at the one place it is visible. Intermediate FP64 denormals are not flushed (no FP64 FTZ mode on the GPU):
// a pair whose true log10 likelihood lies in about [-631, -600] may come out finite here and -inf in the reference
// (tests/test_gpu_parity.py::test_fp64_underflow_window; such a read is removed by filterPoorlyModeledEvidence either way).
__device__ __forceinline__ double flush_denormal(double r) { return r < 2.2250738585072014e-308 ? 0.0 : r; }

template <typename T> struct Num;
template <> struct Num<float> {
    static __device__ __forceinline__ float init_const() { return 1.329227995784916e+36f; }  // 2^120
    template <bool EXACT> static __device__ __forceinline__ float mul(float a, float b) { return EXACT ? __fmul_rn(a, b) : a * b; }
    template <bool EXACT> static __device__ __forceinline__ float add(float a, float b) { return EXACT ? __fadd_rn(a, b) : a + b; }
    // a*b + c, contracted unless EXACT
    template <bool EXACT> static __device__ __forceinline__ float mad(float a, float b, float c) {
        return EXACT ? __fadd_rn(__fmul_rn(a, b), c) : fmaf(a, b, c);
    }
    static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
    static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
    static __device__ __forceinline__ float uxy(const KernelArgs &a) { return a.u_xy_f; }
    static __device__ __forceinline__ float ugapm(const KernelArgs &a) { return a.u_gapm_f; }
};
template <> struct Num<double> {
    static __device__ __forceinline__ double init_const() { return 1.1235582092889474e+307; }  // 2^1020
    template <bool EXACT> static __device__ __forceinline__ double mul(double a, double b) { return EXACT ? __dmul_rn(a, b) : a * b; }
    template <bool EXACT> static __device__ __forceinline__ double add(double a, double b) { return EXACT ? __dadd_rn(a, b) : a + b; }
    template <bool EXACT> static __device__ __forceinline__ double mad(double a, double b, double c) {
        return EXACT ? __dadd_rn(__dmul_rn(a, b), c) : fma(a, b, c);
    }
    static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
    static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
    static __device__ __forceinline__ double uxy(const KernelArgs &a) { return a.u_xy_d; }
    static __device__ __forceinline__ double ugapm(const KernelArgs &a) { return a.u_gapm_d; }
};

// A0 C1 T2 G3 N4 of pairhmm_common.h:69-85; every byte that is not C,T,G,N behaves as 'A'.
__device__ __forceinline__ uint32_t base_index(uint32_t b) {
    uint32_t c = 0u;
    c = (b == 'C') ? 1u : c;
    c = (b == 'T') ? 2u : c;
    c = (b == 'G') ? 3u : c;
    c = (b == 'N') ? 4u : c;
    return c;
}

template <int C> struct TileWord;
template <> struct TileWord<1> { using type = uint32_t; };
template <> struct TileWord<2> { using type = uint2; };
template <> struct TileWord<4> { using type = uint4; };
__device__ __forceinline__ uint32_t tw_get(const uint32_t &t, int) { return t; }
__device__ __forceinline__ uint32_t tw_get(const uint2 &t, int j) { return j == 0 ? t.x : t.y; }
__device__ __forceinline__ uint32_t tw_get(const uint4 &t, int j) { return j == 0 ? t.x : j == 1 ? t.y : j == 2 ? t.z : t.w; }
__device__ __forceinline__ uint32_t tw_or(const uint32_t &t) { return t; }
__device__ __forceinline__ uint32_t tw_or(const uint2 &t) { return t.x | t.y; }
__device__ __forceinline__ uint32_t tw_or(const uint4 &t) { return t.x | t.y | t.z | t.w; }

// per-warp table of the unit's haplotypes (first tile, 2^s/H), filled once per unit so that the SEP
// handler -- which runs divergently, one lane at a time -- needs no global loads and no division
// idx = the haplotype's position in the unit (rescue units sweep only the underflowed ones, see below)
template <typename T> struct HapInfo { uint32_t tix; uint32_t idx; T initY; };

// minimum resident CTAs per SM requested from ptxas (caps registers): the FP32 kernels are issue-bound
// and want >= 5-6 warps per scheduler; the FP64 kernels need their registers
#ifndef PHMM_MINB_F32
#define PHMM_MINB_F32 6
#endif
#define PHMM_MINB(T, W, K) ((sizeof(T) == 4 && (W) == 32 && (K) <= 5) ? PHMM_MINB_F32 : 1)

template <typename T, int W, int K, int KL, bool EXACT, bool UG>
__global__ void __launch_bounds__(kThreads, PHMM_MINB(T, W, K)) forward_kernel(const KernelArgs a) {
    static_assert(W == 32 || W == 16 || W == 8, "lane-group width");
    static_assert(KL >= 0 && KL < K, "last-row slot out of range");
    static_assert(!(EXACT && UG), "the exact mode uses the general path");
    using N = Num<T>;
    constexpr int C = kScalarTile;              // columns per step of these kernels (the word layout is aligned to kTile)
    using Tile = typename TileWord<C>::type;
    constexpr int GB = group_bits(K);           // rows per match-mask register
    constexpr int NG = (K + GB - 1) / GB;
    constexpr int FIELD = code_field(K);
    constexpr int KR = UG ? 1 : K;              // per-row pXY/pGAPM registers only on the general path
    const int lane = threadIdx.x & 31;
    const int sl = lane & (W - 1);              // lane inside its group
    const int seg = lane / W;                   // which of the warp's 32/W reads
    const bool lane0 = sl == 0;
    const T *__restrict__ ph = static_cast<const T *>(a.ph2pr);
    const T *__restrict__ mmt = static_cast<const T *>(a.mm);
    const Tile *__restrict__ tiles = reinterpret_cast<const Tile *>(a.hapw);
    const uint32_t n_units = a.n_units_dev ? *a.n_units_dev : a.n_units;
    __shared__ HapInfo<T> hap_info_all[kThreads / 32][kMaxUnitHaps];
    HapInfo<T> *hap_info = hap_info_all[threadIdx.x >> 5];

    for (;;) {
        uint32_t u = 0;
        if (lane == 0) u = atomicAdd(a.queue, 1u);
        u = __shfl_sync(kFull, u, 0);
        if (u >= n_units) break;
        const DevUnit unit = a.units[u];
        uint32_t unit_nh = unit.shape & 0xffu;
        const uint32_t unit_nreads = (unit.shape >> 8) & 0xffu;
        const bool seg_valid = (uint32_t)seg < unit_nreads;   // a partial unit leaves groups without a read: they
        const uint32_t my_read = unit.read + (seg_valid ? seg : 0);   // recompute read 0 and emit nothing
        const uint32_t my_out = unit.out + (uint32_t)seg * (unit.shape >> 16);
        const DevRead rr = a.reads[my_read];
        const int last_lane = (rr.len - 1) / K;       // owns the read's last row, in slot KL

        // ---- per-slot constants (ReadForPairHMM.cpp:72-80; stripeINITIALIZATION template.h:150-152)
        T pMM[K], pMX[K], pMY[K], d1[K], d3[K], pXYr[KR], pGAPMr[KR];
        uint32_t mask[NG];                            // per haplotype base b: bit GB*b+i = row i of the group matches
#pragma unroll
        for (int i = 0; i < NG; ++i) mask[i] = 0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int g = sl * K + k;                 // 0-based read row
            if (g < rr.len) {
                const uint8_t *p = a.rd + rr.off + g;
                const uint32_t rb = base_index(p[0]);
                const int bq = p[a.plane] & 127, iq = p[2 * a.plane] & 127, dq = p[3 * a.plane] & 127,
                          cq = p[4 * a.plane] & 127;
                const int mx = max(iq, dq), mn = min(iq, dq);
                pMM[k] = mmt[((mx * (mx + 1)) >> 1) + mn];
                pMX[k] = ph[iq];
                pMY[k] = ph[dq];
                if constexpr (!UG) { pXYr[k] = ph[cq]; pGAPMr[k] = N::sub((T)1.0, ph[cq]); }
                const T dist = ph[bq];
                d1[k] = N::sub((T)1.0, dist);
                d3[k] = N::div(dist, (T)3.0);
                // match: same base, or either side is N (template.h:38,48-51)
                const uint32_t one = 1u << (k % GB);
                const uint32_t all = one | (one << GB) | (one << (2 * GB)) | (one << (3 * GB)) | (one << (4 * GB));
                mask[k / GB] |= (rb == 4u) ? all : ((one << (GB * rb)) | (one << (4 * GB)));
            } else {                                   // dead slot below the read
                pMM[k] = pMX[k] = pMY[k] = d1[k] = d3[k] = (T)0;
                if constexpr (!UG) { pXYr[k] = pGAPMr[k] = (T)0; }
            }
        }
        // transition constants of the row below this lane's last slot (= slot 0 of lane+1; 0 past lane 31)
        T nMM = __shfl_down_sync(kFull, pMM[0], 1, W), nMX = __shfl_down_sync(kFull, pMX[0], 1, W);
        T nXXr = (T)0, nGAPMr = (T)0;
        if constexpr (!UG) { nXXr = __shfl_down_sync(kFull, pXYr[0], 1, W); nGAPMr = __shfl_down_sync(kFull, pGAPMr[0], 1, W); }
        if (sl == W - 1) { nMM = nMX = nXXr = nGAPMr = (T)0; }
        // gap-continuation constants: constant bank (UG) or this row's registers
#define PXY(k)   (UG ? N::uxy(a) : pXYr[UG ? 0 : (k)])
#define PGAPM(k) (UG ? N::ugapm(a) : pGAPMr[UG ? 0 : (k)])
#define NXX      (UG ? N::uxy(a) : nXXr)
#define NGAPM    (UG ? N::ugapm(a) : nGAPMr)

        // pipeline steps: W (fill) + sum over haplotypes of their tile count; hap table for the SEP handler
        int steps = 0;
        __syncwarp();                                  // previous unit's readers are done with hap_info
        {
            bool keep = (uint32_t)lane < unit_nh;          // nh <= kMaxUnitHaps == 32 (host guarantees)
            // FP64 launch over rescue units: a unit names the read's whole haplotype run, only the pairs whose FP32 sum
            // fell below the reference's threshold (IntelPairHmm.cc:342) are swept, back to back
            if constexpr (sizeof(T) == 8) if (a.rescue_filter && keep) keep = a.fres[unit.out + lane] < 1e-28f;
            const uint32_t km = __ballot_sync(kFull, keep);
            unit_nh = (uint32_t)__popc(km);
            if (keep) {
                const DevHap hp = a.haps[a.hap_list[unit.hap_list + lane]];
                steps = (int)(hap_tiles(hp.len) * (kTile / C));
                hap_info[__popc(km & ((1u << lane) - 1u))] =
                    HapInfo<T>{hp.off / C, (uint32_t)lane, N::div(N::init_const(), (T)hp.len)};   // template.h:115
            }
        }
        __syncwarp();
#pragma unroll
        for (int o = 16; o; o >>= 1) steps += __shfl_xor_sync(kFull, steps, o);
        steps += W;

        T M[K], X[K], Y[K];
#pragma unroll
        for (int k = 0; k < K; ++k) { M[k] = X[k] = Y[k] = (T)0; }
        T Dprev = (T)0, sumM = (T)0, sumX = (T)0, Dbound = (T)0;
        T Din[C], Uin[C];
#pragma unroll
        for (int j = 0; j < C; ++j) { Din[j] = Uin[j] = (T)0; }
        int cur = -1;
        uint32_t tix = 32 * kTile / C - 1 - sl;        // tile index: lane l sees l PAD tiles, then the lead-in's [PAD..,SEP]
        Tile w = tiles[tix];

        // one DP column for this lane's K rows (computeMXY, template.h:177-192): Dd = diagonal term and
        // Uu = X of the row above, both as delivered by the lane above (or the row-0 boundary)
        auto column = [&](uint32_t cw, T Dd, T Uu) {
            uint32_t mbits[NG];
#pragma unroll
            for (int g = 0; g < NG; ++g) mbits[g] = __funnelshift_r(mask[g], 0u, FIELD ? (cw >> FIELD) : cw);
            T Mn[K], Xn[K], Yn[K];
            {
                const T e = (mbits[0] & 1u) ? d1[0] : d3[0];
                Mn[0] = N::template mul<EXACT>(Dd, e);
                Xn[0] = Uu;
                if (EXACT) Yn[0] = N::template add<true>(N::template mul<true>(M[0], pMY[0]), N::template mul<true>(Y[0], PXY(0)));
                else       Yn[0] = N::template mad<false>(Y[0], PXY(0), M[0] * pMY[0]);
            }
#pragma unroll
            for (int k = 1; k < K; ++k) {
                const T e = ((mbits[k / GB] >> (k % GB)) & 1u) ? d1[k] : d3[k];
                T s;
                if (EXACT) {
                    s = N::template add<true>(N::template add<true>(N::template mul<true>(M[k - 1], pMM[k]),
                                                                      N::template mul<true>(X[k - 1], PGAPM(k))),
                                              N::template mul<true>(Y[k - 1], PGAPM(k)));
                    Mn[k] = N::template mul<true>(s, e);
                    Xn[k] = N::template add<true>(N::template mul<true>(Mn[k - 1], pMX[k]), N::template mul<true>(Xn[k - 1], PXY(k)));
                    Yn[k] = N::template add<true>(N::template mul<true>(M[k], pMY[k]), N::template mul<true>(Y[k], PXY(k)));
                } else {
                    // every FFMA keeps the gap-continuation constant as its multiplier: with UG that is a
                    // constant-bank operand, i.e. two register reads per instruction
                    s = N::template mad<false>(Y[k - 1], PGAPM(k), N::template mad<false>(X[k - 1], PGAPM(k), M[k - 1] * pMM[k]));
                    Mn[k] = s * e;
                    Xn[k] = N::template mad<false>(Xn[k - 1], PXY(k), Mn[k - 1] * pMX[k]);
                    Yn[k] = N::template mad<false>(Y[k], PXY(k), M[k] * pMY[k]);
                }
            }
#pragma unroll
            for (int k = 0; k < K; ++k) { M[k] = Mn[k]; X[k] = Xn[k]; Y[k] = Yn[k]; }
            sumM = N::template add<EXACT>(sumM, M[KL]);                           // template.h:326,329
            sumX = N::template add<EXACT>(sumX, X[KL]);
        };
        // bottom row, pre-combined with the transition constants of the row below it
        auto emit_down = [&](T &Dout, T &Uout) {
            if (EXACT) {
                Dout = N::template add<true>(N::template add<true>(N::template mul<true>(M[K - 1], nMM),
                                                                     N::template mul<true>(X[K - 1], NGAPM)),
                                             N::template mul<true>(Y[K - 1], NGAPM));
                Uout = N::template add<true>(N::template mul<true>(M[K - 1], nMX), N::template mul<true>(X[K - 1], NXX));
            } else {
                Dout = N::template mad<false>(Y[K - 1], NGAPM, N::template mad<false>(X[K - 1], NGAPM, M[K - 1] * nMM));
                Uout = N::template mad<false>(X[K - 1], NXX, M[K - 1] * nMX);
            }
        };
        // flagged word: column 0 of a haplotype (SEP) or fill/drain/alignment (PAD)
        auto flagged = [&](uint32_t cw, uint32_t &next_tix) {
            if (!(cw & 0x40000000u)) {
                if (sl == last_lane && seg_valid && (uint32_t)cur < unit_nh) {
                    const T r = N::template add<EXACT>(sumM, sumX);              // template.h:339-343
                    if constexpr (sizeof(T) == 4) {
                        a.fres[my_out + cur] = r;
                    } else {
                        const uint32_t o = my_out + hap_info[cur].idx;
                        a.out[o] = log10(flush_denormal(r)) - a.log10_init_d;      // IntelPairHmm.cc:344
                        a.flag[o] = 1;
                    }
                }
                ++cur;
                sumM = sumX = (T)0;
                T initY = (T)0;
                next_tix = 32 * kTile / C;                                        // drain PAD tiles
                if ((uint32_t)cur < unit_nh) { const HapInfo<T> hi = hap_info[cur]; next_tix = hi.tix; initY = hi.initY; }
                // row 0 of the matrix seen from row 1: ((0*pMM + 0*pGAPM) + Y0*pGAPM) = Y0*pGAPM[row 1]
                Dbound = N::template mul<EXACT>(initY, PGAPM(0));
                if (lane0) Din[C - 1] = Dbound;        // becomes Dprev of the next step's first column
            }
#pragma unroll
            for (int k = 0; k < K; ++k) { M[k] = (T)0; X[k] = (T)0; Y[k] = (T)0; }
        };

#pragma unroll 1
        for (int t = 0; t < steps; ++t) {
            uint32_t next_tix = tix + 1;
            Tile wn = __ldg(tiles + next_tix);         // next step's tile, in flight during this step
            T Dout[C], Uout[C];
            if ((int)tw_or(w) < 0) {
                const uint32_t seq_tix = next_tix;
#pragma unroll
                for (int j = 0; j < C; ++j) {
                    const uint32_t cw = tw_get(w, j);
                    if ((int)cw < 0) flagged(cw, next_tix);
                    else column(cw, j == 0 ? Dprev : Din[j - 1], Uin[j]);
                    emit_down(Dout[j], Uout[j]);
                }
                if (next_tix != seq_tix) wn = __ldg(tiles + next_tix);
            } else {
#pragma unroll
                for (int j = 0; j < C; ++j) {
                    column(tw_get(w, j), j == 0 ? Dprev : Din[j - 1], Uin[j]);
                    emit_down(Dout[j], Uout[j]);
                }
            }
            Dprev = Din[C - 1];   // D of the last column of the PREVIOUS exchange: diagonal of next step's first column
#pragma unroll
            for (int j = 0; j < C; ++j) {
                const T dn = __shfl_up_sync(kFull, Dout[j], 1, W), un = __shfl_up_sync(kFull, Uout[j], 1, W);
                Din[j] = lane0 ? Dbound : dn;
                Uin[j] = lane0 ? (T)0 : un;
            }
            tix = next_tix;
            w = wn;
        }
#undef PXY
#undef PGAPM
#undef NXX
#undef NGAPM
    }
}

}  // namespace phmm
