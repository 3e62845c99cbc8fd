// PairHMM forward kernels for sm_100a (no tensor cores: the M/X/Y recurrence is a wavefront).
//
// Work decomposition: one WARP sweeps one work unit = one read against a run of haplotypes.
//   * lane l owns K consecutive rows of the DP matrix ("slots" l*K .. l*K+K-1); the read is aligned to
//     the END of the 32*K slots so that its last row is always slot K-1 of lane 31, and the unused
//     slots above it emulate matrix row 0 (M=X=0, Y=2^s/H), which makes the top boundary fall out of
//     the ordinary recurrence (no special-casing of lane 0).
//   * lanes are skewed by one column: at step t lane l works on stream position t-l. The only
//     cross-lane traffic per step is two shuffles: the bottom row of lane l, pre-combined with the
//     transition constants of the row below it (D = diagonal term, U = up term of the X update).
//   * the haplotypes of a unit are streamed back to back through the same pipeline, separated by one
//     SEP word (= column 0 of the next haplotype), so pipeline fill/drain is paid once per unit.
// Numerics follow avx-pairhmm-template.h:177-192 / SURVEY.md Appendix A (association order kept).
#include "kernels.cuh"

namespace phmm {
namespace {

constexpr unsigned kFull = 0xffffffffu;

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
// code word of a haplotype base: b * group_bits(K) for every bucket, each in its own 5-bit field
__device__ __forceinline__ uint32_t code_word(uint32_t b) {
    return (5u * b) | ((4u * b) << 5) | ((6u * b) << 10) | ((3u * b) << 15) | ((2u * b) << 20) | (b << 25);
}

// ---------------------------------------------------------------------------------------------
// haplotype bytes -> code words + SEP (+ PADs up to the tile boundary); block per haplotype; block 0 also
// writes the fill/drain words.
__global__ void prep_haps_kernel(const uint8_t *__restrict__ hap_bytes, const uint64_t *__restrict__ byte_off,
                                 const DevHap *__restrict__ haps, uint32_t n_haps, uint32_t *__restrict__ hapw) {
    if (blockIdx.x == 0)
        for (int i = threadIdx.x; i < kLeadWords; i += blockDim.x) hapw[i] = (i == 32 * kTile - 1) ? kWordSep : kWordPad;
    for (uint32_t h = blockIdx.x; h < n_haps; h += gridDim.x) {
        const DevHap d = haps[h];
        const uint8_t *src = hap_bytes + byte_off[h];
        uint32_t *dst = hapw + d.off;
        const int nw = (d.len / kTile + 1) * kTile;
        for (int c = threadIdx.x; c < nw; c += blockDim.x)
            dst[c] = (c < d.len) ? code_word(base_index(src[c])) : (c == d.len ? kWordSep : kWordPad);
    }
}

// ---------------------------------------------------------------------------------------------
template <int C> struct TileWord;
template <> struct TileWord<1> { using type = uint32_t; };
template <> struct TileWord<2> { using type = uint2; };
template <> struct TileWord<4> { using type = uint4; };
__device__ __forceinline__ uint32_t tw_get(const uint32_t &t, int) { return t; }
__device__ __forceinline__ uint32_t tw_get(const uint2 &t, int j) { return j == 0 ? t.x : t.y; }
__device__ __forceinline__ uint32_t tw_get(const uint4 &t, int j) { return j == 0 ? t.x : j == 1 ? t.y : j == 2 ? t.z : t.w; }
__device__ __forceinline__ uint32_t tw_or(const uint32_t &t) { return t; }  // (kTile=1)
__device__ __forceinline__ uint32_t tw_or(const uint2 &t) { return t.x | t.y; }
__device__ __forceinline__ uint32_t tw_or(const uint4 &t) { return t.x | t.y | t.z | t.w; }

template <typename T> __device__ __forceinline__ T shfl_from(T v, int src) { return __shfl_sync(kFull, v, src); }

template <typename T, int K, bool EXACT>
__global__ void __launch_bounds__(kThreads) forward_kernel(const KernelArgs a) {
    using N = Num<T>;
    constexpr int C = kTile;
    using Tile = typename TileWord<C>::type;
    constexpr int GB = group_bits(K);           // rows per match-mask register
    constexpr int NG = (K + GB - 1) / GB;
    constexpr int FIELD = code_field(K);
    const int lane = threadIdx.x & 31;
    const int src_lane = (lane + 31) & 31;            // lane 0 takes lane 31's (all-zero) output
    const T *__restrict__ ph = static_cast<const T *>(a.ph2pr);
    const T *__restrict__ mmt = static_cast<const T *>(a.mm);
    const Tile *__restrict__ tiles = reinterpret_cast<const Tile *>(a.hapw);
    const uint32_t n_units = a.n_units_dev ? *a.n_units_dev : a.n_units;

    for (;;) {
        uint32_t u = 0;
        if (lane == 0) u = atomicAdd(a.queue, 1u);
        u = __shfl_sync(kFull, u, 0);
        if (u >= n_units) break;
        const DevUnit unit = a.units[u];
        const DevRead rr = a.reads[unit.read];
        const int pad = 32 * K - rr.len;              // >= 1 emulator slots above the read

        // ---- per-slot constants (ReadForPairHMM.cpp:72-80; stripeINITIALIZATION template.h:150-152)
        T pMM[K], pGAPM[K], pMX[K], pMY[K], pXY[K], d1[K], d3[K];
        uint32_t mask[NG];                            // per haplotype base b: bit GB*b+i = row i of the group matches
        uint32_t emu = 0;
#pragma unroll
        for (int i = 0; i < NG; ++i) mask[i] = 0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int g = lane * K + k - pad;         // 0-based read row, <0 = emulator slot
            if (g >= 0) {
                const uint8_t *p = a.rd + rr.off + g;
                const uint32_t rb = base_index(p[0]);
                const int bq = p[a.plane] & 127, iq = p[2 * a.plane] & 127, dq = p[3 * a.plane] & 127,
                          cq = p[4 * a.plane] & 127;
                const int mx = max(iq, dq), mn = min(iq, dq);
                pMM[k] = mmt[((mx * (mx + 1)) >> 1) + mn];
                pMX[k] = ph[iq];
                pMY[k] = ph[dq];
                pXY[k] = ph[cq];
                pGAPM[k] = N::sub((T)1.0, ph[cq]);
                const T dist = ph[bq];
                d1[k] = N::sub((T)1.0, dist);
                d3[k] = N::div(dist, (T)3.0);
                // match: same base, or either side is N (template.h:38,48-51)
                const uint32_t one = 1u << (k % GB);
                const uint32_t all = one | (one << GB) | (one << (2 * GB)) | (one << (3 * GB)) | (one << (4 * GB));
                mask[k / GB] |= (rb == 4u) ? all : ((one << (GB * rb)) | (one << (4 * GB)));
            } else {                                   // row-0 emulator: M=X=0, Y carried unchanged
                pMM[k] = pGAPM[k] = pMX[k] = pMY[k] = (T)0; pXY[k] = (T)1; d1[k] = d3[k] = (T)0;
                emu |= 1u << k;
            }
        }
        // transition constants of the row below this lane's last slot (= slot 0 of lane+1)
        T nMM = __shfl_down_sync(kFull, pMM[0], 1), nGAPM = __shfl_down_sync(kFull, pGAPM[0], 1),
          nMX = __shfl_down_sync(kFull, pMX[0], 1), nXX = __shfl_down_sync(kFull, pXY[0], 1);
        if (lane == 31) { nMM = nGAPM = nMX = nXX = (T)0; }

        // pipeline steps: 32 (fill) + sum over haplotypes of their tile count
        int steps = 0;
        for (uint32_t i = lane; i < unit.nh; i += 32) steps += a.haps[a.hap_list[unit.hap_list + i]].len / C + 1;
#pragma unroll
        for (int o = 16; o; o >>= 1) steps += __shfl_xor_sync(kFull, steps, o);
        steps += 32;

        T M[K], X[K], Y[K];
#pragma unroll
        for (int k = 0; k < K; ++k) { M[k] = X[k] = Y[k] = (T)0; }
        T Dprev = (T)0, sumM = (T)0, sumX = (T)0, initY = (T)0;
        T Din[C], Uin[C];
#pragma unroll
        for (int j = 0; j < C; ++j) { Din[j] = Uin[j] = (T)0; }
        int cur = -1;
        uint32_t tix = 31 - lane;                      // tile index: lane l sees l PAD tiles, then [PAD..,SEP]
        Tile w = tiles[tix];

        // one DP column for this lane's K rows (computeMXY, template.h:177-192): Dd = diagonal term and
        // Uu = X of the row above, both as delivered by the lane above
        auto column = [&](uint32_t cw, T Dd, T Uu) {
            uint32_t mbits[NG];
#pragma unroll
            for (int g = 0; g < NG; ++g) mbits[g] = __funnelshift_r(mask[g], 0u, FIELD ? (cw >> FIELD) : cw);
            T Mn[K], Xn[K], Yn[K];
            {
                const T e = (mbits[0] & 1u) ? d1[0] : d3[0];
                Mn[0] = N::template mul<EXACT>(Dd, e);
                Xn[0] = Uu;
                Yn[0] = N::template mad<EXACT>(M[0], pMY[0], N::template mul<EXACT>(Y[0], pXY[0]));
            }
#pragma unroll
            for (int k = 1; k < K; ++k) {
                const T e = ((mbits[k / GB] >> (k % GB)) & 1u) ? d1[k] : d3[k];
                T s;
                if (EXACT) {
                    s = N::template add<true>(N::template add<true>(N::template mul<true>(M[k - 1], pMM[k]),
                                                                      N::template mul<true>(X[k - 1], pGAPM[k])),
                                              N::template mul<true>(Y[k - 1], pGAPM[k]));
                } else {
                    s = N::template mad<false>(Y[k - 1], pGAPM[k],
                                               N::template mad<false>(X[k - 1], pGAPM[k], M[k - 1] * pMM[k]));
                }
                Mn[k] = N::template mul<EXACT>(s, e);
                Xn[k] = N::template mad<EXACT>(Mn[k - 1], pMX[k], N::template mul<EXACT>(Xn[k - 1], pXY[k]));
                Yn[k] = N::template mad<EXACT>(M[k], pMY[k], N::template mul<EXACT>(Y[k], pXY[k]));
            }
#pragma unroll
            for (int k = 0; k < K; ++k) { M[k] = Mn[k]; X[k] = Xn[k]; Y[k] = Yn[k]; }
            sumM = N::template add<EXACT>(sumM, M[K - 1]);                        // template.h:326,329
            sumX = N::template add<EXACT>(sumX, X[K - 1]);
        };
        // bottom row, pre-combined with the transition constants of the row below it
        auto emit_down = [&](T &Dout, T &Uout) {
            if (EXACT) {
                Dout = N::template add<true>(N::template add<true>(N::template mul<true>(M[K - 1], nMM),
                                                                     N::template mul<true>(X[K - 1], nGAPM)),
                                             N::template mul<true>(Y[K - 1], nGAPM));
                Uout = N::template add<true>(N::template mul<true>(M[K - 1], nMX), N::template mul<true>(X[K - 1], nXX));
            } else {
                Dout = N::template mad<false>(Y[K - 1], nGAPM, N::template mad<false>(X[K - 1], nGAPM, M[K - 1] * nMM));
                Uout = N::template mad<false>(M[K - 1], nMX, X[K - 1] * nXX);
            }
        };
        // flagged word: column 0 of a haplotype (SEP) or fill/drain/alignment (PAD)
        auto flagged = [&](uint32_t cw, uint32_t &next_tix) {
            if (!(cw & 0x40000000u)) {
                if (lane == 31 && (uint32_t)cur < unit.nh) {
                    const T r = N::template add<EXACT>(sumM, sumX);              // template.h:339-343
                    if constexpr (sizeof(T) == 4) {
                        a.fres[unit.out + cur] = r;
                        if (r < 1e-28f) {                                          // IntelPairHmm.cc:342
                            const uint32_t slot = atomicAdd(a.rescue_count, 1u);
                            a.rescue_units[slot] = DevUnit{unit.read, unit.hap_list + (uint32_t)cur, 1u,
                                                           unit.out + (uint32_t)cur};
                        }
                    } else {
                        a.out[unit.out + cur] = log10(r) - a.log10_init_d;        // IntelPairHmm.cc:344
                        a.flag[unit.out + cur] = 1;
                    }
                }
                ++cur;
                sumM = sumX = (T)0;
                if ((uint32_t)cur < unit.nh) {
                    const DevHap h = a.haps[a.hap_list[unit.hap_list + cur]];
                    next_tix = h.off / C;
                    initY = N::div(N::init_const(), (T)h.len);                   // template.h:115
                } else {
                    next_tix = 32;                                                // drain PAD tiles
                    initY = (T)0;
                }
            }
#pragma unroll
            for (int k = 0; k < K; ++k) { M[k] = (T)0; X[k] = (T)0; Y[k] = ((emu >> k) & 1u) ? initY : (T)0; }
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
            for (int j = 0; j < C; ++j) { Din[j] = shfl_from(Dout[j], src_lane); Uin[j] = shfl_from(Uout[j], src_lane); }
            tix = next_tix;
            w = wn;
        }
    }
}

// FP32 results -> log10 (IntelPairHmm.cc:348): the subtraction happens in float.
__global__ void finalize_kernel(const FinalizeArgs a) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const float f = a.fres[i];
    if (f < 1e-28f) return;                           // the FP64 kernel owns this entry
    const float l = (float)log10((double)f);          // correctly rounded float log10
    a.out[i] = (double)__fsub_rn(l, a.log10_init_f);
    a.flag[i] = 0;
}

template <typename T, bool EXACT>
cudaError_t launch_forward(int K, const KernelArgs &a, int grid, cudaStream_t st) {
    switch (K) {
#define PHMM_CASE(KK) case KK: forward_kernel<T, KK, EXACT><<<grid, kThreads, 0, st>>>(a); break;
        PHMM_CASE(1) PHMM_CASE(2) PHMM_CASE(3) PHMM_CASE(4) PHMM_CASE(5) PHMM_CASE(6) PHMM_CASE(7) PHMM_CASE(8)
#undef PHMM_CASE
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

template <typename T, bool EXACT>
int blocks_per_sm(int K) {
    int n = 0;
    switch (K) {
#define PHMM_CASE(KK) case KK: cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, forward_kernel<T, KK, EXACT>, kThreads, 0); break;
        PHMM_CASE(1) PHMM_CASE(2) PHMM_CASE(3) PHMM_CASE(4) PHMM_CASE(5) PHMM_CASE(6) PHMM_CASE(7) PHMM_CASE(8)
#undef PHMM_CASE
        default: break;
    }
    return n;
}

}  // namespace

cudaError_t launch_prep_haps(const uint8_t *hap_bytes, const uint64_t *byte_off, const DevHap *haps,
                             uint32_t n_haps, uint32_t *hapw, cudaStream_t st) {
    const int grid = n_haps == 0 ? 1 : (int)(n_haps < 65535u ? n_haps : 65535u);
    prep_haps_kernel<<<grid, 128, 0, st>>>(hap_bytes, byte_off, haps, n_haps, hapw);
    return cudaGetLastError();
}

cudaError_t launch_forward_float(int K, bool exact, const KernelArgs &a, int grid, cudaStream_t st) {
    return exact ? launch_forward<float, true>(K, a, grid, st) : launch_forward<float, false>(K, a, grid, st);
}

cudaError_t launch_forward_double(int K, const KernelArgs &a, int grid, cudaStream_t st) {
    return launch_forward<double, false>(K, a, grid, st);
}

cudaError_t launch_finalize(const FinalizeArgs &a, cudaStream_t st) {
    if (a.n == 0) return cudaSuccess;
    finalize_kernel<<<(a.n + 255) / 256, 256, 0, st>>>(a);
    return cudaGetLastError();
}

int forward_blocks_per_sm(int K, bool is_double, bool exact) {
    if (is_double) return blocks_per_sm<double, false>(K);
    return exact ? blocks_per_sm<float, true>(K) : blocks_per_sm<float, false>(K);
}

}  // namespace phmm
