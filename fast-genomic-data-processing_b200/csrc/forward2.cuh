// "Wide" PairHMM forward kernel for sm_100a: every value of the recurrence lives in a 64-bit register pair and every
// FMA-pipe instruction works on the whole pair. Two instantiations of the same code (policy P):
//
//   PkF32  two reads per lane in the two halves of the pair; the arithmetic is the sm_100 packed FP32 set FFMA2 /
//          FMUL2 / FADD2 (PTX fma.rn.f32x2, CUDA __ffma2_rn ...). The production FP32 pass.
//   PkF64  one read per lane, the pair is a double (DFMA / DMUL / DADD): the FP64 rescue pass. Same registers, same
//          shared-memory match-factor table, same step schedule.
//
// Why packed FP32: the scalar kernel (forward.cuh) is ISSUE-bound, not FMA-pipe-bound -- 8 FMA-pipe instructions per DP
// cell plus ~3 others (match-bit test, select, shuffles, loop) through a scheduler that issues one warp
// instruction per clock, so the pipe can not be busy more than ~72 % of the time. sm_100 executes a packed
// FP32 instruction on both halves of a register pair in one issue slot (two pipe cycles), and its operands may
// be a pair, a swapped pair or a broadcast scalar (R.F32 / UR.F32), measured with tools/ubench/fma2_rate.cu on
// B200: FFMA2 reg-pair x uniform + reg-pair 0.43-0.47 warp-inst/clk/SMSP = 27-30 of the 32 FP32 lane-ops/clk.
// With 8 packed instructions per PAIR of cells the FMA pipe needs 16 cycles per pair and the scheduler has 16
// issue slots for them: the kernel becomes pipe-bound. The FP64 pipe has the same shape (a warp-wide DFMA holds it
// for two cycles), so what keeps the packed kernel's issue slots free does the same for the rescue pass.
//
// What is packed (PkF32): lane l of a W-lane group owns rows l*K .. l*K+K-1 of TWO reads (A in the low half, B in the
// high half); both sweep the same haplotype stream, so the column word, the pipeline step and the shuffles are
// shared and the two DP problems never interact (packed instructions are two independent IEEE fma.rn.ftz:
// results are bit-identical to the scalar kernels, tests/test_gpu_parity.py compares the raw sums).
//
// Match/mismatch factor without ALU work: per lane and row the factor e = match ? 1-dist : dist/3 depends only
// on the haplotype base b in {A,C,T,G,N}. Each lane writes its 5 x K values to a per-warp table in
// shared memory once per work unit; a column then fetches its K values with K/2 LDS.128 from
// [b][row pair][lane] (16 B per lane, conflict-free) instead of K x 2 x (bit test + FSEL), which would put 4
// instructions per pair of cells on the half-rate ALU pipe. The haplotype words, SEP/PAD protocol, lane skew and
// the rescue queue are those of forward.cuh.
#pragma once
#include <type_traits>

#include "forward.cuh"

namespace phmm {

// ---- policies ---------------------------------------------------------------------------------------------------
struct PkF32 {
    using V = float2;                      // the 64-bit value: (read A, read B)
    using S = float;                       // one DP problem's scalar
    static constexpr int kReads = 2;       // reads per lane
    static __device__ __forceinline__ V pk(S lo, S hi) { return make_float2(lo, hi); }
    static __device__ __forceinline__ V bc(S s) { return make_float2(s, s); }
    static __device__ __forceinline__ S lo(V v) { return v.x; }
    static __device__ __forceinline__ S hi(V v) { return v.y; }
    static __device__ __forceinline__ V zero() { return make_float2(0.f, 0.f); }
    // -ftz=true applies (initNative enables FTZ, IntelPairHmm.cc:230)
    static __device__ __forceinline__ V fma(V a, V b, V c) { return __ffma2_rn(a, b, c); }
    static __device__ __forceinline__ V mul(V a, V b) { return __fmul2_rn(a, b); }
    static __device__ __forceinline__ V add(V a, V b) { return __fadd2_rn(a, b); }
    static __device__ __forceinline__ V sel(bool p, V a, V b) { return make_float2(p ? a.x : b.x, p ? a.y : b.y); }
    static __device__ __forceinline__ V shfl_up(V v, int W) {
        return make_float2(__shfl_up_sync(kFull, v.x, 1, W), __shfl_up_sync(kFull, v.y, 1, W));
    }
    static __device__ __forceinline__ V shfl_down(V v, int W) {
        return make_float2(__shfl_down_sync(kFull, v.x, 1, W), __shfl_down_sync(kFull, v.y, 1, W));
    }
    // Y coefficient of a column: pYY, or 0 under a flagged word. A bit trick instead of a select so that the uniform
    // coefficient itself stays in a uniform register: an FFMA2 with a UR operand reads 4 vector registers, with
    // three vector operands it is register-port bound (tools/ubench/fma2_rate.cu)
    static __device__ __forceinline__ V coef_y(const KernelArgs &a, uint32_t cw) {
        const float cy = __uint_as_float(a.u_xy_bits & ~(uint32_t)((int)cw >> 31));
        return make_float2(cy, cy);
    }
};

struct PkF64 {
    using V = double;
    using S = double;
    static constexpr int kReads = 1;
    static __device__ __forceinline__ V pk(S lo, S) { return lo; }
    static __device__ __forceinline__ V bc(S s) { return s; }
    static __device__ __forceinline__ S lo(V v) { return v; }
    static __device__ __forceinline__ S hi(V v) { return v; }
    static __device__ __forceinline__ V zero() { return 0.0; }
    static __device__ __forceinline__ V fma(V a, V b, V c) { return ::fma(a, b, c); }
    static __device__ __forceinline__ V mul(V a, V b) { return a * b; }
    static __device__ __forceinline__ V add(V a, V b) { return a + b; }
    static __device__ __forceinline__ V sel(bool p, V a, V b) { return p ? a : b; }
    static __device__ __forceinline__ V shfl_up(V v, int W) { return __shfl_up_sync(kFull, v, 1, W); }
    static __device__ __forceinline__ V shfl_down(V v, int W) { return __shfl_down_sync(kFull, v, 1, W); }
    static __device__ __forceinline__ V coef_y(const KernelArgs &a, uint32_t cw) {
        const int keep = ~((int)cw >> 31);
        return __hiloint2double(__double2hiint(a.u_xy_d) & keep, __double2loint(a.u_xy_d) & keep);
    }
};

template <typename V> struct alignas(16) QuadT { V x, y; };

// dynamic shared memory of one CTA: per warp the e-table [6][KP][32] x 16 B (KP = row pairs) + the haplotype table
__host__ __device__ constexpr int pk_row_pairs(int K) { return (K + 1) / 2; }
__host__ __device__ constexpr size_t pk_etab_bytes(int K) { return (size_t)6 * pk_row_pairs(K) * 32 * 16; }
// per-warp table of the unit's haplotypes for the SEP handler: first tile, 2^s/H, number of tiles, and (rescue units:
// only the haplotypes whose FP32 result underflowed are kept) the haplotype's position in the unit
template <typename S> struct PkHapInfo;
template <> struct PkHapInfo<float>  { uint32_t tix; float initY; int32_t tiles; };
template <> struct PkHapInfo<double> { uint32_t tix; int32_t tiles; double initY; uint32_t idx, pad_; };
// prefix-sharing kernel: tix / tiles describe the sweep (first tile after the shared prefix, tiles left); save_tix1 = index
// of the tile AFTER the last one the checkpoint covers (~0: none); slots = restore slot + 1 | save slot << 8 (kernels.cuh)
struct PkHapInfoPfx { uint32_t tix; float initY; int32_t tiles; uint32_t save_tix1, slots; };
template <typename P> __host__ __device__ constexpr size_t pk_smem_bytes_t(int K) {
    return (size_t)(kThreads / 32) * (pk_etab_bytes(K) + sizeof(PkHapInfo<typename P::S>) * kMaxUnitHaps);
}
__host__ __device__ constexpr size_t pk_smem_bytes_pfx(int K) {
    return (size_t)(kThreads / 32) * (pk_etab_bytes(K) + sizeof(PkHapInfoPfx) * kMaxUnitHaps);
}
__host__ __device__ constexpr size_t pk_smem_bytes(int K) { return pk_smem_bytes_t<PkF32>(K); }
__host__ __device__ constexpr size_t pk_smem_bytes_f64(int K) { return pk_smem_bytes_t<PkF64>(K); }

// per-row constants of one read row (ReadForPairHMM.cpp:72-80; stripeINITIALIZATION template.h:150-152)
template <typename S> struct RowConstT { S pMM, pMX, pMY, d1, d3; uint32_t rb; };
template <typename S>
__device__ __forceinline__ RowConstT<S> row_consts(const KernelArgs &a, const DevRead rr, int g, const S *__restrict__ ph,
                                                   const S *__restrict__ mmt) {
    RowConstT<S> c{(S)0, (S)0, (S)0, (S)0, (S)0, 5u};   // dead slot below the read: every constant 0, matches nothing
    if (g < rr.len) {
        const uint8_t *p = a.rd + rr.off + g;
        c.rb = base_index(p[0]);
        const int bq = p[a.plane] & 127, iq = p[2 * a.plane] & 127, dq = p[3 * a.plane] & 127;
        const int mx = max(iq, dq), mn = min(iq, dq);
        c.pMM = mmt[((mx * (mx + 1)) >> 1) + mn];
        c.pMX = ph[iq];
        c.pMY = ph[dq];
        const S dist = ph[bq];
        c.d1 = Num<S>::sub((S)1.0, dist);
        c.d3 = Num<S>::div(dist, (S)3.0);
    }
    return c;
}
// match: same base, or either side is N (template.h:38,48-51)
template <typename S> __device__ __forceinline__ S e_of(const RowConstT<S> &c, uint32_t b) {
    return (c.rb == 5u) ? (S)0 : ((c.rb == b || c.rb == 4u || b == 4u) ? c.d1 : c.d3);
}

#define PHMM_PRAGMA_(x) _Pragma(#x)
#define PHMM_UNROLL(n) PHMM_PRAGMA_(unroll n)
#ifndef PHMM_MINB_PK
#define PHMM_MINB_PK 2
#endif

template <typename P, int W, int K, int KL, bool PFX = false>
__device__ __forceinline__ void wide_forward(const KernelArgs &a) {
    static_assert(!PFX || sizeof(typename P::S) == 4, "prefix sharing is an FP32-pass feature (the rescue is per pair, IntelPairHmm.cc:700-703)");
    static_assert(W == 32 || W == 16 || W == 8, "lane-group width");
    static_assert(KL >= 0 && KL < K, "last-row slot out of range");
    using V = typename P::V;
    using S = typename P::S;
    using Quad = QuadT<V>;
    using HapRow = std::conditional_t<PFX, PkHapInfoPfx, PkHapInfo<S>>;
    constexpr bool F64 = sizeof(S) == 8;
    constexpr int C = kTile;
    using Tile = typename TileWord<C>::type;
    constexpr int KP = pk_row_pairs(K);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sl = lane & (W - 1);              // lane inside its group
    const int seg = lane / W;                   // group inside the warp: reads kReads*seg (low half) [and +1 (high half)]
    const bool lane0 = sl == 0;
    const S *__restrict__ ph = static_cast<const S *>(a.ph2pr);
    const S *__restrict__ mmt = static_cast<const S *>(a.mm);
    const Tile *__restrict__ tiles = reinterpret_cast<const Tile *>(a.hapw);
    const uint32_t n_units = a.n_units_dev ? *a.n_units_dev : a.n_units;
    extern __shared__ __align__(16) unsigned char pk_smem[];
    Quad *etab = reinterpret_cast<Quad *>(pk_smem + (size_t)warp * pk_etab_bytes(K)) + lane;
    constexpr size_t kEtabAll = (size_t)(kThreads / 32) * pk_etab_bytes(K);
    HapRow *hap_info = reinterpret_cast<HapRow *>(pk_smem + kEtabAll) + warp * kMaxUnitHaps;
    const V cXY = P::bc(Num<S>::uxy(a)), cG = P::bc(Num<S>::ugapm(a));   // broadcast uniform operands

    for (;;) {
        uint32_t u = 0;
        if (lane == 0) u = atomicAdd(a.queue, 1u);
        u = __shfl_sync(kFull, u, 0);
        if (u >= n_units) break;
        const DevUnit unit = a.units[u];
        uint32_t unit_nh = unit.shape & 0xffu;
        uint32_t unit_nreads = (unit.shape >> 8) & 0xffu, rstep = 1u;
        const uint32_t stride = unit.shape >> 16;
        // a rescue unit of the W = 16 FP64 instance holds two reads of one region that need not be neighbours: the field is
        // the distance between them (rescue_scan_kernel, kernels.cu)
        if constexpr (F64 && W == 16) if (a.rescue_filter) { rstep = unit_nreads; unit_nreads = 2u; }
        // a partial unit leaves halves without a read: they recompute read 0 and emit nothing
        const uint32_t ra = (uint32_t)P::kReads * (uint32_t)seg, rb = ra + (uint32_t)P::kReads - 1u;
        const bool valid_a = ra < unit_nreads, valid_b = rb < unit_nreads;
        const uint32_t read_a = unit.read + (valid_a ? ra * rstep : 0u), read_b = unit.read + (valid_b ? rb * rstep : 0u);
        const uint32_t out_a = unit.out + ra * rstep * stride, out_b = unit.out + rb * rstep * stride;
        const DevRead rra = a.reads[read_a], rrb = a.reads[read_b];
        const int last_a = (rra.len - 1) / K, last_b = (rrb.len - 1) / K;   // lane that owns the last row (slot KL)

        // ---- per-slot constants, packed (A, B); the match factor goes to the shared table
        V pMM[K], pMX[K], pMY[K];
        __syncwarp();
#pragma unroll
        for (int k = 0; k < K; k += 2) {
            RowConstT<S> c[2][2];
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int g = sl * K + k + q;
                c[q][0] = (k + q < K) ? row_consts<S>(a, rra, g, ph, mmt) : RowConstT<S>{(S)0, (S)0, (S)0, (S)0, (S)0, 5u};
                if constexpr (P::kReads == 2)
                    c[q][1] = (k + q < K) ? row_consts<S>(a, rrb, g, ph, mmt) : RowConstT<S>{(S)0, (S)0, (S)0, (S)0, (S)0, 5u};
                else
                    c[q][1] = c[q][0];
                if (k + q < K) {
                    pMM[k + q] = P::pk(c[q][0].pMM, c[q][1].pMM);
                    pMX[k + q] = P::pk(c[q][0].pMX, c[q][1].pMX);
                    pMY[k + q] = P::pk(c[q][0].pMY, c[q][1].pMY);
                }
            }
#pragma unroll
            for (uint32_t b = 0; b < 5; ++b)
                etab[(b * KP + k / 2) * 32] = Quad{P::pk(e_of(c[0][0], b), e_of(c[0][1], b)),
                                                   P::pk(e_of(c[1][0], b), e_of(c[1][1], b))};
            etab[(5 * KP + k / 2) * 32] = Quad{P::zero(), P::zero()};     // row of the flagged words
        }
        // transition constants of the row below this lane's last slot (= slot 0 of lane+1; 0 past the group)
        V nMM = P::shfl_down(pMM[0], W), nMX = P::shfl_down(pMX[0], W);
        if (sl == W - 1) { nMM = P::zero(); nMX = P::zero(); }

        // haplotype table of the unit for the SEP handler and the step schedule
        if constexpr (PFX) {
            if ((uint32_t)lane < unit_nh) {
                const DevHap hp = a.haps[a.hap_list[unit.hap_list + lane]];
                const uint2 pl = a.hap_plan[unit.hap_list + lane];
                const uint32_t ts = pl.x & 0xffffu, tv = pl.x >> 16;
                hap_info[lane] = HapRow{hp.off / C + ts, hp.len ? __fdiv_rn(Num<float>::init_const(), (float)hp.len) : 0.f,
                                        (int32_t)(hap_tiles(hp.len) - ts), tv ? hp.off / C + tv : 0xffffffffu, pl.y};
            }
        } else if constexpr (!F64) {
            if ((uint32_t)lane < unit_nh) {                // nh <= kMaxUnitHaps == 32 (host guarantees)
                const DevHap hp = a.haps[a.hap_list[unit.hap_list + lane]];
                hap_info[lane] = HapRow{hp.off / C, hp.len ? __fdiv_rn(Num<float>::init_const(), (float)hp.len) : 0.f,   // template.h:115
                                        (int32_t)hap_tiles(hp.len)};
            }
        } else {
            // a rescue unit (rescue_scan_kernel) names the read's whole haplotype run: sweep only those whose FP32 sum
            // fell below the reference's threshold (IntelPairHmm.cc:342), back to back -- one pipeline fill per read
            bool keep = (uint32_t)lane < unit_nh;
            if (a.rescue_filter && keep) {
                keep = a.fres[unit.out + lane] < 1e-28f;
                if constexpr (W == 16) keep |= a.fres[unit.out + rstep * stride + lane] < 1e-28f;   // either read needs it
            }
            const uint32_t km = __ballot_sync(kFull, keep);
            unit_nh = (uint32_t)__popc(km);
            if (keep) {
                const DevHap hp = a.haps[a.hap_list[unit.hap_list + lane]];
                hap_info[__popc(km & ((1u << lane) - 1u))] =
                    HapRow{hp.off / C, (int32_t)hap_tiles(hp.len), hp.len ? __ddiv_rn(Num<double>::init_const(), (double)hp.len) : 0.0,
                           (uint32_t)lane, 0u};
            }
        }
        __syncwarp();

        V M[K], X[K], Y[K];
#pragma unroll
        for (int k = 0; k < K; ++k) { M[k] = X[k] = Y[k] = P::zero(); }
        V Dprev = P::zero(), sumM = P::zero(), sumX = P::zero(), Dbound = P::zero();
        V Din[C], Uin[C];
#pragma unroll
        for (int j = 0; j < C; ++j) { Din[j] = Uin[j] = P::zero(); }

        // prefix sharing: checkpoints of this lane (local memory: the slot index is dynamic), the tile index at which the
        // current sweep writes one, and the slots of the current sweep
        V ck[PFX ? kPfxSlots : 1][3 * K + 3];
        uint32_t save_at = 0xffffffffu, pfx_slots = 0u;
        int cur = -1;
        uint32_t tix1 = 32 - sl;                       // lane l sees l PAD tiles, then [PAD.., SEP] (tile 31)
        const Tile *tp = tiles + tix1;                 // points at w1
        Tile w = tp[-1], w1 = tp[0];                  // current and next tile; tix1 = index of w1

        // one DP column for this lane's K rows (computeMXY, template.h:177-192)
        auto column = [&](uint32_t cw, V Dd, V Uu, auto dirty_tag) {
            const Quad *ep = etab + ((cw >> 25) & 7u) * (KP * 32);   // bits 25..27 of a word = base index (5 if flagged)
            V e[2 * KP];
#pragma unroll
            for (int j = 0; j < KP; ++j) { const Quad q = ep[j * 32]; e[2 * j] = q.x; e[2 * j + 1] = q.y; }
            // a flagged word resets the lane's state arithmetically instead of in a divergent branch: its match
            // factors are 0 (M = 0, hence X = 0) and its Y coefficient is 0; two flagged words in a row (every SEP is
            // followed by a PAD, kernels.cuh) leave M = X = Y = 0, i.e. column 0 of the next haplotype's matrix
            // In a CLEAN step the only flagged words a lane can meet are the PADs that fill a haplotype's last code tile
            // (lead-in, drain and every SEP tile fall into dirty steps). They need no zero Y coefficient: their match
            // factors are 0, so M = X = 0 and nothing reads Y until the SEP tile -- a dirty step -- resets it with its
            // own flagged columns. So the clean instance takes the Y coefficient as the uniform-register operand it is
            // (FFMA2 pair x UR + pair); per lane it would be a third vector operand (5 register reads per FFMA2).
            const V cYY = decltype(dirty_tag)::value ? P::coef_y(a, cw) : cXY;
            V Mn[K], Xn[K], Yn[K];
            Mn[0] = P::mul(Dd, e[0]);
            Xn[0] = Uu;
            Yn[0] = P::fma(Y[0], cYY, P::mul(M[0], pMY[0]));
#pragma unroll
            for (int k = 1; k < K; ++k) {
                const V s = P::fma(Y[k - 1], cG, P::fma(X[k - 1], cG, P::mul(M[k - 1], pMM[k])));
                Mn[k] = P::mul(s, e[k]);
                Xn[k] = P::fma(Xn[k - 1], cXY, P::mul(Mn[k - 1], pMX[k]));
                Yn[k] = P::fma(Y[k], cYY, P::mul(M[k], pMY[k]));
            }
#pragma unroll
            for (int k = 0; k < K; ++k) { M[k] = Mn[k]; X[k] = Xn[k]; Y[k] = Yn[k]; }
            sumM = P::add(sumM, M[KL]);                                           // template.h:326,329
            sumX = P::add(sumX, X[KL]);
        };
        // bottom row, pre-combined with the transition constants of the row below it
        auto emit_down = [&](V &Dout, V &Uout) {
            Dout = P::fma(Y[K - 1], cG, P::fma(X[K - 1], cG, P::mul(M[K - 1], nMM)));
            Uout = P::fma(X[K - 1], cXY, P::mul(M[K - 1], nMX));
        };
        // FP32: the raw scaled sum; finalize_kernel turns it into log10 or queues the read for the FP64 pass
        auto emit_result = [&](S r, uint32_t my_out) {
            if constexpr (!F64) {
                a.fres[my_out + cur] = r;
            } else {
                const uint32_t o = my_out + hap_info[cur].idx;
                // two reads per warp sweep the union of what either needs: a pair whose FP32 result stands keeps it
                if (W == 16 && a.rescue_filter && !(a.fres[o] < 1e-28f)) return;
                a.out[o] = log10(flush_denormal(r)) - a.log10_init_d;              // IntelPairHmm.cc:344
                a.flag[o] = 1;
            }
        };
        // one pipeline step = the C columns of this lane's current tile. DIRTY steps are those in which some lane of the
        // warp may hold a SEP tile (the W steps after a haplotype boundary enters lane 0); all others run the branch-free
        // CLEAN instance, whose body is a single basic block.
        auto step = [&](auto dirty_tag, auto save_tag) {
            constexpr bool DIRTY = decltype(dirty_tag)::value;
            constexpr bool SAVE = PFX && (DIRTY || decltype(save_tag)::value);   // may some lane write a checkpoint in this step?
            bool sep = false;
            if (DIRTY) {
#pragma unroll
                for (int j = 0; j < C; ++j) sep |= (tw_get(w, j) & 0xC0000000u) == 0x80000000u;
                // SEP word = column 0 of the next haplotype: emit the finished one, switch. Done before the columns of the
                // tile: they are all flagged (a SEP sits in slot 0 of its own tile, or closes the all-PAD lead-in), add 0 to
                // the sums and ignore the boundary values, so the order does not matter; PAD words need nothing at all.
                // Exactly one lane of the group holds the SEP in each dirty step, so a divergent handler would run every
                // time with one lane active: the switch is written as selects instead (+1.4 % on the FP32 pass), only the
                // result store of the lane that owns a read's last row is a branch
                if (sep && (uint32_t)cur < unit_nh) {
                    const V r = P::add(sumM, sumX);                                // template.h:339-343
                    if (sl == last_a && valid_a) emit_result(P::lo(r), out_a);
                    if constexpr (P::kReads == 2)
                        if (sl == last_b && valid_b) emit_result(P::hi(r), out_b);
                }
                cur += sep ? 1 : 0;
                sumM = P::sel(sep, P::zero(), sumM);
                sumX = P::sel(sep, P::zero(), sumX);
                const bool has = (uint32_t)cur < unit_nh;
                const HapRow hi = hap_info[has ? cur : 0];
                const uint32_t target = has ? hi.tix : 32u;                        // 32 = drain PAD tiles
                const S initY = has ? hi.initY : (S)0;
                if constexpr (PFX) {
                    save_at = sep ? (has ? hi.save_tix1 : 0xffffffffu) : save_at;
                    pfx_slots = sep ? (has ? hi.slots : 0u) : pfx_slots;
                }
                // row 0 of the matrix seen from row 1: ((0*pMM + 0*pGAPM) + Y0*pGAPM) = Y0*pGAPM[row 1]
                Dbound = P::sel(sep, P::bc(initY * Num<S>::ugapm(a)), Dbound);
                Din[C - 1] = P::sel(sep && lane0, Dbound, Din[C - 1]);   // becomes Dprev of the next step's first column
                // a unit's haplotypes lie back to back in hapw, so the sequential guess is right except after the
                // lead-in, before the drain and between the haplotypes of a rescue unit
                if (sep && target != tix1) { tix1 = target; tp = tiles + tix1; w1 = __ldg(tp); }
            }
            V Dout[C], Uout[C];
#pragma unroll
            for (int j = 0; j < C; ++j) {
                column(tw_get(w, j), j == 0 ? Dprev : Din[j - 1], Uin[j], dirty_tag);
                emit_down(Dout[j], Uout[j]);
            }
            Dprev = Din[C - 1];   // D of the last column of the PREVIOUS exchange: diagonal of next step's first column
#pragma unroll
            for (int j = 0; j < C; ++j) {
                const V dn = P::shfl_up(Dout[j], W), un = P::shfl_up(Uout[j], W);
                Din[j] = P::sel(lane0, Dbound, dn);
                Uin[j] = P::sel(lane0, P::zero(), un);
            }
            if constexpr (PFX) {
                // The SEP tile's flagged columns have wiped the lane's state: a sweep that starts behind a shared prefix takes
                // it from the checkpoint instead. M/X/Y = the lane's rows at the prefix' last column, the sums cover the
                // prefix' columns (same order of additions as a full sweep: the result is bit-identical), Dprev = the
                // lane above's bottom row at that column; the inputs of the first own tile (Din/Uin) arrived in the exchange
                // just above -- the lane above restored one step earlier and has swept that tile in this step
                if (DIRTY && sep && (pfx_slots & 0xffu)) {
                    const V *c = ck[(pfx_slots & 0xffu) - 1u];
#pragma unroll
                    for (int k = 0; k < K; ++k) { M[k] = c[k]; X[k] = c[K + k]; Y[k] = c[2 * K + k]; }
                    sumM = c[3 * K]; sumX = c[3 * K + 1]; Dprev = c[3 * K + 2];
                }
                // the lane has just finished the last tile another haplotype of the unit shares with this one
                if (SAVE && tix1 == save_at) {
                    V *c = ck[(pfx_slots >> 8) & 0xffu];
#pragma unroll
                    for (int k = 0; k < K; ++k) { c[k] = M[k]; c[K + k] = X[k]; c[2 * K + k] = Y[k]; }
                    c[3 * K] = sumM; c[3 * K + 1] = sumX; c[3 * K + 2] = Dprev;
                }
            }
            // rotate the tile queue at the END of the step: w is dead by now, so the load can land in a free register and
            // nothing waits for it here; it is consumed two steps later (a full step of latency hiding)
            w = w1;
            ++tix1; ++tp;
            w1 = __ldg(tp);
        };
        // haplotype boundary h enters lane 0 at the start of segment h and leaves lane W-1 W steps later; segment nh
        // is the drain
        for (uint32_t h = 0; h <= unit_nh; ++h) {
            const int T = h < unit_nh ? hap_info[h].tiles : W;
            const int nd = min(W, T);
PHMM_UNROLL(1)
            for (int i = 0; i < nd; ++i) step(std::true_type{}, std::false_type{});
            if constexpr (PFX) {
                // lane l finishes the checkpoint's last tile in step d + l of the segment (d = tiles between the sweep's
                // start and the checkpoint): only the steps [d, d + W) carry the save test, the rest run the plain clean
                // instance. What sticks out beyond the segment falls into the dirty steps of the next ones (d < T),
                // which always test
                int a_ = T, b_ = T;
                if (h < unit_nh && hap_info[h].save_tix1 != 0xffffffffu) {
                    const int d = (int)(hap_info[h].save_tix1 - hap_info[h].tix);
                    a_ = min(max(d, nd), T); b_ = min(max(d + W, nd), T);
                }
PHMM_UNROLL(1)
                for (int i = nd; i < a_; ++i) step(std::false_type{}, std::false_type{});
PHMM_UNROLL(1)
                for (int i = a_; i < b_; ++i) step(std::false_type{}, std::true_type{});
PHMM_UNROLL(1)
                for (int i = b_; i < T; ++i) step(std::false_type{}, std::false_type{});
            } else {
PHMM_UNROLL(1)
                for (int i = nd; i < T; ++i) step(std::false_type{}, std::false_type{});
            }
        }
    }
}

// FP32 pass: two reads per lane, packed arithmetic
template <int W, int K, int KL>
__global__ void __launch_bounds__(kThreads, PHMM_MINB_PK) forward2_kernel(const KernelArgs a) {
    wide_forward<PkF32, W, K, KL>(a);
}
// FP32 pass with prefix sharing between the same-length haplotypes of a unit (kernels.cuh, Variant::F32_PFX)
template <int W, int K, int KL>
__global__ void __launch_bounds__(kThreads, PHMM_MINB_PK) forward2p_kernel(const KernelArgs a) {
    wide_forward<PkF32, W, K, KL, true>(a);
}
// FP64 rescue pass: one read per lane, same structure
template <int W, int K, int KL>
__global__ void __launch_bounds__(kThreads, PHMM_MINB_PK) forward2d_kernel(const KernelArgs a) {
    wide_forward<PkF64, W, K, KL>(a);
}

}  // namespace phmm
