// Packed FP32 PairHMM forward kernel for sm_100a: two reads per lane in the two halves of 64-bit register
// pairs, every FMA-pipe instruction of the recurrence issued as FFMA2 / FMUL2 / FADD2 (PTX fma.rn.f32x2 ...).
//
// Why: the scalar kernel (forward.cuh) is ISSUE-bound, not FMA-pipe-bound -- 8 FMA-pipe instructions per DP
// cell plus ~3 others (match-bit test, select, shuffles, loop) through a scheduler that issues one warp
// instruction per clock, so the pipe can not be busy more than ~72 % of the time. sm_100 executes a packed
// FP32 instruction on both halves of a register pair in one issue slot (two pipe cycles), and its operands may
// be a pair, a swapped pair or a broadcast scalar (R.F32 / UR.F32), measured with tools/ubench/fma2_rate.cu on
// B200: FFMA2 reg-pair x uniform + reg-pair 0.43-0.47 warp-inst/clk/SMSP = 27-30 of the 32 FP32 lane-ops/clk.
// With 8 packed instructions per PAIR of cells the FMA pipe needs 16 cycles per pair and the scheduler has 16
// issue slots for them: the kernel becomes pipe-bound.
//
// What is packed: lane l of a W-lane group owns rows l*K .. l*K+K-1 of TWO reads (A in the low half, B in the
// high half); both sweep the same haplotype stream, so the column word, the pipeline step and the shuffles are
// shared and the two DP problems never interact (packed instructions are two independent IEEE fma.rn.ftz:
// results are bit-identical to the scalar kernels, tests/test_gpu_parity.py compares the raw sums).
//
// Match/mismatch factor without ALU work: per lane and row the factor e = match ? 1-dist : dist/3 depends only
// on the haplotype base b in {A,C,T,G,N}. Each lane writes its 5 x K pairs (e_A, e_B) to a per-warp table in
// shared memory once per work unit; a column then fetches its K pairs with K/2 LDS.128 from
// [b][row pair][lane] (16 B per lane, conflict-free) instead of K x 2 x (bit test + FSEL), which would put 4
// instructions per pair of cells on the half-rate ALU pipe. The haplotype words, SEP/PAD protocol, lane skew and
// the rescue queue are those of forward.cuh.
#pragma once
#include <type_traits>

#include "forward.cuh"

namespace phmm {

// A packed pair of floats. PHMM_PK_INTRIN selects how the packed instructions are spelled: the sm_100 float2
// intrinsics (__ffma2_rn ...) or inline PTX on 64-bit registers; both become FFMA2 / FMUL2 / FADD2.
#ifndef PHMM_PK_INTRIN
#define PHMM_PK_INTRIN 1
#endif
#if PHMM_PK_INTRIN
using u64 = float2;
__device__ __forceinline__ u64 pk2(float lo, float hi) { return make_float2(lo, hi); }
__device__ __forceinline__ float lo2(u64 v) { return v.x; }
__device__ __forceinline__ float hi2(u64 v) { return v.y; }
__device__ __forceinline__ u64 zero2() { return make_float2(0.f, 0.f); }
// -ftz=true applies (initNative enables FTZ, IntelPairHmm.cc:230)
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { return __fadd2_rn(a, b); }
struct alignas(16) Quad { u64 x, y; };
__device__ __forceinline__ u64 sel2(bool p, u64 a, u64 b) { return make_float2(p ? a.x : b.x, p ? a.y : b.y); }
__device__ __forceinline__ u64 shfl_up2(u64 v, int W) {
    return make_float2(__shfl_up_sync(kFull, v.x, 1, W), __shfl_up_sync(kFull, v.y, 1, W));
}
__device__ __forceinline__ u64 shfl_down2(u64 v, int W) {
    return make_float2(__shfl_down_sync(kFull, v.x, 1, W), __shfl_down_sync(kFull, v.y, 1, W));
}
#else
using u64 = unsigned long long;
__device__ __forceinline__ u64 pk2(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ float lo2(u64 v) { return __uint_as_float((uint32_t)v); }
__device__ __forceinline__ float hi2(u64 v) { return __uint_as_float((uint32_t)(v >> 32)); }
__device__ __forceinline__ u64 zero2() { return 0ull; }
// -ftz=true semantics of the scalar kernels (initNative enables FTZ, IntelPairHmm.cc:230)
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 d; asm("add.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
struct alignas(16) Quad { u64 x, y; };
__device__ __forceinline__ u64 sel2(bool p, u64 a, u64 b) { return p ? a : b; }
__device__ __forceinline__ u64 shfl_up2(u64 v, int W) { return __shfl_up_sync(kFull, v, 1, W); }
__device__ __forceinline__ u64 shfl_down2(u64 v, int W) { return __shfl_down_sync(kFull, v, 1, W); }
#endif

// dynamic shared memory of one CTA: per warp the e-table [5][KP][32] x 16 B (KP = row pairs) + the haplotype table
__host__ __device__ constexpr int pk_row_pairs(int K) { return (K + 1) / 2; }
__host__ __device__ constexpr size_t pk_etab_bytes(int K) { return (size_t)6 * pk_row_pairs(K) * 32 * 16; }
// per-warp table of the unit's haplotypes for the SEP handler: first tile, 2^120/H, number of tiles
struct PkHapInfo { uint32_t tix; float initY; int32_t tiles; };
// PHMM_PK_XCHG = 1: lanes exchange the row boundary through shared memory (STS.128 + WARPSYNC + LDS.128 per column) instead of
// four shuffles + four selects. Per warp 2 (column parity) x 36 slots x 16 B {D pair, U pair}; a group of W lanes owns W+1 slots:
// slot 0 = the matrix's row-0 boundary {Dbound, 0}, lane l writes slot l+1 and reads slot l.
#ifndef PHMM_PK_LDG_END
#define PHMM_PK_LDG_END 1
#endif
#ifndef PHMM_PK_XCHG
#define PHMM_PK_XCHG 0
#endif
constexpr int kXSlots = 36;
__host__ __device__ constexpr size_t pk_xbuf_bytes() { return PHMM_PK_XCHG ? (size_t)2 * kXSlots * 16 : 0; }
__host__ __device__ constexpr size_t pk_smem_bytes(int K) {
    return (size_t)(kThreads / 32) * (pk_etab_bytes(K) + pk_xbuf_bytes() + sizeof(PkHapInfo) * kMaxUnitHaps);
}

// per-row constants of one read row (ReadForPairHMM.cpp:72-80; stripeINITIALIZATION template.h:150-152)
struct RowConst { float pMM, pMX, pMY, d1, d3; uint32_t rb; };
__device__ __forceinline__ RowConst row_consts(const KernelArgs &a, const DevRead rr, int g, const float *__restrict__ ph,
                                               const float *__restrict__ mmt) {
    RowConst c{0.f, 0.f, 0.f, 0.f, 0.f, 5u};           // dead slot below the read: every constant 0, matches nothing
    if (g < rr.len) {
        const uint8_t *p = a.rd + rr.off + g;
        c.rb = base_index(p[0]);
        const int bq = p[a.plane] & 127, iq = p[2 * a.plane] & 127, dq = p[3 * a.plane] & 127;
        const int mx = max(iq, dq), mn = min(iq, dq);
        c.pMM = mmt[((mx * (mx + 1)) >> 1) + mn];
        c.pMX = ph[iq];
        c.pMY = ph[dq];
        const float dist = ph[bq];
        c.d1 = __fsub_rn(1.0f, dist);
        c.d3 = __fdiv_rn(dist, 3.0f);
    }
    return c;
}
// match: same base, or either side is N (template.h:38,48-51)
__device__ __forceinline__ float e_of(const RowConst &c, uint32_t b) {
    return (c.rb == 5u) ? 0.f : ((c.rb == b || c.rb == 4u || b == 4u) ? c.d1 : c.d3);
}

#ifndef PHMM_PK_ORDER
#define PHMM_PK_ORDER 1
#endif
#ifndef PHMM_PK_UNROLL
#define PHMM_PK_UNROLL 1
#endif
#define PHMM_PRAGMA_(x) _Pragma(#x)
#define PHMM_UNROLL(n) PHMM_PRAGMA_(unroll n)
#ifndef PHMM_MINB_PK
#define PHMM_MINB_PK 2
#endif

template <int W, int K, int KL>
__global__ void __launch_bounds__(kThreads, PHMM_MINB_PK) forward2_kernel(const KernelArgs a) {
    static_assert(W == 32 || W == 16 || W == 8, "lane-group width");
    static_assert(KL >= 0 && KL < K, "last-row slot out of range");
    constexpr int C = kTile;
    using Tile = typename TileWord<C>::type;
    constexpr int KP = pk_row_pairs(K);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sl = lane & (W - 1);              // lane inside its group
    const int seg = lane / W;                   // group inside the warp: reads 2*seg (low half) and 2*seg+1 (high half)
    const bool lane0 = sl == 0;
    const float *__restrict__ ph = static_cast<const float *>(a.ph2pr);
    const float *__restrict__ mmt = static_cast<const float *>(a.mm);
    const Tile *__restrict__ tiles = reinterpret_cast<const Tile *>(a.hapw);
    const uint32_t n_units = a.n_units_dev ? *a.n_units_dev : a.n_units;
    extern __shared__ __align__(16) unsigned char pk_smem[];
    Quad *etab = reinterpret_cast<Quad *>(pk_smem + (size_t)warp * pk_etab_bytes(K)) + lane;
    constexpr size_t kEtabAll = (size_t)(kThreads / 32) * pk_etab_bytes(K), kXbufAll = (size_t)(kThreads / 32) * pk_xbuf_bytes();
    Quad *xrd = reinterpret_cast<Quad *>(pk_smem + kEtabAll + (size_t)warp * pk_xbuf_bytes()) + lane + seg;   // slot this lane reads
    PkHapInfo *hap_info = reinterpret_cast<PkHapInfo *>(pk_smem + kEtabAll + kXbufAll) + warp * kMaxUnitHaps;
    const u64 cXY = pk2(a.u_xy_f, a.u_xy_f), cG = pk2(a.u_gapm_f, a.u_gapm_f);   // broadcast uniform operands

    for (;;) {
        uint32_t u = 0;
        if (lane == 0) u = atomicAdd(a.queue, 1u);
        u = __shfl_sync(kFull, u, 0);
        if (u >= n_units) break;
        const DevUnit unit = a.units[u];
        const uint32_t unit_nh = unit.shape & 0xffu, unit_nreads = (unit.shape >> 8) & 0xffu, stride = unit.shape >> 16;
        // a partial unit leaves halves without a read: they recompute read 0 and emit nothing
        const uint32_t ra = 2u * (uint32_t)seg, rb = ra + 1u;
        const bool valid_a = ra < unit_nreads, valid_b = rb < unit_nreads;
        const uint32_t read_a = unit.read + (valid_a ? ra : 0u), read_b = unit.read + (valid_b ? rb : 0u);
        const uint32_t out_a = unit.out + ra * stride, out_b = unit.out + rb * stride;
        const DevRead rra = a.reads[read_a], rrb = a.reads[read_b];
        const int last_a = (rra.len - 1) / K, last_b = (rrb.len - 1) / K;   // lane that owns the last row (slot KL)

        // ---- per-slot constants, packed (A, B); the match factor goes to the shared table
        u64 pMM[K], pMX[K], pMY[K];
        __syncwarp();
#pragma unroll
        for (int k = 0; k < K; k += 2) {
            RowConst c[2][2];
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int g = sl * K + k + q;
                c[q][0] = (k + q < K) ? row_consts(a, rra, g, ph, mmt) : RowConst{0.f, 0.f, 0.f, 0.f, 0.f, 5u};
                c[q][1] = (k + q < K) ? row_consts(a, rrb, g, ph, mmt) : RowConst{0.f, 0.f, 0.f, 0.f, 0.f, 5u};
                if (k + q < K) {
                    pMM[k + q] = pk2(c[q][0].pMM, c[q][1].pMM);
                    pMX[k + q] = pk2(c[q][0].pMX, c[q][1].pMX);
                    pMY[k + q] = pk2(c[q][0].pMY, c[q][1].pMY);
                }
            }
#pragma unroll
            for (uint32_t b = 0; b < 5; ++b)
                etab[(b * KP + k / 2) * 32] = Quad{pk2(e_of(c[0][0], b), e_of(c[0][1], b)),
                                                         pk2(e_of(c[1][0], b), e_of(c[1][1], b))};
            etab[(5 * KP + k / 2) * 32] = Quad{zero2(), zero2()};     // row of the flagged words
        }
        // transition constants of the row below this lane's last slot (= slot 0 of lane+1; 0 past the group)
        u64 nMM = shfl_down2(pMM[0], W), nMX = shfl_down2(pMX[0], W);
        if (sl == W - 1) { nMM = zero2(); nMX = zero2(); }

        // haplotype table of the unit for the SEP handler and the step schedule
        if ((uint32_t)lane < unit_nh) {                // nh <= kMaxUnitHaps == 32 (host guarantees)
            const DevHap hp = a.haps[a.hap_list[unit.hap_list + lane]];
            hap_info[lane] = PkHapInfo{hp.off / C, hp.len ? __fdiv_rn(Num<float>::init_const(), (float)hp.len) : 0.f,   // template.h:115
                                       (int32_t)hap_tiles(hp.len)};
        }
        __syncwarp();

        u64 M[K], X[K], Y[K];
#pragma unroll
        for (int k = 0; k < K; ++k) { M[k] = X[k] = Y[k] = zero2(); }
        u64 Dprev = zero2(), sumM = zero2(), sumX = zero2(), Dbound = zero2();
        u64 Din[C], Uin[C];
#pragma unroll
        for (int j = 0; j < C; ++j) { Din[j] = Uin[j] = zero2(); }
#if PHMM_PK_XCHG
        if (lane0) { xrd[0] = Quad{zero2(), zero2()}; xrd[kXSlots] = Quad{zero2(), zero2()}; }   // row-0 boundary slots
#endif

        int cur = -1;
        uint32_t tix1 = 32 - sl;                       // lane l sees l PAD tiles, then [PAD.., SEP] (tile 31)
        Tile w = tiles[tix1 - 1], w1 = tiles[tix1];    // current and next tile; tix1 = index of w1

        // one DP column for this lane's K row pairs (computeMXY, template.h:177-192)
        auto column = [&](uint32_t cw, u64 Dd, u64 Uu) {
            const Quad *ep = etab + ((cw >> 25) & 7u) * (KP * 32);   // bits 25..27 of a word = base index (5 if flagged)
            u64 e[2 * KP];
#pragma unroll
            for (int j = 0; j < KP; ++j) { const Quad q = ep[j * 32]; e[2 * j] = q.x; e[2 * j + 1] = q.y; }
            // a flagged word resets the lane's state arithmetically instead of in a divergent branch: its match
            // factors are 0 (M = 0, hence X = 0) and its Y coefficient is 0; two flagged words in a row (every SEP is
            // followed by a PAD, kernels.cuh) leave M = X = Y = 0, i.e. column 0 of the next haplotype's matrix
            // (bit trick instead of a select so that the uniform coefficient itself stays in a uniform register:
            // an FFMA2 with a UR operand reads 4 vector registers, with three vector operands it is register-port bound)
            const float cy = __uint_as_float(a.u_xy_bits & ~(uint32_t)((int)cw >> 31));
            const u64 cYY = pk2(cy, cy);
#if PHMM_PK_ORDER == 0
            // rows bottom-up so that every update can be done in place: row k reads the OLD row k-1
#pragma unroll
            for (int k = K - 1; k >= 1; --k) {
                const u64 s = fma2(Y[k - 1], cG, fma2(X[k - 1], cG, mul2(M[k - 1], pMM[k])));
                Y[k] = fma2(Y[k], cYY, mul2(M[k], pMY[k]));
                M[k] = mul2(s, e[k]);
            }
            Y[0] = fma2(Y[0], cYY, mul2(M[0], pMY[0]));
            M[0] = mul2(Dd, e[0]);
            // X of this column depends on the NEW M and X of the row above: top-down, after M
            X[0] = Uu;
#pragma unroll
            for (int k = 1; k < K; ++k) X[k] = fma2(X[k - 1], cXY, mul2(M[k - 1], pMX[k]));
#else
            u64 Mn[K], Xn[K], Yn[K];
            Mn[0] = mul2(Dd, e[0]);
            Xn[0] = Uu;
            Yn[0] = fma2(Y[0], cYY, mul2(M[0], pMY[0]));
#pragma unroll
            for (int k = 1; k < K; ++k) {
                const u64 s = fma2(Y[k - 1], cG, fma2(X[k - 1], cG, mul2(M[k - 1], pMM[k])));
                Mn[k] = mul2(s, e[k]);
                Xn[k] = fma2(Xn[k - 1], cXY, mul2(Mn[k - 1], pMX[k]));
                Yn[k] = fma2(Y[k], cYY, mul2(M[k], pMY[k]));
            }
#pragma unroll
            for (int k = 0; k < K; ++k) { M[k] = Mn[k]; X[k] = Xn[k]; Y[k] = Yn[k]; }
#endif
            sumM = add2(sumM, M[KL]);                                             // template.h:326,329
            sumX = add2(sumX, X[KL]);
        };
        // bottom row, pre-combined with the transition constants of the row below it
        auto emit_down = [&](u64 &Dout, u64 &Uout) {
            Dout = fma2(Y[K - 1], cG, fma2(X[K - 1], cG, mul2(M[K - 1], nMM)));
            Uout = fma2(X[K - 1], cXY, mul2(M[K - 1], nMX));
        };
        auto emit_result = [&](float r, uint32_t my_read, uint32_t my_out, int len) {
            a.fres[my_out + cur] = r;
            if (r < 1e-28f) {                                                      // IntelPairHmm.cc:342
                const int k32 = (len + 31) >> 5;                                   // the FP64 kernels run at W = 32
                const int bin = k32 * (k32 - 1) / 2 + (len - 1) % k32;
                const uint32_t slot = a.rescue_base[bin] + atomicAdd(a.rescue_count + bin, 1u);
                a.rescue_units[slot] = DevUnit{my_read, unit.hap_list + (uint32_t)cur, unit_shape(1u, 1u, 0u), my_out + (uint32_t)cur};
            }
        };
        // SEP word = column 0 of the next haplotype: emit the finished one, switch. Runs before the columns of its
        // tile: they are all flagged (a SEP sits in slot 0 of its own tile, or closes the all-PAD lead-in), add 0 to the
        // sums and ignore the boundary values, so the order does not matter. PAD words need no handler at all.
        auto next_haplotype = [&](uint32_t &next_tix) {
            {
                if ((uint32_t)cur < unit_nh) {
                    const u64 r = add2(sumM, sumX);                                // template.h:339-343
                    if (sl == last_a && valid_a) emit_result(lo2(r), read_a, out_a, rra.len);
                    if (sl == last_b && valid_b) emit_result(hi2(r), read_b, out_b, rrb.len);
                }
                ++cur;
                sumM = sumX = zero2();
                float initY = 0.f;
                next_tix = 32;                                                     // drain PAD tiles
                if ((uint32_t)cur < unit_nh) { const PkHapInfo hi = hap_info[cur]; next_tix = hi.tix; initY = hi.initY; }
                // row 0 of the matrix seen from row 1: ((0*pMM + 0*pGAPM) + Y0*pGAPM) = Y0*pGAPM[row 1]
                const float db = initY * a.u_gapm_f;
                Dbound = pk2(db, db);
                if (lane0) {
                    Din[C - 1] = Dbound;               // becomes Dprev of the next step's first column
#if PHMM_PK_XCHG
                    xrd[0] = Quad{Dbound, zero2()}; xrd[kXSlots] = Quad{Dbound, zero2()};
#endif
                }
            }
        };

        // one pipeline step = the C columns of this lane's current tile. DIRTY steps are those in which some lane of the
        // warp may hold a SEP tile (the W steps after a haplotype boundary enters lane 0); all others run the branch-free
        // CLEAN instance, whose body is a single basic block.
        auto step = [&](auto dirty_tag) {
            constexpr bool DIRTY = decltype(dirty_tag)::value;
            if (DIRTY) {
                bool sep = false;
#pragma unroll
                for (int j = 0; j < C; ++j) sep |= (tw_get(w, j) & 0xC0000000u) == 0x80000000u;
                if (sep) {
                    uint32_t target = tix1;
                    next_haplotype(target);
                    // a unit's haplotypes lie back to back in hapw, so the sequential guess is right except after the
                    // lead-in and before the drain
                    if (target != tix1) { tix1 = target; w1 = __ldg(tiles + tix1); }
                }
            }
#if !PHMM_PK_LDG_END
            const Tile w2 = __ldg(tiles + tix1 + 1);   // two steps ahead: an L2 miss stays hidden
#endif
#if PHMM_PK_XCHG
            static_assert(C == 2, "the exchange buffers alternate with the column parity");
            u64 Dd = Dprev;
#pragma unroll
            for (int j = 0; j < C; ++j) {
                const u64 Dsave = Din[j];              // lane above's D of (this tile, column j): diagonal of column j+1
                column(tw_get(w, j), Dd, Uin[j]);
                u64 Dout, Uout;
                emit_down(Dout, Uout);
                // exchange right away -- its latency hides behind the next column, the values are first read one step
                // later. One barrier per exchange is enough: the buffer of parity j is rewritten only after the barrier of
                // the other parity, which every lane passes after its own read below.
                Quad *xb = xrd + j * kXSlots;
                xb[1] = Quad{Dout, Uout};
                __syncwarp();
                const Quad in = xb[0];
                Din[j] = in.x;
                Uin[j] = in.y;
                Dd = Dsave;
            }
            Dprev = Dd;           // D of the last column of the previous exchange: diagonal of the next step's first column
#else
            u64 Dout[C], Uout[C];
#pragma unroll
            for (int j = 0; j < C; ++j) {
                column(tw_get(w, j), j == 0 ? Dprev : Din[j - 1], Uin[j]);
                emit_down(Dout[j], Uout[j]);
            }
            Dprev = Din[C - 1];   // D of the last column of the PREVIOUS exchange: diagonal of next step's first column
#pragma unroll
            for (int j = 0; j < C; ++j) {
                const u64 dn = shfl_up2(Dout[j], W), un = shfl_up2(Uout[j], W);
                Din[j] = sel2(lane0, Dbound, dn);
                Uin[j] = sel2(lane0, zero2(), un);
            }
#endif
            // rotate the tile queue at the END of the step: w is dead by now, so the load can land in a free register and
            // nothing waits for it here; it is consumed two steps later (a full step of latency hiding)
            w = w1;
            ++tix1;
#if PHMM_PK_LDG_END
            w1 = __ldg(tiles + tix1);
#else
            w1 = w2;
#endif
        };
        // haplotype boundary h enters lane 0 at the start of segment h and leaves lane W-1 W steps later; segment nh
        // is the drain
        for (uint32_t h = 0; h <= unit_nh; ++h) {
            const int T = h < unit_nh ? hap_info[h].tiles : W;
            const int nd = min(W, T), nc = T - nd;
PHMM_UNROLL(1)
            for (int i = 0; i < nd; ++i) step(std::true_type{});
PHMM_UNROLL(1)
            for (int i = 0; i < nc; ++i) step(std::false_type{});
        }
    }
}

}  // namespace phmm
