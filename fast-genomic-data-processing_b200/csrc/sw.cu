// Batched Smith-Waterman with backtrace for sm_100a (include/smithwaterman_cuda.h).
//
// One warp per alignment, claimed from an atomic queue by a persistent grid. The affine-gap recurrence of the reference
// (MAIN_CODE, M2/intel/smithwaterman/PairWiseSW.h:27-62) is swept one anti-diagonal at a time -- the cells of an
// anti-diagonal are independent, lanes take consecutive columns j -- with the last two anti-diagonals of H and the last
// one of E, F kept in shared memory indexed by column. Each cell writes one backtrace byte (move | INSERT_EXT |
// DELETE_EXT, smithwaterman_common.h:44-49) into a per-warp scratch area in anti-diagonal-major order, so the stores
// of a warp are consecutive bytes. The best cell (last row / last column, visited in anti-diagonal order with the
// reference's tie rules, :227-251), the backtrace state machine and the CIGAR text (getCIGAR, :265-447) follow the
// reference step for step; runs of plain matches are walked 32 cells at a time (every lane fetches one cell of the diagonal).
// Integer work, HBM/L2-bound on the backtrace bytes: no tensor cores.
#include <algorithm>
#include <climits>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <type_traits>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/smithwaterman_cuda.h"

namespace sw {

constexpr int kWarps = 4;                       // warps per CTA (fewer when long sequences need the shared memory)
constexpr unsigned kFull = 0xffffffffu;
constexpr int BT_MATCH = 0, BT_INSERT = 1, BT_DELETE = 2, BT_INSERT_EXT = 4, BT_DELETE_EXT = 8;
constexpr int32_t kMinCutoff = -100000000;      // MATRIX_MIN_CUTOFF
constexpr int32_t kLowInit = INT_MIN / 2;       // LOW_INIT_VALUE

struct Args {
    const uint8_t *ref_bases, *alt_bases;
    const int64_t *ref_off, *alt_off;
    int64_t ref_first, alt_first;               // the device copies start at these offsets of the caller's arrays
    const int32_t *ref_len, *alt_len;
    uint32_t n;
    int32_t match, mismatch, open, extend, strategy;
    char *cigars;
    int32_t cigar_stride;
    int32_t *offsets;
    uint8_t *scratch;                           // per resident warp: backtrace bytes, then the CIGAR element list
    uint64_t scratch_per_warp;
    uint32_t bt_width;                          // max alt length + 1
    uint32_t max_ref, max_alt;
    uint32_t *queue;
    const uint32_t *order;                      // queue position -> pair (the pairs of one launch group)
};

// ints of shared memory one warp needs
__host__ __device__ inline size_t warp_smem_ints(uint32_t max_ref, uint32_t max_alt) {
    const size_t c = max_alt + 2, r = max_ref + 2;
    return 7 * c + r + (c + 3) / 4 + (r + 3) / 4;          // H x3, E x2, F, lastrow | lastcol | alt bytes | ref bytes
}

__device__ __forceinline__ int digits_of(int v) { int d = 0; while (v > 0) { v /= 10; ++d; } return d; }

// Best cell, backtrace and CIGAR text of one alignment (shared by both kernels). btget(i, j) returns the backtrace code
// of cell (i, j); every lane runs the same scalar code, lane 0 writes the results.
template <class BtGet>
__device__ __forceinline__ void finish_alignment(const Args &a, uint32_t p, int lane, int nrow, int ncol, const int32_t *lastrow,
                                             const int32_t *lastcol, int32_t *elems, BtGet btget) {
    // ---- best cell: last row (SOFTCLIP / IGNORE only) and last column in anti-diagonal order (:227-251). Every lane
    // runs the same scalar code (shared-memory broadcasts), so there is no divergence and all lanes know the result.
    int32_t max_score = INT_MIN;
    int max_i = 0, max_j = 0;
    const bool row_too = a.strategy == SW_CUDA_SOFTCLIP || a.strategy == SW_CUDA_IGNORE;
    for (int ad = 1; ad <= nrow + ncol; ++ad) {
        if (ad >= nrow + 1 && row_too) {
            const int j = ad - nrow;
            const int32_t s = lastrow[j];
            if (max_score < s || (max_score == s && abs(nrow - j) < abs(max_i - max_j))) { max_score = s; max_i = nrow; max_j = j; }
        }
        if (ad >= ncol + 1) {
            const int i = ad - ncol;
            const int32_t s = lastcol[i];
            if (max_score < s || (max_score == s && (max_j == ncol || abs(i - ncol) <= abs(max_i - max_j)))) { max_score = s; max_i = i; max_j = ncol; }
        }
    }

    // ---- backtrace (getCIGAR, :265-402), elements merged as they are produced (the reference merges equal
    // neighbours afterwards, :404-421 -- same list)
    int n_el = 0, cur_op = -1, cur_len = 0;
    auto push = [&](int op, int len) {
        if (n_el > 0 && op == cur_op) { cur_len += len; return; }
        if (n_el > 0 && lane == 0) { elems[2 * (n_el - 1)] = cur_op; elems[2 * (n_el - 1) + 1] = cur_len; }
        cur_op = op; cur_len = len; ++n_el;
    };
    int i, j;
    if (a.strategy == SW_CUDA_INDEL) { i = nrow; j = ncol; }
    else if (a.strategy == SW_CUDA_LEADING_INDEL) { i = max_i; j = ncol; }
    else { i = max_i; j = max_j; }
    if (j < ncol) push(SW_CUDA_SOFTCLIP, ncol - j);
    int state = 0;
    while (i > 0 && j > 0) {
        if (state == 0) {
            // lane l looks at (i - l, j - l): a run of plain matches is consumed in one step
            const int ii = i - lane, jj = j - lane;
            int btr = 0xff;
            if (ii > 0 && jj > 0) btr = btget(ii, jj);
            const unsigned plain = __ballot_sync(kFull, ii > 0 && jj > 0 && (btr & 3) == BT_MATCH);
            const int run = __ffs(~plain) - 1;                      // leading lanes that are plain matches (32 if all)
            if (run != 0) {
                const int r = run < 0 ? 32 : run;
                push(BT_MATCH, r); i -= r; j -= r;                  // MATCH leaves state 0
                continue;
            }
            btr = __shfl_sync(kFull, btr, 0);
            if ((btr & 3) == BT_INSERT) { --j; push(BT_INSERT, 1); state = btr & BT_INSERT_EXT; }
            else { --i; push(BT_DELETE, 1); state = btr & BT_DELETE_EXT; }       // (btr & 3) == DELETE
        } else {
            const int btr = btget(i, j);
            if (state == BT_INSERT_EXT) { --j; ++cur_len; state = btr & BT_INSERT_EXT; }
            else { --i; ++cur_len; state = btr & BT_DELETE_EXT; }
        }
    }
    int offset;
    if (a.strategy == SW_CUDA_SOFTCLIP) {
        if (j > 0) push(SW_CUDA_SOFTCLIP, j);
        offset = i;
    } else if (a.strategy == SW_CUDA_IGNORE) {
        if (j > 0) push(n_el > 0 ? cur_op : BT_MATCH, j);
        offset = (int16_t)(i - j);
    } else {
        if (i > 0) push(BT_DELETE, i);
        else if (j > 0) push(BT_INSERT, j);
        offset = 0;
    }
    // ---- CIGAR text, last element first (:423-445)
    if (lane == 0) {
        if (n_el > 0) { elems[2 * (n_el - 1)] = cur_op; elems[2 * (n_el - 1) + 1] = cur_len; }
        char *out = a.cigars + (uint64_t)p * a.cigar_stride;
        int cur = 0;
        for (int k = n_el - 1; k >= 0; --k) {
            const int op = elems[2 * k], v = elems[2 * k + 1], d = digits_of(v);
            if (d + 1 > 1 && cur + d + 1 <= a.cigar_stride) {
                int x = v;
                for (int q = d - 1; q >= 0; --q) { out[cur + q] = (char)('0' + x % 10); x /= 10; }
                cur += d;
                out[cur++] = op == BT_MATCH ? 'M' : op == BT_INSERT ? 'I' : op == BT_DELETE ? 'D' : op == SW_CUDA_SOFTCLIP ? 'S' : 'R';
            }
        }
        a.offsets[p] = offset;
    }
}

__global__ void __launch_bounds__(kWarps * 32) align_kernel(const Args a) {
    extern __shared__ int32_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t C = a.max_alt + 2, R = a.max_ref + 2;
    int32_t *base = smem + (size_t)warp * warp_smem_ints(a.max_ref, a.max_alt);
    // H (3 anti-diagonals) and E (2) rotate by OFFSET into `base`, not by pointer: the accesses stay LDS/STS with 32-bit
    // addresses (an array of pointers made them generic 64-bit loads and a third of the inner loop address arithmetic)
    const int oH = 0, oE = 3 * (int)C;
    int32_t *F = base + 5 * C, *lastrow = base + 6 * C, *lastcol = base + 7 * C;
    uint8_t *alt = reinterpret_cast<uint8_t *>(lastcol + R), *ref = alt + ((C + 3) / 4) * 4;
    const uint32_t gwarp = blockIdx.x * (blockDim.x >> 5) + warp;
    uint8_t *bt = a.scratch + (uint64_t)gwarp * a.scratch_per_warp;
    const bool indel = a.strategy == SW_CUDA_INDEL || a.strategy == SW_CUDA_LEADING_INDEL;

    for (;;) {
        uint32_t p = 0;
        if (lane == 0) p = atomicAdd(a.queue, 1u);
        p = __shfl_sync(kFull, p, 0);
        if (p >= a.n) break;
        p = a.order[p];
        const int nrow = a.ref_len[p], ncol = a.alt_len[p];             // rows i = ref, columns j = alt
        const uint8_t *rp = a.ref_bases + (a.ref_off[p] - a.ref_first), *ap = a.alt_bases + (a.alt_off[p] - a.alt_first);
        __syncwarp();
        for (int k = lane; k < nrow; k += 32) ref[k] = rp[k];
        for (int k = lane; k < ncol; k += 32) alt[k] = ap[k];
        for (int j = lane; j <= ncol; j += 32) F[j] = kLowInit;         // F[0][j] (PairWiseSW.h:98-101,224)
        if (lane == 0) base[oH + (int)C] = 0;                           // H[0][0]: anti-diagonal 0 = the "ad-1" buffer of the first step
        __syncwarp();

        // ---- forward sweep over anti-diagonals ad = i + j
        int h0 = oH, h1 = oH + (int)C, h2 = oH + 2 * (int)C, e1 = oE, e2 = oE + (int)C;   // buffers of ad-2, ad-1, ad
        for (int ad = 1; ad <= nrow + ncol; ++ad) {
            int32_t *H0 = base + h0, *H1 = base + h1, *H2 = base + h2, *E1 = base + e1, *E2 = base + e2;
            if (lane == 0) {
                // boundary cells of this anti-diagonal (:214-225): (0, ad) on the top row, (ad, 0) in the left column
                if (ad <= ncol) H2[ad] = indel ? a.open + (ad - 1) * a.extend : 0;
                if (ad <= nrow) { H2[0] = indel ? a.open + (ad - 1) * a.extend : 0; E2[0] = kLowInit; }
            }
            const int jlo = max(1, ad - nrow), jhi = min(ncol, ad - 1);
            uint8_t *btd = bt + (uint64_t)ad * a.bt_width;
            for (int j = jlo + lane; j <= jhi; j += 32) {
                const int i = ad - j;
                const int32_t ext_h = E1[j - 1] + a.extend, open_h = H1[j - 1] + a.open;
                const int32_t e = max(open_h, ext_h);
                int b = open_h > ext_h ? 0 : BT_INSERT_EXT;
                const int32_t ext_v = F[j] + a.extend, open_v = H1[j] + a.open;
                const int32_t f = max(ext_v, open_v);
                if (!(open_v > ext_v)) b |= BT_DELETE_EXT;
                const int32_t m = H0[j - 1] + (ref[i - 1] == alt[j - 1] ? a.match : a.mismatch);
                int32_t h = max(kMinCutoff, m);
                int mv = e > h ? BT_INSERT : BT_MATCH;
                h = max(h, e);
                if (f > h) { mv = BT_DELETE; h = f; }
                E2[j] = e; F[j] = f; H2[j] = h;
                btd[j] = (uint8_t)(mv | b);
                if (i == nrow) lastrow[j] = h;
                if (j == ncol) lastcol[i] = h;
            }
            __syncwarp();
            const int t = h0; h0 = h1; h1 = h2; h2 = t;
            const int u = e1; e1 = e2; e2 = u;
        }

        int32_t *elems = reinterpret_cast<int32_t *>(bt + ((((uint64_t)a.max_ref + a.max_alt + 1) * a.bt_width + 7) & ~7ull));
        const uint32_t btw = a.bt_width;
        finish_alignment(a, p, lane, nrow, ncol, lastrow, lastcol, elems,
                         [=](int i, int j) { return (int)bt[(uint64_t)(i + j) * btw + j]; });
    }
}

// ---- register-blocked variant: lane l owns the K columns l*K+1 .. l*K+K of the alignment matrix (alt positions) and
// sweeps the rows (ref positions) top to bottom, one row per step, lane l one row behind lane l-1 (the same skewed
// wavefront as the PairHMM kernels). H of the row above and F live in registers, the row's H/E chain runs through the
// lane's K cells and leaves for lane l+1 by two shuffles per step; a step's K backtrace codes are one 32-bit word
// (4 bits per cell), stored at [step][lane] -- 128 consecutive bytes per warp and step. ~4x fewer instructions per cell
// than the anti-diagonal kernel; used when the longest alt of the batch fits 32 x 8 columns.
__host__ __device__ inline size_t strip_smem_ints(uint32_t max_ref, uint32_t max_alt) {
    const size_t c = max_alt + 2, r = max_ref + 2;
    return c + r + (r + 3) / 4;                            // lastrow | lastcol | ref bytes
}

template <int K>
__global__ void __launch_bounds__(kWarps * 32) align_strip_kernel(const Args a) {
    extern __shared__ int32_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t C = a.max_alt + 2, R = a.max_ref + 2;
    int32_t *lastrow = smem + (size_t)warp * strip_smem_ints(a.max_ref, a.max_alt), *lastcol = lastrow + C;
    uint8_t *ref = reinterpret_cast<uint8_t *>(lastcol + R);
    const uint32_t gwarp = blockIdx.x * (blockDim.x >> 5) + warp;
    uint32_t *btw = reinterpret_cast<uint32_t *>(a.scratch + (uint64_t)gwarp * a.scratch_per_warp);
    int32_t *elems = reinterpret_cast<int32_t *>(btw + (size_t)(a.max_ref + 33) * 32);
    const bool indel = a.strategy == SW_CUDA_INDEL || a.strategy == SW_CUDA_LEADING_INDEL;
    const int32_t w_open = a.open, w_ext = a.extend, w_match = a.match, w_mis = a.mismatch;

    for (;;) {
        uint32_t p = 0;
        if (lane == 0) p = atomicAdd(a.queue, 1u);
        p = __shfl_sync(kFull, p, 0);
        if (p >= a.n) break;
        p = a.order[p];
        const int nrow = a.ref_len[p], ncol = a.alt_len[p];             // rows i = ref, columns j = alt
        const uint8_t *rp = a.ref_bases + (a.ref_off[p] - a.ref_first), *ap = a.alt_bases + (a.alt_off[p] - a.alt_first);
        __syncwarp();
        for (int k = lane; k < nrow; k += 32) ref[k] = rp[k];
        const int j0 = lane * K;                                        // this lane's columns are j0+1 .. j0+K
        int32_t Hup[K], F[K];
        uint32_t altb[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int j = j0 + k + 1;
            altb[k] = j <= ncol ? ap[j - 1] : 0x100u;                    // a dead column matches nothing
            Hup[k] = indel ? w_open + (j - 1) * w_ext : 0;               // H[0][j] (:214-223)
            F[k] = kLowInit;                                             // F[0][j]
        }
        // what the lane to the left delivers: H[i][j0] and E[i][j0] of this lane's current row i; h_in_prev = H[i-1][j0]
        int32_t h_in_prev = j0 == 0 ? 0 : (indel ? w_open + (j0 - 1) * w_ext : 0);   // H[0][j0]
        int32_t h_out = 0, e_out = kLowInit;
        const int lc = (ncol - 1) / K, kc = (ncol - 1) % K;             // owner of the last column
        const int32_t bnd_step = indel ? w_ext : 0;
        int32_t bnd = indel ? w_open - lane * w_ext : 0;                // H[i][0] of row i = s - lane at step s = 1
        uint32_t *btp = btw + 32 + lane;
        // column ncol's H goes to lastcol[i] every step: picked out of the lane's K registers by multiply-add with 0/1
        // factors the compiler cannot see through (it would turn them into K compares and selects per step), stored by
        // every lane -- the others to slot 0 (H[0][ncol]), which nobody reads
        int32_t pick[K];
#pragma unroll
        for (int k = 0; k < K; ++k) asm("mov.s32 %0, %1;" : "=r"(pick[k]) : "r"(kc == k ? 1 : 0));
        const int lc_step = lane == lc ? 1 : 0;
        int lc_idx = lane == lc ? 1 - lane : 0;                         // i of lane lc at step s = 1
        __syncwarp();

        // one row per lane; `guarded` = some lanes are outside their rows (the ramps at both ends of the sweep)
        auto step = [&](const int s, auto guard) {
            constexpr bool guarded = decltype(guard)::value;
            int32_t h_in = __shfl_up_sync(kFull, h_out, 1), e_in = __shfl_up_sync(kFull, e_out, 1);
            const int i = s - lane;
            if (lane == 0) { h_in = bnd; e_in = kLowInit; }            // H[i][0], E[i][0]
            if (!guarded || (i >= 1 && i <= nrow)) {
                const uint32_t rb = ref[i - 1];
                int32_t h_left = h_in, e_left = e_in, diag = h_in_prev;
                uint32_t word = 0;
#pragma unroll
                for (int k = 0; k < K; ++k) {                            // MAIN_CODE, PairWiseSW.h:27-62
                    // the four comparisons of the cell go into the backtrace word as the SIGN BITS of four differences
                    // (one funnel shift each; every value is within +-1.1e9, no difference overflows). The kernel is
                    // bound by instruction issue: each max is taken once and the flag read off its result
                    const int32_t ext_h = e_left + w_ext;
                    const int32_t e = max(h_left + w_open, ext_h);
                    word = __funnelshift_l((uint32_t)(ext_h - e), word, 1);             // 1: open_h > ext_h
                    const int32_t ext_v = F[k] + w_ext;
                    const int32_t f = max(Hup[k] + w_open, ext_v);
                    word = __funnelshift_l((uint32_t)(ext_v - f), word, 1);             // 1: open_v > ext_v
                    const int32_t m = diag + (rb == altb[k] ? w_match : w_mis);
                    const int32_t h0 = max(kMinCutoff, m);
                    word = __funnelshift_l((uint32_t)(h0 - e), word, 1);                // 1: e > h  (INSERT)
                    const int32_t h1 = max(h0, e);
                    word = __funnelshift_l((uint32_t)(h1 - f), word, 1);                // 1: f > h  (DELETE)
                    const int32_t h = max(h1, f);
                    diag = Hup[k];
                    Hup[k] = h; F[k] = f; h_left = h; e_left = e;
                }
                h_out = h_left; e_out = e_left;
                h_in_prev = h_in;
                *btp = word;
                int32_t last = 0;
#pragma unroll
                for (int k = 0; k < K; ++k) last += Hup[k] * pick[k];
                lastcol[lc_idx] = last;
                if (guarded && i == nrow) {
#pragma unroll
                    for (int k = 0; k < K; ++k) if (j0 + k + 1 <= ncol) lastrow[j0 + k + 1] = Hup[k];
                }
            }
            btp += 32; bnd += bnd_step; lc_idx += lc_step;
        };
        // steps 32 .. nrow - 1 have every lane inside its rows and none on the last row: no range checks there, and two
        // steps per trip so that no register is shuffled between trips
        const int last_step = nrow + 31;
        int s = 1;
        for (; s <= min(31, last_step); ++s) step(s, std::true_type{});
        for (; s + 1 <= nrow - 1; s += 2) { step(s, std::false_type{}); step(s + 1, std::false_type{}); }
        for (; s <= last_step; ++s) step(s, std::true_type{});
        __syncwarp();
        finish_alignment(a, p, lane, nrow, ncol, lastrow, lastcol, elems, [=](int i, int j) {
            const int l = (j - 1) / K, k = (j - 1) % K;
            const uint32_t nib = btw[(size_t)(i + l) * 32 + l] >> (4 * (K - 1 - k));    // open_h> | open_v> | e> | f>
            const int mv = (nib & 1u) ? BT_DELETE : ((nib & 2u) ? BT_INSERT : BT_MATCH);
            return mv | ((nib & 8u) ? 0 : BT_INSERT_EXT) | ((nib & 4u) ? 0 : BT_DELETE_EXT);
        });
    }
}

using KernelFn = void (*)(const Args);
inline KernelFn strip_kernel_for(int K) {
    switch (K) {
        case 1: return align_strip_kernel<1>; case 2: return align_strip_kernel<2>; case 3: return align_strip_kernel<3>;
        case 4: return align_strip_kernel<4>; case 5: return align_strip_kernel<5>; case 6: return align_strip_kernel<6>;
        case 7: return align_strip_kernel<7>; case 8: return align_strip_kernel<8>; default: return nullptr;
    }
}

thread_local std::string g_err;
int fail(int st, const std::string &m) { g_err = m; return st; }

// Grow-only device array: a call reuses what the previous calls of this host thread allocated.
template <class T> struct DevArr {
    T *p = nullptr;
    size_t cap = 0;
    cudaError_t alloc(size_t n) {
        n = std::max<size_t>(1, n);
        if (n <= cap) return cudaSuccess;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        const size_t want = n + n / 4;
        cudaError_t e = cudaMalloc(reinterpret_cast<void **>(&p), want * sizeof(T));
        if (e != cudaSuccess) { cudaGetLastError(); e = cudaMalloc(reinterpret_cast<void **>(&p), n * sizeof(T)); if (e == cudaSuccess) cap = n; return e; }
        cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

// Per host thread (Mutect2Cpp keeps one engine per worker thread, M2/main.cpp:199) and device: stream, events, device
// properties and the device buffers survive between calls, so a per-region call (a few hundred alignments) does not pay
// a dozen cudaMalloc / cudaFree, a stream and two events each time. smithwaterman_cuda_release() frees them.
struct Context {
    int device = -1;
    cudaDeviceProp prop{};
    cudaStream_t st = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    DevArr<uint8_t> d_ref, d_alt, d_scratch;
    DevArr<int64_t> d_roff, d_aoff;
    DevArr<int32_t> d_rlen, d_alen, d_off;
    DevArr<char> d_cig;
    DevArr<uint32_t> d_queue, d_order;
    struct Occ { size_t smem = 0; int warps = 0, per_sm = 0; } occ[10];   // per kernel (index K; 0 = anti-diagonal kernel)
    void release() {
        if (device < 0) return;
        int cur = -1;
        if (cudaGetDevice(&cur) == cudaSuccess && cudaSetDevice(device) == cudaSuccess) {
            d_ref.release(); d_alt.release(); d_scratch.release(); d_roff.release(); d_aoff.release(); d_rlen.release();
            d_alen.release(); d_off.release(); d_cig.release(); d_queue.release(); d_order.release();
            if (ev0) cudaEventDestroy(ev0);
            if (ev1) cudaEventDestroy(ev1);
            if (st) cudaStreamDestroy(st);
            cudaSetDevice(cur);
        }
        cudaGetLastError();
        *this = Context();
    }
    ~Context() { release(); }
};
thread_local Context g_ctx;

}  // namespace sw

extern "C" {

const char *smithwaterman_cuda_last_error(void) { return sw::g_err.c_str(); }

void smithwaterman_cuda_release(void) { sw::g_ctx.release(); }

int smithwaterman_cuda_align_batch(int device, const uint8_t *ref_bases, const int64_t *ref_off, const int32_t *ref_len,
                                   const uint8_t *alt_bases, const int64_t *alt_off, const int32_t *alt_len, int64_t n,
                                   int match, int mismatch, int open, int extend, int strategy, char *cigars,
                                   int cigar_stride, int32_t *offsets, float *kernel_ms) {
    using namespace sw;
    if (kernel_ms) *kernel_ms = 0.f;
    if (n < 0 || n > 0x7fffffff) return fail(SW_CUDA_ERR_ARG, "bad pair count");
    if (n == 0) return SW_CUDA_OK;
    if (!ref_bases || !ref_off || !ref_len || !alt_bases || !alt_off || !alt_len || !cigars || !offsets || cigar_stride < 1)
        return fail(SW_CUDA_ERR_ARG, "null array or cigar_stride < 1");
    if (strategy < SW_CUDA_SOFTCLIP || strategy > SW_CUDA_IGNORE) return fail(SW_CUDA_ERR_ARG, "unknown overhang strategy");
    uint32_t max_ref = 1, max_alt = 1;
    uint64_t ref_bytes = 0, alt_bytes = 0;
    int64_t ref_first = INT64_MAX, alt_first = INT64_MAX;      // only the span the pairs touch is copied to the device
    for (int64_t p = 0; p < n; ++p) {
        if (ref_len[p] < 1 || alt_len[p] < 1 || ref_len[p] > SW_CUDA_MAX_LEN || alt_len[p] > SW_CUDA_MAX_LEN || ref_off[p] < 0 || alt_off[p] < 0)
            return fail(SW_CUDA_ERR_ARG, "sequence length outside [1, SW_CUDA_MAX_LEN] or negative offset");
        max_ref = std::max<uint32_t>(max_ref, (uint32_t)ref_len[p]); max_alt = std::max<uint32_t>(max_alt, (uint32_t)alt_len[p]);
        ref_bytes = std::max<uint64_t>(ref_bytes, (uint64_t)ref_off[p] + (uint64_t)ref_len[p]);
        alt_bytes = std::max<uint64_t>(alt_bytes, (uint64_t)alt_off[p] + (uint64_t)alt_len[p]);
        ref_first = std::min(ref_first, ref_off[p]); alt_first = std::min(alt_first, alt_off[p]);
    }
    ref_bytes -= (uint64_t)ref_first; alt_bytes -= (uint64_t)alt_first;
#define SWCU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return fail(SW_CUDA_ERR_CUDA, std::string(#x ": ") + cudaGetErrorString(e_)); } while (0)
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(SW_CUDA_ERR_NO_DEVICE, "no CUDA device (this aligner has no CPU fallback)"); }
    if (device < 0) SWCU(cudaGetDevice(&device));
    if (device >= ndev) return fail(SW_CUDA_ERR_ARG, "device ordinal out of range");
    sw::Context &cx = sw::g_ctx;
    if (cx.device != device) {
        cx.release();
        SWCU(cudaGetDeviceProperties(&cx.prop, device));
        if (cx.prop.major != 10) return fail(SW_CUDA_ERR_NO_DEVICE, "not an sm_100 device (kernels are built for sm_100a only)");
        SWCU(cudaSetDevice(device));
        SWCU(cudaStreamCreate(&cx.st));
        SWCU(cudaEventCreate(&cx.ev0)); SWCU(cudaEventCreate(&cx.ev1));
        cx.device = device;
    }
    SWCU(cudaSetDevice(device));
    const cudaDeviceProp &prop = cx.prop;

    // Pairs are grouped by alt length: class c = ceil(alt_len / 32) <= 8 runs the register-blocked kernel with K = c
    // columns per lane, longer alts the anti-diagonal kernel (SW_CUDA_KERNEL=diag forces it for all). A class too small to
    // fill the device joins the next wider one. One launch per group, one after the other on the stream: they share
    // the backtrace scratch.
    constexpr int kDiagClass = 9;
    const char *force = std::getenv("SW_CUDA_KERNEL");
    const bool force_diag = force && !std::strcmp(force, "diag");
    std::vector<uint8_t> cls((size_t)n);
    uint64_t cnt[kDiagClass + 1] = {};
    for (int64_t p = 0; p < n; ++p) {
        const int c = force_diag ? kDiagClass : std::min(kDiagClass, (alt_len[p] + 31) / 32);
        cls[(size_t)p] = (uint8_t)c; ++cnt[c];
    }
    int join[kDiagClass + 1];                                 // class -> the class whose launch it rides in
    uint64_t min_group = (uint64_t)prop.multiProcessorCount * 64;
    if (const char *mg = std::getenv("SW_CUDA_MIN_GROUP")) min_group = std::strtoull(mg, nullptr, 10);   // tests: one launch per class
    for (int c = 1; c <= kDiagClass; ++c) join[c] = c;
    for (int c = 1; c < kDiagClass - 1; ++c) {
        if (!cnt[c] || cnt[c] >= min_group) continue;
        int up = c + 1;
        while (up < kDiagClass - 1 && !cnt[up]) ++up;
        if (!cnt[up]) continue;                               // nothing wider to join
        cnt[up] += cnt[c]; cnt[c] = 0; join[c] = up;
    }
    for (int c = 1; c < kDiagClass; ++c) { int t = join[c]; while (join[t] != t) t = join[t]; join[c] = t; }
    struct Group { int K, warps; uint64_t first, count, per_warp, grid; uint32_t max_ref, max_alt; size_t smem; KernelFn kernel; };
    Group groups[kDiagClass + 1];
    int group_of[kDiagClass + 1], ngroups = 0;
    uint64_t first = 0;
    for (int c = 1; c <= kDiagClass; ++c) {
        group_of[c] = -1;
        if (!cnt[c]) continue;
        group_of[c] = ngroups;
        groups[ngroups++] = Group{c == kDiagClass ? 0 : c, kWarps, first, 0, 0, 0, 1, 1, 0, nullptr};
        first += cnt[c];
    }
    std::vector<uint32_t> order((size_t)n);
    for (int64_t p = 0; p < n; ++p) {
        Group &g = groups[group_of[join[cls[(size_t)p]]]];
        order[(size_t)(g.first + g.count++)] = (uint32_t)p;
        g.max_ref = std::max<uint32_t>(g.max_ref, (uint32_t)ref_len[p]); g.max_alt = std::max<uint32_t>(g.max_alt, (uint32_t)alt_len[p]);
    }
    uint64_t scratch_bytes = 0;
    for (int k = 0; k < ngroups; ++k) {
        Group &g = groups[k];
        g.kernel = g.K ? strip_kernel_for(g.K) : align_kernel;
        const size_t warp_smem = (g.K ? strip_smem_ints(g.max_ref, g.max_alt) : warp_smem_ints(g.max_ref, g.max_alt)) * sizeof(int32_t);
        while (g.warps > 1 && g.warps * warp_smem > (size_t)prop.sharedMemPerBlockOptin) --g.warps;
        g.smem = g.warps * warp_smem;
        if (g.smem > (size_t)prop.sharedMemPerBlockOptin) return fail(SW_CUDA_ERR_ARG, "sequences too long for the shared-memory anti-diagonals");
        // the attribute and the occupancy query cost ~20 us each: remembered per kernel for the last shared-memory size, which
        // is rounded up to 4 KB so that per-region calls (whose longest sequences differ a little) hit the entry
        g.smem = std::min<size_t>((g.smem + 4095) & ~(size_t)4095, (size_t)prop.sharedMemPerBlockOptin);
        sw::Context::Occ &oc = cx.occ[g.K];
        if (oc.smem != g.smem || oc.warps != g.warps) {
            SWCU(cudaFuncSetAttribute(g.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem));
            int q = 0;
            SWCU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&q, g.kernel, g.warps * 32, g.smem));
            oc.smem = g.smem; oc.warps = g.warps; oc.per_sm = std::max(1, q);
        }
        const int per_sm = oc.per_sm;
        const uint64_t tail = 8ull * ((uint64_t)g.max_ref + g.max_alt + 6);                 // CIGAR element list
        g.per_warp = g.K ? ((uint64_t)g.max_ref + 33) * 128 + tail
                         : ((uint64_t)g.max_ref + g.max_alt + 1) * ((uint64_t)g.max_alt + 1) + 8 + tail;
        g.per_warp = (g.per_warp + 255) & ~255ull;
        g.grid = std::min<uint64_t>((uint64_t)prop.multiProcessorCount * per_sm, (g.count + g.warps - 1) / g.warps);
        const uint64_t budget = 8ull << 30;                               // backtrace scratch of all resident warps
        g.grid = std::max<uint64_t>(1, std::min<uint64_t>(g.grid, budget / (g.per_warp * g.warps)));
        scratch_bytes = std::max(scratch_bytes, g.per_warp * g.warps * g.grid);
    }

    auto &d_ref = cx.d_ref; auto &d_alt = cx.d_alt; auto &d_scratch = cx.d_scratch; auto &d_roff = cx.d_roff; auto &d_aoff = cx.d_aoff;
    auto &d_rlen = cx.d_rlen; auto &d_alen = cx.d_alen; auto &d_off = cx.d_off; auto &d_cig = cx.d_cig; auto &d_queue = cx.d_queue;
    auto &d_order = cx.d_order;
    SWCU(d_ref.alloc(ref_bytes)); SWCU(d_alt.alloc(alt_bytes)); SWCU(d_roff.alloc(n)); SWCU(d_aoff.alloc(n));
    SWCU(d_rlen.alloc(n)); SWCU(d_alen.alloc(n)); SWCU(d_off.alloc(n)); SWCU(d_cig.alloc((size_t)n * cigar_stride));
    SWCU(d_queue.alloc(ngroups)); SWCU(d_order.alloc(n)); SWCU(d_scratch.alloc(scratch_bytes));
    cudaStream_t st = cx.st;
    cudaEvent_t ev0 = cx.ev0, ev1 = cx.ev1;
    // a failed call drops the thread's context: the next one starts from a clean stream
    auto cleanup = [&]() { cudaStreamSynchronize(st); cudaGetLastError(); cx.release(); };
#define SWCU2(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { cleanup(); return fail(SW_CUDA_ERR_CUDA, std::string(#x ": ") + cudaGetErrorString(e_)); } } while (0)
    SWCU2(cudaMemcpyAsync(d_ref.p, ref_bases + ref_first, ref_bytes, cudaMemcpyHostToDevice, st));
    SWCU2(cudaMemcpyAsync(d_alt.p, alt_bases + alt_first, alt_bytes, cudaMemcpyHostToDevice, st));
    SWCU2(cudaMemcpyAsync(d_roff.p, ref_off, sizeof(int64_t) * n, cudaMemcpyHostToDevice, st));
    SWCU2(cudaMemcpyAsync(d_aoff.p, alt_off, sizeof(int64_t) * n, cudaMemcpyHostToDevice, st));
    SWCU2(cudaMemcpyAsync(d_rlen.p, ref_len, sizeof(int32_t) * n, cudaMemcpyHostToDevice, st));
    SWCU2(cudaMemcpyAsync(d_alen.p, alt_len, sizeof(int32_t) * n, cudaMemcpyHostToDevice, st));
    SWCU2(cudaMemsetAsync(d_cig.p, 0, (size_t)n * cigar_stride, st));
    SWCU2(cudaMemcpyAsync(d_order.p, order.data(), sizeof(uint32_t) * n, cudaMemcpyHostToDevice, st));
    SWCU2(cudaMemsetAsync(d_queue.p, 0, sizeof(uint32_t) * ngroups, st));
    Args a{};
    a.ref_first = ref_first; a.alt_first = alt_first;
    a.ref_bases = d_ref.p; a.alt_bases = d_alt.p; a.ref_off = d_roff.p; a.alt_off = d_aoff.p; a.ref_len = d_rlen.p; a.alt_len = d_alen.p;
    a.match = match; a.mismatch = mismatch; a.open = open; a.extend = extend; a.strategy = strategy;
    a.cigars = d_cig.p; a.cigar_stride = cigar_stride; a.offsets = d_off.p; a.scratch = d_scratch.p;
    SWCU2(cudaEventRecord(ev0, st));
    for (int k = 0; k < ngroups; ++k) {
        const Group &g = groups[k];
        a.n = (uint32_t)g.count; a.order = d_order.p + g.first; a.queue = d_queue.p + k;
        a.scratch_per_warp = g.per_warp; a.bt_width = g.max_alt + 1; a.max_ref = g.max_ref; a.max_alt = g.max_alt;
        g.kernel<<<(unsigned)g.grid, g.warps * 32, g.smem, st>>>(a);
        SWCU2(cudaGetLastError());
    }
    SWCU2(cudaEventRecord(ev1, st));
    SWCU2(cudaMemcpyAsync(cigars, d_cig.p, (size_t)n * cigar_stride, cudaMemcpyDeviceToHost, st));
    SWCU2(cudaMemcpyAsync(offsets, d_off.p, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, st));
    SWCU2(cudaStreamSynchronize(st));
    if (kernel_ms) cudaEventElapsedTime(kernel_ms, ev0, ev1);
    return SW_CUDA_OK;
#undef SWCU
#undef SWCU2
}

int smithwaterman_cuda_align(uint8_t *ref, int ref_length, uint8_t *alt, int alt_length, uint8_t *cigar, int cigar_length,
                             int match, int mismatch, int open, int extend, uint8_t strategy) {
    const int64_t off0 = 0;
    const int32_t rl = ref_length, al = alt_length;
    int32_t offset = 0;
    if (cigar_length < 1 || !cigar) { sw::fail(SW_CUDA_ERR_ARG, "no cigar buffer"); return INT_MIN; }
    std::vector<char> tmp((size_t)cigar_length, 0);
    const int st = smithwaterman_cuda_align_batch(-1, ref, &off0, &rl, alt, &off0, &al, 1, match, mismatch, open, extend, strategy,
                                                  tmp.data(), cigar_length, &offset, nullptr);
    if (st != SW_CUDA_OK) return INT_MIN;
    std::memcpy(cigar, tmp.data(), (size_t)cigar_length);
    return offset;
}

}  // extern "C"
