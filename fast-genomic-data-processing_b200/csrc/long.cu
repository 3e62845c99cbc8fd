// Reads longer than one warp's register block (32 lanes x kMaxRowsPerLane = 384 rows): the DP matrix is swept in ROW
// STRIPES of 32 x kLongRows rows, the last row of a stripe handed to the next one through a per-warp scratch line in
// global memory -- what the reference does with its shiftOutM/X/Y arrays between stripes of one SIMD vector
// (avx-pairhmm-template.h:225,289-293). One warp per (read, haplotype) pair, one column per step, lanes skewed by one
// column; the float pass, the < 1e-28f test, the double pass and the log10 (IntelPairHmm.cc:339-350) all happen here, so
// these pairs bypass finalize / rescue scan. A plain kernel for a rare case (Illumina reads end at 2 x 300): the
// per-row gap-continuation penalty is always read from the row (no uniform-gcp variant), no packing, no tiling.
#include <cmath>

#include "forward.cuh"

namespace phmm {
namespace {

constexpr int kLongRows = 8;                       // rows per lane and stripe
constexpr int kStripe = 32 * kLongRows;

// 2^s-scaled forward sum of one pair in precision T (computeMXY, template.h:177-192)
template <typename T>
__device__ T long_sweep(const LongArgs &a, const DevRead rr, const uint8_t *__restrict__ hap, int H, T *__restrict__ sM,
                        T *__restrict__ sX, T *__restrict__ sY) {
    using N = Num<T>;
    const int lane = threadIdx.x & 31;
    const T *__restrict__ ph, *__restrict__ mmt;
    if constexpr (sizeof(T) == 4) { ph = a.ph2pr_f; mmt = a.mm_f; } else { ph = a.ph2pr_d; mmt = a.mm_d; }
    const T initY = N::div(N::init_const(), (T)H);                                  // template.h:115
    const int R = rr.len, n_stripes = (R + kStripe - 1) / kStripe;
    const int last_stripe = (R - 1) / kStripe, last_lane = ((R - 1) % kStripe) / kLongRows, last_slot = (R - 1) % kLongRows;
    T sumM = (T)0, sumX = (T)0;
    for (int s = 0; s < n_stripes; ++s) {
        // ---- per-row constants of this lane's rows (ReadForPairHMM.cpp:72-80; stripeINITIALIZATION template.h:150-152)
        T pMM[kLongRows], pMX[kLongRows], pMY[kLongRows], pXY[kLongRows], pG[kLongRows], d1[kLongRows], d3[kLongRows];
        uint32_t mm[kLongRows];                                                     // bit b: the row matches haplotype base b (A C T G N)
#pragma unroll
        for (int k = 0; k < kLongRows; ++k) {
            const int g = s * kStripe + lane * kLongRows + k;                       // 0-based read row
            pMM[k] = pMX[k] = pMY[k] = pXY[k] = pG[k] = d1[k] = d3[k] = (T)0; mm[k] = 0u;   // dead slot below the read: e = d3 = 0
            if (g < R) {
                const uint8_t *p = a.rd + rr.off + g;
                const uint32_t rb = base_index(p[0]);
                mm[k] = rb == 4u ? 0x1fu : ((1u << rb) | 0x10u);                    // N on either side matches (template.h:38,48-51)
                const int bq = p[a.plane] & 127, iq = p[2 * a.plane] & 127, dq = p[3 * a.plane] & 127, cq = p[4 * a.plane] & 127;
                const int mx = max(iq, dq), mn = min(iq, dq);
                pMM[k] = mmt[((mx * (mx + 1)) >> 1) + mn];
                pMX[k] = ph[iq]; pMY[k] = ph[dq]; pXY[k] = ph[cq]; pG[k] = N::sub((T)1.0, ph[cq]);
                const T dist = ph[bq];
                d1[k] = N::sub((T)1.0, dist); d3[k] = N::div(dist, (T)3.0);
            }
        }
        T M[kLongRows], X[kLongRows], Y[kLongRows];                                 // this lane's rows at the previous column
#pragma unroll
        for (int k = 0; k < kLongRows; ++k) M[k] = X[k] = Y[k] = (T)0;              // column 0 of rows >= 1
        // the row above this lane's first row, at the column about to be swept (up) and at the one before (diag)
        T upM = (T)0, upX = (T)0, upY = (T)0, dgM = (T)0, dgX = (T)0, dgY = (T)0;
        if (lane == 0) {
            if (s == 0) { upY = initY; dgY = initY; }                                // row 0: M = X = 0, Y = 2^s / H in every column
            else { upM = sM[0]; upX = sX[0]; upY = sY[0]; }                          // last row of the previous stripe, column 1
        }
        const bool sums_here = s == last_stripe && lane == last_lane;
        for (int t = 0; t < H + 31; ++t) {
            const int c = t - lane;                                                  // 0-based haplotype column of this lane
            if (c >= 0 && c < H) {
                const uint32_t hb = base_index(hap[c]);
                T Mn[kLongRows], Xn[kLongRows], Yn[kLongRows];
#pragma unroll
                for (int k = 0; k < kLongRows; ++k) {
                    const T aM = k ? M[k - 1] : dgM, aX = k ? X[k - 1] : dgX, aY = k ? Y[k - 1] : dgY;   // (r-1, c-1)
                    const T uM = k ? Mn[k - 1] : upM, uX = k ? Xn[k - 1] : upX;                           // (r-1, c)
                    const T e = ((mm[k] >> hb) & 1u) ? d1[k] : d3[k];
                    const T sm = N::template mad<false>(aY, pG[k], N::template mad<false>(aX, pG[k], aM * pMM[k]));
                    Mn[k] = sm * e;
                    Xn[k] = N::template mad<false>(uX, pXY[k], uM * pMX[k]);
                    Yn[k] = N::template mad<false>(Y[k], pXY[k], M[k] * pMY[k]);
                }
#pragma unroll
                for (int k = 0; k < kLongRows; ++k) { M[k] = Mn[k]; X[k] = Xn[k]; Y[k] = Yn[k]; }
                if (sums_here) {                                                     // template.h:326,329: increasing columns
#pragma unroll
                    for (int k = 0; k < kLongRows; ++k) if (k == last_slot) { sumM = sumM + M[k]; sumX = sumX + X[k]; }
                }
                if (lane == 31 && s + 1 < n_stripes) {                               // hand the stripe's last row down
                    sM[c] = M[kLongRows - 1]; sX[c] = X[kLongRows - 1]; sY[c] = Y[kLongRows - 1];
                }
                dgM = upM; dgX = upX; dgY = upY;                                     // (r-1, c) becomes the next column's diagonal
            }
            // the lane above has just finished the column this lane sweeps next
            const T nM = __shfl_up_sync(kFull, M[kLongRows - 1], 1), nX = __shfl_up_sync(kFull, X[kLongRows - 1], 1),
                    nY = __shfl_up_sync(kFull, Y[kLongRows - 1], 1);
            if (lane) { upM = nM; upX = nX; upY = nY; }
            else if (s) { const int cn = min(t + 1, H - 1); upM = sM[cn]; upX = sX[cn]; upY = sY[cn]; }
        }
        __syncwarp();                                                                // lane 31's line is read by lane 0 next
    }
    const T r = sumM + sumX;                                                         // template.h:339-343
    return __shfl_sync(kFull, r, last_lane);
}

__global__ void __launch_bounds__(128) forward_long_kernel(const LongArgs a) {
    const int lane = threadIdx.x & 31;
    const uint32_t gwarp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    double *line = reinterpret_cast<double *>(a.scratch + (uint64_t)gwarp * a.scratch_per_warp);
    const uint64_t w = a.scratch_per_warp / (3 * sizeof(double));                   // columns per scratch line
    for (;;) {
        uint32_t u = 0;
        if (lane == 0) u = atomicAdd(a.queue, 1u);
        u = __shfl_sync(kFull, u, 0);
        if (u >= a.n_units) break;
        const DevUnit unit = a.units[u];                                             // one read x one haplotype
        const DevRead rr = a.reads[unit.read];
        const uint32_t hid = a.hap_list[unit.hap_list];
        const int H = a.haps[hid].len;
        const uint8_t *hap = a.hap_bytes + a.hap_byte_off[hid];
        double out;
        uint8_t flag = 0;
        float rf = 0.f;
        if (H <= 0) rf = 0.f;                                                        // no column: the sum over columns is empty
        else if (!a.use_double)
            rf = long_sweep<float>(a, rr, hap, H, reinterpret_cast<float *>(line), reinterpret_cast<float *>(line + w),
                                   reinterpret_cast<float *>(line + 2 * w));
        if (a.use_double || rf < 1e-28f) {                                           // IntelPairHmm.cc:340-344
            const double rd = H <= 0 ? 0.0 : long_sweep<double>(a, rr, hap, H, line, line + w, line + 2 * w);
            out = log10(flush_denormal(rd)) - a.log10_init_d;
            flag = 1;
        } else {
            out = (double)__fsub_rn((float)log10((double)rf), a.log10_init_f);       // as finalize_kernel; IntelPairHmm.cc:348
        }
        if (lane == 0) { a.out[unit.out] = out; a.flag[unit.out] = flag; }
        __syncwarp();
    }
}

}  // namespace

cudaError_t launch_forward_long(const LongArgs &a, int grid_blocks, cudaStream_t st) {
    if (a.n_units == 0) return cudaSuccess;
    forward_long_kernel<<<grid_blocks, 128, 0, st>>>(a);
    return cudaGetLastError();
}

}  // namespace phmm

// ------------------------------------------------------------------------------------------------------------------------
// pairhmm_cuda_forward_pairs_noalloc -- the row-striped kernel above for callers that must not allocate (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE.
// Everything the kernel needs besides the caller's own device arrays lives in file-scope
// __device__ storage that every GPU loading this library holds until exit:
//   g_na_scratch  the work-queue counter (first 256 bytes) and, per resident warp, the three stripe hand-over lines of
//                 long_sweep: kNaWarps x 3 x (PAIRHMM_CUDA_NOALLOC_MAX_HAP_LEN + 8) doubles = 24.1 MiB;
//   g_na_tables   ph2pr / matchToMatchProb in float and double (Context.h:75-89,139,180), 98 KiB, uploaded at set-up.
// In host code those names are host-side stand-ins, not device addresses: at set-up -- once per device, with that device
// current -- cudaGetSymbolAddress gives the real addresses, which are what the kernel is handed. Every call shares the
// arrays; the queue counter is cleared on the caller's stream before every launch (the lines need no clearing: a stripe
// writes a column before the next stripe reads it). Calls that overlap on the GPU (different streams) RACE on both.
// ------------------------------------------------------------------------------------------------------------------------
#include <algorithm>
#include <cstdio>
#include <mutex>

#include "tables.h"

namespace phmm {
namespace {

constexpr int kNaBlocks = 128, kNaWarps = kNaBlocks * 4;                                  // resident warps of the no-alloc launch
constexpr uint64_t kNaLineCols = PAIRHMM_CUDA_NOALLOC_MAX_HAP_LEN + 8;
constexpr uint64_t kNaPerWarp = 3 * kNaLineCols * sizeof(double);                          // 49 344 B, a multiple of 64
constexpr uint64_t kNaHead = 256;                                                          // queue counter
constexpr uint64_t kNaScratchBytes = kNaHead + (uint64_t)kNaWarps * kNaPerWarp;

static __device__ __align__(256) unsigned char g_na_scratch[kNaScratchBytes];
static __device__ __align__(256) HostTables g_na_tables;

struct NoAllocArgs {
    LongArgs base;                  // rd / plane / tables / log10 constants / out / flag / use_double / scratch
    const uint32_t *read_off; const int32_t *read_len;
    const uint64_t *hap_off; const int32_t *hap_len;
    const uint32_t *pair_read, *pair_hap;
    uint32_t n_pairs;
};

__global__ void __launch_bounds__(128) forward_long_noalloc_kernel(const NoAllocArgs n) {
    const LongArgs &a = n.base;
    const int lane = threadIdx.x & 31;
    const uint32_t gwarp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    double *line = reinterpret_cast<double *>(a.scratch + kNaHead + (uint64_t)gwarp * kNaPerWarp);
    const uint64_t w = kNaLineCols;
    for (;;) {
        uint32_t u = 0;
        if (lane == 0) u = atomicAdd(a.queue, 1u);
        u = __shfl_sync(kFull, u, 0);
        if (u >= n.n_pairs) break;
        const uint32_t r = n.pair_read[u], hid = n.pair_hap[u];
        const DevRead rr{n.read_off[r], n.read_len[r]};
        const int H = n.hap_len[hid];
        const uint8_t *hap = a.hap_bytes + n.hap_off[hid];
        double out;
        uint8_t flag = 0;
        if (H > PAIRHMM_CUDA_NOALLOC_MAX_HAP_LEN) {                  // longer than the caller stated: no scratch line for it
            out = __longlong_as_double(0x7ff8000000000000ll); flag = 255;
        } else if (rr.len <= 0) {                                    // zero-length read: no alignment exists (as the engine reports it)
            out = __longlong_as_double(0xfff0000000000000ll); flag = 1;
        } else {
            float rf = 0.f;
            if (H > 0 && !a.use_double)
                rf = long_sweep<float>(a, rr, hap, H, reinterpret_cast<float *>(line), reinterpret_cast<float *>(line + w),
                                       reinterpret_cast<float *>(line + 2 * w));
            if (a.use_double || rf < 1e-28f) {                       // IntelPairHmm.cc:340-344
                const double rd = H <= 0 ? 0.0 : long_sweep<double>(a, rr, hap, H, line, line + w, line + 2 * w);
                out = log10(flush_denormal(rd)) - a.log10_init_d;
                flag = 1;
            } else {
                out = (double)__fsub_rn((float)log10((double)rf), a.log10_init_f);   // IntelPairHmm.cc:348
            }
        }
        if (lane == 0) { a.out[u] = out; if (a.flag) a.flag[u] = flag; }
        __syncwarp();
    }
}

// per device: the device addresses of the two arrays (0 = not set up yet)
std::mutex g_na_mutex;
struct NaDevice { unsigned char *scratch = nullptr; const HostTables *tables = nullptr; };
NaDevice g_na_dev[64];

void na_error(char *buf, size_t len, const char *msg, const char *detail = nullptr) {
    if (!buf || !len) return;
    snprintf(buf, len, "%s%s%s", msg, detail ? ": " : "", detail ? detail : "");
}

}  // namespace
}  // namespace phmm

extern "C" int pairhmm_cuda_forward_pairs_noalloc(int device, void *cuda_stream, const uint8_t *d_read_planes, uint64_t plane_stride,
                                                  const uint32_t *d_read_off, const int32_t *d_read_len,
                                                  const uint8_t *d_hap_bases, const uint64_t *d_hap_off, const int32_t *d_hap_len,
                                                  const uint32_t *d_pair_read, const uint32_t *d_pair_hap, uint32_t n_pairs,
                                                  int32_t max_hap_len, int32_t use_double, double *d_log10_out,
                                                  uint8_t *d_rescued_out, char *error_buf, size_t error_buf_len) {
    using namespace phmm;
    if (error_buf && error_buf_len) error_buf[0] = 0;
    if (n_pairs == 0) return PAIRHMM_OK;
    if (!d_read_planes || !d_read_off || !d_read_len || !d_hap_bases || !d_hap_off || !d_hap_len || !d_pair_read || !d_pair_hap ||
        !d_log10_out) {
        na_error(error_buf, error_buf_len, "pairhmm_cuda_forward_pairs_noalloc: null device pointer");
        return PAIRHMM_ERR_ARG;
    }
    if (max_hap_len < 0 || max_hap_len > PAIRHMM_CUDA_NOALLOC_MAX_HAP_LEN) {
        char msg[160];
        snprintf(msg, sizeof msg, "pairhmm_cuda_forward_pairs_noalloc: max_hap_len %d exceeds the static scratch of this build "
                 "(PAIRHMM_CUDA_NOALLOC_MAX_HAP_LEN = %d); use the allocating entry points", max_hap_len, PAIRHMM_CUDA_NOALLOC_MAX_HAP_LEN);
        na_error(error_buf, error_buf_len, msg);
        return PAIRHMM_ERR_ARG;
    }
    int dev = device;
    cudaError_t e;
    if (dev < 0) { if ((e = cudaGetDevice(&dev)) != cudaSuccess) { cudaGetLastError(); na_error(error_buf, error_buf_len, "cudaGetDevice", cudaGetErrorString(e)); return PAIRHMM_ERR_NO_DEVICE; } }
    else if ((e = cudaSetDevice(dev)) != cudaSuccess) { cudaGetLastError(); na_error(error_buf, error_buf_len, "cudaSetDevice", cudaGetErrorString(e)); return PAIRHMM_ERR_ARG; }
    if (dev >= 64) { na_error(error_buf, error_buf_len, "pairhmm_cuda_forward_pairs_noalloc: device ordinal above 63"); return PAIRHMM_ERR_ARG; }
    NaDevice nd;
    {   // set-up, once per device, with that device current: the arrays' DEVICE addresses, and the lookup tables. (The first
        // call on a device must therefore not sit inside a stream capture: the table upload is a synchronous copy.)
        std::lock_guard<std::mutex> lk(g_na_mutex);
        if (!g_na_dev[dev].scratch) {
            unsigned char *sp = nullptr; HostTables *tp = nullptr;
            if ((e = cudaGetSymbolAddress((void **)&sp, g_na_scratch)) != cudaSuccess ||
                (e = cudaGetSymbolAddress((void **)&tp, g_na_tables)) != cudaSuccess) {
                cudaGetLastError(); na_error(error_buf, error_buf_len, "cudaGetSymbolAddress", cudaGetErrorString(e)); return PAIRHMM_ERR_CUDA;
            }
            const HostTables &t = host_tables();
            if ((e = cudaMemcpy(tp, &t, sizeof t, cudaMemcpyHostToDevice)) != cudaSuccess) {
                cudaGetLastError(); na_error(error_buf, error_buf_len, "table upload", cudaGetErrorString(e)); return PAIRHMM_ERR_CUDA;
            }
            g_na_dev[dev].tables = tp; g_na_dev[dev].scratch = sp;
        }
        nd = g_na_dev[dev];
    }
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    // the work-queue counter must start at zero; the lines keep the last call's contents and need no clearing
    if ((e = cudaMemsetAsync(nd.scratch, 0, kNaHead, st)) != cudaSuccess) { cudaGetLastError(); na_error(error_buf, error_buf_len, "cudaMemsetAsync", cudaGetErrorString(e)); return PAIRHMM_ERR_CUDA; }
    NoAllocArgs n{};
    LongArgs &a = n.base;
    a.rd = d_read_planes; a.plane = plane_stride; a.hap_bytes = d_hap_bases;
    a.queue = reinterpret_cast<uint32_t *>(nd.scratch);
    a.ph2pr_f = nd.tables->ph2pr_f; a.mm_f = nd.tables->mm_f; a.ph2pr_d = nd.tables->ph2pr_d; a.mm_d = nd.tables->mm_d;   // address arithmetic only
    a.log10_init_f = log10f(ldexpf(1.f, 120)); a.log10_init_d = log10(ldexp(1.0, 1020));                                 // Context.h:184,143
    a.scratch = nd.scratch; a.scratch_per_warp = kNaPerWarp;
    a.out = d_log10_out; a.flag = d_rescued_out; a.use_double = use_double ? 1 : 0;
    n.read_off = d_read_off; n.read_len = d_read_len; n.hap_off = d_hap_off; n.hap_len = d_hap_len;
    n.pair_read = d_pair_read; n.pair_hap = d_pair_hap; n.n_pairs = n_pairs;
    const int grid = (int)std::min<uint64_t>((uint64_t)kNaBlocks, ((uint64_t)n_pairs + 3) / 4);
    forward_long_noalloc_kernel<<<grid, 128, 0, st>>>(n);
    if ((e = cudaGetLastError()) != cudaSuccess) { na_error(error_buf, error_buf_len, "kernel launch", cudaGetErrorString(e)); return PAIRHMM_ERR_CUDA; }
    return PAIRHMM_OK;
}
