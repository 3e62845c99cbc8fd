// Auxiliary kernels (haplotype byte -> column-word expansion, FP32 result finalisation) and the
// dispatch from (variant, K, KL) to the forward_kernel instantiations of inst_*.cu. The forward kernel
// itself is in forward.cuh.
#include "forward2.cuh"

namespace phmm {
namespace {

// code word of a haplotype base: b * group_bits(K) for every bucket, each in its own 5-bit field
__device__ __forceinline__ uint32_t code_word(uint32_t b) {
    return (5u * b) | ((4u * b) << 5) | ((6u * b) << 10) | ((3u * b) << 15) | ((2u * b) << 20) | (b << 25);
}

// ---------------------------------------------------------------------------------------------
// haplotype bytes -> code words (+ PADs up to the tile boundary) + the [SEP, PAD..] tile; block per haplotype; block 0 also
// writes the fill/drain words.
__global__ void prep_haps_kernel(const uint8_t *__restrict__ hap_bytes, const uint64_t *__restrict__ byte_off,
                                 const DevHap *__restrict__ haps, uint32_t n_haps, uint32_t *__restrict__ hapw) {
    if (blockIdx.x == 0)
        for (int i = threadIdx.x; i < kLeadWords; i += blockDim.x) hapw[i] = (i == 32 * kTile - 1) ? kWordSep : kWordPad;
    for (uint32_t h = blockIdx.x; h < n_haps; h += gridDim.x) {
        const DevHap d = haps[h];
        const uint8_t *src = hap_bytes + byte_off[h];
        uint32_t *dst = hapw + d.off;
        const int nw = (int)hap_tiles(d.len) * kTile, sep = nw - kTile;
        for (int c = threadIdx.x; c < nw; c += blockDim.x)
            dst[c] = (c < d.len) ? code_word(base_index(src[c])) : (c == sep ? kWordSep : kWordPad);
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

}  // namespace

cudaError_t launch_prep_haps(const uint8_t *hap_bytes, const uint64_t *byte_off, const DevHap *haps,
                             uint32_t n_haps, uint32_t *hapw, cudaStream_t st) {
    const int grid = n_haps == 0 ? 1 : (int)(n_haps < 65535u ? n_haps : 65535u);
    prep_haps_kernel<<<grid, 128, 0, st>>>(hap_bytes, byte_off, haps, n_haps, hapw);
    return cudaGetLastError();
}

#include "inst_dispatch.inc"

static const InstUnit *find_unit(Variant v, int W, int K) {
    for (const InstUnit &u : kInstUnits)
        if (u.v == v && u.W == W && K >= u.k0 && K <= u.k1) return &u;
    return nullptr;
}

bool forward_available(Variant v, int W, int K) { return find_unit(v, W, K) != nullptr; }

cudaError_t launch_forward(Variant v, int W, int K, int KL, const KernelArgs &a, int grid, cudaStream_t st) {
    const InstUnit *u = find_unit(v, W, K);
    return u ? u->launch(K, KL, a, grid, st) : cudaErrorInvalidValue;
}

int forward_blocks_per_sm(Variant v, int W, int K, int KL) {
    const InstUnit *u = find_unit(v, W, K);
    return u ? u->blocks(K, KL) : 0;
}

cudaError_t launch_finalize(const FinalizeArgs &a, cudaStream_t st) {
    if (a.n == 0) return cudaSuccess;
    finalize_kernel<<<(a.n + 255) / 256, 256, 0, st>>>(a);
    return cudaGetLastError();
}

}  // namespace phmm
