// Instantiations of forward_kernel<double, K, KL, EXACT=false, UG=false> for every (K, KL) bin.
#include "forward.cuh"

namespace phmm {

cudaError_t launch_f64_gen(int K, int KL, const KernelArgs &a, int grid, cudaStream_t st) {
    switch (K * 8 + KL) {
#define X(KK, LL) case KK * 8 + LL: forward_kernel<double, KK, LL, false, false><<<grid, kThreads, 0, st>>>(a); break;
#include "bins.inc"
#undef X
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

int blocks_per_sm_f64_gen(int K, int KL) {
    int n = 0;
    switch (K * 8 + KL) {
#define X(KK, LL) case KK * 8 + LL: cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, forward_kernel<double, KK, LL, false, false>, kThreads, 0); break;
#include "bins.inc"
#undef X
        default: break;
    }
    return n;
}

}  // namespace phmm
