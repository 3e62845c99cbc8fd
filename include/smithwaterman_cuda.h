/* Batched Smith-Waterman with backtrace on sm_100a.
 *
 * Replaces the reference's native aligner behind the same C interface:
 *   smithwaterman_initial() / SmithWaterman_align(ref, refLength, alt, altLength, cigar, cigarLength, match, mismatch,
 *   open, extend, strategy)            -- M2/intel/smithwaterman/IntelSmithWaterman.h:33-34, IntelSmithWaterman.cc:69-79
 * (M2 = deepmutect/Mutect2Cpp-master/src), the Intel-GKL AVX2/AVX-512 kernel of PairWiseSW.h. Mutect2Cpp calls it once
 * per read right after the PairHMM (realign every read to its best haplotype, M2/read/AlignmentUtils.cpp:498, parameters
 * (10, -15, -30, -5), SOFTCLIP) and during assembly (M2/graph/ReadThreadingGraph.cpp:478,698; M2/read/CigarUtils.cpp:52).
 * The batch form below is what a GPU needs: thousands of independent alignments per launch.
 *
 * Results are bit-identical to the reference: same scores, same tie-breaking of the best cell (anti-diagonal visiting
 * order, PairWiseSW.h:227-251), same backtrace state machine and CIGAR text (getCIGAR, :265-447), same offset. */
#ifndef SMITHWATERMAN_CUDA_H
#define SMITHWATERMAN_CUDA_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* overhang strategies, numbered as IntelSmithWaterman::getStrategy does (M2/smithwaterman/IntelSmithWaterman.cpp:22-39) */
#define SW_CUDA_SOFTCLIP 9
#define SW_CUDA_INDEL 10
#define SW_CUDA_LEADING_INDEL 11
#define SW_CUDA_IGNORE 12

#define SW_CUDA_OK 0
#define SW_CUDA_ERR_ARG (-1)
#define SW_CUDA_ERR_CUDA (-2)
#define SW_CUDA_ERR_NO_DEVICE (-4)
#define SW_CUDA_MAX_LEN 4096        /* longest ref / alt sequence of this build */

/* n independent alignments. Pair p aligns alt = alt_bases[alt_off[p] .. +alt_len[p]) to ref = ref_bases[ref_off[p] ..
 * +ref_len[p]); its CIGAR text goes to cigars + p * cigar_stride (NUL-terminated when it fits; the reference's caller
 * passes a zeroed buffer of 2 * max(refLength, altLength) bytes, IntelSmithWaterman.cpp:12-14), its alignment offset to
 * offsets[p]. All pointers are host memory. kernel_ms (nullable) receives the CUDA-event time of the kernel alone.
 * device = CUDA ordinal, -1 = current. Returns SW_CUDA_OK or a negative status; smithwaterman_cuda_last_error() has the text
 * (per host thread). */
int smithwaterman_cuda_align_batch(int device, const uint8_t *ref_bases, const int64_t *ref_off, const int32_t *ref_len,
                                   const uint8_t *alt_bases, const int64_t *alt_off, const int32_t *alt_len, int64_t n,
                                   int match, int mismatch, int open, int extend, int strategy, char *cigars,
                                   int cigar_stride, int32_t *offsets, float *kernel_ms);

/* One alignment with the reference's own signature (drop-in for SmithWaterman_align): returns the offset. A failed call
 * leaves cigar untouched and returns INT32_MIN. */
int smithwaterman_cuda_align(uint8_t *ref, int ref_length, uint8_t *alt, int alt_length, uint8_t *cigar, int cigar_length,
                             int match, int mismatch, int open, int extend, uint8_t strategy);

const char *smithwaterman_cuda_last_error(void);

/* The aligner keeps, per host thread, its stream and device buffers between calls (grow-only), so that a per-region call
 * does not pay for allocations; this frees the calling thread's (they are also freed when the thread ends). The
 * reference has no counterpart (its scratch is per call, PairWiseSW.h:108-125). */
void smithwaterman_cuda_release(void);

#ifdef __cplusplus
}
#endif
#endif /* SMITHWATERMAN_CUDA_H */
