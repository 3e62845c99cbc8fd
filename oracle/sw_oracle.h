/* TEST INFRASTRUCTURE ONLY -- see sw_oracle.c. */
#ifndef SW_ORACLE_H
#define SW_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
/* overhang strategies as IntelSmithWaterman::getStrategy maps them (M2/smithwaterman/IntelSmithWaterman.cpp:22-39,
 * smithwaterman_common.h:52-55) */
#define SW_SOFTCLIP 9
#define SW_INDEL 10
#define SW_LEADING_INDEL 11
#define SW_IGNORE 12
/* SmithWaterman_align (M2/intel/smithwaterman/IntelSmithWaterman.cc:69-79): returns the alignment offset, writes the
 * CIGAR string (NUL-terminated if it fits) */
int oracle_sw_align(const uint8_t *ref, int ref_len, const uint8_t *alt, int alt_len, int match, int mismatch, int open,
                    int extend, int strategy, char *cigar, int cigar_len);
void oracle_sw_align_batch(const uint8_t *ref_bases, const int64_t *ref_off, const int32_t *ref_len, const uint8_t *alt_bases,
                           const int64_t *alt_off, const int32_t *alt_len, int64_t n, int match, int mismatch, int open,
                           int extend, int strategy, char *cigars, int cigar_stride, int32_t *offsets);
#ifdef __cplusplus
}
#endif
#endif
