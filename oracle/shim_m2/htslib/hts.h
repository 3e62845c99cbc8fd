/* TEST INFRASTRUCTURE ONLY -- stand-in for htslib's <htslib/hts.h>, which is not installed in this image (htslib is a
 * third-party dependency of Mutect2Cpp, not vendored in the reference tree: CMakeLists.txt:21 points at ../htslib).
 * Declarations only, written from htslib's public API, so that the reference's own translation units on the PairHMM path
 * (SAMRecord.cpp, ReadUtils.cpp, Haplotype.cpp, PairHMMLikelihoodCalculationEngine.cpp, VectorLoglessPairHMM.cpp ...) compile
 * UNMODIFIED for the oracle. No BAM/VCF file is ever opened by the oracle: the I/O entry points are declared, never defined
 * (the linker drops the unreferenced sections that name them, oracle/Makefile --gc-sections). */
#ifndef SHIM_HTSLIB_HTS_H
#define SHIM_HTSLIB_HTS_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef int64_t hts_pos_t;
#define HTS_POS_MAX ((((int64_t)INT_MAX) << 32) | INT_MAX)
#define HTS_POS_MIN INT64_MIN
typedef struct htsFile htsFile;
typedef struct hts_idx_t hts_idx_t;
typedef struct hts_itr_t hts_itr_t;
typedef struct kstring_t { size_t l, m; char *s; } kstring_t;
typedef htsFile samFile;
htsFile *hts_open(const char *fn, const char *mode);
int hts_close(htsFile *fp);
int hts_set_threads(htsFile *fp, int n);
int hts_set_fai_filename(htsFile *fp, const char *fn_aux);
void hts_idx_destroy(hts_idx_t *idx);
void hts_itr_destroy(hts_itr_t *iter);
#ifdef __cplusplus
}
#endif
#endif
