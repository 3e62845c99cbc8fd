/* TEST INFRASTRUCTURE ONLY -- stand-in for htslib's <htslib/faidx.h> (see hts.h beside this file): declarations only. */
#ifndef SHIM_HTSLIB_FAIDX_H
#define SHIM_HTSLIB_FAIDX_H
#include "hts.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef struct faidx_t faidx_t;
faidx_t *fai_load(const char *fn);
faidx_t *fai_load3(const char *fn, const char *fnfai, const char *fngzi, int flags);
void fai_destroy(faidx_t *fai);
char *fai_fetch(const faidx_t *fai, const char *reg, int *len);
char *faidx_fetch_seq(const faidx_t *fai, const char *c_name, int p_beg_i, int p_end_i, int *len);
char *faidx_fetch_seq64(const faidx_t *fai, const char *c_name, hts_pos_t p_beg_i, hts_pos_t p_end_i, hts_pos_t *len);
int faidx_nseq(const faidx_t *fai);
const char *faidx_iseq(const faidx_t *fai, int i);
int faidx_seq_len(const faidx_t *fai, const char *seq);
#ifdef __cplusplus
}
#endif
#endif
