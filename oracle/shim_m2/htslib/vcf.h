/* TEST INFRASTRUCTURE ONLY -- stand-in for htslib's <htslib/vcf.h> (see hts.h beside this file): opaque types only; the
 * oracle never compiles the reference's VCF writer. */
#ifndef SHIM_HTSLIB_VCF_H
#define SHIM_HTSLIB_VCF_H
#include "hts.h"
typedef struct bcf_hdr_t bcf_hdr_t;
typedef struct bcf1_t bcf1_t;
#endif
