/* TEST INFRASTRUCTURE ONLY -- scalar C restatement of the reference PairHMM (see pairhmm_oracle.h).
 * Compile with -ffp-contract=off: the reference's AVX-256 objects contain no FMA (built -mavx only,
 * M2/intel/pairhmm/CMakeLists.txt:7), and this file is meant to agree with them bit for bit.
 * M2 = /root/reference/deepmutect/Mutect2Cpp-master/src. */
#include "pairhmm_oracle.h"
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#if defined(__x86_64__) || defined(__i386__)
#include <xmmintrin.h>
#define FTZ_ON() unsigned _csr = _mm_getcsr(); _mm_setcsr(_csr | 0x8000u) /* IntelPairHmm.cc:230 */
#define FTZ_RESTORE() _mm_setcsr(_csr)
#else
#define FTZ_ON() (void)0
#define FTZ_RESTORE() (void)0
#endif

/* ---- tables: Context.h:30-34 constants ---- */
#define MAX_QUAL 254
#define JAC_TOL 8.0
#define JAC_STEP 0.0001
#define JAC_INV_STEP (1.0 / JAC_STEP)
#define JAC_SIZE 80001 /* (int)(8.0/0.0001)+1 */

static float g_ph2pr_f[128], g_mm_f[ORACLE_MM_COUNT], g_jac_f[JAC_SIZE];
static double g_ph2pr_d[128], g_mm_d[ORACLE_MM_COUNT], g_jac_d[JAC_SIZE];
static pthread_once_t g_once = PTHREAD_ONCE_INIT;

/* Context.h:92-121, NUMBER = float */
static float approx_log10_sum_f(float small, float big) {
    if (small > big) { float t = big; big = small; small = t; }
    if (isinf(small) || isinf(big)) return big;
    float diff = big - small;
    if (diff >= (float)JAC_TOL) return big;
    float d = (float)(diff * ((float)JAC_INV_STEP));
    int ind = (d > 0.0f) ? (int)(d + 0.5f) : (int)(d - 0.5f); /* fastRound, Context.h:87-90 */
    return big + g_jac_f[ind];
}
/* Context.h:92-121, NUMBER = double */
static double approx_log10_sum_d(double small, double big) {
    if (small > big) { double t = big; big = small; small = t; }
    if (isinf(small) || isinf(big)) return big;
    double diff = big - small;
    if (diff >= JAC_TOL) return big;
    double d = diff * JAC_INV_STEP;
    int ind = (d > 0.0) ? (int)(d + 0.5) : (int)(d - 0.5);
    return big + g_jac_d[ind];
}

static void build_tables(void) {
    const double INV_LN10 = 0.434294; /* truncated on purpose, Context.h:76 */
    for (int k = 0; k < JAC_SIZE; k++) { /* Context.h:65-72 */
        double v = log10(1.0 + pow(10.0, -((double)k) * JAC_STEP));
        g_jac_d[k] = v;
        g_jac_f[k] = (float)v;
    }
    for (int i = 0, offset = 0; i <= MAX_QUAL; offset += ++i) /* Context.h:78-88 */
        for (int j = 0; j <= i; j++) {
            double sf = approx_log10_sum_f((float)-0.1 * (float)i, (float)-0.1 * (float)j);
            double lf = log1p(-fmin(1.0, pow(10, sf))) * INV_LN10;
            g_mm_f[offset + j] = (float)pow(10, lf);
            double sd = approx_log10_sum_d((double)-0.1 * (double)i, (double)-0.1 * (double)j);
            double ld = log1p(-fmin(1.0, pow(10, sd))) * INV_LN10;
            g_mm_d[offset + j] = pow(10, ld);
        }
    for (int x = 0; x < 128; x++) { /* Context.h:137-140,178-181 */
        g_ph2pr_d[x] = pow(10.0, -((double)x) / 10.0);
        g_ph2pr_f[x] = powf(10.f, -((float)x) / 10.f);
    }
}

void oracle_tables(float *ph2pr_f, float *mm_f, double *ph2pr_d, double *mm_d, int mm_count) {
    pthread_once(&g_once, build_tables);
    memcpy(ph2pr_f, g_ph2pr_f, sizeof g_ph2pr_f);
    memcpy(ph2pr_d, g_ph2pr_d, sizeof g_ph2pr_d);
    memcpy(mm_f, g_mm_f, (size_t)mm_count * sizeof(float));
    memcpy(mm_d, g_mm_d, (size_t)mm_count * sizeof(double));
}

/* triangular lookup of set_mm_prob (Context.h:156-167): quals are <128 so always the table branch */
static inline int mm_index(int ins, int del) {
    int mn = del, mx = ins;
    if (ins <= del) { mn = ins; mx = del; }
    return ((mx * (mx + 1)) >> 1) + mn;
}

/* pairhmm_common.h:69-85: A0 C1 T2 G3 N4, every other byte behaves as 'A' (zero-initialised table) */
static inline int base_code(uint8_t b) {
    switch (b) { case 'A': return 0; case 'C': return 1; case 'T': return 2; case 'G': return 3; case 'N': return 4; default: return 0; }
}

/* The forward recurrence, one macro body for both precisions (avx-pairhmm-template.h:110-121,
 * 150-169, 177-192, 310-344 restated row-major). */
#define FORWARD_BODY(T, PH2PR, MM, INITC)                                                        \
    pthread_once(&g_once, build_tables);                                                         \
    if (R <= 0 || H <= 0) return (T)0;                                                           \
    T *buf = (T *)malloc(sizeof(T) * 6 * (size_t)(H + 1));                                       \
    T *Mp = buf, *Xp = Mp + (H + 1), *Yp = Xp + (H + 1), *Mc = Yp + (H + 1), *Xc = Mc + (H + 1), \
      *Yc = Xc + (H + 1);                                                                        \
    uint8_t *hc = (uint8_t *)malloc((size_t)H);                                                  \
    for (int c = 0; c < H; c++) hc[c] = (uint8_t)base_code(hap[c]);                              \
    const T initY = (INITC) / (T)H;                                                              \
    for (int c = 0; c <= H; c++) { Mp[c] = (T)0; Xp[c] = (T)0; Yp[c] = initY; }                  \
    for (int r = 1; r <= R; r++) {                                                               \
        const int iq = ins_q[r - 1] & 127, dq = del_q[r - 1] & 127, cq = gcp[r - 1] & 127,       \
                  bq = base_q[r - 1] & 127;                                                      \
        const T pMM = MM[mm_index(iq, dq)], pXX = PH2PR[cq], pYY = PH2PR[cq], pMX = PH2PR[iq],   \
                pMY = PH2PR[dq], pGAPM = (T)1.0 - PH2PR[cq], dist = PH2PR[bq];                   \
        const T d1 = (T)1.0 - dist, d3 = dist / (T)3.0;                                          \
        const int rc = base_code(read[r - 1]);                                                   \
        Mc[0] = Xc[0] = Yc[0] = (T)0;                                                            \
        for (int c = 1; c <= H; c++) {                                                           \
            const int h = hc[c - 1];                                                             \
            const T e = (rc == h || rc == 4 || h == 4) ? d1 : d3;                                \
            T t0 = Mp[c - 1] * pMM;                                                              \
            T t1 = Xp[c - 1] * pGAPM;                                                            \
            T t2 = Yp[c - 1] * pGAPM;                                                            \
            Mc[c] = ((t0 + t1) + t2) * e;                                                        \
            T u0 = Mp[c] * pMX, u1 = Xp[c] * pXX;                                                \
            Xc[c] = u0 + u1;                                                                     \
            T v0 = Mc[c - 1] * pMY, v1 = Yc[c - 1] * pYY;                                        \
            Yc[c] = v0 + v1;                                                                     \
        }                                                                                        \
        T *t;                                                                                    \
        t = Mp; Mp = Mc; Mc = t; t = Xp; Xp = Xc; Xc = t; t = Yp; Yp = Yc; Yc = t;               \
    }                                                                                            \
    T sumM = (T)0, sumX = (T)0;                                                                  \
    for (int c = 1; c <= H; c++) { sumM = sumM + Mp[c]; sumX = sumX + Xp[c]; }                   \
    T res = sumM + sumX;                                                                         \
    free(buf); free(hc);                                                                         \
    return res;

float oracle_forward_float(const uint8_t *read, const uint8_t *base_q, const uint8_t *ins_q,
                           const uint8_t *del_q, const uint8_t *gcp, int R, const uint8_t *hap, int H) {
    FORWARD_BODY(float, g_ph2pr_f, g_mm_f, ldexpf(1.f, 120))
}
double oracle_forward_double(const uint8_t *read, const uint8_t *base_q, const uint8_t *ins_q,
                             const uint8_t *del_q, const uint8_t *gcp, int R, const uint8_t *hap, int H) {
    FORWARD_BODY(double, g_ph2pr_d, g_mm_d, ldexp(1.0, 1020))
}

/* ---- flat pair list driver (IntelPairHmm.cc:332-351 dispatch) ---- */
typedef struct {
    const uint8_t *rb, *bq, *iq, *dq, *gc, *hb;
    const int64_t *roff, *hoff;
    const int32_t *rlen, *hlen;
    const uint32_t *pr, *ph;
    int64_t n_pairs;
    int use_double, tid, nth;
    double *out;
    uint8_t *rescued;
} job_t;

static void *worker(void *arg) {
    job_t *j = (job_t *)arg;
    FTZ_ON();
    const float LOG10_INIT_F = log10f(ldexpf(1.f, 120));   /* Context.h:184 */
    const double LOG10_INIT_D = log10(ldexp(1.0, 1020));   /* Context.h:143 */
    for (int64_t p = j->tid; p < j->n_pairs; p += j->nth) {
        const int64_t ro = j->roff[j->pr[p]], ho = j->hoff[j->ph[p]];
        const int R = j->rlen[j->pr[p]], H = j->hlen[j->ph[p]];
        float f = j->use_double ? 0.0f
                                : oracle_forward_float(j->rb + ro, j->bq + ro, j->iq + ro, j->dq + ro,
                                                       j->gc + ro, R, j->hb + ho, H);
        if (f < 1e-28f) { /* MIN_ACCEPTED, pairhmm_common.h:41 */
            double d = oracle_forward_double(j->rb + ro, j->bq + ro, j->iq + ro, j->dq + ro, j->gc + ro,
                                             R, j->hb + ho, H);
            j->out[p] = log10(d) - LOG10_INIT_D;
            if (j->rescued) j->rescued[p] = 1;
        } else {
            j->out[p] = (double)(log10f(f) - LOG10_INIT_F);
            if (j->rescued) j->rescued[p] = 0;
        }
    }
    FTZ_RESTORE();
    return NULL;
}

double oracle_compute(const uint8_t *read_bases, const uint8_t *base_q, const uint8_t *ins_q,
                      const uint8_t *del_q, const uint8_t *gcp, const int64_t *read_off,
                      const int32_t *read_len, int n_reads, const uint8_t *hap_bases,
                      const int64_t *hap_off, const int32_t *hap_len, int n_haps,
                      const uint32_t *pair_read, const uint32_t *pair_hap, int64_t n_pairs,
                      int use_double, int n_threads, int tables_in_timed_region, double *out,
                      uint8_t *rescued) {
    (void)n_reads; (void)n_haps; (void)tables_in_timed_region; /* the port builds row constants inline */
    pthread_once(&g_once, build_tables);
    if (n_threads < 1) n_threads = 1;
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)n_threads);
    job_t *jobs = (job_t *)malloc(sizeof(job_t) * (size_t)n_threads);
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int t = 0; t < n_threads; t++) {
        job_t j = {read_bases, base_q, ins_q, del_q, gcp, hap_bases, read_off, hap_off, read_len, hap_len,
                   pair_read, pair_hap, n_pairs, use_double, t, n_threads, out, rescued};
        jobs[t] = j;
        pthread_create(&th[t], NULL, worker, &jobs[t]);
    }
    for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    free(th); free(jobs);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

/* raw scaled sums of the FP32 pass for a flat pair list (FTZ on), for bit-exactness tests */
void oracle_forward_float_pairs(const uint8_t *read_bases, const uint8_t *base_q, const uint8_t *ins_q,
                                const uint8_t *del_q, const uint8_t *gcp, const int64_t *read_off,
                                const int32_t *read_len, const uint8_t *hap_bases, const int64_t *hap_off,
                                const int32_t *hap_len, const uint32_t *pair_read, const uint32_t *pair_hap,
                                int64_t n_pairs, float *out) {
    FTZ_ON();
    for (int64_t p = 0; p < n_pairs; p++) {
        const int64_t ro = read_off[pair_read[p]], ho = hap_off[pair_hap[p]];
        out[p] = oracle_forward_float(read_bases + ro, base_q + ro, ins_q + ro, del_q + ro, gcp + ro,
                                      read_len[pair_read[p]], hap_bases + ho, hap_len[pair_hap[p]]);
    }
    FTZ_RESTORE();
}

/* ---- per-read epilogue (see pairhmm_oracle.h) ---- */
void oracle_epilogue(double *m, int64_t n_reads, int32_t n_haps, const int32_t *read_len, double likelihood_cap,
                     double max_error_per_base, int natural_log, double *best_out, double *second_out,
                     int32_t *best_hap, uint8_t *poorly_modeled) {
    const double ln10 = log(10.0);                         /* MathUtils.cpp:17 */
    for (int64_t r = 0; r < n_reads; r++) {
        double *v = m + r * (int64_t)n_haps;
        double best = -INFINITY, second = -INFINITY;
        int bi = -1;
        uint8_t poorly = 1;
        if (n_haps > 0) {
            /* searchBestAllele, AlleleLikelihoods.h:101-117: strict '>' keeps the first maximum */
            best = v[0]; bi = 0;
            for (int a = 1; a < n_haps; a++) {
                const double c = v[a];
                if (c > best) { second = best; best = c; bi = a; }
                else if (c > second) second = c;
            }
            /* normalizeLikelihoods returns early for an infinite cap or <= 1 allele (:375-380); otherwise every value
             * below best + cap is raised to it (:155-165) */
            if (!isinf(likelihood_cap) && n_haps > 1) {
                const double worst = best + likelihood_cap;
                for (int a = 0; a < n_haps; a++) if (v[a] < worst) v[a] = worst;
                if (second < worst) second = worst;
            }
            /* maximumLikelihoodOverAllAlleles(s, e) < log10MinTrueLikelihood(read, rate), :413 */
            {
                double max_err = ceil((double)read_len[r] * max_error_per_base);
                if (max_err > 2.0) max_err = 2.0;          /* min(2.0, ceil(...)), PHLCE.cpp:296 */
                poorly = best < max_err * -4.0 ? 1 : 0;    /* log10QualPerBase = -4.0, PHLCE.cpp:297-298 */
            }
            if (natural_log) {
                for (int a = 0; a < n_haps; a++) v[a] = v[a] * ln10;
                best = best * ln10; second = second * ln10;
            }
        }
        if (best_out) best_out[r] = best;
        if (second_out) second_out[r] = second;
        if (best_hap) best_hap[r] = bi;
        if (poorly_modeled) poorly_modeled[r] = poorly;
    }
}

/* ---- read quality pre-pass (see pairhmm_oracle.h) ---- */
/* GATKVariantContextUtils::findNumberOfRepetitions, GATKVariantContextUtils.cpp:59-100 */
static int count_repetitions(const uint8_t *unit_full, int unit_off, int unit_len, const uint8_t *test_full, int test_off,
                             int test_len, int leading) {
    if (test_len == 0) return 0;
    const int diff = test_len - unit_len;
    int n = 0;
    if (leading) {
        for (int start = 0; start <= diff; start += unit_len) {
            if (memcmp(test_full + test_off + start, unit_full + unit_off, (size_t)unit_len) == 0) n++;   /* Utils::equalRange */
            else return n;
        }
    } else {
        for (int start = diff; start >= 0; start -= unit_len) {
            if (memcmp(test_full + test_off + start, unit_full + unit_off, (size_t)unit_len) == 0) n++;
            else return n;
        }
    }
    return n;
}

/* PairHMMLikelihoodCalculationEngine::findTandemRepeatUnits, PHLCE.cpp:177-254 (MAX_STR_UNIT_LENGTH 8, MAX_REPEAT_LENGTH 20) */
int oracle_tandem_repeat_length(const uint8_t *bases, int len, int offset) {
    int max_bw = 0, bw_off = offset, bw_len = 1;              /* bestBWRepeatUnit = {readBases[offset]} */
    for (int str = 1; str <= 8; str++) {
        if (offset + 1 - str < 0) break;
        max_bw = count_repetitions(bases, offset - str + 1, str, bases, 0, offset + 1, 0);
        if (max_bw > 1) { bw_off = offset - str + 1; bw_len = str; break; }
    }
    int max_rl = max_bw;
    if (offset < len - 1) {
        int max_fw = 0, fw_off = offset + 1, fw_len = 1;      /* bestFWRepeatUnit = {readBases[offset+1]} */
        for (int str = 1; str <= 8; str++) {
            if (offset + str + 1 > len) break;
            max_fw = count_repetitions(bases, offset + 1, str, bases, offset + 1, len - offset - 1, 1);
            if (max_fw > 1) { fw_off = offset + 1; fw_len = str; break; }
        }
        if (fw_len == bw_len && memcmp(bases + fw_off, bases + bw_off, (size_t)fw_len) == 0) {
            max_rl = max_fw + max_bw;                          /* in the middle of a tandem repeat */
        } else {
            /* does the FW unit also repeat backwards from here? (:244-247) */
            max_bw = count_repetitions(bases + fw_off, 0, fw_len, bases, 0, offset + 1, 0);
            max_rl = max_fw + max_bw;
        }
    }
    return max_rl > 20 ? 20 : max_rl;
}

/* getErrorModelAdjustedQual / initializePCRErrorModel, PHLCE.cpp:45-61; MathUtils::fastRound MathUtils.cpp:65-68 */
void oracle_pcr_table(int pcr_rate_factor, uint8_t table[21]) {
    for (int i = 0; i <= 20; i++) {
        const double d = 40.0 - exp((double)i / ((double)pcr_rate_factor * 3.14159265358979323846 /* M_PI */));
        const int r = d > 0.0 ? (int)(d + 0.5) : (int)(d - 0.5);
        const int q = r + 1 > 10 ? r + 1 : 10;                 /* MIN_ADJUSTED_QSCORE */
        table[i] = (uint8_t)(char)q;
    }
}

void oracle_prepare_read(const uint8_t *bases, const uint8_t *base_q, const uint8_t *ins_q, const uint8_t *del_q,
                         int len, int mapq, int pcr_rate_factor, int base_q_threshold, int constant_gcp,
                         uint8_t *out_base_q, uint8_t *out_ins_q, uint8_t *out_del_q, uint8_t *out_gcp) {
    uint8_t table[21];
    for (int i = 0; i < len; i++) {
        out_base_q[i] = base_q[i];
        out_ins_q[i] = ins_q ? ins_q[i] : 45;                  /* ReadUtils.cpp:254-262 */
        out_del_q[i] = del_q ? del_q[i] : 45;
        out_gcp[i] = (uint8_t)constant_gcp;                    /* PHLCE.cpp:284-292 */
    }
    if (pcr_rate_factor > 0) {                                 /* applyPCRErrorModel, :150-158: positions 0 .. len-2 */
        oracle_pcr_table(pcr_rate_factor, table);
        for (int i = 1; i < len; i++) {
            const int rl = oracle_tandem_repeat_length(bases, len, i - 1);
            const int q = 0xff & table[rl];
            if (q < (0xff & out_ins_q[i - 1])) out_ins_q[i - 1] = (uint8_t)q;
            if (q < (0xff & out_del_q[i - 1])) out_del_q[i - 1] = (uint8_t)q;
        }
    }
    for (int i = 0; i < len; i++) {                            /* capMinimumReadQualities, :256-267; MIN_USABLE_Q_SCORE 6 */
        int q = 0xff & out_base_q[i];
        if (mapq < q) q = mapq;
        out_base_q[i] = (uint8_t)q;
        if (out_base_q[i] < (uint8_t)base_q_threshold) out_base_q[i] = 6;
        if (out_ins_q[i] < 6) out_ins_q[i] = 6;
        if (out_del_q[i] < 6) out_del_q[i] = 6;
    }
}
