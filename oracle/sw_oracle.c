/* TEST INFRASTRUCTURE ONLY -- scalar C restatement of the reference's Smith-Waterman with backtrace (Intel GKL,
 * M2/intel/smithwaterman/PairWiseSW.h; M2 = /root/reference/deepmutect/Mutect2Cpp-master/src): the checker of the
 * CUDA aligner. Pinned against the reference's own sources compiled unmodified
 * (oracle/_ref/libref_sw.so, oracle/ref_sw_driver.cpp). Never linked or loaded by the product library. */
#include "sw_oracle.h"
#include <limits.h>
#include <stdlib.h>
#include <string.h>

#define BT_MATCH 0
#define BT_INSERT 1
#define BT_DELETE 2
#define BT_INSERT_EXT 4
#define BT_DELETE_EXT 8
#define MATRIX_MIN_CUTOFF (-100000000)
#define LOW_INIT_VALUE (INT_MIN / 2)

static int digits_of(int v) { int d = 0; while (v > 0) { v /= 10; d++; } return d; }   /* fast_itoa(NULL, v), v >= 0 */

int oracle_sw_align(const uint8_t *ref, int nrow, const uint8_t *alt, int ncol, int match, int mismatch, int open,
                    int extend, int strategy, char *cigar, int cigar_len) {
    /* smithWatermanBackTrack, PairWiseSW.h:66-263: rows i = seq1 = ref, columns j = seq2 = alt */
    const int W = ncol + 1;
    int32_t *H = malloc(sizeof(int32_t) * (size_t)(nrow + 1) * W);
    int32_t *E = malloc(sizeof(int32_t) * (size_t)(nrow + 1) * W);
    int32_t *F = malloc(sizeof(int32_t) * (size_t)(nrow + 1) * W);
    uint8_t *bt = calloc((size_t)(nrow + 1) * W, 1);
    const int indel = strategy == SW_INDEL || strategy == SW_LEADING_INDEL;
    for (int j = 0; j <= ncol; j++) { H[j] = (j && indel) ? open + (j - 1) * extend : 0; F[j] = LOW_INIT_VALUE; E[j] = LOW_INIT_VALUE; }
    for (int i = 1; i <= nrow; i++) {
        H[i * W] = indel ? open + (i - 1) * extend : 0;                       /* :214-223 */
        E[i * W] = LOW_INIT_VALUE; F[i * W] = LOW_INIT_VALUE;                   /* :224-225 */
        for (int j = 1; j <= ncol; j++) {                                       /* MAIN_CODE, :27-62 */
            const int32_t ext_h = E[i * W + j - 1] + extend, open_h = H[i * W + j - 1] + open;
            const int32_t e = open_h > ext_h ? open_h : ext_h;
            int b = open_h > ext_h ? 0 : BT_INSERT_EXT;
            const int32_t ext_v = F[(i - 1) * W + j] + extend, open_v = H[(i - 1) * W + j] + open;
            const int32_t f = ext_v > open_v ? ext_v : open_v;
            if (!(open_v > ext_v)) b |= BT_DELETE_EXT;
            const int32_t m = H[(i - 1) * W + j - 1] + (ref[i - 1] == alt[j - 1] ? match : mismatch);
            int32_t h = m > MATRIX_MIN_CUTOFF ? m : MATRIX_MIN_CUTOFF;
            int mv = e > h ? BT_INSERT : BT_MATCH;
            if (e > h) h = e;
            if (f > h) { mv = BT_DELETE; h = f; }
            E[i * W + j] = e; F[i * W + j] = f; H[i * W + j] = h; bt[i * W + j] = (uint8_t)(mv | b);
        }
    }
    /* best cell: last row (SOFTCLIP / IGNORE only) and last column, visited in anti-diagonal order, :227-251 */
    int32_t max_score = INT_MIN;
    int max_i = 0, max_j = 0;
    for (int ad = 1; ad <= nrow + ncol; ad++) {
        if (ad >= nrow + 1) {
            const int j = ad - nrow;
            const int32_t s = H[nrow * W + j];
            if ((strategy == SW_SOFTCLIP || strategy == SW_IGNORE) &&
                (max_score < s || (max_score == s && abs(nrow - j) < abs(max_i - max_j)))) { max_score = s; max_i = nrow; max_j = j; }
        }
        if (ad >= ncol + 1) {
            const int i = ad - ncol;
            const int32_t s = H[i * W + ncol];
            if (max_score < s || (max_score == s && (max_j == ncol || abs(i - ncol) <= abs(max_i - max_j)))) { max_score = s; max_i = i; max_j = ncol; }
        }
    }
    /* getCIGAR, :265-447 */
    int16_t *ca = malloc(sizeof(int16_t) * 4 * (size_t)(nrow + ncol + 4));
    int id = 0, i, j, offset;
    if (strategy == SW_INDEL) { i = nrow; j = ncol; }
    else if (strategy == SW_LEADING_INDEL) { i = max_i; j = ncol; }
    else { i = max_i; j = max_j; }
    if (j < ncol) { ca[0] = SW_SOFTCLIP; ca[1] = (int16_t)(ncol - j); id = 1; }
    int state = 0;
    while (i > 0 && j > 0) {
        const int btr = bt[i * W + j];
        if (state == BT_INSERT_EXT) { j--; ca[id * 2 - 1]++; state = btr & BT_INSERT_EXT; }
        else if (state == BT_DELETE_EXT) { i--; ca[id * 2 - 1]++; state = btr & BT_DELETE_EXT; }
        else switch (btr & 3) {
            case BT_MATCH: i--; j--; ca[id * 2] = BT_MATCH; ca[id * 2 + 1] = 1; state = 0; id++; break;
            case BT_INSERT: j--; ca[id * 2] = BT_INSERT; ca[id * 2 + 1] = 1; state = btr & BT_INSERT_EXT; id++; break;
            case BT_DELETE: i--; ca[id * 2] = BT_DELETE; ca[id * 2 + 1] = 1; state = btr & BT_DELETE_EXT; id++; break;
        }
    }
    if (strategy == SW_SOFTCLIP) {
        if (j > 0) { ca[id * 2] = SW_SOFTCLIP; ca[id * 2 + 1] = (int16_t)j; id++; }
        offset = i;
    } else if (strategy == SW_IGNORE) {
        if (j > 0) { ca[id * 2] = id > 0 ? ca[(id - 1) * 2] : BT_MATCH; ca[id * 2 + 1] = (int16_t)j; id++; }
        offset = (int16_t)(i - j);
    } else {
        if (i > 0) { ca[id * 2] = BT_DELETE; ca[id * 2 + 1] = (int16_t)i; id++; }
        else if (j > 0) { ca[id * 2] = BT_INSERT; ca[id * 2 + 1] = (int16_t)j; id++; }
        offset = 0;
    }
    int cur_size = 0;
    if (id > 0) {
        int nid = 0;                                                             /* merge equal neighbours, :404-421 */
        int16_t prev = ca[0];
        for (int k = 1; k < id; k++) {
            if (ca[k * 2] == prev) ca[nid * 2 + 1] = (int16_t)(ca[nid * 2 + 1] + ca[k * 2 + 1]);
            else { nid++; ca[nid * 2] = ca[k * 2]; ca[nid * 2 + 1] = ca[k * 2 + 1]; prev = ca[k * 2]; }
        }
        for (int k = nid; k >= 0; k--) {                                         /* printed back to front, :423-445 */
            const char st = ca[2 * k] == BT_MATCH ? 'M' : ca[2 * k] == BT_INSERT ? 'I' : ca[2 * k] == BT_DELETE ? 'D'
                          : ca[2 * k] == SW_SOFTCLIP ? 'S' : 'R';
            const int v = ca[2 * k + 1], need = digits_of(v) + 1;
            if (need > 1 && cur_size + need <= cigar_len) {
                int d = digits_of(v), x = v;
                for (int q = d - 1; q >= 0; q--) { cigar[cur_size + q] = (char)('0' + x % 10); x /= 10; }
                cur_size += d;
                cigar[cur_size++] = st;
            }
        }
    }
    if (cur_size < cigar_len) cigar[cur_size] = 0;
    free(H); free(E); free(F); free(bt); free(ca);
    return offset;
}

void oracle_sw_align_batch(const uint8_t *ref_bases, const int64_t *ref_off, const int32_t *ref_len, const uint8_t *alt_bases,
                           const int64_t *alt_off, const int32_t *alt_len, int64_t n, int match, int mismatch, int open,
                           int extend, int strategy, char *cigars, int cigar_stride, int32_t *offsets) {
    for (int64_t p = 0; p < n; p++) {
        memset(cigars + p * cigar_stride, 0, (size_t)cigar_stride);
        offsets[p] = oracle_sw_align(ref_bases + ref_off[p], ref_len[p], alt_bases + alt_off[p], alt_len[p], match, mismatch,
                                     open, extend, strategy, cigars + p * cigar_stride, cigar_stride);
    }
}
