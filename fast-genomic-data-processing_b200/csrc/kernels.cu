// Auxiliary kernels (haplotype byte -> column-word expansion, FP32 result finalisation, rescue-unit scan) and the
// dispatch from (variant, K, KL) to the forward_kernel instantiations of inst_*.cu. The forward kernel
// itself is in forward.cuh.
#include <algorithm>

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
    const float f = i < a.n ? a.fres[i] : 1.0f;
    const bool resc = f < 1e-28f;                      // the FP64 kernel owns this entry (IntelPairHmm.cc:342)
    const uint32_t m = __ballot_sync(0xffffffffu, resc);
    if (m && (threadIdx.x & 31) == 0) atomicAdd(a.rescued_pairs, (uint32_t)__popc(m));
    if (i >= a.n || resc) return;
    const float l = (float)log10((double)f);          // correctly rounded float log10
    a.out[i] = (double)__fsub_rn(l, a.log10_init_f);
    a.flag[i] = 0;
}

// ---- FP64 rescue lists --------------------------------------------------------------------------------------------
__device__ __forceinline__ int bin_w32(int len) { const int k = (len + 31) >> 5; return k * (k - 1) / 2 + (len - 1) % k; }
__device__ __forceinline__ int bin_w16(int len) { const int k = (len + 15) >> 4; return k * (k - 1) / 2 + (len - 1) % k; }
// a unit of one read for the W = 32 FP64 kernel
__device__ __forceinline__ void emit_single(const RescueScanArgs &a, int bin32, uint32_t read, uint32_t hap_list, uint32_t nh, uint32_t out) {
    const uint32_t slot = a.rescue_base[bin32] + atomicAdd(a.rescue_count + bin32, 1u);
    a.rescue_units[slot] = DevUnit{read, hap_list, unit_shape(nh, 1u, 0u), out};
}
// Candidate `me` (shape = haplotypes | W = 16 bin << 8 | 16-bit hash of its underflow pattern << 16) looks for a read of the same
// bin that needs the same haplotypes of the same run: the table holds, per hash slot, one waiting candidate (index into
// a.pending, + 1). The kernel sweeps the union of what the two reads of a unit need, so two reads with different patterns would
// each pay for the other's haplotypes (BASELINE configs[4]: half a row underflows on average, pairs of any two reads made the
// pass 13 % slower, not faster); a collision of the pattern hash costs time, never a result. Found: both leave as ONE
// unit of the W = 16 kernel, {lower read, run, haplotypes | distance between the reads << 8 | stride << 16}. An occupant that
// does not fit (hash collision, reads too far apart) stays; `me` then leaves alone, as whoever still waits when the scan ends
// does (rescue_flush_kernel). Which reads meet depends on timing; every pair's arithmetic does not.
__device__ void pair_or_wait(const RescueScanArgs &a, const DevUnit me, int len, int neighbourhood, uint32_t stride) {
    const uint32_t idx = atomicAdd(a.pending_count, 1u);
    a.pending[idx] = me;
    __threadfence();
    // partners are sought among the 2^neighbourhood reads around this one: a slot then sees a handful of candidates, not
    // every underflowed read of a deep region (2000 reads x 16 runs on 16 words: the scan took longer than it saved)
    const uint32_t key = me.hap_list * 0x9E3779B1u ^ (me.shape >> 8) * 0x85EBCA6Bu ^ (me.read >> neighbourhood) * 0xC2B2AE35u;
    uint32_t *slot = a.pair_table + ((key >> 12) & (kPairSlots - 1u));
    for (;;) {
        const uint32_t old = atomicCAS(slot, 0u, idx + 1u);
        if (old == 0u) return;                                    // waiting
        const uint4 q = __ldcg(reinterpret_cast<const uint4 *>(a.pending + (old - 1u)));
        const DevUnit c{q.x, q.y, q.z, q.w};
        const uint32_t dist = c.read > me.read ? c.read - me.read : me.read - c.read;
        if (c.hap_list != me.hap_list || c.shape != me.shape || dist == 0u || dist > 255u) {
            emit_single(a, bin_w32(len), me.read, me.hap_list, me.shape & 0xffu, me.out);
            return;
        }
        if (atomicCAS(slot, old, 0u) == old) {
            const DevUnit &lo = c.read < me.read ? c : me;
            const int b16 = (int)((me.shape >> 8) & 0xffu);
            const uint32_t at = a.rescue_base16[b16] + atomicAdd(a.rescue_count16 + b16, 1u);
            a.rescue_units[at] = DevUnit{lo.read, lo.hap_list, (me.shape & 0xffu) | dist << 8 | stride << 16, lo.out};
            return;
        }
    }
}

// One thread per (FP32 work unit, read of the unit): the read's pairs below the threshold -> units of the FP64 pass.
__global__ void rescue_scan_kernel(const RescueScanArgs a) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t u = t / kMaxUnitReads, j = t % kMaxUnitReads;
    if (u >= a.n_units) return;
    const DevUnit unit = a.units[u];
    const uint32_t nh = unit.shape & 0xffu, nreads = (unit.shape >> 8) & 0xffu, stride = unit.shape >> 16;
    if (j >= nreads) return;
    const uint32_t out = unit.out + j * stride;
    const int len = a.reads[unit.read + j].len;
    if (len < 1) return;
    const uint32_t rescued = *a.rescued_pairs;
    // Reads of up to 16 x kMaxKF64W16 rows go to the W = 16 instance of the FP64 kernel two at a time when the chunk has
    // underflowed pairs enough to fill the GPU that way: twice the rows per lane halve the per-cell share of the shuffles,
    // selects and loop overhead (1544 -> 1741 GCUPS with every pair of BASELINE configs[1] through the FP64 kernel). Both reads of
    // a unit sweep the same haplotypes, so a candidate is (read, FIXED run of its FP32 unit's haplotypes, `per` of them) with
    // at least one underflowed pair; the kernel sweeps the union of what the two reads need and writes only the pairs that
    // underflowed. Underflow goes with the read far more than with the haplotype (the haplotypes of a region differ in a few
    // bases), so the union is rarely larger than either part.
    if (a.pair_mode && len <= 16 * kMaxKF64W16 && (a.pair_mode == 2u || rescued >= 4u * a.fp64_warp_slots16)) {
        const uint32_t per = a.max_pairs ? a.max_pairs
                             : min((uint32_t)kMaxUnitHaps, max(1u, rescued / (8u * a.fp64_warp_slots16)));
        // 16 to 32 underflowed reads per neighbourhood (an odd one is left over in every other): 2^6 reads when a third of the
        // pairs underflow, a whole region at 2 %
        const uint32_t span = min(1u << 16, max(16u, (uint32_t)(32ull * a.n_pairs / max(rescued, 1u))));
        const int neighbourhood = 31 - __clz((int)span);
        for (uint32_t first = 0; first < nh; first += per) {
            const uint32_t n = min(per, nh - first);
            uint32_t pattern = 0;
            for (uint32_t i = 0; i < n; ++i) pattern |= (a.fres[out + first + i] < 1e-28f ? 1u : 0u) << i;
            if (pattern)
                pair_or_wait(a, DevUnit{unit.read + j, unit.hap_list + first,
                                        n | (uint32_t)bin_w16(len) << 8 | ((pattern * 0x9E3779B1u) >> 16) << 16, out + first},
                             len, neighbourhood, stride);
        }
        return;
    }
    const int bin = bin_w32(len);                      // everything else: one read per warp, W = 32
    // a run of the read's haplotypes with at most max_pairs underflowed ones becomes one FP64 unit: the whole row in a big
    // chunk (one pipeline fill per read), single pairs in a small one (one region per call: the FP64 pass is as long as
    // its longest unit, and a sweep of one 150 x 400 pair alone takes ~25 us)
    // few underflowed pairs in the chunk: single pairs, so that the FP64 launch is as short as one sweep and its warps are
    // all busy; many: runs of pairs, up to the read's whole row (one pipeline fill per unit) -- about four units per
    // resident warp of the FP64 kernel
    const uint32_t max_pairs = a.max_pairs ? a.max_pairs
                               : min((uint32_t)kMaxUnitHaps, max(1u, rescued / (4u * a.fp64_warp_slots)));
    uint32_t first = 0, cnt = 0;
    for (uint32_t i = 0; i < nh; ++i) {
        cnt += a.fres[out + i] < 1e-28f ? 1u : 0u;
        if (cnt && (cnt == max_pairs || i + 1 == nh)) {
            emit_single(a, bin, unit.read + j, unit.hap_list + first, i + 1 - first, out + first);
            first = i + 1; cnt = 0;
        }
    }
}

// finalize_kernel + rescue_scan_kernel in ONE launch for a chunk so small that the scan would make single-pair units whatever
// underflows (fewer unique pairs than 2 x 4 x the FP64 kernel's resident warps: a region per call). One thread per (FP32 work
// unit, read): a launch less on a path whose kernels run for 50 us each.
__global__ void finalize_scan_small_kernel(const RescueScanArgs a, const FinalizeArgs f) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t u = t / kMaxUnitReads, j = t % kMaxUnitReads;
    if (u >= a.n_units) return;
    const DevUnit unit = a.units[u];
    const uint32_t nh = unit.shape & 0xffu, nreads = (unit.shape >> 8) & 0xffu, stride = unit.shape >> 16;
    if (j >= nreads) return;
    const uint32_t out = unit.out + j * stride;
    const int len = a.reads[unit.read + j].len;
    if (len < 1) return;
    const int bin = bin_w32(len);
    uint32_t cnt = 0;
    for (uint32_t i = 0; i < nh; ++i) {
        const float v = a.fres[out + i];
        if (v < 1e-28f) {                                  // IntelPairHmm.cc:342
            emit_single(a, bin, unit.read + j, unit.hap_list + i, 1u, out + i);
            ++cnt;
        } else {                                           // IntelPairHmm.cc:348, as finalize_kernel
            f.out[out + i] = (double)__fsub_rn((float)log10((double)v), f.log10_init_f);
            f.flag[out + i] = 0;
        }
    }
    if (cnt) atomicAdd(f.rescued_pairs, cnt);
}

// One thread per slot of the pairing table: who still waits leaves alone; the table is all zero again afterwards.
__global__ void rescue_flush_kernel(const RescueScanArgs a) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= kPairSlots) return;
    const uint32_t v = a.pair_table[t];
    if (!v) return;
    a.pair_table[t] = 0u;
    const DevUnit c = a.pending[v - 1u];
    emit_single(a, bin_w32(a.reads[c.read].len), c.read, c.hap_list, c.shape & 0xffu, c.out);
}

// ---------------------------------------------------------------------------------------------
// Read quality pre-pass. Device restatement of GATKVariantContextUtils::findNumberOfRepetitions
// (GATKVariantContextUtils.cpp:59-100) and PairHMMLikelihoodCalculationEngine::findTandemRepeatUnits (PHLCE.cpp:177-254).
__device__ __forceinline__ bool equal_range(const uint8_t *a, const uint8_t *b, int n) {
    for (int i = 0; i < n; ++i) if (a[i] != b[i]) return false;
    return true;
}
__device__ int count_repetitions(const uint8_t *unit, int unit_len, const uint8_t *test, int test_len, bool leading) {
    if (test_len == 0) return 0;
    const int diff = test_len - unit_len;
    int n = 0;
    if (leading) {
        for (int start = 0; start <= diff; start += unit_len) { if (equal_range(test + start, unit, unit_len)) ++n; else return n; }
    } else {
        for (int start = diff; start >= 0; start -= unit_len) { if (equal_range(test + start, unit, unit_len)) ++n; else return n; }
    }
    return n;
}
__device__ int tandem_repeat_length(const uint8_t *bases, int len, int offset) {
    int max_bw = 0, bw_off = offset, bw_len = 1;              // bestBWRepeatUnit = {readBases[offset]}
    for (int str = 1; str <= 8; ++str) {                      // MAX_STR_UNIT_LENGTH
        if (offset + 1 - str < 0) break;
        max_bw = count_repetitions(bases + offset - str + 1, str, bases, offset + 1, false);
        if (max_bw > 1) { bw_off = offset - str + 1; bw_len = str; break; }
    }
    int max_rl = max_bw;
    if (offset < len - 1) {
        int max_fw = 0, fw_off = offset + 1, fw_len = 1;      // bestFWRepeatUnit = {readBases[offset+1]}
        for (int str = 1; str <= 8; ++str) {
            if (offset + str + 1 > len) break;
            max_fw = count_repetitions(bases + offset + 1, str, bases + offset + 1, len - offset - 1, true);
            if (max_fw > 1) { fw_off = offset + 1; fw_len = str; break; }
        }
        if (fw_len == bw_len && equal_range(bases + fw_off, bases + bw_off, fw_len)) {
            max_rl = max_fw + max_bw;                          // in the middle of a tandem repeat
        } else {
            max_bw = count_repetitions(bases + fw_off, fw_len, bases, offset + 1, false);   // PHLCE.cpp:244-247
            max_rl = max_fw + max_bw;
        }
    }
    return max_rl > 20 ? 20 : max_rl;                          // MAX_REPEAT_LENGTH
}

// applyPCRErrorModel (PHLCE.cpp:150-158: positions 0 .. len-2) + capMinimumReadQualities (:256-267), in place
__global__ void prep_reads_kernel(const PrepReadsArgs a) {
    __shared__ uint8_t sb[kMaxReadLen];
    for (uint32_t r = blockIdx.x; r < a.n; r += gridDim.x) {
        const DevRead rr = a.reads[r];
        const int mapq = a.mapq[r];
        uint8_t *p = a.rd + rr.off;
        __syncthreads();
        const bool staged = rr.len <= kMaxReadLen;            // longer reads (long.cu) are scanned in place: this kernel never writes the bases
        if (staged) for (int i = threadIdx.x; i < rr.len; i += blockDim.x) sb[i] = p[i];
        __syncthreads();
        const uint8_t *bases = staged ? sb : p;
        for (int i = threadIdx.x; i < rr.len; i += blockDim.x) {
            int iq = p[2 * a.plane + i], dq = p[3 * a.plane + i];
            if (a.pcr_on && i < rr.len - 1) {
                const int q = a.pcr_table[tandem_repeat_length(bases, rr.len, i)];
                iq = min(iq, q); dq = min(dq, q);
            }
            int bq = min((int)p[a.plane + i], mapq);
            bq = (int)(uint8_t)bq;
            if (bq < (int)(uint8_t)a.base_q_threshold) bq = 6;   // QualityUtils::MIN_USABLE_Q_SCORE
            if (iq < 6) iq = 6;
            if (dq < 6) dq = 6;
            p[a.plane + i] = (uint8_t)bq; p[2 * a.plane + i] = (uint8_t)iq; p[3 * a.plane + i] = (uint8_t)dq;
        }
    }
}

// Per-read epilogue. The operations and their order are those of the reference's host code, so that the result equals
// oracle_epilogue() (oracle/pairhmm_oracle.c) bit for bit on the same input matrix.
__global__ void epilogue_kernel(const EpilogueArgs a) {
    const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= a.n) return;
    const RowInfo row = a.rows[u];
    double *v = a.out + row.off;
    const double ninf = -__longlong_as_double(0x7ff0000000000000ll);
    phmm_read_summary s;
    s.best = ninf; s.second_best = ninf; s.best_hap = -1; s.poorly_modeled = 1; s.pad[0] = s.pad[1] = s.pad[2] = 0;
    if (row.nh) {
        // searchBestAllele(canBeReference = true, priorities = null), AlleleLikelihoods.h:92-117
        double best = v[0], second = ninf;
        int bi = 0;
        for (uint32_t i = 1; i < row.nh; ++i) {
            const double c = v[i];
            if (c > best) { second = best; best = c; bi = (int)i; }
            else if (c > second) second = c;
        }
        // normalizeLikelihoodsPerEvidence, AlleleLikelihoods.h:150-166 (skipped for <= 1 allele or an infinite cap, :375-380)
        if (a.do_cap && row.nh > 1) {
            const double worst = best + a.cap;
            for (uint32_t i = 0; i < row.nh; ++i) if (v[i] < worst) v[i] = worst;
            if (second < worst) second = worst;
        }
        // filterPoorlyModeledEvidence's predicate, AlleleLikelihoods.h:413 with PairHMMLikelihoodCalculationEngine.cpp:294-299
        const double max_err = fmin(2.0, ceil((double)a.reads[u].len * a.err_rate));
        s.poorly_modeled = best < max_err * -4.0 ? 1 : 0;
        if (a.natural_log) {                                // switchToNaturalLog, AlleleLikelihoods.h:462-476
            for (uint32_t i = 0; i < row.nh; ++i) v[i] = __dmul_rn(v[i], a.ln10);
            best = __dmul_rn(best, a.ln10); second = __dmul_rn(second, a.ln10);
        }
        s.best = best; s.second_best = second; s.best_hap = bi;
    }
    a.summ[u] = s;
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

cudaError_t launch_rescue_scan(const RescueScanArgs &a, cudaStream_t st) {
    if (a.n_units == 0) return cudaSuccess;
    const uint64_t threads = (uint64_t)a.n_units * kMaxUnitReads;
    rescue_scan_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(a);
    return cudaGetLastError();
}
cudaError_t launch_finalize_scan_small(const RescueScanArgs &a, const FinalizeArgs &f, cudaStream_t st) {
    if (a.n_units == 0) return cudaSuccess;
    const uint64_t threads = (uint64_t)a.n_units * kMaxUnitReads;
    finalize_scan_small_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(a, f);
    return cudaGetLastError();
}
cudaError_t launch_rescue_flush(const RescueScanArgs &a, cudaStream_t st) {
    if (a.n_units == 0 || !a.pair_mode) return cudaSuccess;
    rescue_flush_kernel<<<kPairSlots / 256, 256, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_prep_reads(const PrepReadsArgs &a, int sm_count, cudaStream_t st) {
    if (a.n == 0) return cudaSuccess;
    const uint32_t grid = std::min<uint32_t>(a.n, (uint32_t)sm_count * 16u);
    prep_reads_kernel<<<grid, 128, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_epilogue(const EpilogueArgs &a, cudaStream_t st) {
    if (a.n == 0) return cudaSuccess;
    epilogue_kernel<<<(a.n + 127) / 128, 128, 0, st>>>(a);
    return cudaGetLastError();
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
