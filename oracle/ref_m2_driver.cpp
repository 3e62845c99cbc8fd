// TEST INFRASTRUCTURE ONLY. C entry points over the reference's OWN caller-side code of the PairHMM path, compiled
// UNMODIFIED from where it lies under /root/reference (oracle/Makefile, target _ref/libref_m2.so):
//   PairHMMLikelihoodCalculationEngine.cpp (modifyReadQualities, applyPCRErrorModel, findTandemRepeatUnits,
//   capMinimumReadQualities, buildGapContinuationPenalties, computeReadLikelihoods, log10MinTrueLikelihood),
//   GATKVariantContextUtils.cpp (findNumberOfRepetitions), VectorLoglessPairHMM.cpp (gather, de-dup, scatter),
//   SAMRecord.cpp, ReadUtils.cpp, Haplotype.cpp, AssemblyResultSet.cpp, AlleleLikelihoods.h ... and the Intel kernels.
// The only stand-ins are for THIRD-PARTY headers that are not in this image (oracle/shim_m2: htslib declarations,
// boost/utility.hpp, boost/multi_array.hpp); every Mutect2Cpp header and source is the reference's own.
//
// Private members of PairHMMLikelihoodCalculationEngine are reached with the usual test trick (`#define private public`
// around the one header, in THIS translation unit only; the reference's own objects are compiled without it -- access
// specifiers change neither the layout GCC produces nor the mangled names).
//
// Built a second time with -DM2_WITH_CUDA_ADAPTER as integration/replay/libm2_replay.so: the same harness where the
// engine's `pairHMM` member (PairHMMLikelihoodCalculationEngine.cpp:35) can be integration/CudaPairHMM instead of
// VectorLoglessPairHMM -- the drop-in, executed on the reference's real SAMRecord / Haplotype / AlleleLikelihoods objects.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

// everything PairHMMLikelihoodCalculationEngine.h includes comes in first (include guards), so the define below touches that
// one class only
#include "utils/pairhmm/PairHMM.h"
#include "AssemblyResultSet.h"
#include "utils/genotyper/AlleleLikelihoods.h"
#include "haplotypecaller/PairHMMNativeArgumentCollection.h"
#include "ReadUtils.h"
#include "ContigMap.h"
#include "utils/pairhmm/VectorLoglessPairHMM.h"
#define private public
#include "haplotypecaller/PairHMMLikelihoodCalculationEngine.h"
#undef private
#ifdef M2_WITH_CUDA_ADAPTER
#include "CudaPairHMM.h"
#endif

namespace {

thread_local std::string g_error;

struct Region {
    std::vector<std::string> samples;
    std::map<std::string, std::vector<std::shared_ptr<SAMRecord>>> per_sample;
    std::vector<std::shared_ptr<SAMRecord>> reads;          // input order
    std::vector<int> read_sample;                           // sample index of every read
    std::vector<std::shared_ptr<Haplotype>> haps;           // input order
    AssemblyResultSet ars;
};

std::shared_ptr<uint8_t[]> dup_bytes(const uint8_t *p, int n) {
    std::shared_ptr<uint8_t[]> q(new uint8_t[n > 0 ? n : 1]);
    if (n > 0) std::memcpy(q.get(), p, (size_t)n);
    return q;
}

// reads as the BAM record holds them: bases, base qualities, mapping quality, optional BI / BD quality strings
std::shared_ptr<SAMRecord> make_read(int idx, const uint8_t *bases, const uint8_t *quals, int len, int mapq, const uint8_t *ins_q,
                                     const uint8_t *del_q) {
    std::string name = "r" + std::to_string(idx);
    auto rec = std::make_shared<SAMRecord>(dup_bytes(bases, len), len, dup_bytes(quals, len), len, name);
    rec->setMappingQuality(mapq);
    // the in-memory constructor (SAMRecord.cpp:259) leaves the CIGAR null, and the copy constructor that
    // ReadUtils::emptyRead uses (SAMRecord.cpp:598-622) dereferences it; a record decoded from a BAM always has one
    std::vector<CigarElement> ce{CigarElement(len, M)};
    rec->setCigar(std::make_shared<Cigar>(ce));
    rec->setReadUnmappedFlag(false);
    std::shared_ptr<SAMRecord> r = rec;
    // ReadUtils::setInsertionBaseQualities stores the BI / BD attribute as a FASTQ string, exactly what a BQSR-ed BAM carries
    if (ins_q) ReadUtils::setInsertionBaseQualities(r, dup_bytes(ins_q, len), len);
    if (del_q) ReadUtils::setDeletionBaseQualities(r, dup_bytes(del_q, len), len);
    return rec;
}

}  // namespace

// the objects are compiled -fvisibility=hidden so that --gc-sections can drop what the path never reaches
#define M2_API __attribute__((visibility("default")))

extern "C" {

M2_API const char *ref_m2_last_error(void) { return g_error.c_str(); }

// modifyReadQualities (PairHMMLikelihoodCalculationEngine.cpp:123-147) + buildGapContinuationPenalties (:284-292) on n reads.
// has_indel_q[i] != 0: read i carries BI / BD arrays (ins_q / del_q at the read's offset); otherwise the default 45 applies
// (ReadUtils.h:40). pcr_model = 1 | 2 | 3 (the reference's enum has no NONE). Outputs are laid out like base_q.
M2_API int ref_m2_prepass(const uint8_t *bases, const uint8_t *base_q, const uint8_t *ins_q, const uint8_t *del_q, const uint8_t *has_indel_q,
                   const int32_t *mapq, const int64_t *read_off, const int32_t *read_len, int64_t n_reads, int pcr_model,
                   int base_q_threshold, int constant_gcp, uint8_t *out_base_q, uint8_t *out_ins_q, uint8_t *out_del_q,
                   uint8_t *out_gcp) {
    try {
        PairHMMNativeArgumentCollection args;
        PairHMMLikelihoodCalculationEngine eng((char)constant_gcp, args, -4.5, (PCRErrorModel)pcr_model, (char)base_q_threshold);
        std::vector<std::shared_ptr<SAMRecord>> reads;
        for (int64_t i = 0; i < n_reads; ++i) {
            const int64_t o = read_off[i];
            const bool own = has_indel_q && has_indel_q[i];
            reads.push_back(make_read((int)i, bases + o, base_q + o, read_len[i], mapq[i], own ? ins_q + o : nullptr, own ? del_q + o : nullptr));
        }
        auto processed = eng.modifyReadQualities(reads);
        auto *gcp = PairHMMLikelihoodCalculationEngine::buildGapContinuationPenalties(*processed, PairHMMLikelihoodCalculationEngine::constantGCP);
        for (int64_t i = 0; i < n_reads; ++i) {
            const int64_t o = read_off[i];
            std::shared_ptr<SAMRecord> &p = (*processed)[i];
            int len = p->getLength();
            if (len != read_len[i]) throw std::runtime_error("processed read length differs");
            // exactly what VectorLoglessPairHMM.cpp:68-77 hands to ReadForPairHMM
            auto iq = ReadUtils::getBaseInsertionQualities(p, len);
            auto dq = ReadUtils::getBaseDeletionQualities(p, len);
            std::memcpy(out_base_q + o, p->getBaseQualitiesNoCopy().get(), (size_t)len);
            std::memcpy(out_ins_q + o, iq.get(), (size_t)len);
            std::memcpy(out_del_q + o, dq.get(), (size_t)len);
            std::memcpy(out_gcp + o, (*gcp)[p.get()].get(), (size_t)len);
        }
        delete gcp;
        return 0;
    } catch (const std::exception &e) { g_error = e.what(); return -1; }
    catch (const char *e) { g_error = e; return -1; }
}

// log10MinTrueLikelihood (PairHMMLikelihoodCalculationEngine.cpp:294-299), the reference's own three lines
M2_API double ref_m2_log10_min_true_likelihood(int read_len, double max_error_per_base) {
    uint8_t b = 'A', q = 30;
    std::vector<uint8_t> bb((size_t)std::max(1, read_len), b), qq((size_t)std::max(1, read_len), q);
    return PairHMMLikelihoodCalculationEngine::log10MinTrueLikelihood(make_read(0, bb.data(), qq.data(), read_len, 60, nullptr, nullptr),
                                                                     max_error_per_base);
}

// One active region through PairHMMLikelihoodCalculationEngine, as Mutect2Engine::callRegion drives it.
//   reads: raw (see ref_m2_prepass), read i belongs to sample read_sample[i] in [0, n_samples);
//   haps: hap_ref = index of the reference haplotype (AssemblyResultSet wants exactly one);
//   stage 0: per sample the private computeReadLikelihoods(SampleMatrix*) only (:63-75: modifyReadQualities, gcp,
//            pairHMM->computeLog10Likelihoods)  -> the raw log10 matrix;
//   stage 1: the public computeReadLikelihoods (:77-97): + normalizeLikelihoods + filterPoorlyModeledEvidence;
//   use_cuda: 0 = the reference's VectorLoglessPairHMM, 1 = integration/CudaPairHMM (replay build only).
// out[i * n_haps + h] = value for read i (input order) and haplotype h (INPUT index, whatever the matrix's allele order
// is); removed[i] = 1 if filterPoorlyModeledEvidence dropped the read (stage 1). allele_order[a] (nullable) = input index of
// the matrix's a-th allele. Returns 0, or -1 with ref_m2_last_error().
M2_API int ref_m2_region(const uint8_t *bases, const uint8_t *base_q, const uint8_t *ins_q, const uint8_t *del_q, const uint8_t *has_indel_q,
                  const int32_t *mapq, const int32_t *read_sample, const int64_t *read_off, const int32_t *read_len, int64_t n_reads,
                  int n_samples, const uint8_t *hap_bases, const int64_t *hap_off, const int32_t *hap_len, int n_haps, int hap_ref,
                  int pcr_model, int base_q_threshold, int constant_gcp, double log10_mismapping, int use_double, int stage,
                  int use_cuda, double *out, uint8_t *removed, int32_t *allele_order) {
    try {
        PairHMMNativeArgumentCollection args;
        args.useDoublePrecision = use_double != 0;
        PairHMMLikelihoodCalculationEngine eng((char)constant_gcp, args, log10_mismapping, (PCRErrorModel)pcr_model, (char)base_q_threshold);
        if (use_cuda) {
#ifdef M2_WITH_CUDA_ADAPTER
            eng.pairHMM = std::make_shared<CudaPairHMM>(args);      // the one-line edit of INTEGRATION.md (PHLCE.cpp:35)
#else
            throw std::runtime_error("built without the CUDA adapter");
#endif
        }
        Region rg;
        for (int s = 0; s < n_samples; ++s) { rg.samples.push_back("sample" + std::to_string(s)); rg.per_sample[rg.samples.back()]; }
        for (int64_t i = 0; i < n_reads; ++i) {
            const int64_t o = read_off[i];
            const bool own = has_indel_q && has_indel_q[i];
            auto r = make_read((int)i, bases + o, base_q + o, read_len[i], mapq[i], own ? ins_q + o : nullptr, own ? del_q + o : nullptr);
            rg.reads.push_back(r);
            rg.per_sample[rg.samples[(size_t)read_sample[i]]].push_back(r);
        }
        std::string contig = "chr20";
        {   // main.cpp fills the process-wide contig table from the BAM header before any region is called
            static std::once_flag once;
            std::call_once(once, [&] { ContigMap::initial(1); ContigMap::insertPair(0, contig); });
        }
        for (int h = 0; h < n_haps; ++h) {
            auto hp = std::make_shared<Haplotype>(dup_bytes(hap_bases + hap_off[h], hap_len[h]), hap_len[h], h == hap_ref);
            hp->setGenomeLocation(std::make_shared<SimpleInterval>(contig, 1000, 1000 + std::max(1, hap_len[h]) - 1));
            rg.haps.push_back(hp);
            if (!rg.ars.add(hp)) throw std::runtime_error("duplicate haplotype: AssemblyResultSet keeps a set");
        }
        AlleleLikelihoods<SAMRecord, Haplotype> *al = nullptr;
        if (stage == 0) {
            // the first half of the public function (:80-91) with the private per-sample overload
            auto list = rg.ars.getHaplotypeList();
            eng.initializePairHMM(*list, rg.per_sample);
            al = new AlleleLikelihoods<SAMRecord, Haplotype>(rg.samples, list, rg.per_sample);
            for (int s = 0; s < n_samples; ++s) eng.computeReadLikelihoods(al->sampleMatrix(s));
        } else {
            al = eng.computeReadLikelihoods(rg.ars, rg.samples, rg.per_sample);
        }
        const double nan = std::nan("");
        for (int64_t i = 0; i < n_reads * n_haps; ++i) out[i] = nan;
        if (removed) std::memset(removed, 1, (size_t)n_reads);
        std::map<const Haplotype *, int> hap_index;
        for (int h = 0; h < n_haps; ++h) hap_index[rg.haps[(size_t)h].get()] = h;
        std::map<const SAMRecord *, int64_t> read_index;
        for (int64_t i = 0; i < n_reads; ++i) read_index[rg.reads[(size_t)i].get()] = i;
        for (int s = 0; s < n_samples; ++s) {
            std::string name = rg.samples[(size_t)s];
            const int si = al->indexOfSamples(name);
            auto *sm = al->sampleMatrix(si);
            auto &ev = al->sampleEvidence(si);
            const int na = al->numberOfAlleles();
            for (int a = 0; a < na; ++a) {
                const int h = hap_index.at(al->getAllele(a).get());
                if (allele_order && s == 0) allele_order[a] = h;
                for (size_t e = 0; e < ev.size(); ++e) out[read_index.at(ev[e].get()) * n_haps + h] = sm->get(a, (int)e);
            }
            if (removed) for (auto &e : ev) removed[read_index.at(e.get())] = 0;
        }
        delete al;
        return 0;
    } catch (const std::exception &e) { g_error = e.what(); return -1; }
    catch (const char *e) { g_error = e; return -1; }
    catch (const std::string &e) { g_error = e; return -1; }
}

}  // extern "C"
