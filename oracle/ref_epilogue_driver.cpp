// TEST INFRASTRUCTURE ONLY. C entry point over the reference's OWN AlleleLikelihoods<E, A> template
// (src/utils/genotyper/AlleleLikelihoods.h, compiled from where it lies; shim_al/ holds stand-ins for the headers it
// includes that would drag in htslib). Pins oracle_epilogue(): normalizeLikelihoods, filterPoorlyModeledEvidence,
// switchToNaturalLog run through the reference's code on the same matrix.
#include <cmath>
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "utils/genotyper/AlleleLikelihoods.h"

template <> double AlleleLikelihoods<SAMRecord, Haplotype>::NATURAL_LOG_INFORMATIVE_THRESHOLD =
    MathUtils::log10ToLog(AlleleLikelihoods<SAMRecord, Haplotype>::LOG_10_INFORMATIVE_THRESHOLD);

// PairHMMLikelihoodCalculationEngine::log10MinTrueLikelihood (PairHMMLikelihoodCalculationEngine.cpp:294-299); that
// translation unit needs the whole Mutect2Cpp build, so its three lines are restated here
static double log10MinTrueLikelihood(std::shared_ptr<SAMRecord> read, double maximumErrorPerBase) {
    double maxErrorsForRead = std::min(2.0, std::ceil(read->getLength() * maximumErrorPerBase));
    double log10QualPerBase = -4.0;
    return maxErrorsForRead * log10QualPerBase;
}

extern "C" int ref_epilogue(double *m, int64_t n_reads, int32_t n_haps, const int32_t *read_len, double likelihood_cap,
                            double max_error_per_base, int natural_log, uint8_t *removed) {
    std::vector<std::string> samples{"s"};
    auto haps = std::make_shared<std::vector<std::shared_ptr<Haplotype>>>();
    for (int a = 0; a < n_haps; ++a) haps->push_back(std::make_shared<Haplotype>(a == 0));
    std::map<std::string, std::vector<std::shared_ptr<SAMRecord>>> reads;
    auto &rl = reads["s"];
    for (int64_t r = 0; r < n_reads; ++r) rl.push_back(std::make_shared<SAMRecord>(read_len[r]));
    std::vector<std::shared_ptr<SAMRecord>> keep = rl;
    AlleleLikelihoods<SAMRecord, Haplotype> al(samples, haps, reads);
    auto *sm = al.sampleMatrix(0);
    for (int64_t r = 0; r < n_reads; ++r)
        for (int a = 0; a < n_haps; ++a) sm->set(a, (int)r, m[r * n_haps + a]);
    al.normalizeLikelihoods(likelihood_cap);
    // values after normalisation, before evidence is removed
    for (int64_t r = 0; r < n_reads; ++r)
        for (int a = 0; a < n_haps; ++a) m[r * n_haps + a] = sm->get(a, (int)r);
    al.filterPoorlyModeledEvidence(&log10MinTrueLikelihood, max_error_per_base);
    // which reads survived
    auto &left = al.sampleEvidence(0);
    size_t k = 0;
    for (int64_t r = 0; r < n_reads; ++r) {
        const bool kept = k < left.size() && left[k] == keep[r];
        removed[r] = kept ? 0 : 1;
        if (kept) ++k;
    }
    if (natural_log) {
        al.switchToNaturalLog();
        // surviving rows come back as natural logs; removed rows are converted with the same reference function
        k = 0;
        for (int64_t r = 0; r < n_reads; ++r) {
            if (removed[r]) { for (int a = 0; a < n_haps; ++a) m[r * n_haps + a] = MathUtils::log10ToLog(m[r * n_haps + a]); continue; }
            for (int a = 0; a < n_haps; ++a) m[r * n_haps + a] = sm->get(a, (int)k);
            ++k;
        }
    }
    return (int)left.size();
}
