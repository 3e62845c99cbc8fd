// See CudaPairHMM.h. Mirrors VectorLoglessPairHMM (src/utils/pairhmm/VectorLoglessPairHMM.cpp) call for call;
// the de-duplication of identical reads, the per-row tables, the float pass, the < 1e-28f double rescue and the
// log10 all happen inside libpairhmm_cuda.
#include "CudaPairHMM.h"

#include <stdexcept>
#include <string>

#include "ReadUtils.h"

CudaPairHMM::CudaPairHMM(PairHMMNativeArgumentCollection &args, int device) {
    // the header this file was compiled against and the library it runs with must agree on the layout of the config struct
    if (pairhmm_cuda_abi_version() != PAIRHMM_CUDA_ABI_VERSION)
        throw std::runtime_error("CudaPairHMM: libpairhmm_cuda ABI version " + std::to_string(pairhmm_cuda_abi_version()) +
                                 ", header version " + std::to_string(PAIRHMM_CUDA_ABI_VERSION));
    // initNative(args.useDoublePrecision, args.pairHmmNativeThreads)   VectorLoglessPairHMM.cpp:14-16
    pairhmm_cuda_config cfg;
    pairhmm_cuda_default_config(&cfg);
    cfg.device = device;
    cfg.use_double = args.useDoublePrecision ? 1 : 0;   // pairHmmNativeThreads has no meaning on the GPU
    // the reference switches to its trie kernels when nHaps / nDistinctLens > 3 (VectorLoglessPairHMM.cpp:35); the engine makes
    // the equivalent choice itself, per launch group, from the prefixes the same-length haplotypes really share (bit-identical
    // results either way)
    cfg.reserved[4] = 1;
    int st = pairhmm_cuda_create(&cfg, &handle);
    if (st != PAIRHMM_OK)
        throw std::runtime_error("CudaPairHMM: pairhmm_cuda_create failed with status " + std::to_string(st) +
                                 " (no sm_100 GPU? the engine has no CPU fallback)");
    is_use_trietree_optimize = false;                   // PairHMM.h:56: the caller then takes the plain path (PHLCE.cpp:69-72)
}

CudaPairHMM::~CudaPairHMM() { pairhmm_cuda_destroy(handle); }

void CudaPairHMM::check(int status, const char *what) {
    // Mutect2Cpp's worker catches std::exception, prints and skips the region (src/main.cpp:303-310)
    if (status != PAIRHMM_OK)
        throw std::runtime_error(std::string("CudaPairHMM::") + what + ": " + pairhmm_cuda_last_error(handle));
}

void CudaPairHMM::initialize(const std::vector<std::shared_ptr<Haplotype>> &haplotypes,
                             const std::map<std::string, std::vector<std::shared_ptr<SAMRecord>>> &perSampleReadList,
                             const int readMaxLength, const int haplotypeMaxLength) {
    (void)perSampleReadList;
    // VectorLoglessPairHMM.cpp:22-34: remember (bases*, length) per haplotype and its index in the list
    hapArray.clear();
    hapArray.reserve(haplotypes.size());
    hapToListIdx.clear();
    hapToListIdx.reserve(haplotypes.size());
    int idx = 0;
    for (const std::shared_ptr<Haplotype> &hap : haplotypes) {
        const int len = hap->getBasesLength();
        const uint8_t *bases = hap->getBases().get();
        hapArray.push_back(phmm_hap{bases, len});
        hapToListIdx.emplace(HapKey{bases, len}, idx++);    // emplace keeps the first of equal haplotypes, like the reference
    }
    is_use_trietree_optimize = false;                      // never ask the caller for the trie entry points
    PairHMM::initialize(readMaxLength, haplotypeMaxLength); // PairHMM.cpp:14-27 (asserts haplotypeMaxLength > 0)
    check(pairhmm_cuda_initialize(handle, readMaxLength, haplotypeMaxLength), "initialize");
}

void CudaPairHMM::computeLog10Likelihoods(SampleMatrix<SAMRecord, Haplotype> *logLikelihoods,
                                          std::vector<std::shared_ptr<SAMRecord>> &processedReads,
                                          phmap::flat_hash_map<SAMRecord *, std::shared_ptr<char[]>> *gcp) {
    if (processedReads.empty())                             // VectorLoglessPairHMM.cpp:46-47
        return;
    const int numReads = (int)processedReads.size();
    const int numHaps = (int)hapArray.size();

    // VectorLoglessPairHMM.cpp:64-77: the five byte arrays of every read, borrowed for the duration of the call
    std::vector<std::shared_ptr<uint8_t[]>> insGops, delGops;
    insGops.reserve(numReads);
    delGops.reserve(numReads);
    std::vector<phmm_read> reads((size_t)numReads);
    for (int r = 0; r < numReads; ++r) {
        int len = processedReads[r]->getLength();
        insGops.emplace_back(ReadUtils::getBaseInsertionQualities(processedReads[r], len));   // default 45
        delGops.emplace_back(ReadUtils::getBaseDeletionQualities(processedReads[r], len));
        reads[r].bases = processedReads[r]->getBasesNoCopy().get();
        reads[r].base_q = processedReads[r]->getBaseQualitiesNoCopy().get();
        reads[r].ins_q = insGops[r].get();
        reads[r].del_q = delGops[r].get();
        reads[r].gcp = reinterpret_cast<const uint8_t *>((*gcp)[processedReads[r].get()].get());
        reads[r].len = len;
    }

    // read-major result array, out[r * numHaps + h] -- the layout of mLogLikelihoodArray_1D (VectorLoglessPairHMM.cpp:123-125)
    mLogLikelihoodArray_1D.resize((size_t)numReads * numHaps);
    check(pairhmm_cuda_compute(handle, reads.data(), numReads, hapArray.data(), numHaps, mLogLikelihoodArray_1D.data(),
                               nullptr),
          "computeLog10Likelihoods");

    // VectorLoglessPairHMM.cpp:132-146: the matrix's allele order may differ from the haplotype list's
    int readIdx = 0;
    for (int r = 0; r < numReads; ++r) {
        int hapIdx = 0;
        for (auto &haplotype : logLikelihoods->alleles()) {
            const int idxInsideHaplotypeList =
                hapToListIdx.at(HapKey{haplotype->getBases().get(), haplotype->getBasesLength()});
            logLikelihoods->set(hapIdx, r, mLogLikelihoodArray_1D[readIdx + idxInsideHaplotypeList]);
            hapIdx++;
        }
        readIdx += numHaps;
    }
}
