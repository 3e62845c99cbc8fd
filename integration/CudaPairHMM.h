// Reference-side adapter: drops libpairhmm_cuda in behind Mutect2Cpp's PairHMM plugin interface.
//
// This file (and CudaPairHMM.cpp) is what a maintainer of deepmutect/Mutect2Cpp-master adds under
// src/utils/pairhmm/ ; the only other edits are the type at
// src/haplotypecaller/PairHMMLikelihoodCalculationEngine.cpp:35
//     this->pairHMM = std::make_shared<VectorLoglessPairHMM>(args);   ->   std::make_shared<CudaPairHMM>(args);
// and `target_link_libraries(Mutect2Cpp_master pairhmm_cuda)` in CMakeLists.txt.
// It implements exactly the virtuals of PairHMM (src/utils/pairhmm/PairHMM.h:13-69) the way
// VectorLoglessPairHMM does (src/utils/pairhmm/VectorLoglessPairHMM.cpp:18-148), minus the CPU kernels.
#ifndef MUTECT2CPP_MASTER_CUDAPAIRHMM_H
#define MUTECT2CPP_MASTER_CUDAPAIRHMM_H

#include <cstring>
#include <memory>
#include <unordered_map>
#include <vector>

#include "PairHMM.h"
#include "haplotypecaller/PairHMMNativeArgumentCollection.h"
#include "pairhmm_cuda.h"

class CudaPairHMM : public PairHMM {
public:
    // device = CUDA ordinal, -1 = the calling thread's current device. One instance per worker thread, as
    // Mutect2Cpp creates one Mutect2Engine (-> one PairHMM) per thread (src/main.cpp:199).
    explicit CudaPairHMM(PairHMMNativeArgumentCollection &args, int device = -1);
    ~CudaPairHMM() override;
    CudaPairHMM(const CudaPairHMM &) = delete;
    CudaPairHMM &operator=(const CudaPairHMM &) = delete;

    void initialize(const std::vector<std::shared_ptr<Haplotype>> &haplotypes,
                    const std::map<std::string, std::vector<std::shared_ptr<SAMRecord>>> &perSampleReadList,
                    int readMaxLength, int haplotypeMaxLength) override;

    void computeLog10Likelihoods(SampleMatrix<SAMRecord, Haplotype> *logLikelihoods,
                                 std::vector<std::shared_ptr<SAMRecord>> &processedReads,
                                 phmap::flat_hash_map<SAMRecord *, std::shared_ptr<char[]>> *gcp) override;

    // the reference's two trie entry points compute bit-identical numbers: same call
    void computeLog10Likelihoods_trie(SampleMatrix<SAMRecord, Haplotype> *logLikelihoods,
                                      std::vector<std::shared_ptr<SAMRecord>> &processedReads,
                                      phmap::flat_hash_map<SAMRecord *, std::shared_ptr<char[]>> *gcp) override {
        computeLog10Likelihoods(logLikelihoods, processedReads, gcp);
    }
    void computeLog10Likelihoods_trie_unique(SampleMatrix<SAMRecord, Haplotype> *logLikelihoods,
                                             std::vector<std::shared_ptr<SAMRecord>> &processedReads,
                                             phmap::flat_hash_map<SAMRecord *, std::shared_ptr<char[]>> *gcp) override {
        computeLog10Likelihoods(logLikelihoods, processedReads, gcp);
    }

private:
    // haplotypes are looked up by CONTENT, as hash_Haplotype / equal_Haplotype do (src/Mutect2/AssemblyResultSet.h:23-35):
    // the alleles of the SampleMatrix need not be the same shared_ptr objects as the list given to initialize()
    struct HapKey {
        const uint8_t *bases; int len;
        bool operator==(const HapKey &o) const { return len == o.len && std::memcmp(bases, o.bases, (size_t)len) == 0; }
    };
    struct HapKeyHash {
        size_t operator()(const HapKey &k) const {
            uint64_t h = 1469598103934665603ull;
            for (int i = 0; i < k.len; ++i) h = (h ^ k.bases[i]) * 1099511628211ull;
            return (size_t)h;
        }
    };
    void check(int status, const char *what);

    pairhmm_cuda_handle *handle = nullptr;
    std::vector<phmm_hap> hapArray;                                  // borrowed pointers, as HaplotypeDataHolder does
    std::unordered_map<HapKey, int, HapKeyHash> hapToListIdx;
};

#endif  // MUTECT2CPP_MASTER_CUDAPAIRHMM_H
