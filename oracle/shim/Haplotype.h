// Minimal stand-in for Mutect2Cpp's Haplotype (the real header pulls in htslib-dependent
// code). Only what avx-pairhmm-template.h:25 and trie/buildTreeUtils.h:8 need to compile.
// Test infrastructure only: lets oracle/Makefile build the reference's own PairHMM sources.
#pragma once
#include <cstdint>
#include <memory>
class Haplotype {
    std::shared_ptr<uint8_t[]> bases_; int len_; bool ref_;
public:
    Haplotype(std::shared_ptr<uint8_t[]> b, int n, bool isRef = false) : bases_(std::move(b)), len_(n), ref_(isRef) {}
    std::shared_ptr<uint8_t[]> getBases() const { return bases_; }
    int getBasesLength() const { return len_; }
    int getLength() const { return len_; }
    bool getIsReference() const { return ref_; }
};
