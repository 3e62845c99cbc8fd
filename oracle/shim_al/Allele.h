// TEST INFRASTRUCTURE ONLY: minimal stand-in for Mutect2Cpp's Allele so that the reference's own
// utils/genotyper/AlleleLikelihoods.h compiles for oracle/ref_epilogue_driver.cpp (the real header pulls in htslib).
#pragma once
#include <cstdint>
#include <memory>
class Allele {
public:
    virtual ~Allele() = default;
    virtual bool getIsReference() const { return false; }
};
// names AlleleLikelihoods.h mentions in members this driver never instantiates (marginalize, contamination, ...)
struct hash_Allele { size_t operator()(const std::shared_ptr<Allele> &a) const { return (size_t)a.get(); } };
struct equal_Allele { bool operator()(const std::shared_ptr<Allele> &a, const std::shared_ptr<Allele> &b) const { return a == b; } };
class SimpleInterval {};
