// TEST INFRASTRUCTURE ONLY: stand-in for Mutect2Cpp's Haplotype as an allele type of AlleleLikelihoods<E, A>.
#pragma once
#include "Allele.h"
class Haplotype : public Allele {
    bool ref_;
public:
    explicit Haplotype(bool isRef = false) : ref_(isRef) {}
    bool getIsReference() const override { return ref_; }
};
