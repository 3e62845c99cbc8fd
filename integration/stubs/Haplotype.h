// COMPILE-CHECK STUB (integration/check.sh): the members of the reference's Haplotype the adapter uses
// (src/haplotype/Haplotype.h; Allele::getBases / getBasesLength).
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <vector>
class Haplotype {
public:
    std::shared_ptr<uint8_t[]> getBases() const;
    int getBasesLength() const;
    int getLength() const;
};
