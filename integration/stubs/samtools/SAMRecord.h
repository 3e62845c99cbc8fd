// COMPILE-CHECK STUB: the members of the reference's SAMRecord the adapter uses (src/samtools/SAMRecord.h).
#pragma once
#include <cstdint>
#include <memory>
class SAMRecord {
public:
    int getLength() const;
    std::shared_ptr<uint8_t[]> getBasesNoCopy();
    std::shared_ptr<uint8_t[]> getBaseQualitiesNoCopy();
    int getBaseQualitiesLength() const;
};
