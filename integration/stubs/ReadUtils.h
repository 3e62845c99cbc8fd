// COMPILE-CHECK STUB: src/read/ReadUtils.h, the two functions VectorLoglessPairHMM.cpp:71-72 calls.
#pragma once
#include <cstdint>
#include <memory>
#include "samtools/SAMRecord.h"
class ReadUtils {
public:
    static std::shared_ptr<uint8_t[]> getBaseInsertionQualities(std::shared_ptr<SAMRecord> &read, int &length);
    static std::shared_ptr<uint8_t[]> getBaseDeletionQualities(std::shared_ptr<SAMRecord> &read, int &length);
};
