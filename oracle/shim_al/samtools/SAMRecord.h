// TEST INFRASTRUCTURE ONLY: stand-in for Mutect2Cpp's SAMRecord -- the epilogue only asks a read for its length.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
class SAMRecord {
    int len_;
public:
    explicit SAMRecord(int len) : len_(len) {}
    int getLength() const { return len_; }
    std::string getName() const { return std::string(); }
};
