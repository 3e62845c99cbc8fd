// TEST INFRASTRUCTURE ONLY: the one MathUtils member AlleleLikelihoods.h uses, with the body of MathUtils.cpp:16-18
// (the real header includes AssemblyRegion.h / ReadCache.h, i.e. htslib).
#pragma once
#include <cmath>
class MathUtils {
public:
    static double log10ToLog(double log10) { return log10 * std::log(10); }
};
