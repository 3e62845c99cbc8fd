// COMPILE-CHECK STUB: SampleMatrix as the adapter sees it (src/utils/genotyper/AlleleLikelihoods.h:986-1002).
// The real header brings `using namespace std` and phmap into PairHMM.h's scope; so does this one.
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>
#include "parallel_hashmap/phmap.h"
using namespace std;
template <typename E, typename A>
class SampleMatrix {
public:
    vector<shared_ptr<A>> &alleles();
    void set(int alleleIndex, int evidenceIndex, double value);
};
