// Host-side construction of the PairHMM lookup tables (product code; restates what
// Context<float>/Context<double> compute at M2/intel/pairhmm/Context.h:65-89,137-183).
// Built once per process with the host libm, uploaded to every device: the per-row constants the
// kernels derive from them (pMM, pMX, ... ) are then single table reads, bit-identical to the
// reference's ReadForPairHMM::initializeData (M2/haplotypecaller/ReadForPairHMM.cpp:72-80).
#pragma once
#include <cstdint>

namespace phmm {

constexpr int kQualCount = 128;                                   // quals are masked with &127
constexpr int kMMCount = (kQualCount * (kQualCount + 1)) / 2;     // 8256 = triangle for q < 128

struct HostTables {
    float ph2pr_f[kQualCount];
    float mm_f[kMMCount];
    double ph2pr_d[kQualCount];
    double mm_d[kMMCount];
};

// Thread-safe, built on first use.
const HostTables &host_tables();

}  // namespace phmm
