#include "tables.h"

#include <algorithm>
#include <cmath>
#include <mutex>
#include <vector>

namespace phmm {
namespace {

// Jacobian-logarithm table: log10(1 + 10^(-k * 1e-4)), k = 0 .. 80000 (Context.h:30-34,65-72).
constexpr double kJacTolerance = 8.0;
constexpr double kJacStep = 0.0001;
constexpr double kJacInvStep = 1.0 / kJacStep;
constexpr int kJacSize = 80001;

template <class N>
struct Jacobian {
    std::vector<N> table;
    Jacobian() : table(kJacSize) {
        for (int k = 0; k < kJacSize; ++k)
            table[k] = (N)(std::log10(1.0 + std::pow(10.0, -((double)k) * kJacStep)));
    }
    // approximateLog10SumLog10 (Context.h:92-121): max + table[round(|diff| * 1e4)], all in N.
    N log10_sum(N lo, N hi) const {
        if (lo > hi) std::swap(lo, hi);
        if (std::isinf(lo) || std::isinf(hi)) return hi;
        N diff = hi - lo;
        if (diff >= (N)kJacTolerance) return hi;
        N scaled = (N)(diff * ((N)kJacInvStep));
        int ind = (scaled > (N)0.0) ? (int)(scaled + (N)0.5) : (int)(scaled - (N)0.5);
        return hi + table[ind];
    }
};

// matchToMatchProb[tri(i) + j], j <= i: 1 - 10^(-i/10) - 10^(-j/10) evaluated the reference's
// way, including its 6-digit 1/ln(10) (Context.h:75-88). Only q < 128 can be looked up.
template <class N>
void fill_match_to_match(N *mm) {
    const Jacobian<N> jac;
    const double inv_ln10 = 0.434294;
    for (int i = 0, offset = 0; i < kQualCount; offset += ++i)
        for (int j = 0; j <= i; ++j) {
            double log10_sum = jac.log10_sum((N)-0.1 * (N)i, (N)-0.1 * (N)j);
            double m2m_log10 = std::log1p(-std::min(1.0, std::pow(10, log10_sum))) * inv_ln10;
            mm[offset + j] = (N)(std::pow(10, m2m_log10));
        }
}

}  // namespace

const HostTables &host_tables() {
    static HostTables t;
    static std::once_flag once;
    std::call_once(once, [] {
        for (int q = 0; q < kQualCount; ++q) {
            t.ph2pr_d[q] = std::pow(10.0, -((double)q) / 10.0);   // Context.h:139
            t.ph2pr_f[q] = powf(10.f, -((float)q) / 10.f);        // Context.h:180
        }
        fill_match_to_match<float>(t.mm_f);
        fill_match_to_match<double>(t.mm_d);
    });
    return t;
}

}  // namespace phmm
