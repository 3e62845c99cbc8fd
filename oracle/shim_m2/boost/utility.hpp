// Shim for the one boost header the reference PairHMM sources include
// (IntelPairHmm.cc:41 uses only BOOST_LIKELY/BOOST_UNLIKELY). Test infrastructure only.
#pragma once
#define BOOST_LIKELY(x)   __builtin_expect(!!(x), 1)
#define BOOST_UNLIKELY(x) __builtin_expect(!!(x), 0)
