// TEST INFRASTRUCTURE ONLY -- stand-in for <boost/multi_array.hpp> (boost is not installed in this image). The reference's
// transformers/recalibration/RecalibrationTables.h, pulled in transitively by Mutect2/AssemblyResultSet.h, only names the
// type and its `index` in six typedefs; the oracle never compiles RecalibrationTables.cpp. The real header also brings in
// boost/config's BOOST_LIKELY / BOOST_UNLIKELY, which PairHMMLikelihoodCalculationEngine.cpp:67 and
// VectorLoglessPairHMM.cpp use without including boost/utility.hpp themselves.
#pragma once
#include <cstddef>
#include "utility.hpp"
namespace boost {
template <class T, std::size_t N> class multi_array { public: typedef std::ptrdiff_t index; };
}
