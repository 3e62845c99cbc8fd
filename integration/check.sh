#!/bin/bash
# Compile-checks the adapter against the reference's REAL PairHMM.h / PairHMMNativeArgumentCollection.h / phmap.h and
# stubs for the htslib-dependent headers (the full Mutect2Cpp tree needs libtorch, boost, jemalloc: not in this image).
set -e
cd "$(dirname "$0")"
M2=${REF:-/root/reference}/deepmutect/Mutect2Cpp-master/src
[ -f "$M2/utils/pairhmm/PairHMM.h" ] || { echo "reference tree absent: nothing to check against"; exit 77; }
g++ -std=gnu++17 -fsyntax-only -Wall -Istubs -I"$M2/utils/pairhmm" -I"$M2" -I../include CudaPairHMM.cpp
echo "CudaPairHMM compiles against $M2/utils/pairhmm/PairHMM.h"
