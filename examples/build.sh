#!/bin/bash
# Builds the C++ example against the C ABI only (no CUDA headers, no torch): the link line a Mutect2Cpp maintainer adds.
set -e
cd "$(dirname "$0")/.."
g++ -O2 -std=c++17 -pthread -Wall -Iinclude examples/region_calls.cpp -L fast-genomic-data-processing_b200 -lpairhmm_cuda \
    -Wl,-rpath,'$ORIGIN/../fast-genomic-data-processing_b200' -o examples/region_calls
