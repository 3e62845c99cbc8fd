#!/bin/bash
# Development helper: run quick_perf against every libpairhmm_cuda_*.so variant in the package dir.
cd "$(dirname "$0")/.."
for lib in fast-genomic-data-processing_b200/libpairhmm_cuda*.so; do
  echo "== $lib"
  PAIRHMM_CUDA_LIB=$PWD/$lib python tools/quick_perf.py ${1:-c2} ${2:-1000} ${3:-3} 2>&1 | grep -E "^\{\"workload|error|Error" | cut -c1-260
done
