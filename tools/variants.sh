#!/bin/bash
# Development helper: run quick_perf against every libpairhmm_cuda_*.so variant in the package dir.
# usage: variants.sh <config> <regions> <reps> [lane groups...]
cd "$(dirname "$0")/.."
cfg=${1:-c2}; n=${2:-1000}; reps=${3:-3}; shift 3
for lib in fast-genomic-data-processing_b200/libpairhmm_cuda*.so; do
  for lg in ${@:-0}; do
    echo "== $lib LANE_GROUP=$lg"
    LANE_GROUP=$lg PAIRHMM_CUDA_LIB=$PWD/$lib python tools/quick_perf.py $cfg $n $reps 2>&1 | grep -E "^\{\"workload|error|Error" | cut -c1-260
  done
done
