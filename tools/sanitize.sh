#!/bin/bash
# compute-sanitizer over every kernel of libpairhmm_cuda.so (memcheck + racecheck for the shared-memory
# tables and the lane exchange). Run on the GPU box: bash tools/sanitize.sh [out_dir]
out=${1:-gpurun_out}
mkdir -p "$out"
cd "$(dirname "$0")/.."
rc=0
for tool in memcheck racecheck synccheck; do
    timeout 900 compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 9 python tools/sanitize_workload.py > "$out/sanitizer_$tool.log" 2>&1
    r=$?
    echo "$tool: exit $r; $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' "$out/sanitizer_$tool.log" | tail -1)"
    [ $r -ne 0 ] && rc=$r
done
exit $rc
