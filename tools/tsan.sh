#!/bin/bash
# SURVEY.md section 5 (race detection): the HOST side of the library -- two slots per handle, the process-wide thread pool,
# T caller threads with a handle each, the shared batcher's dispatcher, the per-device workers of a multi-device handle --
# under ThreadSanitizer. Builds a reduced library (the kernel units configs[1] needs) whose host code is compiled with
# -fsanitize=thread, links examples/region_calls.cpp against it, and runs it on the GPU box:
#     bash tools/tsan.sh build            (here, no GPU)      bash tools/tsan.sh run  (under gpurun; log in gpurun_out/)
# The CUDA runtime / driver are not instrumented: reports whose every frame lies in libcuda / libcudart are theirs.
set -e
cd "$(dirname "$0")/.."
if [ "$1" = build ]; then
  d=/tmp/phmm_builds/tsan
  rm -rf $d && mkdir -p $d/pkg $d/include
  cp -r fast-genomic-data-processing_b200/csrc $d/pkg/csrc && rm -rf $d/pkg/csrc/build $d/pkg/csrc/.gen.stamp
  cp include/*.h $d/include/
  export PHMM_GEN_ONLY=f32_pk_w16_k8_10,f32_pk_w32_k1_6,f64_pk_w32_k1_7,f64_gen_w32_k1_8,f32_gen_w32_k1_8
  make -s -j16 -C $d/pkg/csrc DEFS="-Xcompiler -fsanitize=thread,-g" LDEXTRA="-Xcompiler -fsanitize=thread" \
       OUT=$PWD/tools/tsan_out/libpairhmm_cuda.so
  sed -i 's/g++ -O2 -std=c++17 -fPIC -c tables.cpp/g++ -O2 -std=c++17 -fPIC -fsanitize=thread -c tables.cpp/' $d/pkg/csrc/Makefile || true
  g++ -O1 -g -fsanitize=thread -std=c++17 -pthread -Iinclude examples/region_calls.cpp -L tools/tsan_out -lpairhmm_cuda \
      -Wl,-rpath,'$ORIGIN' -o tools/tsan_out/region_calls_tsan
  ls -la tools/tsan_out
else
  export TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0 history_size=4 second_deadlock_stack=1"
  out=gpurun_out/r2_tsan_region_calls.log
  ( echo "== region_calls 240 (T handles, blocking calls)"; timeout 600 tools/tsan_out/region_calls_tsan 240
    echo "== region_calls 240 --submit"; timeout 600 tools/tsan_out/region_calls_tsan 240 --submit
    echo "== region_calls 240 --batch-window 50 (shared batcher)"; timeout 600 tools/tsan_out/region_calls_tsan 240 --batch-window 50
    echo "== region_calls 240 --devices 1 (multi-device handle, one child)"; timeout 600 tools/tsan_out/region_calls_tsan 240 --devices 1 ) > $out 2>&1
  echo "ThreadSanitizer reports: $(grep -c 'WARNING: ThreadSanitizer' $out)" | tee -a $out
fi
