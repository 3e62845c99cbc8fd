#!/bin/bash
# Development helper: build a reduced library variant (only the kernel units C2 needs) from the working tree plus an
# optional patch command, into fast-genomic-data-processing_b200/libpairhmm_cuda_<name>.so.
# usage: build_variant.sh <name> ["sed/patch command run inside the copied csrc"] [extra nvcc DEFS]
set -e
name=$1; cmd=${2:-true}; defs=${3:-}
root="$(cd "$(dirname "$0")/.." && pwd)"
d=/tmp/phmm_builds/$name
rm -rf $d && mkdir -p $d/pkg $d/include
cp -r $root/fast-genomic-data-processing_b200/csrc $d/pkg/csrc && rm -rf $d/pkg/csrc/build
cp $root/include/*.h $d/include/
cd $d/pkg/csrc
bash -c "$cmd"
export PHMM_GEN_ONLY=${PHMM_GEN_ONLY:-f32_pk_w16_k8_10,f64_pk_w32_k1_7,f64_gen_w32_k1_8,f32_gen_w32_k1_8}
rm -f .gen.stamp     # csrc/Makefile regenerates the instantiation units (only those named above)
make -s -j16 DEFS="$defs" OUT=$root/fast-genomic-data-processing_b200/libpairhmm_cuda_$name.so
ls -la $root/fast-genomic-data-processing_b200/libpairhmm_cuda_$name.so
