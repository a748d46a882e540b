#!/bin/bash
# Experiment build: N = 2^16 / 60-bit kernels only, into lib_exp/<tag>/libb200fhe.so (select with B200FHE_LIB).
# usage: fastbuild.sh <tag> [extra nvcc flags, e.g. -DB200_ROWS_PER_CTA=16]
set -e
cd "$(dirname "$0")"
tag=$1; shift
out=lib_exp/$tag; mkdir -p $out build_exp/$tag
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-O3 -ccbin /usr/bin/g++ -cudart static -DB200_FAST_BUILD $@"
for f in capi.cu ntt_kernels.cu; do $NV -x cu -c csrc/$f -o build_exp/$tag/$f.o & done
wait
$NV -shared -o $out/libb200fhe.so build_exp/$tag/capi.cu.o build_exp/$tag/ntt_kernels.cu.o build/pmult_kernels.cu.o build/bconv_kernels.cu.o build/fourstep_kernels.cu.o build/relia_kernels.cu.o build/keyswitch.cu.o build/evaluator.cu.o build/numth.cpp.o
echo built $out/libb200fhe.so
