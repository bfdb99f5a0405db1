#!/bin/bash
# usage: scripts/build_variant.sh NAME "-DMVD_XY_ZPC=4 ..."  -> build/libmvdecon_NAME.so  (A/B tuning builds)
set -e
cd "$(dirname "$0")/../spim_registration_b200/csrc"
mkdir -p ../../build/$1
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr $2"
nvcc $FLAGS -c conv_kernels.cu -o ../../build/$1/conv_kernels.o
for f in sep_kernels volume_kernels input_kernels engine cabi; do [ -f $f.o ] || nvcc $FLAGS -c $f.cu -o $f.o; done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../build/libmvdecon_$1.so ../../build/$1/conv_kernels.o sep_kernels.o volume_kernels.o input_kernels.o engine.o cabi.o
echo built build/libmvdecon_$1.so
