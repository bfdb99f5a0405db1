#!/bin/bash
# usage: scripts/build_variant.sh NAME "-DMVD_XY_ZPC=4 ..." [FILE]  -> build/libmvdecon_NAME.so  (A/B tuning builds;
# FILE = the source the flags apply to, default conv_kernels)
set -e
cd "$(dirname "$0")/../spim_registration_b200/csrc"
V=${3:-conv_kernels}
mkdir -p ../../build/$1
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr"
nvcc $FLAGS $2 -c $V.cu -o ../../build/$1/$V.o
OBJS=""
for f in conv_kernels z_tma_kernels inv_ring_kernels z_warp_kernels sep_kernels volume_kernels input_kernels engine cabi; do
    if [ $f = $V ]; then OBJS="$OBJS ../../build/$1/$V.o"; else [ -f $f.o ] || nvcc $FLAGS -c $f.cu -o $f.o; OBJS="$OBJS $f.o"; fi
done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../build/libmvdecon_$1.so $OBJS
echo built build/libmvdecon_$1.so
