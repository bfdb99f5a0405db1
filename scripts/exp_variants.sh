#!/bin/bash
# Tuning (run under gpurun): per-kernel times of one loop step for A/B builds of the library.
# usage: scripts/exp_variants.sh NAME...   (NAME = build/libmvdecon_NAME.so, "base" = the in-tree library)
mkdir -p gpurun_out
LOG=gpurun_out/exp_variants.log
: > $LOG
for v in "$@"; do
    echo "== $v" >> $LOG
    if [ "$v" = base ]; then unset MVD_LIB_PATH; else export MVD_LIB_PATH=$PWD/build/libmvdecon_$v.so; fi
    MVD_PROF_MODEL=1 timeout 180 python scripts/prof_conv.py ${SHAPE:-1024 1024 512 31 3} 2>&1 | grep -E "ms/iteration|model|rror" >> $LOG
done
cat $LOG
# parity of the 512-point z window for every variant that was timed
for v in "$@"; do
    if [ "$v" = base ]; then unset MVD_LIB_PATH; else export MVD_LIB_PATH=$PWD/build/libmvdecon_$v.so; fi
    echo "== parity $v" | tee -a $LOG
    timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "test_convolution_matches_oracle and (shape6 or shape7 or shape8)" 2>&1 | tail -2 | tee -a $LOG
done
