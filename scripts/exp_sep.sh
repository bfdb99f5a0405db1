#!/bin/bash
# Tuning (run under gpurun): direct separable path, per convolution, for A/B builds.  usage: scripts/exp_sep.sh NAME...
mkdir -p gpurun_out
LOG=gpurun_out/exp_sep.log
: > $LOG
for v in "$@"; do
    if [ "$v" = base ]; then unset MVD_LIB_PATH; else export MVD_LIB_PATH=$PWD/build/libmvdecon_$v.so; fi
    for K in 31 15 7; do
        echo "== $v K=$K" >> $LOG
        MVD_PROF_SEP=1 MVD_PROF_MODEL=0 timeout 180 python scripts/prof_conv.py 1024 1024 512 $K 3 2>&1 | grep -E "ms/iteration|direct-path|rror" >> $LOG
    done
    echo "== parity $v" >> $LOG
    timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "separable or direct_path" 2>&1 | tail -2 >> $LOG
done
cat $LOG
