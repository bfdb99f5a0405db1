#!/bin/bash
# A/B of the TMA kernels (run under gpurun): parity of the windows that use them, then per-kernel times of one
# loop step with the switches off and on.  usage: scripts/exp_tma.sh [switch ...]   (switch = MVD_Z_TMA, MVD_INV_TMA)
mkdir -p gpurun_out
LOG=gpurun_out/exp_tma.log
: > $LOG
echo "== parity (all switches on)" | tee -a $LOG
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "test_convolution_matches_oracle or test_loop_matches_oracle or test_bricks" 2>&1 | tail -3 | tee -a $LOG
for sw in "$@"; do
    for val in 0 1; do
        echo "== $sw=$val" | tee -a $LOG
        env $sw=$val MVD_PROF_MODEL=1 timeout 240 python scripts/prof_conv.py ${SHAPE:-1024 1024 512 31 3} 2>&1 | grep -E "ms/iteration|model|rror" | tee -a $LOG
    done
done
