#!/bin/bash
# A/B over environment settings (run under gpurun): per-kernel times of one loop step for every "NAME=VALUE[,NAME=VALUE]" argument
# usage: scripts/exp_env.sh LABEL SETTINGS...    ("-" = defaults)
mkdir -p gpurun_out
LOG=gpurun_out/exp_env_$1.log; shift
: > $LOG
for s in "$@"; do
    echo "== $s" | tee -a $LOG
    if [ "$s" = "-" ]; then E=""; else E=$(echo $s | tr ',' ' '); fi
    env $E MVD_PROF_MODEL=1 timeout 240 python scripts/prof_conv.py ${SHAPE:-1024 1024 512 31 3} 2>&1 | grep -E "ms/iteration|model|rror" | tee -a $LOG
done
