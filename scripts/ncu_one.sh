#!/bin/bash
# Full counter set of a few launches of one kernel family (run under gpurun, one GPU, after the same command ran plain).
# usage: [ENV=...] scripts/ncu_one.sh TAG KERNEL_REGEX [SKIP [COUNT]]   -> gpurun_out/ncu_one_TAG_summary.txt (+ source hot spots)
TAG=$1; RE=$2; SKIP=${3:-4}; CNT=${4:-2}
mkdir -p gpurun_out
MVD_PROF_MODEL=0 timeout 300 python scripts/prof_conv.py ${SHAPE:-1024 1024 512 31 1} > gpurun_out/ncu_one_${TAG}_plain.log 2>&1 || { tail -5 gpurun_out/ncu_one_${TAG}_plain.log; exit 1; }
MVD_PROF_MODEL=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$RE" --launch-skip $SKIP -c $CNT -f -o gpurun_out/one_$TAG \
    python scripts/prof_conv.py ${SHAPE:-1024 1024 512 31 1} > gpurun_out/ncu_one_${TAG}.log 2>&1
python scripts/ncu_summary.py gpurun_out/one_$TAG.ncu-rep > gpurun_out/ncu_one_${TAG}_summary.txt 2>&1
ncu -i gpurun_out/one_$TAG.ncu-rep --page source --csv --print-source sass > gpurun_out/one_${TAG}_sass.csv 2>/dev/null
[ $(du -sm gpurun_out | cut -f1) -lt 55 ] || rm -f gpurun_out/one_$TAG.ncu-rep
tail -40 gpurun_out/ncu_one_${TAG}_summary.txt
