#!/bin/bash
# Round profile (run under gpurun, one GPU): plain runs first, then the ncu passes of the same commands.
# usage: scripts/profile_round.sh TAG   -> gpurun_out/{bench,launches,prof}_TAG.*
TAG=${1:-vX}
mkdir -p gpurun_out
set -x
[ -n "$SKIP_TESTS" ] || python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err || { tail -5 gpurun_out/bench_$TAG.err; exit 1; }
# launch list of one timed step (bench.py brackets the timed steps with cudaProfilerStart/Stop)
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_$TAG.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches_$TAG.log 2>&1
# byte model and plain timing of one loop step on two views, then the full counter set of the last view's 12 launches (2 convolutions x 2 tile classes x 3 kernels)
MVD_PROF_MODEL=1 python scripts/prof_conv.py 1024 1024 512 31 1 > gpurun_out/prof_$TAG.log 2>&1 || { tail -5 gpurun_out/prof_$TAG.log; exit 1; }
MVD_PROF_MODEL=0 ncu --set full --clock-control none --import-source on -k regex:"fwd_xy|z_pass|z_warp|z_part|z_tma|inv_xy|inv_ring" --launch-skip 52 -c 12 -f -o gpurun_out/prof_$TAG \
    python scripts/prof_conv.py 1024 1024 512 31 1 > gpurun_out/ncu_full_$TAG.log 2>&1
ncu -i gpurun_out/prof_$TAG.ncu-rep --page raw --csv > gpurun_out/prof_${TAG}_raw.csv 2>/dev/null
python scripts/ncu_summary.py gpurun_out/prof_$TAG.ncu-rep > gpurun_out/ncu_full_${TAG}_summary.txt 2>&1
# gpurun brings back at most 64 MiB: the summaries matter more than the report
[ $(du -sm gpurun_out | cut -f1) -lt 60 ] || rm -f gpurun_out/prof_$TAG.ncu-rep
ls -la gpurun_out/
cat gpurun_out/prof_$TAG.log
