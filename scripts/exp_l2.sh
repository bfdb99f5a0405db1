#!/bin/bash
# Tuning experiment (run under gpurun): does batch scratch that stays in L2 between the three kernels
# beat whole-class batches?  One line per setting in gpurun_out/exp_l2.log.
mkdir -p gpurun_out
LOG=gpurun_out/exp_l2.log
: > $LOG
run() {  # label, env...
    local label="$1"; shift
    echo "== $label: $*" >> $LOG
    env "$@" MVD_PROF_MODEL=${MODEL:-0} timeout 180 python scripts/prof_conv.py 1024 1024 512 31 3 2>&1 | grep -E "ms/iteration|model|rror" >> $LOG
}
MODEL=1 run base X=1
run z512_b2_l1 MVD_SCRATCH_MB=70 MVD_LANES=1
run z512_b2_l1_persist MVD_SCRATCH_MB=70 MVD_LANES=1 MVD_L2_PERSIST=80
run z512_b1_l3 MVD_SCRATCH_MB=35 MVD_LANES=3
run z256_b3_l1 MVD_ZMAX=256 MVD_SCRATCH_MB=52 MVD_LANES=1
run z256_b3_l1_persist MVD_ZMAX=256 MVD_SCRATCH_MB=52 MVD_LANES=1 MVD_L2_PERSIST=64
run z256_b2_l3 MVD_ZMAX=256 MVD_SCRATCH_MB=35 MVD_LANES=3
run z256_all MVD_ZMAX=256
run z128_b6_l1 MVD_ZMAX=128 MVD_SCRATCH_MB=52 MVD_LANES=1
run z128_b6_l1_persist MVD_ZMAX=128 MVD_SCRATCH_MB=52 MVD_LANES=1 MVD_L2_PERSIST=64
run z128_b3_l4 MVD_ZMAX=128 MVD_SCRATCH_MB=26 MVD_LANES=4
run z128_b3_l4_persist MVD_ZMAX=128 MVD_SCRATCH_MB=26 MVD_LANES=4 MVD_L2_PERSIST=90
MODEL=1 run z128_all MVD_ZMAX=128
run z128_b6_l1_zpc1 MVD_ZMAX=128 MVD_SCRATCH_MB=52 MVD_LANES=1 MVD_LIB_PATH=$PWD/build/libmvdecon_zpc1.so
run z128_b3_l4_zpc1 MVD_ZMAX=128 MVD_SCRATCH_MB=26 MVD_LANES=4 MVD_LIB_PATH=$PWD/build/libmvdecon_zpc1.so
run z256_b3_l1_zpc1 MVD_ZMAX=256 MVD_SCRATCH_MB=52 MVD_LANES=1 MVD_LIB_PATH=$PWD/build/libmvdecon_zpc1.so
run z256_b2_l3_zpc2 MVD_ZMAX=256 MVD_SCRATCH_MB=35 MVD_LANES=3 MVD_LIB_PATH=$PWD/build/libmvdecon_zpc2.so
cat $LOG
