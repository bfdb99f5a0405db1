#!/bin/bash
# Multi-GPU bench lines (run under gpurun --gpus 8): cfg3 at N = 8, 4, 2, cfg5 at N = 8, the single-process handle at N = 8.
TAG=${1:-vX}
mkdir -p gpurun_out
run() {  # N config steps extra...
    local n=$1 c=$2 s=$3; shift 3
    timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n + c)) \
        bench.py --gpus $n --config $c --steps $s --warmup 3 "$@" > gpurun_out/bench_${TAG}_cfg${c}_n$n.json 2> gpurun_out/bench_${TAG}_cfg${c}_n$n.err \
        || { echo "cfg$c n$n FAILED"; tail -5 gpurun_out/bench_${TAG}_cfg${c}_n$n.err; }
}
timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi_device" 2>&1 | tail -2
if [ -n "$FULL" ]; then run 4 3 10 --no-parity; run 2 3 10 --no-parity; run 8 5 5 --no-parity; fi
if [ -n "$LEAN" ]; then   # the two headline lines only: 20 steps without the parity leg, 10 steps with it
    run 8 3 20 --no-parity && mv gpurun_out/bench_${TAG}_cfg3_n8.json gpurun_out/bench_${TAG}_cfg3_n8_20steps.json
    run 8 3 10
    for f in gpurun_out/bench_${TAG}_cfg*_n*.json; do python -c "import json,sys; b=json.load(open(sys.argv[1])); print(sys.argv[1], round(b['ms_per_step'],2), round(b['value'],2), round(b['e2e']['value'],2), (b.get('parity_multi_gpu') or {}).get('ok'))" $f; done
    exit 0
fi
run 8 3 20 --no-parity && mv gpurun_out/bench_${TAG}_cfg3_n8.json gpurun_out/bench_${TAG}_cfg3_n8_20steps.json
MVD_BARRIER=nccl timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29599 \
    bench.py --gpus 8 --steps 10 --warmup 3 --no-parity > gpurun_out/bench_${TAG}_cfg3_n8_ncclbarrier.json 2> gpurun_out/bench_${TAG}_cfg3_n8_ncclbarrier.err \
    || tail -3 gpurun_out/bench_${TAG}_cfg3_n8_ncclbarrier.err
run 8 3 10
timeout 420 python bench.py --gpus 8 --single-process --steps 10 --warmup 3 > gpurun_out/bench_${TAG}_cfg3_n8_single.json 2> gpurun_out/bench_${TAG}_cfg3_n8_single.err || tail -5 gpurun_out/bench_${TAG}_cfg3_n8_single.err
for f in gpurun_out/bench_${TAG}_cfg*_n*.json; do python - "$f" <<'PY'
import json, sys
try:
    b = json.load(open(sys.argv[1]))
    print(sys.argv[1].split("bench_")[1], "%.2f ms/step, %.2f Gvoxel-it/s, e2e %.2f" % (b["ms_per_step"], b["value"], b["e2e"]["value"]), b["config"].get("brick_grid_xyz"),
          (b.get("parity_multi_gpu") or {}).get("ok"), {k: round(v["gbs"]) for k, v in b.get("roofline", {}).get("kernels", {}).items()})
except Exception as e:
    print(sys.argv[1], "failed:", e)
PY
done
