#!/bin/bash
# One multi-GPU bench line of cfg3 (run under gpurun --gpus N): usage scripts/measure_n.sh N TAG [STEPS]
N=$1; TAG=$2; S=${3:-10}
mkdir -p gpurun_out
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
    bench.py --gpus $N --steps $S --warmup 3 > gpurun_out/bench_${TAG}_cfg3_n$N.json 2> gpurun_out/bench_${TAG}_cfg3_n$N.err || tail -5 gpurun_out/bench_${TAG}_cfg3_n$N.err
python -c "import json,sys; b=json.load(open(sys.argv[1])); print(sys.argv[1], round(b['ms_per_step'],2), round(b['value'],2), round(b['e2e']['value'],2), (b.get('parity_multi_gpu') or {}).get('ok'))" gpurun_out/bench_${TAG}_cfg3_n$N.json
