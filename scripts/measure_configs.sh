#!/bin/bash
# Bench lines of the named configurations on one GPU (run under gpurun): cfg1, cfg2, cfg4 and the default cfg3.
# usage: scripts/measure_configs.sh TAG   -> gpurun_out/bench_TAG_cfgN.json
TAG=${1:-vX}
mkdir -p gpurun_out
for c in 1 2 4; do
    python bench.py --config $c --steps 5 --warmup 3 > gpurun_out/bench_${TAG}_cfg$c.json 2> gpurun_out/bench_${TAG}_cfg$c.err || tail -3 gpurun_out/bench_${TAG}_cfg$c.err
done
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_${TAG}_cfg3.json 2> gpurun_out/bench_${TAG}_cfg3.err || tail -3 gpurun_out/bench_${TAG}_cfg3.err
for c in 1 2 4 3; do python - <<PY
import json
try:
    b = json.load(open("gpurun_out/bench_${TAG}_cfg$c.json"))
    r = b["roofline"]
    print("cfg$c: %.2f ms/step, %.3f Gvoxel-it/s, e2e %.3f, top %s frac %.3f, loop frac %.3f, cpu %s" % (b["ms_per_step"], b["value"], b["e2e"]["value"], r["kernel"], r["frac"], r["loop"]["frac"], b.get("cpu_baseline", {}).get("value")))
    print("   ", {k: (round(v["ms"], 2), round(v["gbs"])) for k, v in r["kernels"].items()})
except Exception as e:
    print("cfg$c: failed", e)
PY
done
