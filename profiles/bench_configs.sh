#!/bin/bash
# Device-resident PD iteration rate at the other BASELINE.json configs (single-GPU shares).
# usage: bash profiles/bench_configs.sh <tag>   -> gpurun_out/<tag>_configs.jsonl
TAG=${1:-r01}
OUT=gpurun_out/${TAG}_configs.jsonl
: > $OUT
run() { timeout 600 python bench.py --no-e2e --no-cpu-baseline --steps 10 --warmup 3 "$@" 2>>gpurun_out/${TAG}_configs.err >> $OUT || echo "{\"failed\": \"$*\"}" >> $OUT; }
run --nband 4 --nx 2048 --dtype f32      # config 2 fp32
run --nband 4 --nx 2048 --dtype f64      # config 2 fp64
run --nband 8 --nx 4096 --dtype f64      # config 3 in the reference's default precision
run --nband 2 --nx 8192 --dtype f32      # config 4: one GPU's share (16 bands / 8 GPUs)
run --nband 8 --nx 10240 --dtype f32     # config 5: one GPU's share (64 bands / 8 GPUs), PSF 20480^2 = 2^12 * 5
python - <<PY
import json
for l in open("$OUT"):
    d = json.loads(l)
    if "failed" in d: print(d); continue
    print(d["config"]["workload"][:40], d["dtype"], d["value"], "it/s", d["ms_per_step"], "ms",
          {k: v["ms_per_step"] for k, v in d["kernels"].items() if v["ms_per_step"] > 0.05})
PY
