#!/bin/bash
# BASELINE.json configs 3, 4 and 5 on the 8 GPUs of one box (gpurun --gpus 8), each bounded by its own
# timeout; lines land in gpurun_out/r02w_config<N>_n8.json.  Also the NCCL sharded test (world 2 and 4).
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 200 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q > gpurun_out/r02w_sharded_test.log 2>&1
echo "sharded test rc=$?" >> gpurun_out/r02w_sharded_test.log; tail -2 gpurun_out/r02w_sharded_test.log
timeout 300 $TR --master-port 29621 bench.py --gpus 8 --config 3 --steps 20 --warmup 3 \
    > gpurun_out/r02w_config3_n8.json 2> gpurun_out/r02w_config3_n8.err; echo "config3 rc=$?"
timeout 400 $TR --master-port 29622 bench.py --gpus 8 --config 4 --steps 10 --warmup 3 \
    > gpurun_out/r02w_config4_n8.json 2> gpurun_out/r02w_config4_n8.err; echo "config4 rc=$?"
timeout 600 $TR --master-port 29623 bench.py --gpus 8 --config 5 --steps 10 --warmup 3 \
    > gpurun_out/r02w_config5_n8.json 2> gpurun_out/r02w_config5_n8.err; echo "config5 rc=$?"
for c in 3 4 5; do
python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r02w_config${c}_n8.json").read().strip().splitlines()[-1])
    print("config ${c}:", d["value"], "it/s", d["ms_per_step"], "ms; e2e", (d["e2e"] or {}).get("value"), "parity", d.get("parity_vs_n1"), d.get("memory"))
    for k, v in list(d["kernels"].items())[:9]:
        print("    ", k, v)
except Exception as e:
    print("config ${c}: no line", e)
PY
tail -4 gpurun_out/r02w_config${c}_n8.err
done
