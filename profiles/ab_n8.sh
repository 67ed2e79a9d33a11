#!/bin/bash
# A/B on one 8-GPU box: band-sharded loop with the fused primal pass (default) and without it.
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
for rep in 1 2; do
for v in fuse nofuse; do
  if [ $v = nofuse ]; then export PFB_PAR_NOFUSE=1; else unset PFB_PAR_NOFUSE; fi
  timeout 200 $TR --master-port 2963$rep bench.py --gpus 8 --steps 40 --warmup 5 --no-cpu-baseline --no-e2e --no-parity \
      > gpurun_out/r02x_n8_${v}_$rep.json 2> gpurun_out/r02x_n8_${v}_$rep.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/r02x_n8_${v}_$rep.json").read().strip().splitlines()[-1])
print("$v $rep", d["value"], d["ms_per_step"], {k: x["ms_per_step"] for k, x in list(d["kernels"].items())[:7]})
PY
done
done
