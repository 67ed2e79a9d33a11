#!/bin/bash
# N-GPU step vs number of all-reduce groups of the band-sum exchange
N=${1:-8}
for c in 1 2 5; do
  echo "== PFB_PAR_CHUNKS=$c"
  PFB_PAR_CHUNKS=$c python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], {k:v['ms_per_step'] for k,v in d['kernels'].items()})"
done
