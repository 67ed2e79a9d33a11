#!/bin/bash
# where the N-GPU step goes: all-reduce alone, step without the exchange, step without the overlap
N=${1:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 profiles/nccl_probe.py 2>&1 | grep "all_reduce"
for env in "X=1" "PFB_PAR_NOOVERLAP=1"; do
  echo "== $env"
  env $env python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], {k:v['ms_per_step'] for k,v in d['kernels'].items()})"
done
