#!/bin/bash
# multi-GPU bench line: profiles/bench_n.sh N [extra bench.py flags]
N=$1; shift
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline "$@" 2>gpurun_out/b$N.err | tee gpurun_out/bench_n$N.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'] if d['e2e'] else None)
for k,v in d['kernels'].items(): print(k, v)"
tail -2 gpurun_out/b$N.err
