#!/bin/bash
# Hessian parity subset + kernel-timing bench line (iteration helper)
python -m pytest tests -m gpu -x -q -k "hess or minor" 2>&1 | tail -3
python bench.py --no-e2e --no-cpu-baseline "$@" 2>gpurun_out/b.err | tee gpurun_out/b_quick.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])
for k,v in d['kernels'].items(): print(k, v)"
tail -3 gpurun_out/b.err
