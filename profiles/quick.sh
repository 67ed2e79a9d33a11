#!/bin/bash
# quick GPU check used while iterating: psi/minor parity tests + one kernel-timing bench line
python -m pytest tests -m gpu -x -q -k "psi or minor or solvers or gap" 2>&1 | tail -3
python bench.py --no-e2e --no-cpu-baseline "$@" 2>gpurun_out/b.err | tee gpurun_out/b_quick.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])
for k,v in d['kernels'].items(): print(k, v)
print(d['ops'])"
tail -5 gpurun_out/b.err
