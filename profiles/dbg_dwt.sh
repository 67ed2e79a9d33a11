#!/bin/bash
# experiments on the finest-level analysis walk (k_dwt_stream_all): strip width and store suppression
for cfg in "29 0" "24 0" "28 0" "16 0" "29 1" "29 2" "29 3" "24 1"; do
  set -- $cfg
  echo "== SW=$1 DBG=$2"
  PFB_DWT_SW=$1 PFB_DWT_DBG=$2 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['kernels'].get('k_dwt_stream'))"
done
