# experiment: stream-DWT with flush on/off (PFB_DWT_DBG bit 0 = skip slab stores)
for d in 0 1; do echo "dbg=$d"; PFB_DWT_DBG=$d python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import json,sys;d=json.load(sys.stdin);print(d['ms_per_step'],[ (k[:24],v['ms_per_step']) for k,v in d['kernels'].items() if 'dwt' in k or 'self' in k])"; done
python -m pytest tests -m gpu -q -k "psi or minor" 2>&1 | tail -2
