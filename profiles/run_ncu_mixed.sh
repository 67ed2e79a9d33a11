#!/bin/bash
# ncu evidence for the run-time radix Hessian kernels at a typical good_size padding (6000^2 / 12000^2).
TAG=${1:-r02zf}
python profiles/hess_probe.py 2 6000 f32 5 > gpurun_out/${TAG}_mixed_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_hessm -c 3 \
    -o gpurun_out/${TAG}_mixed -f python profiles/hess_probe.py 1 6000 f32 1 > gpurun_out/${TAG}_ncu.log 2>&1
REP=gpurun_out/${TAG}_mixed.ncu-rep
python profiles/ncu_summary.py $REP > gpurun_out/${TAG}_mixed_summary.txt 2>&1
for k in k_hessm_rows_fwd k_hessm_cols k_hessm_rows_inv; do
    python profiles/ncu_stalls.py $REP $k > gpurun_out/${TAG}_stalls_${k}.txt 2>&1
done
rm -f $REP
