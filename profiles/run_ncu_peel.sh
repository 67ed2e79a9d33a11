#!/bin/bash
# ncu evidence for the peeled (odd factor x 2^k) Hessian kernels at config 5's size:
# $1 = tag.  Plain run first, then one --set full capture of the three passes (one band).
TAG=${1:-r02q}
python profiles/hess_probe.py 2 10240 f32 5 > gpurun_out/${TAG}_peel_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_hessp -c 3 \
    -o gpurun_out/${TAG}_peel -f python profiles/hess_probe.py 1 10240 f32 1 > gpurun_out/${TAG}_ncu.log 2>&1
REP=gpurun_out/${TAG}_peel.ncu-rep
python profiles/ncu_summary.py $REP > gpurun_out/${TAG}_peel_summary.txt 2>&1
for k in k_hessp_rows_fwd k_hessp_cols k_hessp_rows_inv; do
    python profiles/ncu_stalls.py $REP $k > gpurun_out/${TAG}_stalls_${k}.txt 2>&1
done
rm -f $REP
