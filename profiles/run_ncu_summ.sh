#!/bin/bash
# As run_ncu.sh, but the (large) report is condensed ON THE BOX and removed: gpurun returns at most
# 64 MiB.  $1 = tag, $2 = kernel regex of the --set full capture.
TAG=${1:-r02}
KRE=${2:-k_hessf|k_dwt|k_idwt|k_dual|k_pd}
bash profiles/run_ncu.sh $TAG "$KRE"
REP=gpurun_out/${TAG}_full.ncu-rep
python profiles/ncu_summary.py $REP > gpurun_out/${TAG}_full_summary.txt 2>&1
for k in k_hessf_rows_inv_tma k_hessf_cols1 k_hessf_rows_fwd k_dwt_stream_all k_idwt_stream; do
    python profiles/ncu_stalls.py $REP $k > gpurun_out/${TAG}_stalls_${k}.txt 2>&1
done
cp profiles/traffic.json /tmp/traffic_before.json
python profiles/make_traffic.py $TAG $REP > /dev/null 2>&1 && cp profiles/traffic.json gpurun_out/${TAG}_traffic.json
rm -f $REP
