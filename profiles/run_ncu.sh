#!/bin/bash
# ncu evidence for one primal-dual step (run under gpurun from the repo root).
#   $1 = tag (e.g. r01a), $2 = kernel regex (default: every library kernel, k_*)
# Writes gpurun_out/<tag>_launches.csv (per-launch time + DRAM bytes) and <tag>_full.ncu-rep
set -e
TAG=${1:-r01}
KRE=${2:-k_}
CMD="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --ncu-step"
$CMD > gpurun_out/${TAG}_plain.log 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    --profile-from-start off -k regex:k_ --csv --log-file gpurun_out/${TAG}_launches.csv $CMD \
    > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:${KRE} \
    -o gpurun_out/${TAG}_full -f $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
