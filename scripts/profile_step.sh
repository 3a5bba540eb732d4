#!/bin/bash
# ncu evidence for one fine-tune step (run under gpurun, ONE GPU).  Usage: scripts/profile_step.sh <tag> [gemm_skip]
# 1. plain run must exit 0; 2. launch list (device time per launch); 3. ncu --set full on a few backward GEMMs.
set -u
TAG=${1:-r01}
SKIP=${2:-745}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-roofline"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 1300 -c 1000 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s $SKIP -c 6 -o gpurun_out/${TAG}_gemm $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attn_ -s 200 -c 4 -o gpurun_out/${TAG}_attn $CMD > gpurun_out/${TAG}_ncu3.log 2>&1
tail -2 gpurun_out/${TAG}_plain.log | cut -c1-300
tail -3 gpurun_out/${TAG}_ncu1.log gpurun_out/${TAG}_ncu2.log gpurun_out/${TAG}_ncu3.log
