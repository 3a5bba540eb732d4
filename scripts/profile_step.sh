#!/bin/bash
# ncu evidence for one fine-tune step (run under gpurun, ONE GPU).  Usage: scripts/profile_step.sh <tag>
# 1. plain run must exit 0; 2. launch list (device time per launch); 3. ncu --set full on the dominant GEMM and the attention kernels.
set -u
TAG=${1:-r01}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-roofline --no-infer"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -s 1300 -c 1000 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_pair_kernel -s 700 -c 8 -o gpurun_out/${TAG}_gemm $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attn_ -s 120 -c 5 -o gpurun_out/${TAG}_attn $CMD > gpurun_out/${TAG}_ncu3.log 2>&1
tail -2 gpurun_out/${TAG}_plain.log | cut -c1-300
for f in 1 2 3; do tail -2 gpurun_out/${TAG}_ncu$f.log | cut -c1-200; done
