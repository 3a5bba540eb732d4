#!/bin/bash
# ncu evidence for one fine-tune step (run under gpurun, ONE GPU).  Usage: scripts/profile_step.sh <tag> [workload]
# 1. the plain run must exit 0; 2. launch list (device time per launch, cold-cache and serialised: shares, not absolutes);
# 3. ncu --set full on the dominant GEMM (its DRAM bytes feed profiles/traffic.json through scripts/capture_traffic.py) and on the
#    attention kernels.  Never quote a throughput printed under ncu.
set -u
TAG=${1:-r02}
WL=${2:-adapter}
CMD="python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu-baseline --no-roofline --no-infer --no-lora-sample --no-e2e"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_pair_kernel -s 500 -c 24 -o gpurun_out/${TAG}_gemm $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attn_ -s 130 -c 6 -o gpurun_out/${TAG}_attn $CMD > gpurun_out/${TAG}_ncu3.log 2>&1
tail -2 gpurun_out/${TAG}_plain.log | cut -c1-300
for f in 1 2 3; do tail -2 gpurun_out/${TAG}_ncu$f.log | cut -c1-200; done
