#!/bin/bash
# Round-end evidence run (under gpurun, ONE GPU): parity tests, smoke, ncu captures of the final build, the four bench workloads.
# Usage: bash scripts/final_round.sh <tag>
set -u
TAG=${1:-r03f}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_gpu_tests.txt 2>&1; echo "gpu tests rc=$?"; tail -2 gpurun_out/${TAG}_gpu_tests.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.txt 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${TAG}_smoke.txt
bash scripts/profile_step.sh ${TAG} adapter
timeout 300 python scripts/decode_traffic.py 256 > gpurun_out/${TAG}_decode_plain.log 2>&1 || { echo "decode_traffic failed"; tail -3 gpurun_out/${TAG}_decode_plain.log; }
timeout 600 ncu --profile-from-start off --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv \
  --log-file gpurun_out/${TAG}_decode.csv python scripts/decode_traffic.py 256 > gpurun_out/${TAG}_ncu4.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:attn_decode_stream -c 2 -o gpurun_out/${TAG}_dattn \
  python scripts/decode_traffic.py 256 > gpurun_out/${TAG}_ncu5.log 2>&1
for f in 4 5; do tail -1 gpurun_out/${TAG}_ncu$f.log | cut -c1-160; done
timeout 120 python scripts/attn_decode_probe.py 576 256 > gpurun_out/${TAG}_attn_decode_probe.txt 2>&1
timeout 120 python scripts/attn_decode_probe.py 576 128 >> gpurun_out/${TAG}_attn_decode_probe.txt 2>&1
timeout 120 python scripts/attn_decode_probe.py 576 64 >> gpurun_out/${TAG}_attn_decode_probe.txt 2>&1
DH_DEC_ATTN=mma timeout 120 python scripts/attn_decode_probe.py 576 256 >> gpurun_out/${TAG}_attn_decode_probe.txt 2>&1
grep sequences gpurun_out/${TAG}_attn_decode_probe.txt
