"""DRAM traffic of ONE decode token step: run under
  ncu --profile-from-start off --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv
The profiled range is two eager token steps of `n` sequences (prompts U[330,480], about 16 tokens generated so far).
Usage: python scripts/decode_traffic.py [n]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import build_model  # noqa: E402
from dualhyp_b200 import synthetic  # noqa: E402
from dualhyp_b200.generate import generate_batch  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
dev = torch.device("cuda", 0)
model = build_model("adapter", dev)
prompts = [p.to(dev) for p in synthetic.synth_prompts(n, 32000, 4242)]
generate_batch(model, prompts, 16, return_device=True, use_graph=False)
loop = model._engine._bufs["decode_loops"][n]
with torch.inference_mode():
    loop.body()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    loop.body()
    loop.body()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
print("kernels per token step:", loop.kernels_per_token)
