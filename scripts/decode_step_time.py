"""Token-step time of the decode CUDA graph alone (graph replays between events), for A/B runs of decode knobs.  B200 only.
Usage: [DH_DECODE_SPLIT_MAX=..] python scripts/decode_step_time.py [utterances ...]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import build_model  # noqa: E402
from dualhyp_b200 import synthetic  # noqa: E402
from dualhyp_b200.generate import generate_batch  # noqa: E402

dev = torch.device("cuda", 0)
model = build_model("adapter", dev)
for n in [int(a) for a in sys.argv[1:]] or [256]:
    prompts = [p.to(dev) for p in synthetic.synth_prompts(n, 32000, 4242)]
    generate_batch(model, prompts, 8, return_device=True)
    loop = model._engine._bufs["decode_loops"][n]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(3):
        torch.cuda.synchronize()
        e0.record()
        for _ in range(16):
            loop.graph.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 16)
    print(f"n={n}: token step {best:.3f} ms (best of 3 x 16 graph replays)", flush=True)
