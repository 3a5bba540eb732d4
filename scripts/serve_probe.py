"""Continuous batching (dualhyp_b200.serve.DecodeService) against static batches (generate_batch) on TinyLlama-1.1B with
answers of different lengths: 256 synthetic DualHyp prompts U[330,480], per-utterance token budgets U[lo,hi] standing in for
EOS (random-init weights never emit it).  Static = batches of 64 that each run to their longest answer; service = 64 slots
refilled as answers end.  Both host prompts -> host ids.  Usage: python scripts/serve_probe.py [lo hi]"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from dualhyp_b200 import synthetic  # noqa: E402
from dualhyp_b200.generate import generate_batch  # noqa: E402
from dualhyp_b200.serve import DecodeService  # noqa: E402

lo, hi = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (8, 64)
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
model = bench.build_model("adapter", dev)
N, SLOTS = 256, 64
prompts = synthetic.synth_prompts(N, 32000, 777)
g = torch.Generator().manual_seed(5)
budgets = torch.randint(lo, hi + 1, (N,), generator=g).tolist()


def static():
    out = []
    for i in range(0, N, SLOTS):
        b = budgets[i: i + SLOTS]
        full = generate_batch(model, prompts[i: i + SLOTS], max(b))
        out += [f[: p.numel() + k] for f, p, k in zip(full, prompts[i: i + SLOTS], b)]
    return out


svc = DecodeService(model, SLOTS, hi, poll_every=8)


def timed(fn):
    fn()
    torch.cuda.synchronize()
    best, res = 1e9, None
    for _ in range(3):
        t0 = time.perf_counter()
        res = fn()
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return best, res


ts, rs = timed(static)
tv, rv = timed(lambda: svc.run(prompts, budgets))
same = sum(a.tolist() == b.tolist() for a, b in zip(rs, rv))
print(f"budgets U[{lo},{hi}] (mean {sum(budgets) / N:.1f}), {N} utterances, {SLOTS} slots")
print(f"static batches : {ts * 1e3:8.1f} ms  {N / ts:7.1f} utt/s")
print(f"service        : {tv * 1e3:8.1f} ms  {N / tv:7.1f} utt/s   stats {svc.stats}")
print(f"identical outputs: {same} / {N}")
