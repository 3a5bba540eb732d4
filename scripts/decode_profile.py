"""Per-op device time of ONE eager decode token step (CUDA events around every C-ABI call), B200 only.
Usage: python scripts/decode_profile.py [utterances] ."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import build_model  # noqa: E402
from dualhyp_b200 import ops, synthetic  # noqa: E402
from dualhyp_b200.generate import generate_batch  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda", 0)
model = build_model("adapter", dev)
prompts = [p.to(dev) for p in synthetic.synth_prompts(n, 32000, 4242)]
generate_batch(model, prompts, 8, return_device=True)
loop = model._engine._bufs["decode_loops"][n]
with torch.inference_mode():
    for _ in range(3):
        loop.body()
    torch.cuda.synchronize()
    with ops.profile() as recs:
        for _ in range(4):
            loop.body()
        torch.cuda.synchronize()
agg = {}
for name, meta, e0, e1 in recs:
    key = (name, meta[:5] if meta else None)
    a = agg.setdefault(key, [0, 0.0])
    a[0] += 1
    a[1] += e0.elapsed_time(e1)
tot = sum(v[1] for v in agg.values()) / 4
print(f"n={n}: sum of per-op event times {tot:.3f} ms per token step")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"  {k[0]:18s} {str(k[1]):38s} x{v[0] // 4:3d}  {1e3 * v[1] / v[0]:8.1f} us each  {v[1] / 4:7.3f} ms/step")

# ---- prefill: per-op device time of one batched prompt pass (same prompts) ----
eng = model._engine
ids = torch.nn.utils.rnn.pad_sequence(prompts, batch_first=True, padding_value=0)
with torch.inference_mode():
    eng.prefill(ids, loop.packed)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    eng.prefill(ids, loop.packed)
    e1.record()
    torch.cuda.synchronize()
    print(f"prefill of {tuple(ids.shape)}: {e0.elapsed_time(e1):.2f} ms (events around the whole pass)")
    with ops.profile() as recs:
        eng.prefill(ids, loop.packed)
        torch.cuda.synchronize()
agg = {}
for name, meta, a, b in recs:
    key = (name, meta[:5] if meta else None)
    v = agg.setdefault(key, [0, 0.0])
    v[0] += 1
    v[1] += a.elapsed_time(b)
print(f"sum of per-op event times {sum(v[1] for v in agg.values()):.2f} ms")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:12]:
    print(f"  {k[0]:18s} {str(k[1]):38s} x{v[0]:3d}  {1e3 * v[1] / v[0]:8.1f} us each  {v[1]:7.3f} ms")
