"""Per-op device time of one PACKED prompt pass (engine.prefill_packed) on TinyLlama-1.1B.  Usage: python scripts/prefill_profile.py [n_prompts]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from dualhyp_b200 import ops, synthetic  # noqa: E402
from dualhyp_b200.generate import generate_batch  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
model = bench.build_model("adapter", dev)
prompts = [p.to(dev) for p in synthetic.synth_prompts(n, 32000, 4242)]
generate_batch(model, prompts, 4)                       # allocates the cache and the loop
eng = model._engine
loop = eng._bufs["decode_loops"][n]
lens = torch.tensor([p.numel() for p in prompts], dtype=torch.int32)
cu = torch.zeros(n + 1, dtype=torch.int32)
cu[1:] = lens.cumsum(0)
pos = torch.cat([torch.arange(int(L), dtype=torch.int32) for L in lens.tolist()]).to(dev)
seq = torch.repeat_interleave(torch.arange(n, dtype=torch.int32), lens.long()).to(dev)
ids = torch.cat(prompts)
cu_d = cu.to(dev)
Tmax = int(lens.max())
with torch.inference_mode():
    for _ in range(2):
        eng.prefill_packed(ids, cu_d, pos, seq, Tmax, loop.packed)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        eng.prefill_packed(ids, cu_d, pos, seq, Tmax, loop.packed)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    s = eng.spec
    flops = 2.0 * ids.numel() * s.n_layer * (s.qkv_out * s.n_embd + s.n_embd * s.n_embd + 3 * s.n_embd * s.intermediate_size)
    print(f"packed prefill of {n} prompts, {ids.numel()} tokens: {ms:.2f} ms  ({flops / ms / 1e9:.0f} TF/s on the linear layers alone)")
    with ops.profile() as recs:
        eng.prefill_packed(ids, cu_d, pos, seq, Tmax, loop.packed)
        torch.cuda.synchronize()
agg = {}
for name, meta, a, b in recs:
    key = (name, meta[:5] if meta else None)
    v = agg.setdefault(key, [0, 0.0])
    v[0] += 1
    v[1] += a.elapsed_time(b)
print(f"sum of per-op event times {sum(v[1] for v in agg.values()):.2f} ms")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
    fl = 2.0 * k[1][0] * k[1][1] * (k[1][2] + k[1][3]) if (k[0] == "gemm" and k[1]) else 0
    print(f"  {k[0]:18s} {str(k[1]):38s} x{v[0]:3d}  {1e3 * v[1] / v[0]:8.1f} us each  {v[1]:7.3f} ms" + (f"  {fl * v[0] / v[1] / 1e9:7.0f} TF/s" if fl else ""))
