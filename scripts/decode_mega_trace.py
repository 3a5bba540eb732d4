"""Stage timeline of the persistent decode kernel (csrc/decode.cu): globaltimer stamps the first and the last CTA's compute
warps leave at every stage boundary (DH_DECODE_TRACE=1), averaged over the layers of one replayed token step.  B200 only.
Usage: DH_DECODE_TRACE=1 python scripts/decode_mega_trace.py [utterances]"""
import os
import sys

os.environ["DH_DECODE_TRACE"] = "1"
os.environ["DH_DECODE_MEGA"] = "1"
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
generate_batch(model, prompts, 8, return_device=True)
eng = model._engine
loop = eng._bufs["decode_loops"][n]
assert loop.plan is not None, "the persistent decode kernel is not in use"
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for _ in range(16):
    loop.graph.replay()
e1.record()
torch.cuda.synchronize()
print(f"n={n}: token step {e0.elapsed_time(e1) / 16:.3f} ms (graph replay)")
from dualhyp_b200 import ops  # noqa: E402
raw = eng._bufs["dec_trace"].cpu().numpy().astype("int64")
L, P = eng.spec.n_layer, ops.sm_count()
tr = raw[: 3 * L * 24].reshape(3, L, 24)
bars = raw[3 * L * 24:].reshape(8, P, 2)
names = {0: "layer start", 1: "bar A (n ready)", 2: "QKV first tile MMA done", 3: "QKV drained", 4: "bar B", 5: "first item qkv reduce+rope",
         6: "attention items done", 7: "bar C", 8: "proj first tile MMA done", 9: "proj drained", 10: "bar D", 11: "resid+norm row",
         12: "bar E", 13: "fc12 first tile MMA done", 14: "fc12 drained", 15: "bar F", 16: "swiglu", 17: "bar G",
         18: "mlp first tile MMA done", 19: "mlp drained", 20: "bar H", 21: "resid+norm row (layer end)"}
order = sorted(names)
for c, who in enumerate(("CTA 0", f"CTA {P // 2}", f"CTA {P - 1}")):
    t = tr[c]
    print(f"{who}: {(t[2:, 0] - t[1:-1, 0]).mean() / 1e3:.1f} us per layer")
    prev = None
    for i in order:
        if prev is not None:
            d = (t[2:-1, i] - t[2:-1, prev]).mean() / 1e3
            print(f"   {names[i]:>30s}  +{d:6.2f} us")
        prev = i
print("grid barriers of layer 5: last arrival -> release seen (mean over CTAs), arrival spread (max - min), mean wait")
for b in range(8):
    arr, rel = bars[b, :, 0], bars[b, :, 1]
    if arr.min() == 0:
        continue
    print(f"   barrier {'ABCDEFGH'[b]}: mechanics {(rel.mean() - arr.max()) / 1e3:5.2f} us, release spread {(rel.max() - rel.min()) / 1e3:5.2f} us, "
          f"arrival spread {(arr.max() - arr.min()) / 1e3:5.2f} us, mean wait {(rel - arr).mean() / 1e3:5.2f} us; latest CTA {int(arr.argmax())}, earliest {int(arr.argmin())}")
