"""Decode attention alone on cold KV caches: a CUDA graph of 22 calls over 22 different caches (one per layer, as in the
token step), events around the replay.  Usage: [DH_DEC_CUDA_CORES=1] python scripts/attn_decode_probe.py [max_len [sequences]]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dualhyp_b200 import ops  # noqa: E402

max_len = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
dev = torch.device("cuda", 0)
nseq = int(sys.argv[2]) if len(sys.argv) > 2 else 64
G, qpk, L = 4, 8, 22
torch.manual_seed(0)
lens = torch.randint(394, 545, (nseq,), dtype=torch.int32, device=dev)
ck = [torch.randn(nseq, G, max_len, 64, device=dev).bfloat16() for _ in range(L)]
cv = [torch.randn(nseq, G, max_len, 64, device=dev).bfloat16() for _ in range(L)]
qkv = torch.randn(nseq, G * (qpk + 2) * 64, device=dev).bfloat16()
pk = torch.randn(10, G * (qpk + 2) * 64, device=dev).bfloat16()
gate = torch.randn(G * qpk, device=dev)
out = torch.empty(nseq, G * qpk * 64, device=dev, dtype=torch.bfloat16)


def run():
    for i in range(L):
        ops.attention_decode(qkv, ck[i], cv[i], lens, out, G, qpk, pk, gate, 10)


run()
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    run()
g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
best = 1e9
for _ in range(5):
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
mb = float(lens.sum()) * G * 64 * 2 * 2 / 1e6
print(f"{nseq} sequences, cache rows of {max_len}: {best / L * 1e3:.2f} us per call, K+V {mb:.1f} MB -> {mb / (best / L) / 1e3:.2f} TB/s")

if os.environ.get("DH_DEC_ATTN_DBG") == "3":
    import ctypes as C
    from dualhyp_b200 import _lib
    buf = (C.c_ulonglong * 8)()
    span = (C.c_ulonglong * (2 * 512 * 4))()
    _lib.load().dh_debug_ds_trace(buf, span)
    import numpy as np
    sp = np.array(span, dtype=np.int64).reshape(2, 512, 4)
    nb = min(512, nseq * G, (ops.sm_count() // G) * G)
    a, b = sp[0, :nb], sp[1, :nb]                      # calls 20 and 21 of the graph (22 calls: slots alternate)
    if a[:, 0].min() > b[:, 0].min():
        a, b = b, a
    t0 = a[:, 0].min()
    def show(name, x):
        print(f"   {name}: blocks enter {x[:, 0].min() - t0} .. {x[:, 0].max() - t0} ns, reach the item loop {x[:, 1].min() - t0} .. {x[:, 1].max() - t0} "
              f"(entry -> loop {np.mean(x[:, 1] - x[:, 0]):.0f} ns mean), leave it {x[:, 2].min() - t0} .. {x[:, 2].max() - t0} (loop {np.mean(x[:, 2] - x[:, 1]):.0f} ns mean, "
              f"{(x[:, 2] - x[:, 1]).max()} max)")
    show("call N  ", a)
    show("call N+1", b)
    t = [int(x) for x in buf]
    warps, chunks, items = t[7], t[5], t[6]
    print(f"per compute warp (cycles): total {t[4] / warps:.0f}; wait-full {t[0] / warps:.0f}, chunk math {t[1] / warps:.0f}, part stores + barrier wait {t[2] / warps:.0f}, "
          f"merge + store {t[3] / warps:.0f}; {chunks / warps:.2f} chunks and {items / warps:.2f} items per warp; {t[1] / max(chunks, 1):.0f} cycles per chunk, "
          f"{t[2] / max(items, 1):.0f} per barrier wait, {t[3] / max(items, 1):.0f} per merge + store")
