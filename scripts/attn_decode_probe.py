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
