"""Decode-shaped GEMMs (M = 64 rows) on cold weights: this library's skinny kernel vs torch.matmul (cuBLAS) as a yardstick.
Each measurement replays a CUDA graph of 24 calls over 24 DIFFERENT weight tensors (> 126 MB L2 in total), events around
the replay.  Usage: python scripts/skinny_probe.py [M]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dualhyp_b200 import ops  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda", 0)
REP = 24


def timed(fn):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
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
    return best / REP * 1e3      # us per call


for N, K in [(2048, 2048), (2560, 2048), (11264, 2048), (2048, 5632), (32000, 2048)]:
    reps = REP if N < 32000 else 4
    Ws = [torch.randn(N, K, device=dev, dtype=torch.bfloat16) * 0.02 for _ in range(reps)]
    x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)

    def ours():
        for W in Ws:
            ops.gemm(x, W, out)

    def cublas():
        for W in Ws:
            torch.matmul(x, W.t(), out=out)

    scale = REP / reps
    t_o, t_c = timed(ours) * scale, timed(cublas) * scale
    mb = N * K * 2 / 1e6
    print(f"M={M} N={N:6d} K={K:5d}  weights {mb:6.1f} MB   ours {t_o:7.2f} us ({mb / t_o * 1e3 / 1e3:5.2f} TB/s)   cuBLAS {t_c:7.2f} us ({mb / t_c * 1e3 / 1e3:5.2f} TB/s)")
