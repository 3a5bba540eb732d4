"""Run the attention kernels alone (for ncu): python scripts/attn_only.py [B] [T] [prefix]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from dualhyp_b200 import ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
T = int(sys.argv[2]) if len(sys.argv) > 2 else 512
prefix = int(sys.argv[3]) if len(sys.argv) > 3 else 0
G, qpk = 4, 8
H, N = G * qpk, G * (qpk + 2) * 64
dev = torch.device("cuda")
qkv = torch.randn(B * T, N, device=dev, dtype=torch.bfloat16)
out = torch.empty(B * T, H * 64, device=dev, dtype=torch.bfloat16)
lse = torch.empty(B, H, T, device=dev)
pk = torch.randn(prefix, N, device=dev, dtype=torch.bfloat16) if prefix else None
gate = torch.randn(H, device=dev) if prefix else None
dout = torch.randn_like(out)
dqkv = torch.empty_like(qkv)
ws = torch.empty(ops.attention_bwd_workspace(B, T, G, qpk), dtype=torch.uint8, device=dev)
dpk = torch.zeros(prefix, N, device=dev) if prefix else None
dg = torch.zeros(H, device=dev) if prefix else None


def run():
    ops.attention_fwd(qkv, out, lse, B, T, G, qpk, pk, gate, prefix)
    ops.attention_bwd(qkv, out, dout, lse, dqkv, ws, B, T, G, qpk, pk, gate, dpk, dg, prefix)


for _ in range(3):
    run()
torch.cuda.synchronize()
e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
n = 20
e[0].record()
for _ in range(n):
    ops.attention_fwd(qkv, out, lse, B, T, G, qpk, pk, gate, prefix)
e[1].record()
for _ in range(n):
    ops.attention_bwd(qkv, out, dout, lse, dqkv, ws, B, T, G, qpk, pk, gate, dpk, dg, prefix)
e[2].record()
torch.cuda.synchronize()
print(f"B={B} T={T} prefix={prefix} fwd {e[0].elapsed_time(e[1]) / n * 1e3:.1f} us  bwd {e[1].elapsed_time(e[2]) / n * 1e3:.1f} us")
