"""The bar BASELINE.md section 3 names: the same DualHyp adapter fine-tune step on the STOCK stack a user of the reference gets on
this box -- cuBLAS through F.linear, flash-attn 2.x `flash_attn_func` for the causal attention (ger/model.py:280-286 takes it
when the module imports and the tensors are bf16), unfused ATen for RMSNorm / RoPE / SwiGLU / cross entropy, torch autograd for
the backward -- restated here in plain torch (bf16-true, ger/model.py:163-316, ger/adapter.py:116-146, finetune/ger.py:278-285)
on random-init TinyLlama-1.1B weights and the bench's synthetic batch.  Self-contained on purpose (no import of oracle/ or of
the product kernels).  B200 only:
    python scripts/stock_stack_step.py [prompts=16] [seq=512] [steps=5]
Prints ms per step (CUDA events, after 3 warm-up steps) and tokens/s over the B x T positions of the right-padded batch."""
import math
import os
import sys
import time

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dualhyp_b200 import synthetic  # noqa: E402  (inputs only: the counter-hash batch bench.py uses)

try:
    from flash_attn import flash_attn_func
except Exception as e:  # noqa: BLE001
    flash_attn_func = None
    print("flash_attn not importable:", e)

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
T = int(sys.argv[2]) if len(sys.argv) > 2 else 512
STEPS = int(sys.argv[3]) if len(sys.argv) > 3 else 5
L, NH, G, C, I, V, HS, AT, START, EPS = 22, 32, 4, 2048, 5632, 32000, 64, 10, 2, 1e-5
QPK = NH // G
dev = torch.device("cuda", 0)
bf = torch.bfloat16
gen = torch.Generator(device=dev).manual_seed(1337)


def lin(o, i):
    return ((torch.rand((o, i), device=dev, generator=gen) * 2 - 1) / math.sqrt(i)).to(bf)


W = dict(wte=torch.randn((V, C), device=dev, generator=gen).to(bf), ln_f=torch.ones(C, device=dev, dtype=bf), head=lin(V, C), layers=[])
for i in range(L):
    lw = dict(n1=torch.ones(C, device=dev, dtype=bf), n2=torch.ones(C, device=dev, dtype=bf), qkv=lin((NH + 2 * G) * HS, C), proj=lin(C, C),
              fc1=lin(I, C), fc2=lin(I, C), mlp=lin(C, I))
    if i >= START:
        lw["awte"] = torch.randn((AT, C), device=dev, generator=gen).to(bf).requires_grad_(True)
        lw["gate"] = (torch.randn((1, 1, NH, 1), device=dev, generator=gen) * 0.5).to(bf).requires_grad_(True)
    W["layers"].append(lw)
theta = 1.0 / (10000 ** (torch.arange(0, HS, 2, device=dev).float() / HS))
ang = torch.outer(torch.arange(T, device=dev).float(), theta).repeat(1, 2)
cos, sin = ang.cos().to(bf), ang.sin().to(bf)                                   # ger/model.py:349-355 (bf16 tables, quirk Q1)


def rmsnorm(x, w):                                                               # ger/rmsnorm.py:17-21
    xf = x.float()
    return (xf * torch.rsqrt(torch.mean(xf * xf, dim=-1, keepdim=True) + EPS)).to(x.dtype) * w


def rope(x):                                                                     # ger/model.py:358-365
    x1, x2 = x[..., : HS // 2], x[..., HS // 2:]
    return ((x * cos) + (torch.cat((-x2, x1), dim=-1) * sin)).type_as(x)


def split_heads(qkv, t):                                                         # ger/model.py:216-231
    qkv = qkv.view(qkv.shape[0], t, G, QPK + 2, HS).permute(0, 2, 3, 1, 4)
    q, k, v = qkv.split((QPK, 1, 1), dim=2)
    k = k.expand(-1, G, QPK, t, HS)
    v = v.expand(-1, G, QPK, t, HS)
    return q.reshape(-1, NH, t, HS), k.reshape(-1, NH, t, HS), v.reshape(-1, NH, t, HS)


def attention(x, lw):
    q, k, v = split_heads(F.linear(x, lw["qkv"]), T)
    q, k = rope(q), rope(k)
    scale = 1.0 / math.sqrt(HS)
    if flash_attn_func is not None:                                              # ger/model.py:280-286
        y = flash_attn_func(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), dropout_p=0.0, softmax_scale=scale, causal=True).transpose(1, 2)
    else:
        y = F.scaled_dot_product_attention(q, k, v, is_causal=True, scale=scale)
    if "awte" in lw:                                                             # ger/adapter.py:129-146
        _, ak, av = split_heads(F.linear(lw["awte"].view(1, AT, C), lw["qkv"]), AT)
        ay = F.scaled_dot_product_attention(q, ak.expand(q.shape[0], -1, -1, -1), av.expand(q.shape[0], -1, -1, -1), attn_mask=None, scale=scale)
        y = y + lw["gate"].view(1, NH, 1, 1) * ay          # gating_factor is (1, 1, n_head, 1) against y as (B, T, n_head, hs) in the reference
    return F.linear(y.transpose(1, 2).reshape(x.shape[0], T, C), lw["proj"])


def step(idx, targets):
    x = F.embedding(idx, W["wte"])
    for lw in W["layers"]:
        x = x + attention(rmsnorm(x, lw["n1"]), lw)
        n2 = rmsnorm(x, lw["n2"])
        x = x + F.linear(F.silu(F.linear(n2, lw["fc1"])) * F.linear(n2, lw["fc2"]), lw["mlp"])
    h = rmsnorm(x, W["ln_f"])
    # finetune/ger.py:278-281 + ger/utils.py:424-463: lm_head in chunks of 128 positions, CE summed over chunks / all positions
    loss, n = 0.0, 0
    for c0 in range(0, T, 128):
        lg = F.linear(h[:, c0:c0 + 128], W["head"])
        tg = targets[:, c0:c0 + 128]
        if c0 + 128 >= T:
            lg = lg[:, :-1]
        loss = loss + F.cross_entropy(lg.reshape(-1, V).float(), tg[:, : lg.shape[1]].reshape(-1), ignore_index=-1, reduction="sum")
        n += lg.shape[0] * lg.shape[1]
    loss = loss / n
    loss.backward()
    return loss


idx, labels = synthetic.synth_batch(B, T, V, 1337)
idx, targets = idx.to(dev), labels[:, 1:].to(dev)
targets = F.pad(targets, (0, 1), value=-1)
params = [p for lw in W["layers"] for p in (lw.get("awte"), lw.get("gate")) if p is not None]
for _ in range(3):
    step(idx, targets)
    for p in params:
        p.grad = None
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(STEPS):
    loss = step(idx, targets)
    for p in params:
        p.grad = None
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / STEPS
print(f"stock stack (cuBLAS + {'flash-attn ' + __import__('flash_attn').__version__ if flash_attn_func else 'torch SDPA'} + ATen + autograd), adapter fine-tune step, "
      f"{B} x {T} right-padded, bf16-true: {ms:.2f} ms per step = {B * T / ms * 1e3:.0f} padded tokens/s (loss {float(loss):.4f}); "
      f"peak memory {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB")
