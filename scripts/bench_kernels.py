"""Micro-benchmarks of the individual kernels at the TinyLlama shapes (CUDA events, L2 flushed between
iterations), next to cuBLAS / flash-attn on the same shapes for context.  Development tool, not bench.py."""
import json
import sys
import os

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dualhyp_b200 import ops  # noqa: E402

DEV = "cuda"
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=DEV)


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def gemm_case(M, N, K, b_layout=0, name=""):
    A = torch.randn(M, K, device=DEV, dtype=torch.bfloat16)
    B = torch.randn((N, K) if b_layout == 0 else (K, N), device=DEV, dtype=torch.bfloat16)
    out = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    t = timeit(lambda: ops.gemm(A, B, out, b_layout=b_layout))
    Bt = B.t() if b_layout == 0 else B
    tc = timeit(lambda: torch.matmul(A, Bt, out=out))
    fl = 2.0 * M * N * K
    print(json.dumps({"kernel": "gemm", "name": name, "M": M, "N": N, "K": K, "b_layout": b_layout, "ms": round(t, 4),
                      "tflops": round(fl / t / 1e9, 1), "cublas_ms": round(tc, 4), "cublas_tflops": round(fl / tc / 1e9, 1)}), flush=True)


def gemm_epilogue_cases(M, C=2048, I=5632, G=4, qpk=8):
    """The decoder's GEMMs with their real fused epilogues (what the training step launches)."""
    bf = torch.bfloat16
    n = torch.randn(M, C, device=DEV, dtype=bf)
    x32 = torch.randn(M, C, device=DEV)
    out32 = torch.empty(M, C, device=DEV)
    cos = torch.randn(2048, 64, device=DEV, dtype=bf)
    NQ = G * (qpk + 2) * 64
    wqkv = torch.randn(NQ, C, device=DEV, dtype=bf) * 0.02
    qkv = torch.empty(M, NQ, device=DEV, dtype=bf)
    w12 = torch.randn(2 * I, C, device=DEV, dtype=bf) * 0.02
    h12 = torch.empty(M, 2 * I, device=DEV, dtype=bf)
    act = torch.empty(M, I, device=DEV, dtype=bf)
    wmlp = torch.randn(C, I, device=DEV, dtype=bf) * 0.02
    wmlp_t = wmlp.t().contiguous()
    w12_t = w12.t().contiguous()
    wproj = torch.randn(C, C, device=DEV, dtype=bf) * 0.02
    dh12 = torch.empty(M, 2 * I, device=DEV, dtype=bf)
    dn = torch.empty(M, C, device=DEV, dtype=bf)
    cases = [
        ("qkv+rope", NQ, C, lambda: ops.gemm(n, wqkv, qkv, epilogue=ops.EPI_QKV_ROPE, rope_cos=cos, rope_sin=cos, seq_len=512, q_per_kv=qpk)),
        ("proj+resid_f32", C, C, lambda: ops.gemm(n, wproj, out32, epilogue=ops.EPI_RESID_F32, aux=x32)),
        ("fc12+swiglu", 2 * I, C, lambda: ops.gemm(n, w12, h12, epilogue=ops.EPI_SWIGLU, out2=act)),
        ("fc12+swiglu(no h12)", 2 * I, C, lambda: ops.gemm(n, w12, None, epilogue=ops.EPI_SWIGLU, out2=act)),
        ("mlp_proj+resid_f32", C, I, lambda: ops.gemm(act, wmlp, out32, epilogue=ops.EPI_RESID_F32, aux=x32)),
        ("d_act+swiglu_bwd", I, C, lambda: ops.gemm(n, wmlp_t, dh12, epilogue=ops.EPI_SWIGLU_BWD, aux=h12)),
        ("d_n2", C, 2 * I, lambda: ops.gemm(dh12, w12_t, dn)),
    ]
    for name, N, K, fn in cases:
        t = timeit(fn)
        fl = 2.0 * M * N * K
        print(json.dumps({"kernel": "gemm+epi", "name": name, "M": M, "N": N, "K": K, "ms": round(t, 4), "tflops": round(fl / t / 1e9, 1)}), flush=True)


def attn_case(B, T, G=4, qpk=8):
    N, H = G * (qpk + 2) * 64, G * qpk
    qkv = torch.randn(B * T, N, device=DEV, dtype=torch.bfloat16)
    out = torch.empty(B * T, H * 64, device=DEV, dtype=torch.bfloat16)
    lse = torch.empty(B, H, T, device=DEV)
    t = timeit(lambda: ops.attention_fwd(qkv, out, lse, B, T, G, qpk))
    dout = torch.randn_like(out)
    dqkv = torch.empty_like(qkv)
    ws = torch.empty(ops.attention_bwd_workspace(B, T, G, qpk), dtype=torch.uint8, device=DEV)
    tb = timeit(lambda: ops.attention_bwd(qkv, out, dout, lse, dqkv, ws, B, T, G, qpk))
    fl = 4.0 * B * H * T * T * 64 / 2
    res = {"kernel": "attention", "B": B, "T": T, "fwd_ms": round(t, 4), "fwd_tflops": round(fl / t / 1e9, 1),
           "bwd_ms": round(tb, 4), "bwd_tflops": round(2.5 * fl / tb / 1e9, 1)}
    try:
        from flash_attn import flash_attn_func
        q = torch.randn(B, T, H, 64, device=DEV, dtype=torch.bfloat16, requires_grad=True)
        k = torch.randn(B, T, G, 64, device=DEV, dtype=torch.bfloat16, requires_grad=True)
        v = torch.randn(B, T, G, 64, device=DEV, dtype=torch.bfloat16, requires_grad=True)
        res["fa2_fwd_ms"] = round(timeit(lambda: flash_attn_func(q, k, v, causal=True)), 4)
        o = flash_attn_func(q, k, v, causal=True)
        do = torch.randn_like(o)
        res["fa2_bwd_ms"] = round(timeit(lambda: torch.autograd.grad(o, (q, k, v), do, retain_graph=True)), 4)
    except Exception as ex:  # noqa: BLE001
        res["fa2"] = repr(ex)[:80]
    print(json.dumps(res), flush=True)


def leaf_cases(M=8192, C=2048):
    x = torch.randn(M, C, device=DEV)
    w = torch.ones(C, device=DEV)
    y = torch.empty(M, C, device=DEV, dtype=torch.bfloat16)
    rstd = torch.empty(M, device=DEV)
    t = timeit(lambda: ops.rmsnorm_fwd(x, w, y, rstd, 1e-5))
    by = M * C * 6
    print(json.dumps({"kernel": "rmsnorm_fwd", "ms": round(t, 4), "GBs": round(by / t / 1e6, 1)}), flush=True)
    dy = torch.randn(M, C, device=DEV, dtype=torch.bfloat16)
    dx = torch.zeros(M, C, device=DEV)
    dxb = torch.empty(M, C, device=DEV, dtype=torch.bfloat16)
    t = timeit(lambda: ops.rmsnorm_bwd(x, w, rstd, dy, dx, dx, dxb))
    by = M * C * (4 + 2 + 4 + 4 + 2)
    print(json.dumps({"kernel": "rmsnorm_bwd", "ms": round(t, 4), "GBs": round(by / t / 1e6, 1)}), flush=True)
    V = 32000
    logits = torch.randn(M, V, device=DEV, dtype=torch.bfloat16)
    tg = torch.randint(0, V, (M,), device=DEV)
    rl = torch.empty(M, device=DEV)
    t = timeit(lambda: ops.cross_entropy(logits, tg, rl, 1.0, True), iters=5)
    print(json.dumps({"kernel": "cross_entropy", "ms": round(t, 4), "GBs": round(M * V * 4 / t / 1e6, 1)}), flush=True)


if __name__ == "__main__":
    M = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    for name, N, K, bl in [("qkv", 2560, 2048, 0), ("proj", 2048, 2048, 0), ("fc12", 11264, 2048, 0), ("mlp_proj", 2048, 5632, 0),
                           ("lm_head", 32000, 2048, 0), ("d_act", 5632, 2048, 1), ("d_n2", 2048, 11264, 1), ("d_n1", 2048, 2560, 1),
                           ("d_hidden", 2048, 32000, 1)]:
        gemm_case(M, N, K, bl, name)
    gemm_epilogue_cases(M)
    attn_case(M // 512, 512)
    leaf_cases(M)
