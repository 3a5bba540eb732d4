"""GPU parity tests of the individual sm_100a kernels, called through the C ABI (dualhyp_b200.ops ->
ctypes -> libdualhyp.so).  The checker is fp32 torch arithmetic on the same bf16-rounded inputs (the same
formulas `oracle/oracle.py` restates); tolerances are stated per test and are bf16-output tolerances:
one bf16 ulp is 2^-8 relative, so `rel_err` (Frobenius) < 1e-2 means "within a few bf16 ulps"."""
import math

import pytest
import torch
import torch.nn.functional as F

from dualhyp_b200 import ops, synthetic
from oracle import oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def rnd(shape, stream, bound=1.0, dtype=torch.bfloat16, seed=11):
    return synthetic.hash_uniform(shape, seed, stream, bound, device=DEV).to(dtype)


def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def report(name, got, ref):
    got, ref = got.float(), ref.float()
    err = (got - ref).abs()
    i = int(err.argmax())
    return (f"{name}: rel {rel(got, ref):.3e} max_abs {float(err.max()):.3e} at flat {i} "
            f"got {float(got.flatten()[i]):.5f} ref {float(ref.flatten()[i]):.5f} "
            f"nan {int(torch.isnan(got).sum())}")


# ----------------------------------------------------------------------------------------------
# GEMM
# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (128, 128, 64), (256, 256, 128), (1000, 512, 256), (77, 320, 192),
                                   (4096, 2560, 2048), (1024, 256, 64), (1100, 768, 320), (2048, 512, 264), (8192, 2048, 2048)])
def test_gemm_kmajor(M, N, K):
    A, B = rnd((M, K), 1), rnd((N, K), 2)
    out = torch.full((M, N), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.gemm(A, B, out)
    torch.cuda.synchronize()
    ref = A.float() @ B.float().t()
    assert rel(out, ref) < 4e-3, report("gemm", out, ref)


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (256, 256, 128), (300, 512, 320), (2048, 2048, 2560)])
def test_gemm_b_mn_major(M, N, K):
    """dgrad form: D = A @ B with B stored [K, N] (the forward weight as stored)."""
    A, B = rnd((M, K), 3), rnd((K, N), 4)
    out = torch.full((M, N), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.gemm(A, B, out, b_layout=1)
    torch.cuda.synchronize()
    ref = A.float() @ B.float()
    assert rel(out, ref) < 4e-3, report("gemm b_mn", out, ref)


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (64, 256, 1024), (24, 2048, 4096), (64, 256, 571)])
def test_gemm_wgrad_form(M, N, K):
    """wgrad form: D[M,N] = A^T B with A stored [K, M], B stored [K, N]; fp32 output, accumulate."""
    A, B = rnd((K, M), 5), rnd((K, N), 6)
    out = torch.ones((M, N), device=DEV, dtype=torch.float32)
    ops.gemm(A, B, out, a_layout=1, b_layout=1, epilogue=ops.EPI_F32, alpha=0.5, accumulate=True)
    torch.cuda.synchronize()
    ref = 1.0 + 0.5 * (A.float().t() @ B.float())
    assert rel(out, ref) < 1e-4, report("gemm wgrad", out, ref)


@pytest.mark.parametrize("M,N,K,S", [(64, 2048, 8192, 4), (2560, 64, 8192, 7), (64, 256, 1000, 3), (304, 128, 4096, 16), (64, 2048, 6659, 8)])
def test_gemm_wgrad_split_k(M, N, K, S):
    """LoRA factor gradients: out[M, N] = A^T B with A [K, M], B [K, N] (K = tokens).  split_k leaves S fp32 slices whose
    sum is the product (ragged K ranges included); the slices are deterministic."""
    A, B = rnd((K, M), 331, bound=0.5), rnd((K, N), 332, bound=0.5)
    ws = torch.full((S, M, N), float("nan"), device=DEV, dtype=torch.float32)
    ops.gemm(A, B, ws, a_layout=1, b_layout=1, epilogue=ops.EPI_F32, alpha=0.5, split_k=S)
    ws2 = torch.empty_like(ws)
    ops.gemm(A, B, ws2, a_layout=1, b_layout=1, epilogue=ops.EPI_F32, alpha=0.5, split_k=S)
    torch.cuda.synchronize()
    ref = 0.5 * (A.float().t() @ B.float())
    assert rel(ws.sum(0), ref) < 2e-3, report("wgrad split-k", ws.sum(0), ref)
    assert torch.equal(ws, ws2)


def test_gemm_dual_source_lora():
    """pretrained + LoRA delta folded in as a K extension (ger/lora.py:388-402)."""
    M, N, K, K2 = 512, 640, 256, 64
    A, B, A2, B2 = rnd((M, K), 7), rnd((N, K), 8), rnd((M, K2), 9), rnd((N, K2), 10)
    out = torch.empty((M, N), device=DEV, dtype=torch.bfloat16)
    ops.gemm(A, B, out, A2=A2, B2=B2)
    torch.cuda.synchronize()
    ref = A.float() @ B.float().t() + A2.float() @ B2.float().t()
    assert rel(out, ref) < 4e-3, report("gemm dual", out, ref)


def test_gemm_residual_and_colscale():
    M, N, K = 384, 256, 128
    A, B = rnd((M, K), 11), rnd((N, K), 12)
    res = rnd((M, N), 13, dtype=torch.float32)
    out = torch.empty((M, N), device=DEV, dtype=torch.float32)
    ops.gemm(A, B, out, epilogue=ops.EPI_RESID_F32, aux=res)
    torch.cuda.synchronize()
    ref = res + (A.float() @ B.float().t()).bfloat16().float()
    assert rel(out, ref) < 1e-3, report("resid", out, ref)
    scale, bias = rnd((N,), 14, dtype=torch.float32) + 1.5, rnd((N,), 15, dtype=torch.float32)
    o2 = torch.empty((M, N), device=DEV, dtype=torch.bfloat16)
    ops.gemm(A, B, o2, col_scale=scale, col_bias=bias)
    torch.cuda.synchronize()
    ref2 = scale * ((A.float() @ B.float().t()) + bias)
    assert rel(o2, ref2) < 6e-3, report("adapter_v2 head", o2, ref2)


def test_gemm_qkv_rope_epilogue():
    """QKV linear + RoPE on the group-interleaved row (ger/model.py:213-246) against the oracle's split + apply_rope."""
    B_, T, G, qpk, C = 2, 96, 2, 2, 256
    cfg = O.OracleConfig(n_head=G * qpk, n_query_groups=G, n_embd=C, block_size=128)
    x, W = rnd((B_ * T, C), 16), rnd((cfg.qkv_out, C), 17, bound=1 / 16)
    cos, sin = O.build_rope_cache(128, 64)
    out = torch.empty((B_ * T, cfg.qkv_out), device=DEV, dtype=torch.bfloat16)
    ops.gemm(x, W, out, epilogue=ops.EPI_QKV_ROPE, rope_cos=cos.to(DEV), rope_sin=sin.to(DEV), seq_len=T, q_per_kv=qpk)
    torch.cuda.synchronize()
    qkv = (x.float() @ W.float().t()).cpu().view(B_, T, -1)
    q, k, v = O.split_qkv(qkv, cfg)
    q, k = O.apply_rope(q, cos[:T], sin[:T]), O.apply_rope(k, cos[:T], sin[:T])
    got = out.float().cpu().view(B_, T, G, qpk + 2, 64)
    gq = got[:, :, :, :qpk].permute(0, 2, 3, 1, 4).reshape(B_, G * qpk, T, 64)
    gk = got[:, :, :, qpk].permute(0, 2, 1, 3)
    gv = got[:, :, :, qpk + 1].permute(0, 2, 1, 3)
    assert rel(gq, q) < 6e-3, report("rope q", gq, q)
    assert rel(gk, k[:, ::qpk]) < 6e-3, report("rope k", gk, k[:, ::qpk])
    assert rel(gv, v[:, ::qpk]) < 6e-3, report("v", gv, v[:, ::qpk])


@pytest.mark.parametrize("M", [1, 5, 64, 100, 128, 129, 200, 256])
def test_gemm_skinny_decode_shapes(M):
    """M <= 256 (decode: one token per sequence; above 128 rows a stage carries two activation boxes and the weight tile feeds two
    TMEM accumulators) takes the cluster split-K weight-streaming kernel: every epilogue the decode
    step uses, at the TinyLlama shapes (split counts 1, 2, 8, 9, with and without the LoRA second source), is checked
    against fp32 math and must be bit-identical run to run (partials are summed in split order, not arrival order)."""
    C, I, G, qpk = 2048, 5632, 4, 8
    x = rnd((M, C), 301)
    # plain bf16 + column scale / bias (lm_head shape class, many tiles => no split) and a 2048-wide one (8 splits)
    for N in (4096, 2048):
        W = rnd((N, C), 302, bound=1 / 32)
        o1 = torch.full((M, N), float("nan"), device=DEV, dtype=torch.bfloat16)
        ops.gemm(x, W, o1)
        o2 = torch.full((M, N), float("nan"), device=DEV, dtype=torch.bfloat16)
        ops.gemm(x, W, o2)
        torch.cuda.synchronize()
        ref = x.float() @ W.float().t()
        assert rel(o1, ref) < 4e-3, report(f"splitk bf16 N={N}", o1, ref)
        assert torch.equal(o1, o2), "split-K result differs between two runs"
    # residual epilogue, K = 5632 (9 splits) + LoRA K-extension
    a = rnd((M, I), 303)
    Wp = rnd((C, I), 304, bound=1 / 64)
    a2, b2 = rnd((M, 64), 305, bound=0.1), rnd((C, 64), 306, bound=0.1)
    res = rnd((M, C), 307, dtype=torch.float32)
    out = torch.full((M, C), float("nan"), device=DEV, dtype=torch.float32)
    ops.gemm(a, Wp, out, A2=a2, B2=b2, epilogue=ops.EPI_RESID_F32, aux=res)
    torch.cuda.synchronize()
    ref = res + (a.float() @ Wp.float().t() + a2.float() @ b2.float().t()).bfloat16().float()
    assert rel(out, ref) < 1e-3, report("splitk resid", out, ref)
    # SwiGLU (2 splits at N = 11264)
    W1, W2 = rnd((I, C), 308, bound=1 / 32), rnd((I, C), 309, bound=1 / 32)
    act = torch.full((M, I), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.gemm(x, _interleave(W1, W2).contiguous(), None, epilogue=ops.EPI_SWIGLU, out2=act)
    torch.cuda.synchronize()
    ref = F.silu(x.float() @ W1.float().t()) * (x.float() @ W2.float().t())
    assert rel(act, ref) < 8e-3, report("splitk swiglu", act, ref)
    # QKV + RoPE with explicit positions (decode)
    cfg = O.OracleConfig(n_head=G * qpk, n_query_groups=G, n_embd=C, block_size=2048)
    Wq = rnd((cfg.qkv_out, C), 310, bound=1 / 32)
    cos, sin = O.build_rope_cache(2048, 64)
    pos = synthetic.hash_ints((M,), 5, 311, 0, 2048).to(torch.int32)
    qkv = torch.full((M, cfg.qkv_out), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.gemm(x, Wq, qkv, epilogue=ops.EPI_QKV_ROPE, rope_cos=cos.to(DEV), rope_sin=sin.to(DEV), positions=pos.to(DEV), seq_len=1, q_per_kv=qpk)
    torch.cuda.synchronize()
    raw = (x.float() @ Wq.float().t()).bfloat16().float().cpu().view(M, G, qpk + 2, 64)
    c, s_ = cos.float()[pos.long()][:, None, None, :], sin.float()[pos.long()][:, None, None, :]
    rot = torch.cat((-raw[..., 32:], raw[..., :32]), dim=-1)
    want = raw.clone()
    want[:, :, : qpk + 1] = (raw * c + rot * s_)[:, :, : qpk + 1]
    assert rel(qkv.cpu().view(M, G, qpk + 2, 64), want) < 6e-3, report("splitk rope", qkv.cpu().view(M, G, qpk + 2, 64), want)


@pytest.mark.parametrize("M,S", [(1, 9), (64, 9), (64, 3), (100, 16), (128, 8), (200, 9), (256, 8)])
def test_gemm_split_k_partials_and_fused_consumer(M, S):
    """dh_gemm split_k: slice s holds the product over K range s (their sum is the GEMM); dh_splitk_resid_rmsnorm sums
    the slices in order, adds the residual (bf16-rounded product, like DH_EPI_RESID_F32) and applies the next RMSNorm --
    bit-identical to the unsplit residual GEMM followed by dh_rmsnorm_fwd is not required (other summation order), so
    the check is against fp32 math; run-to-run it must be bit-identical."""
    C, I = 2048, 5632
    a, W = rnd((M, I), 321), rnd((C, I), 322, bound=1 / 64)
    a2, b2 = rnd((M, 64), 323, bound=0.1), rnd((C, 64), 324, bound=0.1)
    res, g = rnd((M, C), 325, dtype=torch.float32), rnd((C,), 326, dtype=torch.float32) + 1.5
    ws = torch.full((S, M, C), float("nan"), device=DEV, dtype=torch.float32)
    ops.gemm(a, W, ws, A2=a2, B2=b2, epilogue=ops.EPI_F32, split_k=S)
    x1 = torch.empty((M, C), device=DEV, dtype=torch.float32)
    n1 = torch.empty((M, C), device=DEV, dtype=torch.bfloat16)
    ops.splitk_resid_rmsnorm(ws, res, g, x1, n1, 1e-5)
    x2, n2 = torch.empty_like(x1), torch.empty_like(n1)
    ops.gemm(a, W, ws, A2=a2, B2=b2, epilogue=ops.EPI_F32, split_k=S)
    ops.splitk_resid_rmsnorm(ws, res, g, x2, n2, 1e-5)
    torch.cuda.synchronize()
    prod = a.float() @ W.float().t() + a2.float() @ b2.float().t()
    assert rel(ws.sum(0), prod) < 1e-3, report("split-k slices", ws.sum(0), prod)
    xr = res + prod.bfloat16().float()
    nr = xr * torch.rsqrt((xr * xr).mean(-1, keepdim=True) + 1e-5) * g
    assert rel(x1, xr) < 1e-3, report("split-k resid", x1, xr)
    assert rel(n1, nr) < 4e-3, report("split-k norm", n1, nr)
    assert torch.equal(x1, x2) and torch.equal(n1, n2)


@pytest.mark.parametrize("M,G,qpk,S", [(64, 4, 8, 7), (3, 2, 2, 2), (128, 1, 1, 4), (250, 4, 8, 7)])
def test_splitk_qkv_rope_cache_consumer(M, G, qpk, S):
    """Split-K QKV partial products -> dh_splitk_qkv_rope_cache must equal the fused DH_EPI_QKV_ROPE GEMM followed by
    dh_kv_cache_write up to the summation order (checked against fp32 math at bf16 tolerance; the cache rows must be the
    k / v slots of the produced qkv row bit for bit)."""
    C, N, max_len = 2048, G * (qpk + 2) * 64, 96
    x, W = rnd((M, C), 341), rnd((N, C), 342, bound=1 / 32)
    cos, sin = O.build_rope_cache(max_len, 64)
    cos, sin = cos.to(DEV), sin.to(DEV)
    pos = synthetic.hash_ints((M,), 5, 343, 0, max_len).to(torch.int32).to(DEV)
    seq = torch.arange(M, device=DEV, dtype=torch.int32)
    ws = torch.full((S, M, N), float("nan"), device=DEV, dtype=torch.float32)
    qkv = torch.full((M, N), float("nan"), device=DEV, dtype=torch.bfloat16)
    ck = torch.zeros((M, G, max_len, 64), device=DEV, dtype=torch.bfloat16)
    cv = torch.zeros_like(ck)
    ops.gemm(x, W, ws, epilogue=ops.EPI_F32, split_k=S)
    ops.splitk_qkv_rope_cache(ws, qkv, ck, cv, seq, pos, cos, sin, G, qpk)
    ref = torch.empty((M, N), device=DEV, dtype=torch.bfloat16)
    ops.gemm(x, W, ref, epilogue=ops.EPI_QKV_ROPE, rope_cos=cos, rope_sin=sin, positions=pos, seq_len=1, q_per_kv=qpk)
    torch.cuda.synchronize()
    assert rel(qkv, ref) < 6e-3, report("qkv consumer", qkv, ref)
    got = qkv.view(M, G, qpk + 2, 64)
    rows = torch.arange(M, device=DEV)
    for g in range(G):
        assert torch.equal(ck[rows, g, pos.long()], got[:, g, qpk]) and torch.equal(cv[rows, g, pos.long()], got[:, g, qpk + 1])
    assert int((ck != 0).any(-1).sum()) <= M * G and int((cv != 0).any(-1).sum()) <= M * G       # nothing else was written


def _interleave(w1, w2):
    I = w1.shape[0]
    return torch.stack((w1.view(I // 32, 32, -1), w2.view(I // 32, 32, -1)), dim=1).reshape(2 * I, -1)


def test_gemm_swiglu_fwd_bwd():
    M, C, I = 256, 128, 256
    x, W1, W2 = rnd((M, C), 18), rnd((I, C), 19, bound=0.2), rnd((I, C), 20, bound=0.2)
    W12 = _interleave(W1, W2).contiguous()
    h12 = torch.empty((M, 2 * I), device=DEV, dtype=torch.bfloat16)
    act = torch.empty((M, I), device=DEV, dtype=torch.bfloat16)
    ops.gemm(x, W12, h12, epilogue=ops.EPI_SWIGLU, out2=act)
    torch.cuda.synchronize()
    h1, h2 = x.float() @ W1.float().t(), x.float() @ W2.float().t()
    ref = F.silu(h1) * h2
    assert rel(act, ref) < 8e-3, report("swiglu act", act, ref)
    h12v = h12.float().view(M, I // 32, 2, 32)
    assert rel(h12v[:, :, 0].reshape(M, I), h1) < 4e-3 and rel(h12v[:, :, 1].reshape(M, I), h2) < 4e-3
    # backward epilogue: d(act) = dy @ Wp with Wp [Cout, I] stored as the forward weight (MN-major B)
    Cout = 128
    dy, Wp = rnd((M, Cout), 21), rnd((Cout, I), 22, bound=0.2)
    dh12 = torch.empty((M, 2 * I), device=DEV, dtype=torch.bfloat16)
    ops.gemm(dy, Wp, dh12, b_layout=1, epilogue=ops.EPI_SWIGLU_BWD, aux=h12)
    torch.cuda.synchronize()
    h1b, h2b = h12v[:, :, 0].reshape(M, I), h12v[:, :, 1].reshape(M, I)
    dact = (dy.float() @ Wp.float()).bfloat16().float()
    sig = torch.sigmoid(h1b)
    rdh1 = dact * h2b * (sig * (1 + h1b * (1 - sig)))
    rdh2 = dact * (h1b * sig)
    g = dh12.float().view(M, I // 32, 2, 32)
    assert rel(g[:, :, 0].reshape(M, I), rdh1) < 8e-3, report("dh1", g[:, :, 0].reshape(M, I), rdh1)
    assert rel(g[:, :, 1].reshape(M, I), rdh2) < 8e-3, report("dh2", g[:, :, 1].reshape(M, I), rdh2)


@pytest.mark.parametrize("M", [1024, 1100, 2304])
def test_gemm_pair_kernel_epilogues(M):
    """Every fused epilogue at shapes that dispatch to the CTA-pair kernel (M >= 1024, N % 256 == 0, K-major operands),
    including a ragged last row tile; same references as the small-shape tests above."""
    N, K = 512, 320
    A, B = rnd((M, K), 40), rnd((N, K), 41, bound=0.25)
    acc = A.float() @ B.float().t()
    # bf16 out with column bias / scale (AdapterV2Linear form)
    scale, bias = rnd((N,), 42, dtype=torch.float32) + 1.5, rnd((N,), 43, dtype=torch.float32)
    o = torch.full((M, N), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.gemm(A, B, o, col_scale=scale, col_bias=bias)
    assert rel(o, scale * (acc + bias)) < 6e-3, report("pair bf16+colscale", o, scale * (acc + bias))
    # fp32 out, accumulate
    o32 = torch.ones((M, N), device=DEV, dtype=torch.float32)
    ops.gemm(A, B, o32, epilogue=ops.EPI_F32, alpha=0.5, accumulate=True)
    assert rel(o32, 1.0 + 0.5 * acc) < 1e-4, report("pair f32 acc", o32, 1.0 + 0.5 * acc)
    # residual add into the fp32 stream / bf16 stream
    res = rnd((M, N), 44, dtype=torch.float32)
    ops.gemm(A, B, o32, epilogue=ops.EPI_RESID_F32, aux=res)
    assert rel(o32, res + acc.bfloat16().float()) < 1e-3, report("pair resid f32", o32, res + acc.bfloat16().float())
    resb = res.bfloat16()
    ops.gemm(A, B, o, epilogue=ops.EPI_RESID_BF16, aux=resb)
    assert rel(o, resb.float() + acc.bfloat16().float()) < 6e-3, report("pair resid bf16", o, resb.float() + acc)
    # dual source (LoRA K extension)
    A2, B2 = rnd((M, 64), 45), rnd((N, 64), 46)
    ops.gemm(A, B, o, A2=A2, B2=B2)
    assert rel(o, acc + A2.float() @ B2.float().t()) < 4e-3, report("pair dual", o, acc + A2.float() @ B2.float().t())
    torch.cuda.synchronize()


def test_gemm_pair_kernel_rope_and_swiglu():
    B_, T, G, qpk, C = 3, 400, 2, 2, 256          # M = 1200 (ragged), N = 512
    cfg = O.OracleConfig(n_head=G * qpk, n_query_groups=G, n_embd=C, block_size=512)
    M = B_ * T
    x, W = rnd((M, C), 47), rnd((cfg.qkv_out, C), 48, bound=1 / 16)
    assert cfg.qkv_out % 256 == 0
    cos, sin = O.build_rope_cache(512, 64)
    out = torch.full((M, cfg.qkv_out), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.gemm(x, W, out, epilogue=ops.EPI_QKV_ROPE, rope_cos=cos.to(DEV), rope_sin=sin.to(DEV), seq_len=T, q_per_kv=qpk)
    torch.cuda.synchronize()
    qkv = (x.float() @ W.float().t()).cpu().view(B_, T, -1)
    q, k, v = O.split_qkv(qkv, cfg)
    q, k = O.apply_rope(q, cos[:T], sin[:T]), O.apply_rope(k, cos[:T], sin[:T])
    got = out.float().cpu().view(B_, T, G, qpk + 2, 64)
    gq = got[:, :, :, :qpk].permute(0, 2, 3, 1, 4).reshape(B_, G * qpk, T, 64)
    assert rel(gq, q) < 6e-3, report("pair rope q", gq, q)
    assert rel(got[:, :, :, qpk].permute(0, 2, 1, 3), k[:, ::qpk]) < 6e-3
    assert rel(got[:, :, :, qpk + 1].permute(0, 2, 1, 3), v[:, ::qpk]) < 6e-3
    # explicit positions (decode form) give the same rotation
    out2 = torch.empty_like(out)
    pos = torch.arange(T, device=DEV, dtype=torch.int32).repeat(B_)
    ops.gemm(x, W, out2, epilogue=ops.EPI_QKV_ROPE, rope_cos=cos.to(DEV), rope_sin=sin.to(DEV), positions=pos, seq_len=1, q_per_kv=qpk)
    assert torch.equal(out, out2)
    # SwiGLU forward + backward epilogues
    I = 256
    W1, W2 = rnd((I, C), 49, bound=0.2), rnd((I, C), 50, bound=0.2)
    W12 = _interleave(W1, W2).contiguous()
    h12 = torch.empty((M, 2 * I), device=DEV, dtype=torch.bfloat16)
    act = torch.empty((M, I), device=DEV, dtype=torch.bfloat16)
    ops.gemm(x, W12, h12, epilogue=ops.EPI_SWIGLU, out2=act)
    h1, h2 = x.float() @ W1.float().t(), x.float() @ W2.float().t()
    assert rel(act, F.silu(h1) * h2) < 8e-3, report("pair swiglu act", act, F.silu(h1) * h2)
    act2 = torch.empty_like(act)
    ops.gemm(x, W12, None, epilogue=ops.EPI_SWIGLU, out2=act2)      # inference form: h12 not kept
    assert torch.equal(act, act2)
    h12v = h12.float().view(M, I // 32, 2, 32)
    h1b, h2b = h12v[:, :, 0].reshape(M, I), h12v[:, :, 1].reshape(M, I)
    assert rel(h1b, h1) < 4e-3 and rel(h2b, h2) < 4e-3
    Cout = 128
    dy, WpT = rnd((M, Cout), 51), rnd((I, Cout), 52, bound=0.2)      # transposed proj weight: K-major B of the dgrad
    dh12 = torch.empty((M, 2 * I), device=DEV, dtype=torch.bfloat16)
    ops.gemm(dy, WpT, dh12, epilogue=ops.EPI_SWIGLU_BWD, aux=h12)
    torch.cuda.synchronize()
    dact = (dy.float() @ WpT.float().t()).bfloat16().float()
    sig = torch.sigmoid(h1b)
    g = dh12.float().view(M, I // 32, 2, 32)
    assert rel(g[:, :, 0].reshape(M, I), dact * h2b * (sig * (1 + h1b * (1 - sig)))) < 8e-3
    assert rel(g[:, :, 1].reshape(M, I), dact * (h1b * sig)) < 8e-3


# ----------------------------------------------------------------------------------------------
# leaf ops
# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("M,C,dt", [(1037, 2048, torch.float32), (520, 256, torch.bfloat16), (4096, 1024, torch.float32)])
def test_rmsnorm_fwd_many_rows(M, C, dt):
    """M >= 512 takes the warp-per-row kernel (ragged last block of 8 rows); same tolerance as the block-per-row one, and
    the rstd it leaves feeds the backward."""
    x = rnd((M, C), 36, 2.0, dt)
    w = rnd((C,), 37, 0.5, torch.float32) + 1.0
    y = torch.full((M, C), float("nan"), device=DEV, dtype=torch.bfloat16)
    rstd = torch.full((M,), float("nan"), device=DEV)
    ops.rmsnorm_fwd(x, w, y, rstd, 1e-5)
    ref = O.rmsnorm(x.float(), w, 1e-5)
    assert rel(y, ref) < 4e-3, report("rmsnorm", y, ref)
    want = torch.rsqrt((x.float() ** 2).mean(-1) + 1e-5)
    assert rel(rstd, want) < 1e-5
    ys = torch.empty((300, C), device=DEV, dtype=torch.bfloat16)
    ops.rmsnorm_fwd(x[:300].contiguous(), w, ys, None, 1e-5)                   # block-per-row kernel on the same rows
    assert (ys.float() - y[:300].float()).abs().max() <= 2 ** -7 * ref[:300].abs().max()   # at most one bf16 ulp apart


def test_rmsnorm_fwd_bwd():
    M, C = 300, 2048
    x = rnd((M, C), 30, 2.0, torch.float32)
    w = rnd((C,), 31, 0.5, torch.float32) + 1.0
    y = torch.empty((M, C), device=DEV, dtype=torch.bfloat16)
    rstd = torch.empty(M, device=DEV)
    ops.rmsnorm_fwd(x, w, y, rstd, 1e-5)
    ref = O.rmsnorm(x, w, 1e-5)
    assert rel(y, ref) < 4e-3, report("rmsnorm", y, ref)
    dy = rnd((M, C), 32)
    dres = rnd((M, C), 33, dtype=torch.float32)
    xr = x.clone().requires_grad_(True)
    O.rmsnorm(xr, w, 1e-5).backward(dy.float())
    dx = torch.empty_like(x)
    dxb = torch.empty((M, C), device=DEV, dtype=torch.bfloat16)
    ops.rmsnorm_bwd(x, w, rstd, dy, dres, dx, dxb)
    torch.cuda.synchronize()
    assert rel(dx, dres + xr.grad) < 1e-5, report("rmsnorm bwd", dx, dres + xr.grad)
    assert rel(dxb, dres + xr.grad) < 4e-3


def test_embedding_and_rope_apply():
    V, C = 1000, 256
    wte = rnd((V, C), 34, dtype=torch.float32)
    idx = synthetic.hash_ints((4, 50), 3, 35, 0, V).to(DEV)
    out = torch.empty((200, C), device=DEV)
    ops.embedding_fwd(idx, wte, out)
    assert torch.equal(out, wte[idx.flatten()])
    # apply_rotary_emb_func semantics (ger/fused_rotary_embedding.py): (B,T,H,D), cos/sin (T, rot/2)
    B_, T, H, D, rot = 2, 40, 3, 64, 32
    x = rnd((B_, T, H, D), 36)
    ang = torch.outer(torch.arange(T, dtype=torch.float32), 1.0 / (10000 ** (torch.arange(0, rot, 2).float() / rot))).to(DEV)
    cos, sin = ang.cos(), ang.sin()
    o = torch.empty_like(x)
    ops.rope_apply(x, cos, sin, o)
    xf = x.float()
    x1, x2 = xf[..., : rot // 2], xf[..., rot // 2: rot]
    c, s = cos[None, :, None, :], sin[None, :, None, :]
    ref = torch.cat((x1 * c - x2 * s, x1 * s + x2 * c, xf[..., rot:]), dim=-1)
    assert rel(o, ref) < 4e-3, report("rope_apply", o, ref)
    back = torch.empty_like(x)
    ops.rope_apply(o, cos, sin, back, conj=True)          # backward rotation undoes the forward one
    assert rel(back[..., :rot], x[..., :rot]) < 1e-2


def test_cross_entropy_and_argmax():
    M, V = 257, 32000
    logits = rnd((M, V), 40, 6.0)
    tg = synthetic.hash_ints((M,), 5, 41, -1, V).to(DEV)
    tg[::7] = -1
    ref_rows = F.cross_entropy(logits.float(), tg, ignore_index=-1, reduction="none")
    lf = logits.float().requires_grad_(True)
    (F.cross_entropy(lf, tg, ignore_index=-1, reduction="none").sum() * 0.37).backward()
    ids = torch.empty(M, dtype=torch.int64, device=DEV)
    ops.argmax(logits, ids)
    assert torch.equal(ids, logits.float().argmax(dim=-1))
    row = torch.empty(M, device=DEV)
    ops.cross_entropy(logits, tg, row, 0.37, True)
    out2 = torch.empty(2, device=DEV)
    ops.loss_reduce(row, tg, out2)
    torch.cuda.synchronize()
    assert rel(row, ref_rows) < 1e-5, report("ce rows", row, ref_rows)
    assert abs(float(out2[0]) - float(ref_rows.sum())) < 1e-2 and int(out2[1]) == int((tg >= 0).sum())
    assert rel(logits, lf.grad) < 8e-3, report("ce grad", logits, lf.grad)
    assert float(logits[::7].abs().max()) == 0.0


def test_adamw_matches_torch():
    n = 10000
    p0, g = rnd((n,), 50, dtype=torch.float32), rnd((n,), 51, dtype=torch.float32)
    p = p0.clone()
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    tp = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([tp], lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.02)
    for step in range(1, 4):
        tp.grad = g * step
        opt.step()
        ops.adamw_step(p, g * step, m, v, 1e-3, 0.9, 0.999, 1e-8, 0.02, step)
    torch.cuda.synchronize()
    assert rel(p, tp.data) < 1e-6, report("adamw", p, tp.data)


# ----------------------------------------------------------------------------------------------
# attention
# ----------------------------------------------------------------------------------------------
def _attn_ref(qkv, B_, T, G, qpk, prefix_kv=None, gate=None):
    """fp32 reference through the oracle's split/sdpa (+ prefix) on an already-roped interleaved row."""
    cfg = O.OracleConfig(n_head=G * qpk, n_query_groups=G)
    v5 = qkv.view(B_, T, G, qpk + 2, 64).permute(0, 2, 3, 1, 4)
    q, k, v = v5.split((qpk, 1, 1), dim=2)
    k = k.expand(B_, G, qpk, T, 64).reshape(B_, -1, T, 64)
    v = v.expand(B_, G, qpk, T, 64).reshape(B_, -1, T, 64)
    q = q.reshape(B_, -1, T, 64)
    y = O.sdpa(q, k, v, None, 0.125)
    if prefix_kv is not None:
        aT = prefix_kv.shape[0]
        a5 = prefix_kv.view(1, aT, G, qpk + 2, 64).permute(0, 2, 3, 1, 4)
        _, ak, av = a5.split((qpk, 1, 1), dim=2)
        ak = ak.repeat_interleave(qpk, dim=2).reshape(1, -1, aT, 64)
        av = av.repeat_interleave(qpk, dim=2).reshape(1, -1, aT, 64)
        ay = O.sdpa(q, ak, av, torch.ones(T, aT, dtype=torch.bool), 0.125)
        y = y + gate.view(1, 1, -1, 1) * ay
    return y.reshape(B_, T, -1)


@pytest.mark.parametrize("B_,T,G,qpk,prefix", [(1, 128, 1, 1, 0), (2, 256, 2, 2, 0), (2, 160, 2, 2, 10), (1, 512, 4, 8, 10),
                                               (3, 77, 1, 4, 0)])
def test_attention_fwd(B_, T, G, qpk, prefix):
    N = G * (qpk + 2) * 64
    qkv = rnd((B_ * T, N), 60, 2.0)
    pk = rnd((prefix, N), 61, 2.0) if prefix else None
    gate = rnd((G * qpk,), 62, 1.0, torch.float32) if prefix else None
    out = torch.full((B_ * T, G * qpk * 64), float("nan"), device=DEV, dtype=torch.bfloat16)
    lse = torch.empty((B_, G * qpk, T), device=DEV)
    ops.attention_fwd(qkv, out, lse, B_, T, G, qpk, pk, gate, prefix)
    torch.cuda.synchronize()
    ref = _attn_ref(qkv.float().cpu(), B_, T, G, qpk, None if pk is None else pk.float().cpu(),
                    None if gate is None else gate.cpu())
    got = out.float().cpu().view(B_, T, -1)
    assert rel(got, ref) < 1e-2, report("attn fwd", got, ref)


@pytest.mark.parametrize("B_,T,G,qpk,prefix,rope", [(1, 128, 1, 1, 0, False), (2, 256, 2, 2, 0, False), (2, 160, 2, 2, 10, False),
                                                    (1, 384, 2, 4, 10, True)])
def test_attention_bwd(B_, T, G, qpk, prefix, rope):
    N = G * (qpk + 2) * 64
    H = G * qpk
    qkv = rnd((B_ * T, N), 63, 2.0)
    pk = rnd((prefix, N), 64, 2.0) if prefix else None
    gate = rnd((H,), 65, 1.0, torch.float32) if prefix else None
    dout = rnd((B_ * T, H * 64), 66, 1.0)
    out = torch.empty((B_ * T, H * 64), device=DEV, dtype=torch.bfloat16)
    lse = torch.empty((B_, H, T), device=DEV)
    ops.attention_fwd(qkv, out, lse, B_, T, G, qpk, pk, gate, prefix)
    dqkv = torch.full((B_ * T, N), float("nan"), device=DEV, dtype=torch.bfloat16)
    ws = torch.empty(ops.attention_bwd_workspace(B_, T, G, qpk), dtype=torch.uint8, device=DEV)
    dpk = torch.zeros((prefix, N), device=DEV) if prefix else None
    dg = torch.zeros(H, device=DEV) if prefix else None
    cos, sin = O.build_rope_cache(512, 64)
    ops.attention_bwd(qkv, out, dout, lse, dqkv, ws, B_, T, G, qpk, pk, gate, dpk, dg, prefix,
                      rope_cos=cos.to(DEV) if rope else None, rope_sin=sin.to(DEV) if rope else None)
    torch.cuda.synchronize()
    # reference by autograd.  With rope=True the kernel returns the gradient w.r.t. the PRE-rotation row, so the
    # reference treats `qkv` as the rotated value of a leaf `pre` (rotation is orthogonal: pre = R^-1 qkv).
    x = qkv.float().cpu().requires_grad_(True)
    pkr = pk.float().cpu().requires_grad_(True) if prefix else None
    gr = gate.cpu().clone().requires_grad_(True) if prefix else None
    y = _attn_ref(x, B_, T, G, qpk, pkr, gr)
    y.backward(dout.float().cpu().view(B_, T, -1))
    ref = x.grad.view(B_, T, G, qpk + 2, 64).clone()
    if rope:
        c, s = cos[:T].float()[None, :, None, None, :], sin[:T].float()[None, :, None, None, :]
        qk = ref[:, :, :, : qpk + 1]
        rot = torch.cat((qk[..., 32:], -qk[..., :32]), dim=-1)      # conjugate rotation
        ref[:, :, :, : qpk + 1] = qk * c + rot * s
    got = dqkv.float().cpu().view(B_, T, G, qpk + 2, 64)
    assert rel(got[:, :, :, qpk + 1], ref[:, :, :, qpk + 1]) < 1.5e-2, report("dv", got[:, :, :, qpk + 1], ref[:, :, :, qpk + 1])
    assert rel(got[:, :, :, qpk], ref[:, :, :, qpk]) < 1.5e-2, report("dk", got[:, :, :, qpk], ref[:, :, :, qpk])
    assert rel(got[:, :, :, :qpk], ref[:, :, :, :qpk]) < 1.5e-2, report("dq", got[:, :, :, :qpk], ref[:, :, :, :qpk])
    if prefix:
        assert rel(dg.cpu(), gr.grad) < 1e-2, report("dgate", dg.cpu(), gr.grad)
        assert rel(dpk.cpu(), pkr.grad) < 1e-2, report("dprefix", dpk.cpu(), pkr.grad)


@pytest.mark.parametrize("lens,G,qpk,prefix", [([200, 77, 128, 1, 333], 2, 2, 10), ([480, 330, 415], 4, 8, 10), ([129], 1, 1, 0)])
def test_attention_fwd_varlen_equals_per_sequence(lens, G, qpk, prefix):
    """Packed variable-length forward (cu_seqlens, no pad rows) == the fixed-length kernel run on each sequence alone,
    bit for bit (same tiles relative to the sequence start), for sequences that start at arbitrary row offsets."""
    N, H = G * (qpk + 2) * 64, G * qpk
    M = sum(lens)
    qkv = rnd((M, N), 351, 2.0)
    pk, gate = (rnd((prefix, N), 352, 2.0), rnd((H,), 353, 1.0, torch.float32)) if prefix else (None, None)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=DEV)
    out = torch.full((M, H * 64), float("nan"), device=DEV, dtype=torch.bfloat16)
    lse = torch.full((H, M), float("nan"), device=DEV)
    ops.attention_fwd_varlen(qkv, out, lse, cu, max(lens), G, qpk, pk, gate, prefix)
    torch.cuda.synchronize()
    o = 0
    for L in lens:
        one = torch.empty((L, H * 64), device=DEV, dtype=torch.bfloat16)
        lse1 = torch.empty((1, H, L), device=DEV)
        ops.attention_fwd(qkv[o:o + L].contiguous(), one, lse1, 1, L, G, qpk, pk, gate, prefix)
        torch.cuda.synchronize()
        assert torch.equal(out[o:o + L], one), report(f"varlen seq at {o}", out[o:o + L], one)
        assert torch.equal(lse[:, o:o + L], lse1[0])
        o += L
    assert not torch.isnan(out.float()).any()


@pytest.mark.parametrize("lens,G,qpk,prefix,rope", [([200, 77, 128, 1, 333], 2, 2, 10, True), ([480, 330, 415], 4, 8, 10, True),
                                                    ([129], 1, 1, 0, False), ([256, 100, 300], 2, 4, 0, True)])
def test_attention_bwd_varlen_equals_per_sequence(lens, G, qpk, prefix, rope):
    """Packed variable-length backward (cu_seqlens, no pad rows) == the fixed-length kernels run on each sequence alone:
    dqkv bit for bit (same tiles relative to the sequence start; rows beyond a sequence -- the next sequence's rows inside
    the last 128-row box -- must contribute nothing), prefix / gate gradients to fp32 accumulation-order tolerance."""
    N, H = G * (qpk + 2) * 64, G * qpk
    M = sum(lens)
    qkv = rnd((M, N), 451, 2.0)
    dout = rnd((M, H * 64), 454, 1.0)
    pk, gate = (rnd((prefix, N), 452, 2.0), rnd((H,), 453, 1.0, torch.float32)) if prefix else (None, None)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=DEV)
    cos, sin = (t.to(DEV) for t in O.build_rope_cache(512, 64))
    out = torch.empty((M, H * 64), device=DEV, dtype=torch.bfloat16)
    lse = torch.empty((H, M), device=DEV)
    ops.attention_fwd_varlen(qkv, out, lse, cu, max(lens), G, qpk, pk, gate, prefix)
    dqkv = torch.full((M, N), float("nan"), device=DEV, dtype=torch.bfloat16)
    ws = torch.empty(M * H * 4, dtype=torch.uint8, device=DEV)
    dpk = torch.zeros((prefix, N), device=DEV) if prefix else None
    dg = torch.zeros(H, device=DEV) if prefix else None
    ops.attention_bwd_varlen(qkv, out, dout, lse, dqkv, ws, cu, max(lens), G, qpk, pk, gate, dpk, dg, prefix,
                             rope_cos=cos if rope else None, rope_sin=sin if rope else None)
    torch.cuda.synchronize()
    assert not torch.isnan(dqkv.float()).any()
    dpk1 = torch.zeros_like(dpk) if prefix else None
    dg1 = torch.zeros_like(dg) if prefix else None
    o = 0
    for L in lens:
        q1, do1 = qkv[o:o + L].contiguous(), dout[o:o + L].contiguous()
        out1 = torch.empty((L, H * 64), device=DEV, dtype=torch.bfloat16)
        lse1 = torch.empty((1, H, L), device=DEV)
        ops.attention_fwd(q1, out1, lse1, 1, L, G, qpk, pk, gate, prefix)
        d1 = torch.full((L, N), float("nan"), device=DEV, dtype=torch.bfloat16)
        ws1 = torch.empty(ops.attention_bwd_workspace(1, L, G, qpk), dtype=torch.uint8, device=DEV)
        ops.attention_bwd(q1, out1, do1, lse1, d1, ws1, 1, L, G, qpk, pk, gate, dpk1, dg1, prefix,
                          rope_cos=cos if rope else None, rope_sin=sin if rope else None)
        torch.cuda.synchronize()
        assert torch.equal(dqkv[o:o + L], d1), report(f"varlen bwd seq at {o}", dqkv[o:o + L], d1)
        o += L
    if prefix:
        assert rel(dg, dg1) < 1e-5, report("dgate", dg, dg1)
        assert rel(dpk, dpk1) < 1e-5, report("dprefix", dpk, dpk1)


def test_attention_decode_matches_prefill():
    B_, T, G, qpk, prefix = 3, 200, 2, 4, 10
    N, H = G * (qpk + 2) * 64, G * qpk
    qkv = rnd((B_ * T, N), 67, 2.0)
    pk, gate = rnd((prefix, N), 68, 2.0), rnd((H,), 69, 1.0, torch.float32)
    out = torch.empty((B_ * T, H * 64), device=DEV, dtype=torch.bfloat16)
    lse = torch.empty((B_, H, T), device=DEV)
    ops.attention_fwd(qkv, out, lse, B_, T, G, qpk, pk, gate, prefix)
    max_len = 256
    ck = torch.zeros((B_, G, max_len, 64), device=DEV, dtype=torch.bfloat16)
    cv = torch.zeros_like(ck)
    seq = torch.arange(B_, device=DEV, dtype=torch.int32).repeat_interleave(T)
    pos = torch.arange(T, device=DEV, dtype=torch.int32).repeat(B_)
    ops.kv_cache_write(qkv, ck, cv, seq, pos, G, qpk)
    lens = torch.tensor([T, T - 37, 1], device=DEV, dtype=torch.int32)
    rows = torch.stack([qkv[b * T + int(lens[b]) - 1] for b in range(B_)]).contiguous()
    dec = torch.empty((B_, H * 64), device=DEV, dtype=torch.bfloat16)
    ops.attention_decode(rows, ck, cv, lens, dec, G, qpk, pk, gate, prefix)
    torch.cuda.synchronize()
    ref = torch.stack([out[b * T + int(lens[b]) - 1] for b in range(B_)])
    assert rel(dec, ref) < 1e-2, report("decode attn", dec, ref)


@pytest.mark.parametrize("stream_kv", [False, True])
@pytest.mark.parametrize("nseq,G,qpk,prefix,max_len", [(5, 4, 8, 10, 640), (3, 2, 2, 0, 640), (2, 1, 1, 16, 640), (4, 4, 8, 10, 2048),
                                                     (200, 4, 8, 10, 640), (301, 2, 4, 0, 320)])
def test_attention_decode_against_fp32(nseq, G, qpk, prefix, max_len, stream_kv):
    """The decode kernel against fp32 softmax attention over the valid cache slots (lengths from 1 to the whole row, not
    multiples of the 16-key tiles), with and without the gated prefix; rows of 640 slots take the 4-tiles-ahead V prefetch,
    rows of 2048 x 8 heads the 2-tiles-ahead build (the score buffer takes the shared memory).  `stream_kv`: the same cases
    through the streamed kernel (dh_attention_decode_stream: TMA ring of 32-key chunks; with 200 / 301 sequences every resident
    block walks two to three (sequence, kv group) items as one chunk stream)."""
    H = G * qpk
    N = G * (qpk + 2) * 64
    lens = synthetic.hash_ints((nseq,), 3, 401, 1, max_len + 1).to(torch.int32)
    lens[0], lens[-1] = 1, max_len
    ck, cv = rnd((nseq, G, max_len, 64), 402, 2.0), rnd((nseq, G, max_len, 64), 403, 2.0)
    row = rnd((nseq, N), 404, 2.0)
    pk = rnd((max(prefix, 1), N), 405, 2.0)
    gate = rnd((H,), 406, 1.0, torch.float32)
    out = torch.full((nseq, H * 64), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.attention_decode(row, ck, cv, lens.to(DEV), out, G, qpk, pk if prefix else None, gate if prefix else None, prefix, stream_kv=stream_kv)
    torch.cuda.synchronize()
    r = row.float().view(nseq, G, qpk + 2, 64)
    ref = torch.zeros((nseq, G, qpk, 64), device=DEV)
    for b in range(nseq):
        L = int(lens[b])
        q = r[b, :, :qpk]                                                  # [G, qpk, 64]
        sc = torch.einsum("gqd,gkd->gqk", q, ck[b, :, :L].float()) / 8.0
        ref[b] = torch.einsum("gqk,gkd->gqd", sc.softmax(-1), cv[b, :, :L].float())
        if prefix:
            pkv = pk.float().view(-1, G, qpk + 2, 64)[:prefix]
            sa = torch.einsum("gqd,agd->gqa", q, pkv[:, :, qpk]) / 8.0
            ref[b] += gate.view(G, qpk, 1) * torch.einsum("gqa,agd->gqd", sa.softmax(-1), pkv[:, :, qpk + 1])
    assert rel(out, ref.reshape(nseq, -1)) < 6e-3, report("decode attn fp32", out, ref.reshape(nseq, -1))


def test_relpred_fwd_bwd():
    B_, T, C, pool = 3, 103, 96, 10
    x = synthetic.synth_features(B_, T, C, 3, 70, DEV)
    p = "audio_noise_classifier."
    shapes = {"conv1.weight": (256, C, 3), "conv1.bias": (256,), "conv2.weight": (256, 256, 3), "conv2.bias": (256,),
              "classifier.weight": (3, 256), "classifier.bias": (3,)}
    names = list(shapes)
    sd = {p + n: synthetic.hash_uniform(shapes[n], 5, 80 + i, 0.08, DEV, bf16_exact=False) for i, n in enumerate(names)}
    n_out = -(-T // pool)
    h1, h2 = torch.empty((B_, T, 256), device=DEV), torch.empty((B_, T, 256), device=DEV)
    pooled, logits = torch.empty((B_, n_out, 256), device=DEV), torch.empty((B_, n_out, 3), device=DEV)
    args = [sd[p + n] for n in names]
    ops.relpred_fwd(x, *args, h1, h2, pooled, logits, pool)
    # reference on the CPU in true fp32 (cuDNN would use TF32 convolutions on the GPU)
    work = {k: v.cpu().clone().requires_grad_(True) for k, v in sd.items()}
    ref = O.noise_mask_classifier(x.cpu(), work, p, pool)
    assert rel(logits.cpu(), ref.detach()) < 1e-4, report("relpred fwd", logits.cpu(), ref.detach())
    dl = rnd((B_, n_out, 3), 71, dtype=torch.float32)
    ref.backward(dl.cpu())
    grads = [torch.zeros_like(a) for a in args]
    scratch = torch.empty(2 * B_ * T * 256, device=DEV)
    ops.relpred_bwd(x, args[0], args[2], args[4], h1, h2, pooled, dl, *grads, scratch, pool)
    torch.cuda.synchronize()
    for n, g in zip(names, grads):
        assert rel(g.cpu(), work[p + n].grad) < 1e-4, report("relpred " + n, g.cpu(), work[p + n].grad)


def test_prefix_project_fwd_bwd():
    """Batched adapter prefix K/V projection (ger/adapter.py:129-137) and its backward, all layers in one launch."""
    L, aT, C, G, qpk = 3, 10, 512, 2, 4
    N = G * (qpk + 2) * 64
    prefix = rnd((L, aT, C), 90, 1.0)
    w = rnd((L, N, C), 91, 0.05)
    rows = torch.cat([torch.arange((g * (qpk + 2) + qpk) * 64, (g * (qpk + 2) + qpk + 2) * 64) for g in range(G)]).to(DEV)
    wkv = w.index_select(1, rows).contiguous()
    out = torch.zeros((L, aT, N), device=DEV, dtype=torch.bfloat16)
    ops.prefix_project_fwd(prefix, wkv, out, G, qpk)
    ref = torch.einsum("lac,lnc->lan", prefix.float(), w.float())
    got = out.float().index_select(2, rows)
    assert rel(got, ref.index_select(2, rows)) < 1e-2, report("prefix fwd", got, ref.index_select(2, rows))
    qcols = torch.ones(N, dtype=torch.bool, device=DEV)
    qcols[rows] = False
    assert float(out[:, :, qcols].abs().max()) == 0.0          # q slots untouched
    d_pkv = torch.zeros((L, aT, N), device=DEV)
    d_pkv[:, :, rows] = rnd((L, aT, rows.numel()), 92, 1.0, torch.float32)
    d_prefix = torch.empty((L, aT, C), device=DEV)
    ops.prefix_project_bwd(d_pkv, wkv, d_prefix, G, qpk)
    refb = torch.einsum("lan,lnc->lac", d_pkv, w.float())
    assert rel(d_prefix, refb) < 1e-3, report("prefix bwd", d_prefix, refb)


def test_valid_row_compaction_and_bounded_gemm():
    """dh_compact_targets / dh_gather_rows / dh_scatter_rows and dh_gemm's device-side m_limit (fused loss path)."""
    M, C, N = 1500, 256, 512
    g = torch.Generator().manual_seed(5)
    tg = torch.full((M,), -1, dtype=torch.int64)
    valid = torch.randperm(M, generator=g)[:300].sort().values
    tg[valid] = torch.randint(0, N, (300,), generator=g)
    tg = tg.to(DEV)
    count = torch.zeros(1, dtype=torch.int32, device=DEV)
    rows = torch.empty(M, dtype=torch.int32, device=DEV)
    tg_c = torch.empty(M, dtype=torch.int64, device=DEV)
    ops.compact_targets(tg, count, rows, tg_c)
    assert int(count) == 300
    assert torch.equal(rows[:300].cpu().long(), valid) and bool((rows[300:] == -1).all())
    assert torch.equal(tg_c[:300], tg[valid.to(DEV)]) and bool((tg_c[300:] == -1).all())
    h = rnd((M, C), 95, 1.0)
    h_c = torch.zeros((M, C), device=DEV, dtype=torch.bfloat16)
    ops.gather_rows(h, rows, count, h_c)
    assert torch.equal(h_c[:300], h[valid.to(DEV)]) and float(h_c[300:].abs().max()) == 0.0
    w = rnd((N, C), 96, 0.1)
    out = torch.full((M, N), 7.0, device=DEV, dtype=torch.bfloat16)
    ops.gemm(h_c, w, out, m_limit=count)
    torch.cuda.synchronize()
    ref = h_c[:300].float() @ w.float().t()
    assert rel(out[:300], ref) < 1e-2, report("bounded gemm", out[:300], ref)
    assert bool((out[1280:] == 7.0).all())                 # tiles beyond the limit are never touched
    back = torch.full((M, C), 3.0, device=DEV, dtype=torch.bfloat16)
    ops.scatter_rows(h_c, rows, count, back)
    exp = torch.zeros_like(h)
    exp[valid.to(DEV)] = h[valid.to(DEV)]
    assert torch.equal(back, exp)
    # all targets ignored: count 0, nothing runs, nothing hangs
    tg0 = torch.full((M,), -1, dtype=torch.int64, device=DEV)
    ops.compact_targets(tg0, count, rows, tg_c)
    ops.gemm(h_c, w, out, m_limit=count)
    torch.cuda.synchronize()
    assert int(count) == 0


@pytest.mark.parametrize("p", [0.05, 0.5])
def test_dropout_bf16_statistics_and_determinism(p):
    """dh_dropout_bf16 (the LoRA branch's nn.Dropout, ger/lora.py:94-98,165): keep rate within 5 sigma of 1 - p, kept values =
    x / (1 - p) exactly (p quantised to 16 bits), the mean is preserved, the same (seed, offset) regenerates the mask bit for
    bit, another offset or seed gives an independent one, and `accumulate` adds the masked tensor onto the output."""
    n = 1 << 22
    x = torch.ones(n, device=DEV, dtype=torch.bfloat16)
    a, b, c, d = (torch.empty_like(x) for _ in range(4))
    ops.dropout_bf16(x, a, p, 1234, 7)
    ops.dropout_bf16(x, b, p, 1234, 7)
    ops.dropout_bf16(x, c, p, 1234, 8)
    ops.dropout_bf16(x, d, p, 99, 7)
    torch.cuda.synchronize()
    assert torch.equal(a, b)
    keep = float((a != 0).float().mean())
    sigma = (p * (1 - p) / n) ** 0.5
    assert abs(keep - (1 - p)) < 5 * sigma + 1e-4, (keep, 1 - p)
    q = round(p * 65536) / 65536
    kept = a[a != 0].float()
    assert float((kept - 1 / (1 - q)).abs().max()) < 2 ** -7 / (1 - q)          # one bf16 rounding of 1 / (1 - p)
    assert abs(float(a.float().mean()) - 1.0) < 5 * sigma / (1 - p) + 4e-3
    for other in (c, d):                                                        # independent masks: joint keep rate = (1 - p)^2
        both = float(((a != 0) & (other != 0)).float().mean())
        assert abs(both - (1 - p) ** 2) < 6 * sigma + 1e-3, both
    y = rnd((4096, 512), 901, 2.0)
    base = rnd((4096, 512), 902, 2.0)
    m = torch.empty_like(y)
    ops.dropout_bf16(y, m, p, 5, 3)
    acc = base.clone()
    ops.dropout_bf16(y, acc, p, 5, 3, accumulate=True)
    torch.cuda.synchronize()
    assert rel(acc, base.float() + m.float()) < 4e-3


@pytest.mark.parametrize("M,I,C", [(1, 5632, 2048), (100, 5632, 2048), (128, 5632, 2048), (200, 5632, 2048), (256, 5632, 2048), (37, 512, 2048),
                                   (256, 704, 2048), (5, 512, 256), (130, 128, 128)])
def test_gemm_streamk_swiglu_decode(M, I, C):
    """dh_gemm with a workspace (ABI v4): the decode fc_1|fc_2 GEMM as stream-K over all SMs, partial tiles in the workspace, the
    CTA that completes a tile sums its slices in CTA order and applies SwiGLU.  Against fp32 math at bf16 tolerance and against the
    cluster split-K kernel (same math, other summation order); three launches on the same workspace must agree bit for bit (the
    arrival counters reset themselves, the slice sum does not depend on who arrives last).  The last two shapes have fewer units than
    SMs (some CTAs idle)."""
    x = rnd((M, C), 701)
    W1, W2 = rnd((I, C), 702, bound=1 / 32), rnd((I, C), 703, bound=1 / 32)
    W12 = _interleave(W1, W2).contiguous()
    nb = ops.gemm_streamk_workspace_bytes(M, 2 * I, C)
    assert nb > 0
    ws = torch.zeros(nb, device=DEV, dtype=torch.uint8)
    outs = []
    for _ in range(3):
        act = torch.full((M, I), float("nan"), device=DEV, dtype=torch.bfloat16)
        ops.gemm(x, W12, None, epilogue=ops.EPI_SWIGLU, out2=act, workspace=ws)
        outs.append(act)
    plain = torch.full((M, I), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.gemm(x, W12, None, epilogue=ops.EPI_SWIGLU, out2=plain)
    torch.cuda.synchronize()
    ref = F.silu(x.float() @ W1.float().t()) * (x.float() @ W2.float().t())
    assert rel(outs[0], ref) < 8e-3, report("stream-K swiglu", outs[0], ref)
    assert rel(outs[0], plain) < 8e-3
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    n_tiles = 2 * I // 128
    assert int(ws[: 4 * n_tiles].view(torch.int32).abs().sum()) == 0          # every tile's counter is back at zero
