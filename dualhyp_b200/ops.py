"""Tensor-level wrappers over the C ABI (include/dualhyp.h).  PyTorch is used for device memory and the
current stream only; every function here enqueues hand-written sm_100a kernels and nothing else."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from . import _lib
from ._lib import (EPI_BF16, EPI_F32, EPI_QKV_ROPE, EPI_RESID_BF16, EPI_RESID_F32, EPI_SWIGLU, EPI_SWIGLU_BWD,
                   GemmArgs, check)

Tensor = torch.Tensor
_launches = 0           # kernels launched through this module (bench.py reports it as gpu_launches)


def launches() -> int:
    return _launches


# ---- optional per-launch timing (bench.py's roofline pass): CUDA events on the launching stream ----
_prof = None


class profile:
    """`with ops.profile() as recs:` records (name, meta, start_event, end_event) for every op launched inside."""

    def __enter__(self):
        global _prof
        _prof = []
        return _prof

    def __exit__(self, *exc):
        global _prof
        _prof = None
        return False


_region_depth = 0


class region:
    """`with ops.region("relpred_fwd"):` -- a composite op (several C-ABI launches plus tensor glue) shows up in `profile()` as ONE
    record timed around the whole block; the ops launched inside are not recorded a second time."""

    def __init__(self, name: str):
        self.name, self.e0 = name, None

    def __enter__(self):
        global _region_depth
        if _prof is not None and _region_depth == 0:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        _region_depth += 1
        return self

    def __exit__(self, *exc):
        global _region_depth
        _region_depth -= 1
        if self.e0 is not None and _prof is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            _prof.append((self.name, None, self.e0, e1))
        self.e0 = None
        return False


def _profiled(name, meta_fn=None):
    def deco(fn):
        def wrapper(*args, **kwargs):
            if _prof is None or _region_depth > 0:
                return fn(*args, **kwargs)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn(*args, **kwargs)
            e1.record()
            _prof.append((name, meta_fn(*args, **kwargs) if meta_fn else None, e0, e1))
            return out
        wrapper.__name__, wrapper.__doc__ = fn.__name__, fn.__doc__
        return wrapper
    return deco


B_FROZEN = 0x100


def _gemm_meta(A, B, out, *, epilogue=EPI_BF16, b_layout=0, a_layout=0, A2=None, **kw):
    M, K = (A.shape if a_layout == 0 else (A.shape[1], A.shape[0]))
    N = B.shape[0] if b_layout == 0 else B.shape[1]
    K2 = 0 if A2 is None else (A2.shape[1] if kw.get("a2_layout", 0) == 0 else A2.shape[0])
    if kw.get("m_limit") is not None:
        M = 0          # row count only known on the device: no flops are credited to bounded GEMMs in the profile
    return (int(M), int(N), int(K), int(K2), int(epilogue), int(a_layout), int(b_layout))


def _count(n: int = 1) -> None:
    global _launches
    _launches += n


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _dev(t: Tensor) -> None:
    if not t.is_cuda:
        raise _lib.DualHypError("dualhyp_b200 ops need CUDA tensors (there is no CPU fallback)")
    _lib.init(t.device.index or 0)


def _rowmajor(t: Tensor, what: str) -> int:
    """Leading dimension of a 2-D row-major view (last stride 1)."""
    if t.dim() != 2 or t.stride(1) != 1:
        raise _lib.DualHypError(f"{what}: expected a 2-D row-major tensor, got shape {tuple(t.shape)} strides {t.stride()}")
    return t.stride(0)


@_profiled("gemm", _gemm_meta)
def gemm(A: Tensor, B: Tensor, out: Optional[Tensor], *, epilogue: int = EPI_BF16, b_layout: int = 0, a_layout: int = 0,
         A2: Optional[Tensor] = None, B2: Optional[Tensor] = None, a2_layout: int = 0, b2_layout: int = 0,
         out2: Optional[Tensor] = None, aux: Optional[Tensor] = None, alpha: float = 1.0, accumulate: bool = False,
         col_scale: Optional[Tensor] = None, col_bias: Optional[Tensor] = None, rope_cos: Optional[Tensor] = None,
         rope_sin: Optional[Tensor] = None, positions: Optional[Tensor] = None, seq_len: int = 0, q_per_kv: int = 0,
         m_limit: Optional[Tensor] = None, split_k: int = 0, b_frozen: bool = False, workspace: Optional[Tensor] = None) -> None:
    """D = A op(B) (+ A2 op(B2)) with a fused epilogue; see `dh_gemm` in include/dualhyp.h.  split_k > 1 (decode, M <= 128,
    EPI_F32): `out` is fp32 [split_k, M, N] of partial products for `splitk_resid_rmsnorm` to sum."""
    _dev(A)
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16
    a = GemmArgs()
    a.A, a.lda = A.data_ptr(), _rowmajor(A, "A")
    a.B, a.ldb = B.data_ptr(), _rowmajor(B, "B")
    if a_layout == 0:
        M, K = A.shape
    else:
        K, M = A.shape
    if b_layout == 0:
        N, Kb = B.shape
    else:
        Kb, N = B.shape
    assert K == Kb, f"K mismatch: A gives {K}, B gives {Kb}"
    a.M, a.N, a.K1, a.K2 = M, N, K, 0
    a.a_layout, a.b_layout = a_layout, b_layout | (B_FROZEN if b_frozen else 0)      # DH_B_FROZEN: see include/dualhyp.h
    if A2 is not None:
        assert B2 is not None and A2.dtype == torch.bfloat16 and B2.dtype == torch.bfloat16
        a.A2, a.lda2 = A2.data_ptr(), _rowmajor(A2, "A2")
        a.B2, a.ldb2 = B2.data_ptr(), _rowmajor(B2, "B2")
        a.K2 = A2.shape[1] if a2_layout == 0 else A2.shape[0]
        a.a2_layout, a.b2_layout = a2_layout, b2_layout
    a.epilogue, a.accumulate, a.alpha = epilogue, int(accumulate), alpha
    a.seq_len, a.q_per_kv = seq_len, q_per_kv
    if split_k > 1:
        assert out is not None and out.dim() == 3 and out.shape[0] == split_k and out.is_contiguous() and out.dtype == torch.float32
        a.split_k, a.out, a.ldo = split_k, out.data_ptr(), out.shape[2]
    elif out is not None:
        a.out, a.ldo = out.data_ptr(), _rowmajor(out, "out")
    if out2 is not None:
        a.out2, a.ldo2 = out2.data_ptr(), _rowmajor(out2, "out2")
    if aux is not None:
        a.aux, a.ldaux = aux.data_ptr(), _rowmajor(aux, "aux")
    a.col_scale, a.col_bias = _ptr(col_scale), _ptr(col_bias)
    a.rope_cos, a.rope_sin, a.positions = _ptr(rope_cos), _ptr(rope_sin), _ptr(positions)
    a.m_limit = _ptr(m_limit)
    if workspace is not None:          # stream-K decode GEMM (see dh_gemm_args.workspace): zeroed once by the owner, one per GEMM shape
        a.workspace, a.workspace_bytes = workspace.data_ptr(), workspace.numel() * workspace.element_size()
    check(_lib.load().dh_gemm(C.byref(a), _stream()), "dh_gemm")
    _count()


@_profiled("rmsnorm_fwd")
def rmsnorm_fwd(x: Tensor, weight: Tensor, y: Tensor, rstd: Optional[Tensor], eps: float) -> None:
    _dev(x)
    M, Cdim = x.shape
    assert x.is_contiguous() and y.is_contiguous() and weight.dtype == torch.float32 and y.dtype == torch.bfloat16
    xd = 0 if x.dtype == torch.float32 else 1
    check(_lib.load().dh_rmsnorm_fwd(x.data_ptr(), xd, weight.data_ptr(), y.data_ptr(), _ptr(rstd), M, Cdim, eps, _stream()),
          "dh_rmsnorm_fwd")
    _count()


@_profiled("splitk_resid_rmsnorm")
def splitk_resid_rmsnorm(partial: Tensor, resid: Tensor, weight: Tensor, x_out: Tensor, y: Tensor, eps: float) -> None:
    """x_out = resid + bf16(sum of the split-K slices); y = rmsnorm(x_out) * weight (bf16)."""
    _dev(partial)
    S, M, Cdim = partial.shape
    assert partial.is_contiguous() and partial.dtype == torch.float32 and resid.dtype == torch.float32 and x_out.dtype == torch.float32
    assert resid.is_contiguous() and x_out.is_contiguous() and y.is_contiguous() and y.dtype == torch.bfloat16
    assert resid.shape == (M, Cdim) and x_out.shape == (M, Cdim) and y.shape == (M, Cdim) and weight.dtype == torch.float32
    check(_lib.load().dh_splitk_resid_rmsnorm(partial.data_ptr(), S, M * Cdim, resid.data_ptr(), weight.data_ptr(), x_out.data_ptr(),
                                              y.data_ptr(), M, Cdim, eps, _stream()), "dh_splitk_resid_rmsnorm")
    _count()


@_profiled("splitk_qkv_rope_cache")
def splitk_qkv_rope_cache(partial: Tensor, qkv: Tensor, cache_k: Tensor, cache_v: Tensor, seq: Tensor, pos: Tensor, cos: Tensor,
                          sin: Tensor, n_groups: int, q_per_kv: int) -> None:
    """qkv = RoPE(bf16(sum of the split-K slices)) at pos; k / v slots also into the KV cache at (seq, pos)."""
    _dev(partial)
    S, M, N = partial.shape
    assert partial.is_contiguous() and partial.dtype == torch.float32 and qkv.is_contiguous() and qkv.shape == (M, N)
    assert N == n_groups * (q_per_kv + 2) * 64 and seq.dtype == torch.int32 and pos.dtype == torch.int32 and cos.dtype == torch.bfloat16
    assert cache_k.is_contiguous() and cache_v.is_contiguous()
    check(_lib.load().dh_splitk_qkv_rope_cache(partial.data_ptr(), S, M * N, qkv.data_ptr(), cache_k.data_ptr(), cache_v.data_ptr(),
                                               seq.data_ptr(), pos.data_ptr(), cos.data_ptr(), sin.data_ptr(), M, cache_k.shape[2],
                                               n_groups, q_per_kv, _stream()), "dh_splitk_qkv_rope_cache")
    _count()


def sm_count() -> int:
    return int(_lib.load().dh_sm_count()) or 148


def decode_split_k(N: int, K: int) -> int:
    """Split count of a decode GEMM whose consumer sums the slices: every SM gets one CTA (N / 128 tiles x splits <= SM count),
    at least two 64-wide k-blocks per CTA."""
    sms = int(_lib.load().dh_sm_count()) or 148
    tiles = max(1, N // 128)
    cap = int(os.environ.get("DH_DECODE_SPLIT_MAX", "8"))          # developer knob for A/B runs
    return max(1, min(sms // tiles, (K // 64) // 2, cap))


@_profiled("rmsnorm_bwd")
def rmsnorm_bwd(x: Tensor, weight: Tensor, rstd: Tensor, dy: Tensor, dres: Optional[Tensor], dx_out: Tensor,
                dx_bf16: Optional[Tensor]) -> None:
    _dev(x)
    M, Cdim = x.shape
    assert x.is_contiguous() and dy.is_contiguous() and dx_out.is_contiguous() and dy.dtype == torch.bfloat16
    xd = 0 if x.dtype == torch.float32 else 1
    check(_lib.load().dh_rmsnorm_bwd(x.data_ptr(), xd, weight.data_ptr(), rstd.data_ptr(), dy.data_ptr(), _ptr(dres),
                                     dx_out.data_ptr(), _ptr(dx_bf16), M, Cdim, _stream()), "dh_rmsnorm_bwd")
    _count()


@_profiled("embedding_fwd")
def embedding_fwd(idx: Tensor, wte: Tensor, out: Tensor) -> None:
    _dev(wte)
    assert idx.dtype == torch.int64 and idx.is_contiguous() and wte.is_contiguous() and out.is_contiguous()
    M = idx.numel()
    dt = {torch.float32: 0, torch.bfloat16: 1}
    check(_lib.load().dh_embedding_fwd(idx.data_ptr(), wte.data_ptr(), dt[wte.dtype], out.data_ptr(), dt[out.dtype], M,
                                       wte.shape[1], wte.shape[0], _stream()), "dh_embedding_fwd")
    _count()


def rope_apply(x: Tensor, cos: Tensor, sin: Tensor, out: Tensor, interleaved: bool = False, conj: bool = False) -> None:
    """x (B,T,H,D) bf16, cos/sin (T, rot/2) f32|bf16 -- ger/fused_rotary_embedding.py:10-90."""
    _dev(x)
    B, T, H, D = x.shape
    assert x.is_contiguous() and out.is_contiguous() and x.dtype == torch.bfloat16 and cos.is_contiguous() and sin.is_contiguous()
    assert cos.shape[0] >= T and cos.shape == sin.shape and cos.dtype == sin.dtype
    csd = 0 if cos.dtype == torch.float32 else 1
    check(_lib.load().dh_rope_apply(x.data_ptr(), out.data_ptr(), cos.data_ptr(), sin.data_ptr(), csd, B, T, H, D,
                                    2 * cos.shape[1], int(interleaved), int(conj), _stream()), "dh_rope_apply")
    _count(2 if (out.data_ptr() != x.data_ptr() and 2 * cos.shape[1] < D) else 1)


def rope_qkv_inplace(qkv: Tensor, cos: Tensor, sin: Tensor, seq_len: int, n_groups: int, q_per_kv: int, conj: bool,
                     positions: Optional[Tensor] = None) -> None:
    _dev(qkv)
    assert qkv.is_contiguous() and qkv.dtype == torch.bfloat16 and cos.dtype == torch.bfloat16
    check(_lib.load().dh_rope_qkv_inplace(qkv.data_ptr(), cos.data_ptr(), sin.data_ptr(), _ptr(positions), qkv.shape[0], seq_len,
                                          n_groups, q_per_kv, int(conj), _stream()), "dh_rope_qkv_inplace")
    _count()


@_profiled("attention_fwd")
def attention_fwd(qkv: Tensor, out: Tensor, lse: Tensor, B: int, T: int, n_groups: int, q_per_kv: int,
                  prefix_kv: Optional[Tensor] = None, gate: Optional[Tensor] = None, prefix_len: int = 0) -> None:
    _dev(qkv)
    assert qkv.is_contiguous() and out.is_contiguous() and lse.is_contiguous() and qkv.dtype == torch.bfloat16
    assert qkv.shape == (B * T, n_groups * (q_per_kv + 2) * 64) and out.shape == (B * T, n_groups * q_per_kv * 64)
    if prefix_kv is not None:
        assert prefix_kv.is_contiguous() and prefix_kv.shape[1] == qkv.shape[1] and gate.dtype == torch.float32
    check(_lib.load().dh_attention_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), _ptr(prefix_kv), _ptr(gate), B, T, n_groups,
                                       q_per_kv, prefix_len, _stream()), "dh_attention_fwd")
    _count()


@_profiled("attention_fwd")
def attention_fwd_varlen(qkv: Tensor, out: Tensor, lse: Tensor, cu_seqlens: Tensor, max_len: int, n_groups: int, q_per_kv: int,
                         prefix_kv: Optional[Tensor] = None, gate: Optional[Tensor] = None, prefix_len: int = 0) -> None:
    """Causal attention over packed sequences: rows [cu_seqlens[b], cu_seqlens[b + 1]) of qkv / out are sequence b."""
    _dev(qkv)
    M = qkv.shape[0]
    assert qkv.is_contiguous() and out.is_contiguous() and lse.is_contiguous() and qkv.dtype == torch.bfloat16
    assert cu_seqlens.dtype == torch.int32 and cu_seqlens.is_cuda and lse.numel() >= n_groups * q_per_kv * M
    assert qkv.shape[1] == n_groups * (q_per_kv + 2) * 64 and out.shape == (M, n_groups * q_per_kv * 64)
    check(_lib.load().dh_attention_fwd_varlen(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), _ptr(prefix_kv), _ptr(gate),
                                              cu_seqlens.data_ptr(), cu_seqlens.numel() - 1, max_len, M, n_groups, q_per_kv, prefix_len,
                                              _stream()), "dh_attention_fwd_varlen")
    _count()


def attention_bwd_workspace(B: int, T: int, n_groups: int, q_per_kv: int) -> int:
    return int(_lib.load().dh_attention_bwd_workspace(B, T, n_groups, q_per_kv))


@_profiled("attention_bwd")
def attention_bwd(qkv: Tensor, out: Tensor, dout: Tensor, lse: Tensor, dqkv: Optional[Tensor], workspace: Tensor, B: int, T: int,
                  n_groups: int, q_per_kv: int, prefix_kv: Optional[Tensor] = None, gate: Optional[Tensor] = None,
                  d_prefix_kv: Optional[Tensor] = None, d_gate: Optional[Tensor] = None, prefix_len: int = 0,
                  need_dqkv: bool = True, rope_cos: Optional[Tensor] = None, rope_sin: Optional[Tensor] = None) -> None:
    _dev(qkv)
    assert qkv.is_contiguous() and out.is_contiguous() and dout.is_contiguous() and dout.dtype == torch.bfloat16
    if d_prefix_kv is not None:
        assert d_prefix_kv.dtype == torch.float32 and d_prefix_kv.is_contiguous() and d_gate.dtype == torch.float32
    check(_lib.load().dh_attention_bwd(qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(), _ptr(prefix_kv), _ptr(gate),
                                       _ptr(dqkv), _ptr(d_prefix_kv), _ptr(d_gate), int(need_dqkv), _ptr(rope_cos), _ptr(rope_sin),
                                       B, T, n_groups, q_per_kv, prefix_len, workspace.data_ptr(),
                                       workspace.numel() * workspace.element_size(), _stream()), "dh_attention_bwd")
    _count(3 if need_dqkv else 1)


@_profiled("attention_bwd")
def attention_bwd_varlen(qkv: Tensor, out: Tensor, dout: Tensor, lse: Tensor, dqkv: Optional[Tensor], workspace: Tensor,
                         cu_seqlens: Tensor, max_len: int, n_groups: int, q_per_kv: int, prefix_kv: Optional[Tensor] = None,
                         gate: Optional[Tensor] = None, d_prefix_kv: Optional[Tensor] = None, d_gate: Optional[Tensor] = None,
                         prefix_len: int = 0, need_dqkv: bool = True, rope_cos: Optional[Tensor] = None,
                         rope_sin: Optional[Tensor] = None) -> None:
    """Backward of `attention_fwd_varlen`: rows [cu_seqlens[b], cu_seqlens[b + 1]) of qkv / out / dout / dqkv are sequence b."""
    _dev(qkv)
    M = qkv.shape[0]
    assert qkv.is_contiguous() and out.is_contiguous() and dout.is_contiguous() and dout.dtype == torch.bfloat16
    assert cu_seqlens.dtype == torch.int32 and cu_seqlens.is_cuda and lse.numel() >= n_groups * q_per_kv * M
    if d_prefix_kv is not None:
        assert d_prefix_kv.dtype == torch.float32 and d_prefix_kv.is_contiguous() and d_gate.dtype == torch.float32
    check(_lib.load().dh_attention_bwd_varlen(qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(), _ptr(prefix_kv), _ptr(gate),
                                              _ptr(dqkv), _ptr(d_prefix_kv), _ptr(d_gate), int(need_dqkv), _ptr(rope_cos), _ptr(rope_sin),
                                              cu_seqlens.data_ptr(), cu_seqlens.numel() - 1, max_len, M, n_groups, q_per_kv, prefix_len,
                                              workspace.data_ptr(), workspace.numel() * workspace.element_size(), _stream()),
          "dh_attention_bwd_varlen")
    _count(3 if need_dqkv else 1)


@_profiled("attention_decode")
def attention_decode(qkv: Tensor, cache_k: Tensor, cache_v: Tensor, lens: Tensor, out: Tensor, n_groups: int, q_per_kv: int,
                     prefix_kv: Optional[Tensor] = None, gate: Optional[Tensor] = None, prefix_len: int = 0, stream_kv: Optional[bool] = None) -> None:
    """`stream_kv=True` calls the streamed kernel (csrc/decode_attn.cu) directly, False the whole-row mma.sync kernel; None (default):
    `dh_attention_decode` picks (the streamed one for row lengths that are a multiple of 32 keys)."""
    _dev(qkv)
    nseq, max_len = cache_k.shape[0], cache_k.shape[2]
    assert cache_k.is_contiguous() and cache_v.is_contiguous() and lens.dtype == torch.int32 and qkv.is_contiguous()
    lib = _lib.load()
    fn, name = ((lib.dh_attention_decode, "dh_attention_decode") if stream_kv is None else
                (lib.dh_attention_decode_stream, "dh_attention_decode_stream") if stream_kv else (lib.dh_attention_decode_rows, "dh_attention_decode_rows"))
    check(fn(qkv.data_ptr(), cache_k.data_ptr(), cache_v.data_ptr(), lens.data_ptr(), out.data_ptr(),
             _ptr(prefix_kv), _ptr(gate), nseq, max_len, n_groups, q_per_kv, prefix_len, _stream()), name)
    _count()


@_profiled("kv_cache_write")
def kv_cache_write(qkv: Tensor, cache_k: Tensor, cache_v: Tensor, seq: Tensor, pos: Tensor, n_groups: int, q_per_kv: int) -> None:
    _dev(qkv)
    assert seq.dtype == torch.int32 and pos.dtype == torch.int32 and qkv.is_contiguous()
    check(_lib.load().dh_kv_cache_write(qkv.data_ptr(), cache_k.data_ptr(), cache_v.data_ptr(), seq.data_ptr(), pos.data_ptr(),
                                        qkv.shape[0], cache_k.shape[2], n_groups, q_per_kv, _stream()), "dh_kv_cache_write")
    _count()


@_profiled("cross_entropy")
def cross_entropy(logits: Tensor, targets: Tensor, row_loss: Tensor, grad_scale: float, write_grad: bool) -> None:
    _dev(logits)
    assert logits.dtype == torch.bfloat16 and targets.dtype == torch.int64 and targets.is_contiguous()
    M, V = logits.shape
    check(_lib.load().dh_cross_entropy(logits.data_ptr(), _rowmajor(logits, "logits"), targets.data_ptr(), row_loss.data_ptr(), M, V,
                                       grad_scale, int(write_grad), _stream()), "dh_cross_entropy")
    _count()


@_profiled("loss_reduce")
def loss_reduce(row_loss: Tensor, targets: Tensor, out2: Tensor) -> None:
    _dev(row_loss)
    check(_lib.load().dh_loss_reduce(row_loss.data_ptr(), targets.data_ptr(), out2.data_ptr(), row_loss.numel(), _stream()),
          "dh_loss_reduce")
    _count()


@_profiled("argmax")
def argmax(logits: Tensor, ids: Tensor) -> None:
    _dev(logits)
    M, V = logits.shape
    assert logits.dtype == torch.bfloat16 and ids.dtype == torch.int64
    check(_lib.load().dh_argmax(logits.data_ptr(), _rowmajor(logits, "logits"), ids.data_ptr(), M, V, _stream()), "dh_argmax")
    _count()


@_profiled("cast_f32_to_bf16")
def cast_f32_to_bf16(src: Tensor, dst: Tensor) -> None:
    _dev(src)
    assert src.dtype == torch.float32 and dst.dtype == torch.bfloat16 and src.shape == dst.shape
    check(_lib.load().dh_cast_f32_to_bf16(src.data_ptr(), _rowmajor(src, "src"), dst.data_ptr(), _rowmajor(dst, "dst"), src.shape[0],
                                          src.shape[1], _stream()), "dh_cast_f32_to_bf16")
    _count()


@_profiled("adamw_step")
def adamw_step(param: Tensor, grad: Tensor, exp_avg: Tensor, exp_avg_sq: Tensor, lr: float, beta1: float, beta2: float, eps: float,
               weight_decay: float, step: int, grad_scale: float = 1.0) -> None:
    _dev(param)
    for t in (param, grad, exp_avg, exp_avg_sq):
        assert t.dtype == torch.float32 and t.is_contiguous()
    check(_lib.load().dh_adamw_step(param.data_ptr(), grad.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(), param.numel(), lr,
                                    beta1, beta2, eps, weight_decay, step, grad_scale, _stream()), "dh_adamw_step")
    _count()


@_profiled("relpred_fwd")
def relpred_fwd(x: Tensor, w1, b1, w2, b2, wc, bc, h1: Tensor, h2: Tensor, pooled: Tensor, logits: Tensor, pool: int) -> None:
    _dev(x)
    B, T, Cd = x.shape
    for t in (x, w1, b1, w2, b2, wc, bc, h1, h2, pooled, logits):
        assert t.dtype == torch.float32 and t.is_contiguous()
    check(_lib.load().dh_relpred_fwd(x.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(), wc.data_ptr(),
                                     bc.data_ptr(), h1.data_ptr(), h2.data_ptr(), pooled.data_ptr(), logits.data_ptr(), B, T, Cd,
                                     pool, _stream()), "dh_relpred_fwd")
    _count(4)


@_profiled("relpred_bwd")
def relpred_bwd(x, w1, w2, wc, h1, h2, pooled, d_logits, dw1, db1, dw2, db2, dwc, dbc, scratch, pool: int) -> None:
    _dev(x)
    B, T, Cd = x.shape
    for t in (x, w1, w2, wc, h1, h2, pooled, d_logits, dw1, db1, dw2, db2, dwc, dbc, scratch):
        assert t.dtype == torch.float32 and t.is_contiguous()
    assert scratch.numel() >= 2 * B * T * 256
    check(_lib.load().dh_relpred_bwd(x.data_ptr(), w1.data_ptr(), w2.data_ptr(), wc.data_ptr(), h1.data_ptr(), h2.data_ptr(),
                                     pooled.data_ptr(), d_logits.data_ptr(), dw1.data_ptr(), db1.data_ptr(), dw2.data_ptr(),
                                     db2.data_ptr(), dwc.data_ptr(), dbc.data_ptr(), scratch.data_ptr(), B, T, Cd, pool, _stream()),
          "dh_relpred_bwd")
    _count(7)


@_profiled("prefix_project_fwd")
def prefix_project_fwd(prefix: Tensor, wkv: Tensor, prefix_kv: Tensor, n_groups: int, q_per_kv: int) -> None:
    """prefix bf16 [L, aT, C], wkv bf16 [L, n_groups*128, C] -> k / v slots of prefix_kv bf16 [L, aT, n_groups*(qpk+2)*64]."""
    _dev(prefix)
    L, aT, C = prefix.shape
    assert prefix.is_contiguous() and wkv.is_contiguous() and prefix_kv.is_contiguous()
    assert wkv.shape == (L, n_groups * 128, C) and prefix_kv.shape == (L, aT, n_groups * (q_per_kv + 2) * 64)
    assert prefix.dtype == wkv.dtype == prefix_kv.dtype == torch.bfloat16
    check(_lib.load().dh_prefix_project_fwd(prefix.data_ptr(), wkv.data_ptr(), prefix_kv.data_ptr(), L, aT, C, n_groups, q_per_kv,
                                            _stream()), "dh_prefix_project_fwd")
    _count()


@_profiled("prefix_project_bwd")
def prefix_project_bwd(d_prefix_kv: Tensor, wkv: Tensor, d_prefix: Tensor, n_groups: int, q_per_kv: int) -> None:
    _dev(d_prefix_kv)
    L, aT, C = d_prefix.shape
    assert d_prefix_kv.is_contiguous() and wkv.is_contiguous() and d_prefix.is_contiguous()
    assert d_prefix_kv.dtype == torch.float32 and d_prefix.dtype == torch.float32 and wkv.dtype == torch.bfloat16
    assert d_prefix_kv.shape == (L, aT, n_groups * (q_per_kv + 2) * 64) and wkv.shape == (L, n_groups * 128, C)
    check(_lib.load().dh_prefix_project_bwd(d_prefix_kv.data_ptr(), wkv.data_ptr(), d_prefix.data_ptr(), L, aT, C, n_groups, q_per_kv,
                                            _stream()), "dh_prefix_project_bwd")
    _count()


@_profiled("compact_targets")
def compact_targets(targets: Tensor, count: Tensor, rows: Tensor, targets_c: Tensor) -> None:
    """targets int64 [M] (-1 = ignore) -> count int32 [1], rows int32 [M] (ascending valid rows, -1 beyond), targets_c int64 [M]."""
    _dev(targets)
    M = targets.numel()
    assert targets.dtype == torch.int64 and targets_c.dtype == torch.int64 and rows.dtype == torch.int32 and count.dtype == torch.int32
    assert targets.is_contiguous() and rows.numel() == M and targets_c.numel() == M
    check(_lib.load().dh_compact_targets(targets.data_ptr(), M, count.data_ptr(), rows.data_ptr(), targets_c.data_ptr(), _stream()),
          "dh_compact_targets")
    _count()


@_profiled("gather_rows")
def gather_rows(src: Tensor, rows: Tensor, count: Tensor, dst: Tensor) -> None:
    _dev(src)
    assert src.dtype == torch.bfloat16 and dst.dtype == torch.bfloat16 and src.is_contiguous() and dst.is_contiguous()
    check(_lib.load().dh_gather_rows(src.data_ptr(), rows.data_ptr(), count.data_ptr(), dst.data_ptr(), src.shape[0], src.shape[1],
                                     _stream()), "dh_gather_rows")
    _count()


@_profiled("scatter_rows")
def scatter_rows(src: Tensor, rows: Tensor, count: Tensor, dst: Tensor) -> None:
    _dev(src)
    assert src.dtype == torch.bfloat16 and dst.dtype == torch.bfloat16 and src.is_contiguous() and dst.is_contiguous()
    check(_lib.load().dh_scatter_rows(src.data_ptr(), rows.data_ptr(), count.data_ptr(), dst.data_ptr(), dst.shape[0], dst.shape[1],
                                      _stream()), "dh_scatter_rows")
    _count()


def relpred_head_fwd(h2: Tensor, wc: Tensor, bc: Tensor, pooled: Tensor, logits: Tensor, pool: int) -> None:
    """AvgPool1d(pool, ceil_mode) + Linear(256 -> 3) over h2 f32 [B, T, 256] (ger/relprompt.py:143-146)."""
    _dev(h2)
    B, T, H = h2.shape
    assert H == 256 and all(t.dtype == torch.float32 and t.is_contiguous() for t in (h2, wc, bc, pooled, logits))
    check(_lib.load().dh_relpred_head_fwd(h2.data_ptr(), wc.data_ptr(), bc.data_ptr(), pooled.data_ptr(), logits.data_ptr(), B, T, pool, _stream()),
          "dh_relpred_head_fwd")
    _count(2)


def relpred_head_bwd(h2: Tensor, pooled: Tensor, d_logits: Tensor, wc: Tensor, dz2: Tensor, dwc: Tensor, dbc: Tensor, pool: int) -> None:
    _dev(h2)
    B, T, H = h2.shape
    assert H == 256 and all(t.dtype == torch.float32 and t.is_contiguous() for t in (h2, pooled, d_logits, wc, dz2, dwc, dbc))
    check(_lib.load().dh_relpred_head_bwd(h2.data_ptr(), pooled.data_ptr(), d_logits.data_ptr(), wc.data_ptr(), dz2.data_ptr(), dwc.data_ptr(),
                                          dbc.data_ptr(), B, T, pool, _stream()), "dh_relpred_head_bwd")
    _count(2)


@_profiled("dropout")
def dropout_bf16(x: Tensor, out: Tensor, p: float, seed: int, offset: int, accumulate: bool = False) -> None:
    """out (+)= dropout(x, p) with the Philox mask of (seed, offset); see include/dualhyp.h."""
    _dev(x)
    assert x.dtype == torch.bfloat16 and out.dtype == torch.bfloat16 and x.is_contiguous() and out.is_contiguous() and x.numel() == out.numel()
    check(_lib.load().dh_dropout_bf16(x.data_ptr(), out.data_ptr(), x.numel(), float(p), int(seed) & (2 ** 64 - 1), int(offset), int(accumulate), _stream()),
          "dh_dropout_bf16")
    _count()


def gemm_streamk_workspace_bytes(M: int, N: int, K: int) -> int:
    return int(_lib.load().dh_gemm_streamk_workspace_bytes(int(M), int(N), int(K)))


class DecodePlan:
    """Host + device copies of the stage table / TMA descriptors of the persistent decode kernel (`dh_decode_layers`)."""

    def __init__(self, desc, keep):
        lib = _lib.load()
        n = int(lib.dh_decode_plan_bytes(desc.n_layer))
        self.host = (C.c_uint8 * n)()
        check(lib.dh_decode_plan_build(C.byref(desc), self.host), "dh_decode_plan_build")
        blob = torch.frombuffer(self.host, dtype=torch.uint8)
        self.dev = torch.empty(n + 64, dtype=torch.uint8, device=keep[0].device)
        off = (-self.dev.data_ptr()) % 64                      # tensor maps need 64-byte alignment
        self.dev_view = self.dev[off:off + n]
        self.dev_view.copy_(blob)
        self.keep = keep                                       # the tensors whose addresses are baked into the plan
        self.kernels = 1


def decode_workspace_bytes(desc) -> int:
    return int(_lib.load().dh_decode_workspace_bytes(C.byref(desc)))


@_profiled("decode_layers")
def decode_layers(plan: DecodePlan) -> None:
    """All decoder layers of one decode token step as one persistent kernel (see include/dualhyp.h)."""
    check(_lib.load().dh_decode_layers(plan.dev_view.data_ptr(), C.addressof(plan.host), _stream()), "dh_decode_layers")
    _count()
