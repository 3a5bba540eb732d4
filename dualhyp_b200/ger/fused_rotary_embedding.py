"""`ger.fused_rotary_embedding` mirror (reference ger/fused_rotary_embedding.py:10-90).

The reference wraps flash-attn's `rotary_emb.apply_rotary` CUDA extension (not installed anywhere here and
dead code in the reference: its call sites are commented out, ger/model.py:237-238).  This keeps the same
autograd.Function surface -- `apply_rotary_emb_func(x, cos, sin, interleaved=False, inplace=False)` with
x (batch, seqlen, nheads, headdim), cos/sin (seqlen, rotary_dim/2), backward = conjugate rotation -- on the
`dh_rope_apply` kernel.  On the hot path itself RoPE is fused into the QKV GEMM epilogue instead."""
import torch

from .. import ops


class ApplyRotaryEmb(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cos, sin, interleaved=False, inplace=False):
        batch, seqlen, nheads, headdim = x.shape
        rotary_seqlen, rotary_dim = cos.shape
        rotary_dim *= 2
        assert rotary_dim <= headdim
        assert seqlen <= rotary_seqlen
        assert sin.shape == (rotary_seqlen, rotary_dim // 2)
        if x.dtype != torch.bfloat16:
            raise NotImplementedError("the B200 rotary kernel takes bf16 activations")
        xc = x if x.is_contiguous() else x.contiguous()
        if inplace and xc is not x:
            raise ValueError("inplace rotary needs a contiguous input")
        out = xc if inplace else torch.empty_like(xc)
        ops.rope_apply(xc, cos.contiguous(), sin.contiguous(), out, interleaved=interleaved, conj=False)
        ctx.save_for_backward(cos, sin)
        ctx.interleaved = interleaved
        ctx.inplace = inplace
        return out

    @staticmethod
    def backward(ctx, do):
        cos, sin = ctx.saved_tensors
        doc = do.contiguous()
        dx = doc if ctx.inplace else torch.empty_like(doc)
        ops.rope_apply(doc, cos.contiguous(), sin.contiguous(), dx, interleaved=ctx.interleaved, conj=True)
        return dx, None, None, None, None


apply_rotary_emb_func = ApplyRotaryEmb.apply
