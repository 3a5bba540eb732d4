"""`ger.relprompt` mirror (reference ger/relprompt.py): the LoRA GPT plus the two reliability predictors
(`NoiseMaskClassifier`) and `resize_token_embeddings`, with the reference's forward signature
(`audio_query` / `lip_query` / `max_seq_length` are accepted and, as in the reference, unused: RelPrompt
conditions the LLM only through the <<C>>/<<M>>/<<N>> tokens in the prompt)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Type, Union

import torch
import torch.nn as nn

from .. import ops
from . import lora as _lora
from .config import Config as BaseConfig
from .lora import (AdapterV2Linear, LoRALayer, LoRALinear, LoRAQKVLinear, lora_filter)  # noqa: F401
from .lora import Block as _LoRABlock
from .lora import CausalSelfAttention as _LoRACausalSelfAttention
from .lora import LLaMAMLP as _LoRALLaMAMLP
from .model import Block as BaseBlock
from .utils import map_old_state_dict_weights  # noqa: F401

Tensor = torch.Tensor


def mark_only_lora_as_trainable(model: nn.Module, bias: str = "none") -> None:
    """ger/relprompt.py:79-119: LoRA tensors and the noise classifiers train, everything else is frozen."""
    for n, p in model.named_parameters():
        if "audio_proj" in n or "visual_proj" in n or "noise_classifier" in n:
            p.requires_grad = True
            continue
        if "lora_" not in n:
            p.requires_grad = False
    if bias == "none":
        return
    if bias == "all":
        for n, p in model.named_parameters():
            if "bias" in n:
                p.requires_grad = True
    elif bias == "lora_only":
        for m in model.modules():
            if isinstance(m, LoRALayer) and hasattr(m, "bias") and m.bias is not None:
                m.bias.requires_grad = True
    else:
        raise NotImplementedError


class _RelPredFp32Fn(torch.autograd.Function):
    """The whole reliability head as one node on the fp32 `dh_relpred_*` kernels (CUDA cores; DH_RELPRED=fp32): the arithmetic of
    the reference on CPU.  No dropout."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, wc, bc, pool):
        xf = x.detach().float().contiguous()
        B, T, _ = xf.shape
        n_out = -(-T // pool)
        dev = xf.device
        h1 = torch.empty((B, T, 256), device=dev)
        h2 = torch.empty((B, T, 256), device=dev)
        pooled = torch.empty((B, n_out, 256), device=dev)
        logits = torch.empty((B, n_out, 3), device=dev)
        ws = [t.detach().float().contiguous() for t in (w1, b1, w2, b2, wc, bc)]
        ops.relpred_fwd(xf, *ws, h1, h2, pooled, logits, pool)
        ctx.save_for_backward(xf, ws[0], ws[2], ws[4], h1, h2, pooled)
        ctx.pool = pool
        ctx.dtypes = [t.dtype for t in (w1, b1, w2, b2, wc, bc)]
        return logits

    @staticmethod
    def backward(ctx, d_logits):
        xf, w1, w2, wc, h1, h2, pooled = ctx.saved_tensors
        B, T, _ = xf.shape
        grads = [torch.zeros_like(w1), torch.zeros(256, device=xf.device), torch.zeros_like(w2), torch.zeros(256, device=xf.device),
                 torch.zeros_like(wc), torch.zeros(3, device=xf.device)]
        scratch = torch.empty(2 * B * T * 256, device=xf.device)
        ops.relpred_bwd(xf, w1, w2, wc, h1, h2, pooled, d_logits.float().contiguous(), *grads, scratch, ctx.pool)
        return (None, *[g.to(dt) for g, dt in zip(grads, ctx.dtypes)], None)


def _frames3(h: Tensor) -> Tensor:
    """[B, T, C] -> [B * T, 3 C]: row (b, t) = frames t-1, t, t+1 (zeros outside the clip) -- the A operand that turns
    Conv1d(k=3, padding=1) into one GEMM against the weight reshaped to [out, (tap, c)]."""
    B, T, C = h.shape
    hp = torch.nn.functional.pad(h, (0, 0, 1, 1))
    parts = [hp[:, 0:T], hp[:, 1:T + 1], hp[:, 2:T + 2]]
    if (3 * C) % 64:            # dh_gemm wants the weight gradient's N = 3 C in whole 64-column tiles: zero columns up to it
        parts.append(hp.new_zeros((B, T, -(3 * C) % 64)))
    return torch.cat(parts, dim=-1).reshape(B * T, -1)


def _taps_first(w: Tensor) -> Tensor:
    """Conv1d weight [out, C, 3] -> bf16 [out, 3 C (+ zero columns to a multiple of 64)] with k = tap * C + c (the column order of
    `_frames3`)."""
    wr = w.detach().permute(0, 2, 1).reshape(w.shape[0], -1).to(torch.bfloat16)
    if wr.shape[1] % 64:
        wr = torch.nn.functional.pad(wr, (0, -wr.shape[1] % 64))
    return wr.contiguous()


class _RelPredFn(torch.autograd.Function):
    """The reliability head on the tensor cores (ger/relprompt.py:137-147; the reference runs these contractions as TF32
    cuDNN / cuBLAS calls under `set_float32_matmul_precision("high")`, finetune/ger.py:431).  Both convolutions, conv2's input
    gradient and both weight gradients are `dh_gemm` calls (tcgen05, bf16 operands, fp32 accumulation; K = 3 * 1280, 3 * 256 and
    the frame count for the weight gradients); bias + ReLU, the Philox dropout between the convolutions
    (`dh_dropout_bf16`, ger/relprompt.py:141), pooling and the classifier are small launches around them."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, wc, bc, pool, p_drop, seed):
        with ops.region("relpred_fwd"):
            B, T, C = x.shape
            dev = x.device
            n_out = -(-T // pool)
            A1 = _frames3(x.detach().to(torch.bfloat16))
            W1r, W2r = _taps_first(w1), _taps_first(w2)
            z1 = torch.empty((B * T, 256), device=dev, dtype=torch.float32)
            ops.gemm(A1, W1r, z1, epilogue=ops.EPI_F32)
            h1 = torch.relu_(z1.add_(b1.detach().float())).to(torch.bfloat16)
            inv_keep = 1.0
            if p_drop > 0.0 and seed is not None:
                h1d = torch.empty_like(h1)
                ops.dropout_bf16(h1, h1d, p_drop, seed, 0)
                thresh = round(p_drop * 65536)
                inv_keep = 65536.0 / (65536 - thresh)
                h1 = h1d
            A2 = _frames3(h1.view(B, T, 256))
            h2 = torch.empty((B, T, 256), device=dev, dtype=torch.float32)
            ops.gemm(A2, W2r, h2.view(B * T, 256), epilogue=ops.EPI_F32)
            torch.relu_(h2.add_(b2.detach().float()))
            pooled = torch.empty((B, n_out, 256), device=dev)
            logits = torch.empty((B, n_out, 3), device=dev)
            wcf = wc.detach().float().contiguous()
            ops.relpred_head_fwd(h2, wcf, bc.detach().float().contiguous(), pooled, logits, pool)
            ctx.save_for_backward(A1, A2, W2r, h2, pooled, wcf)
            ctx.pool, ctx.inv_keep, ctx.shape = pool, inv_keep, (B, T, C)
            ctx.dtypes = [t.dtype for t in (w1, b1, w2, b2, wc, bc)]
        return logits

    @staticmethod
    def backward(ctx, d_logits):
        with ops.region("relpred_bwd"):
            A1, A2, W2r, h2, pooled, wcf = ctx.saved_tensors
            B, T, C = ctx.shape
            dev = h2.device
            bf, f32 = torch.bfloat16, torch.float32
            dz2 = torch.empty((B, T, 256), device=dev, dtype=f32)
            dwc, dbc = torch.zeros((3, 256), device=dev), torch.zeros(3, device=dev)
            ops.relpred_head_bwd(h2, pooled, d_logits.float().contiguous(), wcf, dz2, dwc, dbc, ctx.pool)
            dz2 = dz2.view(B * T, 256)
            db2 = dz2.sum(0)
            dz2b = dz2.to(bf)
            dW2r = torch.empty((256, 3 * 256), device=dev, dtype=f32)
            ops.gemm(dz2b, A2, dW2r, a_layout=1, b_layout=1, epilogue=ops.EPI_F32)            # [256, (tap, c)] = dz2^T A2, K = frames
            dA2 = torch.empty((B * T, 3 * 256), device=dev, dtype=bf)
            ops.gemm(dz2b, W2r, dA2, b_layout=1)                                                # d(frames3(h1)) = dz2 W2r
            P = torch.nn.functional.pad(dA2.view(B, T, 3, 256), (0, 0, 0, 0, 1, 1))             # frame t of tap k came from h1[t + k - 1]
            dh1 = P[:, 2:, 0].float() + P[:, 1:-1, 1].float() + P[:, :-2, 2].float()
            h1 = A2.view(B, T, 3, 256)[:, :, 1]                                                 # the centre tap IS h1 (after dropout)
            dz1 = torch.where(h1 > 0, dh1 * ctx.inv_keep, torch.zeros((), device=dev)).reshape(B * T, 256)   # ReLU and dropout mask in one test
            db1 = dz1.sum(0)
            dW1r = torch.empty((256, A1.shape[1]), device=dev, dtype=f32)
            ops.gemm(dz1.to(bf), A1, dW1r, a_layout=1, b_layout=1, epilogue=ops.EPI_F32)
            dw1 = dW1r[:, : 3 * C].reshape(256, 3, C).permute(0, 2, 1)
            dw2 = dW2r.view(256, 3, 256).permute(0, 2, 1)
            grads = (dw1, db1, dw2, db2, dwc, dbc)
        return (None, *[g.to(dt) for g, dt in zip(grads, ctx.dtypes)], None, None, None)


class NoiseMaskClassifier(nn.Module):
    """ger/relprompt.py:126-147.  (B, T, C) -> (B, ceil(T / pool_size), 3)."""

    def __init__(self, input_dim, hidden_dim=256, dropout=0.1, pool_size=10):
        super().__init__()
        if hidden_dim != 256:
            raise NotImplementedError("the reliability-head kernel is specialised for hidden_dim=256 (the reference's only setting)")
        self.pool_size = pool_size
        self.conv1 = nn.Conv1d(input_dim, hidden_dim, kernel_size=3, padding=1)
        self.conv2 = nn.Conv1d(hidden_dim, hidden_dim, kernel_size=3, padding=1)
        self.pool = nn.AvgPool1d(kernel_size=pool_size, stride=pool_size, ceil_mode=True)
        self.classifier = nn.Linear(hidden_dim, 3)
        self.dropout = nn.Dropout(dropout)

    def forward(self, x):
        import os
        args = (x, self.conv1.weight, self.conv1.bias, self.conv2.weight, self.conv2.bias, self.classifier.weight, self.classifier.bias,
                self.pool_size)
        p = float(self.dropout.p) if self.training else 0.0
        if os.environ.get("DH_RELPRED", "tc") == "fp32":
            if p > 0:
                raise NotImplementedError("DH_RELPRED=fp32 (the CUDA-core head) has no dropout; call eval() or use the default tensor-core head")
            return _RelPredFp32Fn.apply(*args)
        # dropout seed from torch's CPU generator, like the LoRA dropout (GPT._dropout_seed)
        seed = int(torch.randint(0, 2 ** 62, (1,), dtype=torch.int64).item()) if p > 0 else None
        return _RelPredFn.apply(*args, p, seed)


@dataclass
class Config(BaseConfig):
    """ger/relprompt.py:150-179."""
    r: int = 16
    alpha: int = 1
    dropout: float = 0.0
    to_query: bool = True
    to_key: bool = True
    to_value: bool = True
    to_projection: bool = True
    to_mlp: bool = False
    to_head: bool = False
    lora_start_layer: int = 0
    whisper_dim: int = 1280
    raven_dim: int = 1024
    pool_size: int = 10

    @property
    def mlp_class(self) -> Type:
        return _LoRALLaMAMLP


class GPT(_lora.GPT):
    _variant = "lora"

    def __init__(self, config: Config) -> None:
        nn.Module.__init__(self)
        assert config.padded_vocab_size is not None
        self.config = config
        if config.to_head or config.to_mlp:
            raise NotImplementedError("to_head / to_mlp LoRA are not used by the DualHyp scripts")
        self.lm_head = AdapterV2Linear(config.n_embd, config.padded_vocab_size, bias=config.lm_head_bias)
        self.transformer = nn.ModuleDict(dict(
            wte=nn.Embedding(config.padded_vocab_size, config.n_embd),
            h=nn.ModuleList(Block(config, i) for i in range(config.n_layer)),
            ln_f=config.norm_class(config.n_embd, eps=config.norm_eps)))
        self._init_runtime()
        self.audio_noise_classifier = NoiseMaskClassifier(config.whisper_dim, pool_size=2 * config.pool_size)
        self.visual_noise_classifier = NoiseMaskClassifier(config.raven_dim, pool_size=config.pool_size)

    def resize_token_embeddings(self, new_vocab_size: int):
        """ger/relprompt.py:215-230: ADDS `new_vocab_size` rows to wte (new rows ~ N(0, std(old))); lm_head keeps
        its size (quirk Q10)."""
        old_embed: nn.Embedding = self.transformer.wte
        old_num, dim = old_embed.weight.shape
        if new_vocab_size == old_num:
            return
        new_embed = nn.Embedding(old_num + new_vocab_size, dim, device=old_embed.weight.device, dtype=old_embed.weight.dtype)
        new_embed.weight.data[:old_num].copy_(old_embed.weight.data)
        nn.init.normal_(new_embed.weight.data[old_num:], mean=0.0, std=old_embed.weight.data.float().std().item())
        self.transformer.wte = new_embed

    def forward(self, idx: Tensor, audio_query: Tensor = None, lip_query: Tensor = None, max_seq_length: Optional[int] = None,
                input_pos: Optional[Tensor] = None, lm_head_chunk_size: int = 0):
        """ger/relprompt.py:232-294."""
        return self._forward_impl(idx, input_pos, lm_head_chunk_size)

    @classmethod
    def from_name(cls, name: str, **kwargs: Any):
        return cls(Config.from_name(name, **kwargs))


class Block(_LoRABlock):
    pass


CausalSelfAttention = _LoRACausalSelfAttention
LLaMAMLP = _LoRALLaMAMLP


def merge_lora_weights(model: GPT) -> None:
    for module in model.modules():
        if isinstance(module, LoRALinear):
            module.merge()
