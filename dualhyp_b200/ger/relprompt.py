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


class _RelPredFn(torch.autograd.Function):
    """The whole reliability head as one node on the `dh_relpred_*` kernels (fp32, as the reference runs it)."""

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
        if self.training and self.dropout.p > 0:
            import warnings
            warnings.warn("classifier dropout is not applied by the B200 reliability-head kernel (parity runs use eval())")
        return _RelPredFn.apply(x, self.conv1.weight, self.conv1.bias, self.conv2.weight, self.conv2.bias,
                                self.classifier.weight, self.classifier.bias, self.pool_size)


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
