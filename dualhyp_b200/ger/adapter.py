"""`ger.adapter` mirror (reference ger/adapter.py): LLaMA-Adapter zero-init gated prefix attention.

The reference's own `GPT.forward` is broken against its ger/model.py (SURVEY headline 2); this one works and
computes what the reference's attention math defines (ger/adapter.py:116-146): for layers >=
adapter_start_layer, `y = causal_attention(q, k, v) + gating_factor * softmax(q ak^T / sqrt(hs)) av` with
ak, av = the k, v slots of the SAME qkv linear applied to `adapter_wte.weight`, no RoPE on them, a separate
softmax over the aT prefix keys, no tanh on the gate (quirk Q9).  The prefix K/V are recomputed every forward
(the reference's `adapter_kv_cache` is a stale-graph bug in training, quirk Q8)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Dict, Optional, Tuple

import torch
import torch.nn as nn

from ..engine import EngineSpec
from .config import Config as BaseConfig
from .model import GPT as BaseModel
from .model import Block as BaseBlock
from .model import CausalSelfAttention as BaseCausalSelfAttention

Tensor = torch.Tensor


@dataclass
class Config(BaseConfig):
    adapter_prompt_length: int = 10
    adapter_start_layer: int = 2


class GPT(BaseModel):
    _variant = "adapter"

    def __init__(self, config: Config) -> None:
        nn.Module.__init__(self)
        assert config.padded_vocab_size is not None
        self.config = config
        self.lm_head = nn.Linear(config.n_embd, config.padded_vocab_size, bias=config.lm_head_bias)
        self.transformer = nn.ModuleDict(dict(
            wte=nn.Embedding(config.padded_vocab_size, config.n_embd),
            h=nn.ModuleList(Block(config, i) for i in range(config.n_layer)),
            ln_f=config.norm_class(config.n_embd, eps=config.norm_eps)))
        self._init_runtime()

    def _engine_spec(self) -> EngineSpec:
        c = self.config
        return EngineSpec(n_layer=c.n_layer, n_head=c.n_head, n_query_groups=c.n_query_groups, n_embd=c.n_embd,
                          intermediate_size=c.intermediate_size, vocab=c.padded_vocab_size, block_size=c.block_size,
                          norm_eps=c.norm_eps, variant="adapter", adapter_prompt_length=c.adapter_prompt_length,
                          adapter_start_layer=c.adapter_start_layer)

    def _trainable_filter(self, name: str) -> bool:
        return adapter_filter(name, None)

    def forward(self, idx: Tensor, input_pos: Optional[Tensor] = None, lm_head_chunk_size: int = 0):
        """Signature of ger/adapter.py:47-49."""
        T = idx.size(1)
        if self.max_seq_length < T:
            raise ValueError(f"Cannot forward sequence of length {T}, max seq length is only {self.max_seq_length}.")
        return self._forward_impl(idx, input_pos, lm_head_chunk_size)

    @classmethod
    def from_name(cls, name: str, **kwargs: Any):
        return cls(Config.from_name(name, **kwargs))


class Block(BaseBlock):
    def __init__(self, config: Config, block_idx: int) -> None:
        nn.Module.__init__(self)
        self.norm_1 = config.norm_class(config.n_embd, eps=config.norm_eps)
        self.attn = CausalSelfAttention(config, block_idx)
        if not config.shared_attention_norm:
            self.norm_2 = config.norm_class(config.n_embd, eps=config.norm_eps)
        self.mlp = config.mlp_class(config)
        self.config = config


class CausalSelfAttention(BaseCausalSelfAttention):
    def __init__(self, config: Config, block_idx: int) -> None:
        super().__init__(config)
        if block_idx >= config.adapter_start_layer:
            self.adapter_wte = nn.Embedding(config.adapter_prompt_length, config.n_embd)
            self.gating_factor = nn.Parameter(torch.zeros(1, 1, config.n_head, 1))
            self.adapter_kv_cache: Optional[Tuple[Tensor, Tensor]] = None
        self.block_idx = block_idx

    def reset_parameters(self) -> None:
        nn.init.zeros_(self.gating_factor)

    def _load_from_state_dict(self, state_dict: Dict, prefix: str, *args: Any, **kwargs: Any) -> None:
        """Older checkpoints stored the gate as (1, n_head, 1, 1) (ger/adapter.py:151-155)."""
        if (key := prefix + "gating_factor") in state_dict and state_dict[key].size(1) == self.config.n_head:
            state_dict[key] = state_dict[key].permute(0, 2, 1, 3)
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)


def mark_only_adapter_as_trainable(model: GPT) -> None:
    for name, param in model.named_parameters():
        param.requires_grad = adapter_filter(name, param)


def adapter_filter(key: str, value: Any) -> bool:
    return "adapter_wte" in key or "gating_factor" in key
