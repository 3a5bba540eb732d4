"""`ger.lora` mirror (reference ger/lora.py): LoRA GPT with the same Config fields, module tree, parameter
names and base->LoRA checkpoint key remapping.  This is what finetune/ger.py trains for GER and DualHyp.

On the B200 the LoRA branches are not separate small GEMMs + adds: `after_A = x A^T` is one skinny GEMM and
the `B` product rides inside the main QKV / proj GEMM as a K extension of the accumulator (engine.py); the
QKV delta keeps the reference's [q | k | v] column order on top of the group-interleaved pretrained
columns."""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Tuple, Type, Union

import torch
import torch.nn as nn

from ..engine import EngineSpec
from . import model as _model
from .config import Config as BaseConfig
from .model import GPT as BaseModel
from .model import Block as BaseBlock
from .model import CausalSelfAttention as BaseCausalSelfAttention
from .model import LLaMAMLP as BaseLLaMAMLP
from .model import KVCache, LazyLogits, RoPECache
from .utils import map_old_state_dict_weights

Tensor = torch.Tensor


class AdapterV2Linear(nn.Module):
    """ger/lora.py:63-75: adapter_scale * (linear(x) + adapter_bias), both frozen at 1 / 0."""

    def __init__(self, in_features: int, out_features: int, **kwargs) -> None:
        super().__init__()
        self.linear = nn.Linear(in_features, out_features, **kwargs)
        self.adapter_bias = nn.Parameter(torch.zeros(out_features), requires_grad=False)
        self.adapter_scale = nn.Parameter(torch.ones(out_features), requires_grad=False)

    def reset_parameters(self) -> None:
        nn.init.zeros_(self.adapter_bias)
        nn.init.ones_(self.adapter_scale)


class LoRALayer(nn.Module):
    def __init__(self, r: int, lora_alpha: int, lora_dropout: float):
        super().__init__()
        assert r >= 0
        self.r = r
        self.lora_alpha = lora_alpha
        self.lora_dropout_p = float(lora_dropout)
        self.merged = False


class LoRALinear(LoRALayer):
    """ger/lora.py:103-166: `linear` (frozen) + lora_A [r, in], lora_B [out, r], scaling = alpha / r."""

    def __init__(self, in_features: int, out_features: int, r: int = 0, lora_alpha: int = 1, lora_dropout: float = 0.0, **kwargs):
        super().__init__(r=r, lora_alpha=lora_alpha, lora_dropout=lora_dropout)
        self.linear = nn.Linear(in_features, out_features, **kwargs)
        if r > 0:
            self.lora_A = nn.Parameter(self.linear.weight.new_zeros((r, in_features)))
            self.lora_B = nn.Parameter(self.linear.weight.new_zeros((out_features, r)))
            self.scaling = self.lora_alpha / self.r
            self.reset_parameters()

    def reset_parameters(self):
        if hasattr(self, "lora_A"):
            nn.init.kaiming_uniform_(self.lora_A, a=math.sqrt(5))
            nn.init.zeros_(self.lora_B)

    def merge(self):
        """W += (B A) * scaling  (ger/lora.py:152-157); tiny host-side update of the frozen weight."""
        if self.r > 0 and not self.merged:
            delta = ((self.lora_B.data.float() @ self.lora_A.data.float()) * self.scaling).to(self.linear.weight.dtype)
            with torch.no_grad():
                self.linear.weight.add_(delta)      # in place THROUGH the parameter: bumps its version, so the engine re-packs W
            self.merged = True                      # the engine drops the LoRA K-extension of merged modules (GPT._trainables)


class LoRAQKVLinear(LoRALinear):
    """ger/lora.py:169-402 for the combined QKV projection with GQA.  lora_A [r*n_enabled, in];
    lora_B [sum(enabled q/k/v widths), r]."""

    def __init__(self, in_features: int, out_features: int, n_head: int, n_query_groups: int, r: int = 0, lora_alpha: int = 1,
                 lora_dropout: float = 0.0, enable_lora: Union[bool, Tuple[bool, bool, bool]] = False, **kwargs):
        super(LoRALinear, self).__init__(r=r, lora_alpha=lora_alpha, lora_dropout=lora_dropout)
        self.linear = nn.Linear(in_features, out_features, **kwargs)
        self.n_head = n_head
        self.n_query_groups = n_query_groups
        if isinstance(enable_lora, bool):
            enable_lora = [enable_lora] * 3
        assert len(enable_lora) == 3
        self.enable_lora = enable_lora
        if r > 0 and any(enable_lora):
            self.lora_A = nn.Parameter(self.linear.weight.new_zeros((r * sum(enable_lora), in_features)))
            enable_q, enable_k, enable_v = enable_lora
            self.kv_embd_size = self.linear.in_features // (n_head // n_query_groups)
            qkv_shapes = (self.linear.in_features * enable_q, self.kv_embd_size * enable_k, self.kv_embd_size * enable_v)
            self.qkv_shapes = [s for s in qkv_shapes if s]
            self.lora_B = nn.Parameter(self.linear.weight.new_zeros(sum(self.qkv_shapes), r))
            self.scaling = self.lora_alpha / self.r
            self.reset_parameters()

    def merge(self):
        """ger/lora.py:349-365 with all of q, k, v enabled: delta rows [q | k | v] each from its own A slice."""
        if self.r > 0 and any(self.enable_lora) and not self.merged:
            assert all(self.enable_lora)
            r = self.r
            A, Bm = self.lora_A.data.float(), self.lora_B.data.float()
            parts, off = [], 0
            for i, s in enumerate(self.qkv_shapes):
                parts.append(Bm[off:off + s] @ A[i * r:(i + 1) * r])
                off += s
            with torch.no_grad():
                self.linear.weight.add_((torch.cat(parts, dim=0) * self.scaling).to(self.linear.weight.dtype))
            self.merged = True


def mark_only_lora_as_trainable(model: nn.Module, bias: str = "none") -> None:
    """ger/lora.py:405-439."""
    for n, p in model.named_parameters():
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


def lora_filter(key: str, value: Any) -> bool:
    return "lora_" in key


@dataclass
class Config(BaseConfig):
    """ger/lora.py:446-472."""
    r: int = 0
    alpha: int = 1
    dropout: float = 0.0
    to_query: bool = False
    to_key: bool = False
    to_value: bool = False
    to_projection: bool = False
    to_mlp: bool = False
    to_head: bool = False
    lora_start_layer: int = 0

    @property
    def mlp_class(self) -> Type:
        import sys
        return getattr(sys.modules[__name__], self._mlp_class)


class GPT(BaseModel):
    _variant = "lora"

    def __init__(self, config: Config) -> None:
        nn.Module.__init__(self)
        assert config.padded_vocab_size is not None
        self.config = config
        if config.to_head:
            raise NotImplementedError("LoRA on lm_head (to_head=True) is not used by the DualHyp scripts (finetune/ger.py:407)")
        if config.to_mlp:
            raise NotImplementedError("LoRA on the MLP (to_mlp=True) is not used by the DualHyp scripts (finetune/ger.py:406)")
        self.lm_head = AdapterV2Linear(config.n_embd, config.padded_vocab_size, bias=config.lm_head_bias)
        self.transformer = nn.ModuleDict(dict(
            wte=nn.Embedding(config.padded_vocab_size, config.n_embd),
            h=nn.ModuleList(Block(config, i) for i in range(config.n_layer)),
            ln_f=config.norm_class(config.n_embd, eps=config.norm_eps)))
        self._init_runtime()

    def lm_head_weight(self) -> Tensor:
        return self.lm_head.linear.weight

    def _engine_spec(self) -> EngineSpec:
        c = self.config
        qkv_flags = (c.to_query, c.to_key, c.to_value)
        if c.r > 0 and any(qkv_flags) and not all(qkv_flags):
            raise NotImplementedError("the B200 LoRA-QKV path implements the all-enabled case the DualHyp scripts use (q, k, v together)")
        return EngineSpec(n_layer=c.n_layer, n_head=c.n_head, n_query_groups=c.n_query_groups, n_embd=c.n_embd,
                          intermediate_size=c.intermediate_size, vocab=c.padded_vocab_size, block_size=c.block_size,
                          norm_eps=c.norm_eps, variant="lora", r=c.r, alpha=float(c.alpha), lora_qkv=all(qkv_flags) and c.r > 0,
                          lora_proj=bool(c.to_projection) and c.r > 0, lora_start_layer=c.lora_start_layer,
                          lora_dropout=float(c.dropout))

    def _trainable_filter(self, name: str) -> bool:
        return "lora_" in name

    def forward(self, idx: Tensor, input_pos: Optional[Tensor] = None, lm_head_chunk_size: int = 0):
        """ger/lora.py:504-549."""
        return self._forward_impl(idx, input_pos, lm_head_chunk_size)

    @classmethod
    def from_name(cls, name: str, **kwargs: Any):
        return cls(Config.from_name(name, **kwargs))

    def _load_from_state_dict(self, state_dict: Dict, prefix: str, *args: Any, **kwargs: Any) -> None:
        mapping = {"lm_head.weight": "lm_head.linear.weight", "lm_head.bias": "lm_head.linear.bias"}
        state_dict = map_old_state_dict_weights(state_dict, mapping, prefix)
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)


class Block(BaseBlock):
    def __init__(self, config: Config, block_idx: int) -> None:
        nn.Module.__init__(self)
        self.norm_1 = config.norm_class(config.n_embd, eps=config.norm_eps)
        self.attn = CausalSelfAttention(config) if block_idx >= config.lora_start_layer else BaseCausalSelfAttention(config)
        if not config.shared_attention_norm:
            self.norm_2 = config.norm_class(config.n_embd, eps=config.norm_eps)
        self.mlp = config.mlp_class(config) if block_idx >= config.lora_start_layer else BaseLLaMAMLP(config)
        self.config = config


class CausalSelfAttention(BaseCausalSelfAttention):
    def __init__(self, config: Config) -> None:
        nn.Module.__init__(self)
        shape = (config.n_head + 2 * config.n_query_groups) * config.head_size
        self.attn = LoRAQKVLinear(in_features=config.n_embd, out_features=shape, r=config.r, lora_alpha=config.alpha,
                                  lora_dropout=config.dropout, enable_lora=(config.to_query, config.to_key, config.to_value),
                                  bias=config.bias, n_head=config.n_head, n_query_groups=config.n_query_groups)
        self.proj = LoRALinear(config.n_embd, config.n_embd, bias=config.bias, r=(config.r if config.to_projection else 0),
                               lora_alpha=config.alpha, lora_dropout=config.dropout)
        self.kv_cache: Optional[KVCache] = None
        self.config = config

    def _load_from_state_dict(self, state_dict: Dict, prefix: str, *args: Any, **kwargs: Any) -> None:
        mapping = {"attn.weight": "attn.linear.weight", "attn.bias": "attn.linear.bias",
                   "proj.weight": "proj.linear.weight", "proj.bias": "proj.linear.bias"}
        state_dict = map_old_state_dict_weights(state_dict, mapping, prefix)
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)


class LLaMAMLP(BaseLLaMAMLP):
    def __init__(self, config: Config) -> None:
        nn.Module.__init__(self)
        kw = dict(bias=config.bias, r=(config.r if config.to_mlp else 0), lora_alpha=config.alpha,
                  lora_dropout=(config.dropout if config.to_mlp else 0))
        self.fc_1 = LoRALinear(config.n_embd, config.intermediate_size, **kw)
        self.fc_2 = LoRALinear(config.n_embd, config.intermediate_size, **kw)
        self.proj = LoRALinear(config.intermediate_size, config.n_embd, **kw)

    def _load_from_state_dict(self, state_dict: Dict, prefix: str, *args: Any, **kwargs: Any) -> None:
        mapping = {"fc_1.weight": "fc_1.linear.weight", "fc_1.bias": "fc_1.linear.bias", "fc_2.weight": "fc_2.linear.weight",
                   "fc_2.bias": "fc_2.linear.bias", "proj.weight": "proj.linear.weight", "proj.bias": "proj.linear.bias"}
        state_dict = map_old_state_dict_weights(state_dict, mapping, prefix)
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)


def merge_lora_weights(model: GPT) -> None:
    """ger/lora.py:707-711."""
    for module in model.modules():
        if isinstance(module, LoRALinear):
            module.merge()
