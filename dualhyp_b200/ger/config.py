"""`ger.config` mirror: the `Config` dataclass and name registry the drivers construct models from
(reference ger/config.py:16-157, TinyLlama entry :1542-1567).

Only the decoder families the DualHyp scripts use are registered (TinyLlama-1.1B, plus the Llama-2-7B
shape the reference's scripts mention); other lit-gpt entries of the reference's 1.5k-line table are not
part of the hot path.  `register_config()` adds more.  Fields the reference's model code ignores
(`rope_base`, `rope_condense_ratio`, `gelu_approximate`, `scale_embeddings`) are kept
for signature compatibility and ignored here too.
"""
from __future__ import annotations

import json
from copy import deepcopy
from dataclasses import dataclass, field
from pathlib import Path
from typing import Any, Dict, List, Literal, Optional, Type, Union


def find_multiple(n: int, k: int) -> int:
    """ger/utils.py `find_multiple`."""
    assert k > 0
    return n if n % k == 0 else n + k - (n % k)


@dataclass
class Config:
    name: str = ""
    hf_config: dict = field(default_factory=dict)
    scale_embeddings: bool = False
    block_size: int = 4096
    vocab_size: int = 50254
    padding_multiple: int = 512
    padded_vocab_size: Optional[int] = None
    n_layer: int = 16
    n_head: int = 32
    head_size: Optional[int] = None
    n_embd: int = 4096
    rotary_percentage: float = 0.25
    parallel_residual: bool = True
    bias: bool = True
    lm_head_bias: bool = False
    n_query_groups: Optional[int] = None
    shared_attention_norm: bool = False
    _norm_class: Literal["LayerNorm", "RMSNorm"] = "LayerNorm"
    norm_eps: float = 1e-5
    _mlp_class: Literal["GptNeoxMLP", "LLaMAMLP", "GemmaMLP", "LLaMAMoE"] = "GptNeoxMLP"
    gelu_approximate: str = "none"
    intermediate_size: Optional[int] = None
    rope_condense_ratio: int = 1
    rope_base: int = 10000
    n_expert: int = 0
    n_expert_per_token: int = 0

    def __post_init__(self):
        if not self.name:
            self.name = self.hf_config.get("name", self.name)
        if self.head_size is None:
            assert self.n_embd % self.n_head == 0
            self.head_size = self.n_embd // self.n_head
        if self.padded_vocab_size is None:
            self.padded_vocab_size = find_multiple(self.vocab_size, self.padding_multiple)
        else:
            self.vocab_size = min(self.vocab_size, self.padded_vocab_size)
        if self.n_query_groups is not None:
            assert self.n_head % self.n_query_groups == 0
        else:
            self.n_query_groups = self.n_head
        if self.intermediate_size is None:
            if self._mlp_class == "LLaMAMLP":
                raise ValueError("The config needs to set the `intermediate_size`")
            self.intermediate_size = 4 * self.n_embd
        self.rope_n_elem = int(self.rotary_percentage * self.head_size)

    # -- constructors (ger/config.py:94-133) --------------------------------------------------------
    @classmethod
    def from_name(cls, name: str, **kwargs: Any):
        if name in name_to_config:
            conf_dict = name_to_config[name]
        else:
            try:
                conf_dict = next(c for c in configs if name == c["hf_config"]["name"])
            except StopIteration:
                raise ValueError(f"{name!r} is not a supported config name")
        conf_dict = conf_dict.copy()
        if "condense_ratio" in kwargs:
            kwargs["rope_condense_ratio"] = kwargs.pop("condense_ratio")
        conf_dict.update(kwargs)
        return cls(**conf_dict)

    @classmethod
    def from_json(cls, path: Union[str, Path], **kwargs: Any):
        with open(path, encoding="utf-8") as fp:
            json_kwargs = json.load(fp)
        if "condense_ratio" in json_kwargs:
            json_kwargs["rope_condense_ratio"] = json_kwargs.pop("condense_ratio")
        if "condense_ratio" in kwargs:
            kwargs["rope_condense_ratio"] = kwargs.pop("condense_ratio")
        if "org" in json_kwargs:
            json_kwargs["hf_config"] = {"name": json_kwargs["name"], "org": json_kwargs.pop("org")}
        if "org" in kwargs:
            kwargs["hf_config"] = {"name": kwargs.get("name", json_kwargs["name"]), "org": kwargs.pop("org")}
        json_kwargs.update(kwargs)
        return cls(**json_kwargs)

    @classmethod
    def from_checkpoint(cls, path: Path, **kwargs: Any):
        if (config_path := path / "lit_config.json").is_file():
            return cls.from_json(config_path, **kwargs)
        if (model_name := path.name) in name_to_config:
            return cls.from_name(model_name, **kwargs)
        raise FileNotFoundError(f"For {str(path)!r} neither 'lit_config.json' nor matching config exists.")

    @property
    def mlp_class(self) -> Type:
        from . import model
        return getattr(model, self._mlp_class)

    @property
    def norm_class(self) -> Type:
        if self._norm_class == "RMSNorm":
            from .rmsnorm import RMSNorm
            return RMSNorm
        import torch
        return getattr(torch.nn, self._norm_class)

    # -- what the B200 engine supports ---------------------------------------------------------------
    def check_engine_supported(self) -> None:
        """The sm_100a kernels are specialised for the LLaMA-style decoder the DualHyp scripts train."""
        problems = []
        if self.head_size != 64:
            problems.append(f"head_size {self.head_size} != 64")
        if self._norm_class != "RMSNorm" or self._mlp_class != "LLaMAMLP":
            problems.append("needs RMSNorm + LLaMAMLP")
        if self.parallel_residual or self.shared_attention_norm or self.bias or self.lm_head_bias:
            problems.append("needs sequential residual and bias-free linears")
        if self.rotary_percentage != 1.0:
            problems.append("rotary_percentage must be 1.0")
        if self.n_embd % 256 or self.intermediate_size % 64 or self.padded_vocab_size % 64:
            problems.append("n_embd % 256, intermediate_size % 64, padded_vocab_size % 64 must be 0")
        if problems:
            raise NotImplementedError("config not supported by the B200 engine: " + "; ".join(problems))


configs: List[Dict[str, Any]] = []

tiny_llama = dict(
    name="tiny-llama-1.1b{}", hf_config=dict(org="TinyLlama", name="TinyLlama-1.1B{}"), block_size=2048, vocab_size=32000,
    padding_multiple=64, n_layer=22, n_head=32, n_embd=2048, rotary_percentage=1.0, parallel_residual=False, bias=False,
    _norm_class="RMSNorm", norm_eps=1e-5, _mlp_class="LLaMAMLP", intermediate_size=5632, n_query_groups=4)
for kind, hf_postfix in (("", "-intermediate-step-1431k-3T"), ("-chat", "-Chat-v1.0")):
    c = deepcopy(tiny_llama)
    c["name"] = tiny_llama["name"].format(kind)
    c["hf_config"]["name"] = tiny_llama["hf_config"]["name"].format(hf_postfix)
    configs.append(c)

# a TinyLlama-shaped toy used by the golden fixtures and the smoke test
configs.append(dict(
    name="tiny-llama-test", hf_config=dict(org="dualhyp-b200", name="tiny-llama-test"), block_size=256, vocab_size=512,
    padded_vocab_size=512, n_layer=3, n_head=4, n_embd=256, rotary_percentage=1.0, parallel_residual=False, bias=False,
    _norm_class="RMSNorm", norm_eps=1e-5, _mlp_class="LLaMAMLP", intermediate_size=512, n_query_groups=2))

name_to_config: Dict[str, Dict[str, Any]] = {c["name"]: c for c in configs}


def register_config(conf: Dict[str, Any]) -> None:
    configs.append(conf)
    name_to_config[conf["name"]] = conf
