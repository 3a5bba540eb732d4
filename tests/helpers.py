"""Shared helpers for the parity tests: the tiny TinyLlama-shaped configs used by the golden fixtures."""
import os

import numpy as np
import torch

from oracle.oracle import OracleConfig

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

TINY = dict(n_layer=3, n_head=4, n_query_groups=2, head_size=64, n_embd=256, intermediate_size=512,
            padded_vocab_size=512, block_size=256, norm_eps=1e-5)
LORA = dict(r=8, alpha=16, to_query=True, to_key=True, to_value=True, to_projection=True)


def tiny_cfg(variant: str) -> OracleConfig:
    if variant == "lora":
        return OracleConfig(**TINY, **LORA, variant="lora")
    if variant == "adapter":
        return OracleConfig(**TINY, variant="adapter", adapter_prompt_length=10, adapter_start_layer=1)
    if variant == "relprompt":
        return OracleConfig(**TINY, **LORA, variant="relprompt", whisper_dim=48, raven_dim=32, pool_size=5)
    raise ValueError(variant)


def full_cfg(variant: str, **over) -> OracleConfig:
    """TinyLlama-1.1B (ger/config.py:1542-1560)."""
    base = dict(variant=variant)
    if variant in ("lora", "relprompt"):
        base.update(r=8, alpha=16, to_query=True, to_key=True, to_value=True, to_projection=True)
    base.update(over)
    return OracleConfig(**base)


def load_golden(name: str):
    z = np.load(os.path.join(GOLDEN, name))
    return {k: z[k] for k in z.files}


def t(a):
    return torch.from_numpy(np.asarray(a))


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))
