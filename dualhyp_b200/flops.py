"""Algorithmic FLOP counts of the decoder path: only the dense contractions are credited
(linears + attention), per token of a length-T causal sequence; LoRA/adapter terms (<0.5 %) and elementwise
work are not.  These are the numerators of bench.py's tensor-pipe roofline."""
from __future__ import annotations


def linear_params_per_layer(cfg) -> int:
    qkv = (cfg.n_head + 2 * cfg.n_query_groups) * cfg.head_size * cfg.n_embd
    return qkv + cfg.n_embd * cfg.n_embd + 3 * cfg.n_embd * cfg.intermediate_size


def forward_flops_per_token(cfg, T: int, rho: float = 1.0) -> float:
    L, C = cfg.n_layer, cfg.n_embd
    head = cfg.padded_vocab_size * C
    return 2.0 * (L * linear_params_per_layer(cfg) + rho * head) + L * 2.0 * T * C


def backward_flops_per_token(cfg, variant: str, T: int, rho: float = 1.0) -> float:
    """dgrad only (frozen weights).  LoRA from layer 0: every layer except the QKV dgrad of the first one;
    adapter from `adapter_start_layer`: only the layers at or above it, again without the first QKV dgrad."""
    L, C = cfg.n_layer, cfg.n_embd
    head = cfg.padded_vocab_size * C
    qkv = (cfg.n_head + 2 * cfg.n_query_groups) * cfg.head_size * C
    if variant == "adapter":
        n = L - getattr(cfg, "adapter_start_layer", 2)
    else:
        n = L - getattr(cfg, "lora_start_layer", 0)
    return 2.0 * (n * linear_params_per_layer(cfg) - qkv + rho * head) + 2.5 * n * 2.0 * T * C


def step_flops(cfg, variant: str, B: int, T: int, training: bool = True, rho: float = 1.0) -> float:
    f = forward_flops_per_token(cfg, T, rho)
    if training:
        f += backward_flops_per_token(cfg, variant, T, rho)
    return f * B * T
