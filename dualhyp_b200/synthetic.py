"""Deterministic synthetic weights, prompts and feature frames.

There is no network for checkpoints or the DualHyp N-best dataset, so every parity test and every
bench run is driven by inputs generated here.  Values come from a counter-based integer hash
(splitmix64 finaliser) evaluated with int64 tensor arithmetic, so the *same bits* are produced on the
CPU and on the GPU, in this container and on the GPU box, independent of torch's RNG streams.

Prompt structure mirrors the reference collate (`data/av_dataset.py:258-292`): each sample is a
prompt span (labels -1), a response span (labels = ids) and right padding (id 0, label -1).
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import torch

Tensor = torch.Tensor

_M1 = -0x61C8864680B583EB            # 0x9E3779B97F4A7C15 as signed int64
_M2 = -0x40A7B892E31B1A47            # 0xBF58476D1CE4E5B9
_M3 = -0x6B2FB644ECCEEE15            # 0x94D049BB133111EB


def _lsr(x: Tensor, n: int) -> Tensor:
    """Logical shift right on int64 (torch's >> is arithmetic)."""
    return (x >> n) & ((1 << (64 - n)) - 1)


def _mix(z: Tensor) -> Tensor:
    z = (z ^ _lsr(z, 30)) * _M2
    z = (z ^ _lsr(z, 27)) * _M3
    return z ^ _lsr(z, 31)


def hash_u24(n: int, seed: int, stream: int, device="cpu", chunk: int = 1 << 26) -> Tensor:
    """n integers in [0, 2^24), int64, from (seed, stream, counter)."""
    out = torch.empty(n, dtype=torch.int64, device=device)
    key = (seed * 0x100000001B3 + stream * 0x1000193 + 0x632BE59BD9B4E019) & 0x7FFFFFFFFFFFFFFF
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        ctr = torch.arange(s, e, dtype=torch.int64, device=device)
        z = _mix(ctr * _M1 + key)
        out[s:e] = _lsr(z, 40)
    return out


def hash_uniform(shape, seed: int, stream: int, bound: float = 1.0, device="cpu", bf16_exact: bool = True) -> Tensor:
    """fp32 tensor, uniform in [-bound, bound); rounded to bf16-representable values by default so the
    bf16 device copy of a weight is bit-identical to what the fp32 oracle multiplies with."""
    n = int(math.prod(shape)) if len(shape) else 1
    u = hash_u24(n, seed, stream, device).to(torch.float32) * (1.0 / (1 << 24)) - 0.5
    w = (u * (2.0 * bound)).reshape(shape)
    return w.bfloat16().float() if bf16_exact else w


def hash_ints(shape, seed: int, stream: int, lo: int, hi: int, device="cpu") -> Tensor:
    n = int(math.prod(shape)) if len(shape) else 1
    return (hash_u24(n, seed, stream, device) % (hi - lo) + lo).reshape(shape)


_SQRT3 = math.sqrt(3.0)
PEAKY_MIX = 0.12
PEAKY_A, PEAKY_B = 7919, 104729          # the `peaky` head permutation v -> (A v + B) mod V (A odd and not a multiple of 5: a bijection for V = 2^a 5^b ...)


def synth_state_dict(cfg, variant: str, seed: int = 1337, device="cpu", lora_b_std: float = 0.02,
                     gate_std: float = 0.5, dtype=torch.float32, extra_wte_rows: int = 0, peaky: bool = False) -> Dict[str, Tensor]:
    """A full state dict with the reference's key names for `variant` in
    {"base", "lora", "adapter", "relprompt"}.  `cfg` needs n_layer, n_head, n_query_groups, head_size,
    n_embd, intermediate_size, padded_vocab_size and the PEFT fields of the variant.

    Distributions follow PyTorch's defaults in spirit (what `empty_init=False` yields in the reference,
    finetune/ger.py:115): Linear ~ U(+-1/sqrt(fan_in)), Embedding unit variance, norms = 1; then, so the
    PEFT paths are live rather than zero-initialised: lora_B ~ U with std `lora_b_std`, gating_factor
    ~ U with std `gate_std`.

    `peaky`: a random-init decoder has near-uniform logits, so greedy ids flip on bf16 noise.  With peaky=True the output head is
    tied to a PERMUTATION of the embedding (lm_head row v = wte row (PEAKY_A * v + PEAKY_B) mod V, scaled by 1 / sqrt(C)): the
    residual stream is dominated by the current token's embedding, so the arg-max is the permutation's pre-image of that token
    with a top-1 / top-2 margin far above bf16 noise, and a greedy decode walks a non-trivial, fully determined id sequence
    through every layer, the KV cache and the adapter / LoRA paths ("make logits peaky")."""
    sd: Dict[str, Tensor] = {}
    C, I, V = cfg.n_embd, cfg.intermediate_size, cfg.padded_vocab_size
    qkv_out = (cfg.n_head + 2 * cfg.n_query_groups) * cfg.head_size
    lora_keys = variant in ("lora", "relprompt")
    lin = ".linear.weight" if lora_keys else ".weight"
    stream = [0]

    def u(shape, bound):
        stream[0] += 1
        return hash_uniform(shape, seed, stream[0], bound, device).to(dtype)

    def linear(out_f, in_f):
        return u((out_f, in_f), 1.0 / math.sqrt(in_f))

    # RelPrompt grows wte by 3 rows for <<C>>/<<M>>/<<N>> while lm_head stays at V (ger/relprompt.py:215-230)
    sd["transformer.wte.weight"] = u((V + extra_wte_rows, C), _SQRT3)
    r = getattr(cfg, "r", 0)
    for i in range(cfg.n_layer):
        p = f"transformer.h.{i}."
        sd[p + "norm_1.weight"] = torch.ones(C, device=device, dtype=dtype) + u((C,), 0.25)
        sd[p + "attn.attn" + lin] = linear(qkv_out, C)
        sd[p + "attn.proj" + lin] = linear(C, C)
        sd[p + "norm_2.weight"] = torch.ones(C, device=device, dtype=dtype) + u((C,), 0.25)
        sd[p + "mlp.fc_1" + lin] = linear(I, C)
        sd[p + "mlp.fc_2" + lin] = linear(I, C)
        sd[p + "mlp.proj" + lin] = linear(C, I)
        if lora_keys and r > 0 and i >= getattr(cfg, "lora_start_layer", 0):
            if cfg.to_query or cfg.to_key or cfg.to_value:
                n_en = int(cfg.to_query) + int(cfg.to_key) + int(cfg.to_value)
                kv = C // (cfg.n_head // cfg.n_query_groups)
                n_b = C * int(cfg.to_query) + kv * int(cfg.to_key) + kv * int(cfg.to_value)
                sd[p + "attn.attn.lora_A"] = linear(r * n_en, C)
                sd[p + "attn.attn.lora_B"] = u((n_b, r), lora_b_std * _SQRT3)
            if cfg.to_projection:
                sd[p + "attn.proj.lora_A"] = linear(r, C)
                sd[p + "attn.proj.lora_B"] = u((C, r), lora_b_std * _SQRT3)
        if variant == "adapter" and i >= cfg.adapter_start_layer:
            sd[p + "attn.adapter_wte.weight"] = u((cfg.adapter_prompt_length, C), _SQRT3)
            sd[p + "attn.gating_factor"] = u((1, 1, cfg.n_head, 1), gate_std * _SQRT3)
    sd["transformer.ln_f.weight"] = torch.ones(C, device=device, dtype=dtype) + u((C,), 0.25)
    if lora_keys:
        sd["lm_head.linear.weight"] = linear(V, C)
        sd["lm_head.adapter_bias"] = torch.zeros(V, device=device, dtype=dtype)
        sd["lm_head.adapter_scale"] = torch.ones(V, device=device, dtype=dtype)
    else:
        sd["lm_head.weight"] = linear(V, C)
    if peaky:
        perm = (PEAKY_A * torch.arange(V, device=device, dtype=torch.int64) + PEAKY_B) % V
        key = "lm_head.linear.weight" if lora_keys else "lm_head.weight"
        head = sd["transformer.wte.weight"][:V].index_select(0, perm) / math.sqrt(C)
        # a blend with the random head keeps the other logits (and the layers' contribution to them) alive: the margin stays
        # ~10x the bf16 noise of the logit scale instead of ~90x
        sd[key] = (PEAKY_MIX * head + (1.0 - PEAKY_MIX) * sd[key]).contiguous()
    if variant == "relprompt":
        for name, dim in (("audio_noise_classifier", cfg.whisper_dim), ("visual_noise_classifier", cfg.raven_dim)):
            q = name + "."
            sd[q + "conv1.weight"] = u((256, dim, 3), 1.0 / math.sqrt(3 * dim))
            sd[q + "conv1.bias"] = u((256,), 1.0 / math.sqrt(3 * dim))
            sd[q + "conv2.weight"] = u((256, 256, 3), 1.0 / math.sqrt(3 * 256))
            sd[q + "conv2.bias"] = u((256,), 1.0 / math.sqrt(3 * 256))
            sd[q + "classifier.weight"] = u((3, 256), 1.0 / math.sqrt(256))
            sd[q + "classifier.bias"] = u((3,), 1.0 / math.sqrt(256))
    return sd


def synth_batch(B: int, T: int, vocab: int, seed: int = 1337, prompt_range: Tuple[int, int] = (330, 480),
                response_range: Tuple[int, int] = (8, 24), extra_token_ids: Optional[Tuple[int, ...]] = None,
                device="cpu") -> Tuple[Tensor, Tensor]:
    """(input_ids int64 [B,T], labels int64 [B,T]).  DualHyp default spans: prompt U[330,480], response
    U[8,24] of T=512 (5 ASR + 5 VSR hypotheses, data/prompts.py:9-13); GER (cfg1): pass
    prompt_range=(150,230) with T=256.  `extra_token_ids` (RelPrompt <<C>>/<<M>>/<<N>> = vocab..vocab+2)
    are sprinkled into the prompt span."""
    ids = hash_ints((B, T), seed, 101, 0, vocab)
    plo, phi = prompt_range
    rlo, rhi = response_range
    plen = hash_ints((B,), seed, 102, plo, phi + 1)
    rlen = hash_ints((B,), seed, 103, rlo, rhi + 1)
    pos = torch.arange(T).unsqueeze(0)
    plen = torch.clamp(plen, max=T - 2).unsqueeze(1)
    end = torch.clamp(plen + rlen.unsqueeze(1), max=T)
    in_resp = (pos >= plen) & (pos < end)
    pad = pos >= end
    if extra_token_ids:
        sel = hash_ints((B, T), seed, 104, 0, 8) == 0
        which = hash_ints((B, T), seed, 105, 0, len(extra_token_ids))
        extra = torch.tensor(extra_token_ids, dtype=torch.int64)[which]
        ids = torch.where(sel & (pos < plen), extra, ids)
    ids = torch.where(pad, torch.zeros_like(ids), ids)
    labels = torch.where(in_resp, ids, torch.full_like(ids, -1))
    return ids.to(device), labels.to(device)


def synth_prompts(n: int, vocab: int, seed: int = 1337, prompt_range: Tuple[int, int] = (330, 480)):
    """n variable-length prompts (list of 1-D int64 tensors) for the decode workload (cfg5)."""
    lens = hash_ints((n,), seed, 201, prompt_range[0], prompt_range[1] + 1).tolist()
    flat = hash_ints((sum(lens),), seed, 202, 3, vocab)          # avoid 0/1/2 (pad/bos/eos) inside prompts
    out, o = [], 0
    for L in lens:
        out.append(flat[o:o + L].clone())
        o += L
    return out


def synth_features(B: int, frames: int, dim: int, seed: int, stream: int, device="cpu") -> Tensor:
    """Unit-variance fp32 encoder-feature frames (stand-in for Whisper / BRAVEn encoder outputs,
    finetune/relprompt.py:347-351)."""
    return hash_uniform((B, frames, dim), seed, stream, _SQRT3, device, bf16_exact=False)
