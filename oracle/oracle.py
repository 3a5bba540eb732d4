"""TEST INFRASTRUCTURE ONLY - CPU restatement of the DualHyp hot path (the oracle).

This file restates, in plain fp32 torch tensor arithmetic on the CPU, what the reference computes on
its decoder hot path.  It is the checker for the CUDA product path; it is never the thing measured or
shipped.  Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import it.  The product package (`dualhyp_b200/`) never imports anything from `oracle/`.

Parity status: **pinned** against outputs of the reference itself.  The reference ships no tests or
golden vectors, so `oracle/make_golden.py` imports the unmodified reference in the
build container (through `oracle/refshim.py`), runs it on deterministic hash-generated weights and
token ids, and commits the results under `tests/golden/`.  `tests/test_oracle_golden.py` checks this
restatement against those fixtures everywhere, and against the live reference where it is mounted.

Every function cites the reference file:line it follows (paths relative to `/root/reference`).
All functions are functional: weights come in as a flat ``dict[str, Tensor]`` with the reference's
state-dict key names.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
StateDict = Dict[str, Tensor]


@dataclass
class OracleConfig:
    """The subset of ger/config.py:16-92 the hot path reads, plus the PEFT fields of
    ger/lora.py:459-468, ger/adapter.py:23-24 and ger/relprompt.py:163-175."""

    n_layer: int = 22
    n_head: int = 32
    n_query_groups: int = 4
    head_size: int = 64
    n_embd: int = 2048
    intermediate_size: int = 5632
    padded_vocab_size: int = 32000
    block_size: int = 2048
    norm_eps: float = 1e-5
    # variant: "base" | "lora" | "adapter" | "relprompt"
    variant: str = "lora"
    # LoRA
    r: int = 0
    alpha: int = 1
    to_query: bool = False
    to_key: bool = False
    to_value: bool = False
    to_projection: bool = False
    lora_start_layer: int = 0
    # adapter
    adapter_prompt_length: int = 10
    adapter_start_layer: int = 2
    # relprompt
    whisper_dim: int = 1280
    raven_dim: int = 1024
    pool_size: int = 10

    @property
    def q_per_kv(self) -> int:
        return self.n_head // self.n_query_groups

    @property
    def qkv_out(self) -> int:
        return (self.n_head + 2 * self.n_query_groups) * self.head_size


# --------------------------------------------------------------------------------------------
# leaf ops
# --------------------------------------------------------------------------------------------

def rmsnorm(x: Tensor, weight: Tensor, eps: float) -> Tensor:
    """ger/rmsnorm.py:17-21 -- mean(x*x) in the input dtype, rsqrt, scale by weight."""
    norm_x = torch.mean(x * x, dim=-1, keepdim=True)
    return weight * (x * torch.rsqrt(norm_x + eps))


def build_rope_cache(seq_len: int, n_elem: int, base: int = 10000) -> Tuple[Tensor, Tensor]:
    """ger/model.py:319-346 with the call site's hard-wired dtype=bfloat16 (ger/model.py:120-126):
    the tables are rounded to bf16 in *every* precision mode; `rope_base` from the
    config is ignored (quirk Q12)."""
    theta = 1.0 / (base ** (torch.arange(0, n_elem, 2) / n_elem))
    seq_idx = torch.arange(seq_len) / 1
    idx_theta = torch.outer(seq_idx, theta).repeat(1, 2)
    return torch.cos(idx_theta).bfloat16(), torch.sin(idx_theta).bfloat16()


def apply_rope(x: Tensor, cos: Tensor, sin: Tensor) -> Tensor:
    """ger/model.py:349-355 -- NeoX half rotation; result cast back to x's dtype."""
    hs = x.size(-1)
    x1, x2 = x[..., : hs // 2], x[..., hs // 2:]
    rotated = torch.cat((-x2, x1), dim=-1)
    return ((x * cos) + (rotated * sin)).type_as(x)


# nn.Dropout on the LoRA branch input (ger/lora.py:94-98,165,373): the reference draws the mask from torch's generator.  To compare
# the CUDA path (Philox masks) with this restatement on the SAME mask, a test may install the multiplicative masks here:
# {(layer, "attn" | "proj"): tensor broadcastable to the branch input, already scaled by 1 / (1 - p)}.  None = identity (p = 0 / eval).
LORA_DROPOUT_MASKS: Optional[Dict[Tuple[int, str], Tensor]] = None


def _dropped(x: Tensor, layer: Optional[int], which: str) -> Tensor:
    if LORA_DROPOUT_MASKS is None or layer is None or (layer, which) not in LORA_DROPOUT_MASKS:
        return x
    return x * LORA_DROPOUT_MASKS[(layer, which)].view_as(x)


def lora_linear(x: Tensor, W: Tensor, A: Optional[Tensor], B: Optional[Tensor], scaling: float, layer: Optional[int] = None) -> Tensor:
    """ger/lora.py:159-166: pretrained(x) + (dropout(x) @ A^T @ B^T) * scaling."""
    pre = F.linear(x, W)
    if A is None:
        return pre
    return pre + (_dropped(x, layer, "proj") @ A.transpose(0, 1) @ B.transpose(0, 1)) * scaling


def lora_qkv_linear(x: Tensor, W: Tensor, A: Optional[Tensor], B: Optional[Tensor], scaling: float,
                    cfg: OracleConfig, layer: Optional[int] = None) -> Tensor:
    """ger/lora.py:367-402 with all of q,k,v enabled and n_head != n_query_groups (the split path of
    `conv1d`, :343-347).  The delta columns are laid out [q | k | v] contiguously (qkv_shapes,
    :228-234) and are added as-is onto the *group-interleaved* pretrained output (quirk Q3)."""
    pre = F.linear(x, W)
    if A is None:
        return pre
    r = A.size(0) // 3
    after_A = F.linear(_dropped(x, layer, "attn"), A)          # (..., 3r); ger/lora.py:373 `self.lora_dropout(x)`
    kv = cfg.n_embd // cfg.q_per_kv                            # kv_embd_size, ger/lora.py:226
    shapes = (cfg.n_embd, kv, kv)
    parts = []
    off = 0
    for i, s in enumerate(shapes):                             # 1x1 conv == per-part linear
        parts.append(F.linear(after_A[..., i * r:(i + 1) * r], B[off:off + s]))
        off += s
    return pre + torch.cat(parts, dim=-1) * scaling


def split_qkv(qkv: Tensor, cfg: OracleConfig) -> Tuple[Tensor, Tensor, Tensor]:
    """ger/model.py:216-231 -- group-interleaved rows [q_per_kv q-heads | k | v] per query group;
    k, v expanded to n_head heads (quirk Q4; mathematically a broadcast)."""
    B, T, _ = qkv.shape
    qpk = cfg.q_per_kv
    qkv = qkv.view(B, T, cfg.n_query_groups, qpk + 2, cfg.head_size).permute(0, 2, 3, 1, 4)
    q, k, v = qkv.split((qpk, 1, 1), dim=2)
    k = k.expand(B, cfg.n_query_groups, qpk, T, cfg.head_size)
    v = v.expand(B, cfg.n_query_groups, qpk, T, cfg.head_size)
    q = q.reshape(B, -1, T, cfg.head_size)
    k = k.reshape(B, -1, T, cfg.head_size)
    v = v.reshape(B, -1, T, cfg.head_size)
    return q, k, v


def sdpa(q: Tensor, k: Tensor, v: Tensor, mask: Optional[Tensor], scale: float) -> Tensor:
    """ger/model.py:270-290, CPU branch: softmax(q k^T * scale [+mask]) v, causal iff mask is None.
    Written out explicitly (no fused kernel) so every step is visible.  Returns (B, T, nh, hs)."""
    att = (q @ k.transpose(-2, -1)) * scale
    Tq, Tk = q.size(-2), k.size(-2)
    if mask is None:
        causal = torch.ones(Tq, Tk, dtype=torch.bool).tril()
        att = att.masked_fill(~causal, float("-inf"))
    else:
        att = att.masked_fill(~mask, float("-inf"))
    y = torch.softmax(att, dim=-1) @ v
    return y.transpose(1, 2)


def adapter_prefix_attention(q: Tensor, prefix: Tensor, W_qkv: Tensor, gating: Tensor, cfg: OracleConfig) -> Tensor:
    """ger/adapter.py:129-146 -- K/V of the learned prefix through the *same* qkv linear, no RoPE,
    its own softmax over the aT prefix keys (all-ones mask), gate (1,1,nh,1) without tanh (quirk Q9).
    Returns gating * ay in (B, T, nh, hs)."""
    aT = cfg.adapter_prompt_length
    aqkv = F.linear(prefix.reshape(1, aT, cfg.n_embd), W_qkv)
    qpk = cfg.q_per_kv
    aqkv = aqkv.view(1, aT, cfg.n_query_groups, qpk + 2, cfg.head_size).permute(0, 2, 3, 1, 4)
    _, ak, av = aqkv.split((qpk, 1, 1), dim=2)
    ak = ak.repeat_interleave(qpk, dim=2).reshape(1, -1, aT, cfg.head_size)
    av = av.repeat_interleave(qpk, dim=2).reshape(1, -1, aT, cfg.head_size)
    amask = torch.ones(q.size(2), aT, dtype=torch.bool)
    ay = sdpa(q, ak, av, amask, 1.0 / math.sqrt(cfg.head_size))
    return gating * ay


def llama_mlp(x: Tensor, W1: Tensor, W2: Tensor, Wp: Tensor) -> Tensor:
    """ger/model.py:312-316 -- proj(silu(fc_1 x) * fc_2 x)."""
    return F.linear(F.silu(F.linear(x, W1)) * F.linear(x, W2), Wp)


# --------------------------------------------------------------------------------------------
# the decoder
# --------------------------------------------------------------------------------------------

def _w(sd: StateDict, base: str, lora: bool) -> Tensor:
    return sd[base + (".linear.weight" if lora else ".weight")]


def _lora_keys(cfg: OracleConfig) -> bool:
    return cfg.variant in ("lora", "relprompt")


def attention(x: Tensor, sd: StateDict, i: int, cfg: OracleConfig, cos: Tensor, sin: Tensor,
              kv_cache: Optional[Tuple[Tensor, Tensor]] = None, input_pos: Optional[Tensor] = None,
              mask: Optional[Tensor] = None):
    """ger/model.py:202-268 (CausalSelfAttention.forward) with the LoRA linears of ger/lora.py:586-617
    or the adapter prefix of ger/adapter.py:116-146 plugged in according to cfg.variant."""
    p = f"transformer.h.{i}.attn."
    lk = _lora_keys(cfg)
    B, T, C = x.shape
    Wqkv = _w(sd, p + "attn", lk)
    use_lora = lk and cfg.r > 0 and i >= cfg.lora_start_layer
    scaling = cfg.alpha / cfg.r if cfg.r > 0 else 0.0
    if use_lora and (cfg.to_query or cfg.to_key or cfg.to_value):
        assert cfg.to_query and cfg.to_key and cfg.to_value, "oracle restates the all-enabled path only"
        qkv = lora_qkv_linear(x, Wqkv, sd[p + "attn.lora_A"], sd[p + "attn.lora_B"], scaling, cfg, layer=i)
    else:
        qkv = F.linear(x, Wqkv)
    q, k, v = split_qkv(qkv, cfg)
    q = apply_rope(q, cos, sin)          # rotary_percentage == 1.0 => whole head (ger/model.py:240-246)
    k = apply_rope(k, cos, sin)
    if kv_cache is not None:             # ger/model.py:248-259
        ck, cv = kv_cache
        k = ck.index_copy_(2, input_pos, k)
        v = cv.index_copy_(2, input_pos, v)
        kv_cache = (k, v)
    scale = 1.0 / math.sqrt(cfg.head_size)
    y = sdpa(q, k, v, mask, scale)
    if cfg.variant == "adapter" and i >= cfg.adapter_start_layer:
        y = y + adapter_prefix_attention(q, sd[p + "adapter_wte.weight"], Wqkv, sd[p + "gating_factor"], cfg)
    y = y.reshape(B, T, C)
    Wp = _w(sd, p + "proj", lk)
    if use_lora and cfg.to_projection:
        y = lora_linear(y, Wp, sd[p + "proj.lora_A"], sd[p + "proj.lora_B"], scaling, layer=i)
    else:
        y = F.linear(y, Wp)
    return y, kv_cache


def block(x: Tensor, sd: StateDict, i: int, cfg: OracleConfig, cos, sin, kv_cache=None, input_pos=None, mask=None):
    """ger/model.py:163-188, the non-parallel-residual branch (:185-186)."""
    p = f"transformer.h.{i}."
    lk = _lora_keys(cfg)
    h, kv_cache = attention(rmsnorm(x, sd[p + "norm_1.weight"], cfg.norm_eps), sd, i, cfg, cos, sin,
                            kv_cache, input_pos, mask)
    x = x + h
    n2 = rmsnorm(x, sd[p + "norm_2.weight"], cfg.norm_eps)
    x = x + llama_mlp(n2, _w(sd, p + "mlp.fc_1", lk), _w(sd, p + "mlp.fc_2", lk), _w(sd, p + "mlp.proj", lk))
    return x, kv_cache


def lm_head(x: Tensor, sd: StateDict, cfg: OracleConfig) -> Tensor:
    """ger/lora.py:70-71 (AdapterV2Linear when to_head=False) or plain nn.Linear (ger/adapter.py:36)."""
    if _lora_keys(cfg):
        return sd["lm_head.adapter_scale"] * (F.linear(x, sd["lm_head.linear.weight"]) + sd["lm_head.adapter_bias"])
    return F.linear(x, sd["lm_head.weight"])


def hidden_states(sd: StateDict, cfg: OracleConfig, idx: Tensor) -> Tensor:
    """Everything of GPT.forward up to and including ln_f (ger/lora.py:518-545, no KV cache)."""
    B, T = idx.shape
    assert cfg.block_size >= T
    cos, sin = build_rope_cache(cfg.block_size, cfg.head_size)
    cos, sin = cos[:T], sin[:T]
    x = F.embedding(idx, sd["transformer.wte.weight"])
    for i in range(cfg.n_layer):
        x, _ = block(x, sd, i, cfg, cos, sin)
    return rmsnorm(x, sd["transformer.ln_f.weight"], cfg.norm_eps)


def gpt_forward(sd: StateDict, cfg: OracleConfig, idx: Tensor, lm_head_chunk_size: int = 0):
    """ger/lora.py:504-549 / ger/relprompt.py:232-294 (which reduces to the same thing when no cache
    is used) / the patched adapter composition of SURVEY section 8c (ger/model.py:65-114 driving
    ger/adapter.py blocks, chunking as ger/adapter.py:69-72)."""
    x = hidden_states(sd, cfg, idx)
    if lm_head_chunk_size > 0:
        return [lm_head(x_i, sd, cfg) for x_i in x.split(lm_head_chunk_size, dim=1)]
    return lm_head(x, sd, cfg)


def chunked_cross_entropy(logits, targets: Tensor, chunk_size: int = 128) -> Tensor:
    """ger/utils.py:424-463.  List input + chunk_size>0 => sum_valid CE / (all positions) (quirk Q2);
    chunk_size == 0 => mean over valid positions."""
    if isinstance(logits, list):
        if chunk_size == 0:
            logits = torch.cat(logits, dim=1)
            return F.cross_entropy(logits.reshape(-1, logits.size(-1)), targets.reshape(-1), ignore_index=-1)
        lc = [l.reshape(-1, l.size(-1)) for l in logits]
        tc = [t.reshape(-1) for t in targets.split(logits[0].size(1), dim=1)]
        losses = [F.cross_entropy(a, b, ignore_index=-1, reduction="none") for a, b in zip(lc, tc)]
        return torch.cat(losses).mean()
    logits = logits.reshape(-1, logits.size(-1))
    targets = targets.reshape(-1)
    if chunk_size == 0:
        return F.cross_entropy(logits, targets, ignore_index=-1)
    losses = [F.cross_entropy(a, b, ignore_index=-1, reduction="none")
              for a, b in zip(logits.split(chunk_size), targets.split(chunk_size))]
    return torch.cat(losses).mean()


def train_loss(sd: StateDict, cfg: OracleConfig, idx: Tensor, labels: Tensor) -> Tensor:
    """The driver's micro-step, finetune/ger.py:278-281: chunked lm_head, drop the last position,
    shift the targets, chunked CE (train reduction)."""
    logits = gpt_forward(sd, cfg, idx, lm_head_chunk_size=128)
    logits[-1] = logits[-1][..., :-1, :]
    return chunked_cross_entropy(logits, labels[..., 1:])


def val_loss(sd: StateDict, cfg: OracleConfig, idx: Tensor, labels: Tensor) -> Tensor:
    """finetune/ger.py:345-346: full logits, mean over valid targets."""
    logits = gpt_forward(sd, cfg, idx)
    return chunked_cross_entropy(logits[..., :-1, :], labels[..., 1:], chunk_size=0)


def trainable_keys(sd: StateDict, cfg: OracleConfig) -> List[str]:
    """ger/lora.py:405-439, ger/relprompt.py:79-119, ger/adapter.py:158-165."""
    if cfg.variant == "adapter":
        return [k for k in sd if "adapter_wte" in k or "gating_factor" in k]
    if cfg.variant == "relprompt":
        return [k for k in sd if "lora_" in k or "noise_classifier" in k]
    return [k for k in sd if "lora_" in k]


def loss_and_grads(sd: StateDict, cfg: OracleConfig, idx: Tensor, labels: Tensor, reduction: str = "train"):
    """Loss and d(loss)/d(trainable) by torch autograd over this restatement."""
    keys = trainable_keys(sd, cfg)
    work = {k: (v.detach().clone().requires_grad_(True) if k in keys else v.detach()) for k, v in sd.items()}
    loss = (train_loss if reduction == "train" else val_loss)(work, cfg, idx, labels)
    grads = torch.autograd.grad(loss, [work[k] for k in keys]) if keys else ()
    return loss.detach(), dict(zip(keys, (g.detach() for g in grads)))


# --------------------------------------------------------------------------------------------
# greedy decode with the KV cache
# --------------------------------------------------------------------------------------------

@torch.no_grad()
def generate_greedy(sd: StateDict, cfg: OracleConfig, prompt: Tensor, max_returned_tokens: int,
                    eos_id: Optional[int] = None, return_logits: bool = False):
    """generate/base.py:19-82 with the reference's temperature>0, top_k=1 setting, which keeps only
    the arg-max logit before `multinomial` (:65-70) => arg-max decoding (quirk Q7: exact ties are
    broken towards the lowest index here).  KV cache as ger/model.py:132-151,248-259: caches span
    max_seq_length (= block_size) slots of n_head *expanded* heads and attention runs over all slots
    under the tril mask row (ger/lora.py:523-531)."""
    T = prompt.size(0)
    assert max_returned_tokens > T
    assert cfg.block_size >= max_returned_tokens - 1
    S = cfg.block_size
    cos_all, sin_all = build_rope_cache(S, cfg.head_size)
    mask_cache = torch.ones(S, S, dtype=torch.bool).tril()[None, None]
    caches = [(torch.zeros(1, cfg.n_head, S, cfg.head_size), torch.zeros(1, cfg.n_head, S, cfg.head_size))
              for _ in range(cfg.n_layer)]
    idx = torch.empty(max_returned_tokens, dtype=prompt.dtype)
    idx[:T] = prompt
    input_pos = torch.arange(0, T)
    step_logits = []
    for _ in range(max_returned_tokens - T):
        x_ids = idx.index_select(0, input_pos).view(1, -1)
        cos, sin = cos_all.index_select(0, input_pos), sin_all.index_select(0, input_pos)
        mask = mask_cache.index_select(2, input_pos)
        x = F.embedding(x_ids, sd["transformer.wte.weight"])
        for i in range(cfg.n_layer):
            x, caches[i] = block(x, sd, i, cfg, cos, sin, caches[i], input_pos, mask)
        x = rmsnorm(x, sd["transformer.ln_f.weight"], cfg.norm_eps)
        logits = lm_head(x, sd, cfg)[0, -1]
        if return_logits:
            step_logits.append(logits.clone())
        nxt = torch.argmax(logits)
        input_pos = input_pos[-1:] + 1
        idx[input_pos] = nxt
        if eos_id is not None and int(nxt) == eos_id:
            out = idx[: int(input_pos) + 1]
            return (out, step_logits) if return_logits else out
    return (idx, step_logits) if return_logits else idx


# --------------------------------------------------------------------------------------------
# RelPrompt reliability predictors
# --------------------------------------------------------------------------------------------

def noise_mask_classifier(x: Tensor, sd: StateDict, prefix: str, pool_size: int, dropout_mask: Optional[Tensor] = None) -> Tensor:
    """ger/relprompt.py:126-147.  x (B, T, C) -> logits (B, ceil(T/pool), 3).  `dropout_mask` (B, T, 256), already scaled by
    1 / (1 - p): the nn.Dropout between the convolutions (:141) with a mask supplied by the test; None = identity (eval).
    AvgPool1d(ceil_mode=True): the last, partial window averages over its valid frames only."""
    h = x.transpose(1, 2)
    h = F.relu(F.conv1d(h, sd[prefix + "conv1.weight"], sd[prefix + "conv1.bias"], padding=1))
    if dropout_mask is not None:
        h = h * dropout_mask.transpose(1, 2)
    h = F.relu(F.conv1d(h, sd[prefix + "conv2.weight"], sd[prefix + "conv2.bias"], padding=1))
    B, C, T = h.shape
    n_out = -(-T // pool_size)
    pooled = []
    for j in range(n_out):
        pooled.append(h[:, :, j * pool_size:min(T, (j + 1) * pool_size)].mean(dim=2))
    h = torch.stack(pooled, dim=1)                              # (B, n_out, C)
    return F.linear(h, sd[prefix + "classifier.weight"], sd[prefix + "classifier.bias"])


def mask_loss(audio_logits: Tensor, visual_logits: Tensor, audio_targets: Tensor, visual_targets: Tensor) -> Tensor:
    """finetune/relprompt.py:360-387 -- trim logits/targets to the common length, CE(audio)+CE(visual)."""
    def trim(lg, tg):
        n = min(lg.size(1), tg.size(1))
        return lg[:, :n, :], tg[:, :n]
    al, at = trim(audio_logits, audio_targets)
    vl, vt = trim(visual_logits, visual_targets)
    return F.cross_entropy(al.reshape(-1, 3), at.reshape(-1)) + F.cross_entropy(vl.reshape(-1, 3), vt.reshape(-1))


def relprompt_loss(sd: StateDict, cfg: OracleConfig, idx, labels, audio_feats, visual_feats, audio_t, visual_t,
                   mask_loss_weight: float = 0.02) -> Tuple[Tensor, Tensor, Tensor]:
    """finetune/relprompt.py:356-403: llm_loss + weight * mask_loss."""
    al = noise_mask_classifier(audio_feats, sd, "audio_noise_classifier.", 2 * cfg.pool_size)
    vl = noise_mask_classifier(visual_feats, sd, "visual_noise_classifier.", cfg.pool_size)
    ml = mask_loss(al, vl, audio_t, visual_t)
    llm = train_loss(sd, cfg, idx, labels)
    return llm + mask_loss_weight * ml, llm, ml
