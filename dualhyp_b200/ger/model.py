"""`ger.model` mirror (reference ger/model.py): GPT, Block, CausalSelfAttention, LLaMAMLP, build_rope_cache,
apply_rope -- same names, constructor arguments, parameter names (state-dict keys) and forward()
signatures, with every device op executed by libdualhyp.so through `dualhyp_b200.engine`.

The whole decoder is one autograd node (`_DecoderFn`): its inputs are the few trainable tensors
(LoRA A/B, adapter prefix/gate), its output is the ln_f hidden state; the frozen 1.1 B weights never enter
autograd.  lm_head (+ cross entropy) is a second node so that callers who only need a loss never
materialise [B, T, V] logits (`LazyLogits`, recognised by `ger.utils.chunked_cross_entropy`).
"""
from __future__ import annotations

import math
from typing import Any, Dict, List, Optional, Tuple, Union

import torch
import torch.nn as nn

from .. import ops
from ..engine import DecoderEngine, EngineSpec
from .config import Config

RoPECache = Tuple[torch.Tensor, torch.Tensor]
KVCache = Tuple[torch.Tensor, torch.Tensor]
Tensor = torch.Tensor


# ---------------------------------------------------------------------------------------------------
# autograd nodes
# ---------------------------------------------------------------------------------------------------
class _DecoderFn(torch.autograd.Function):
    """idx -> ln_f hidden [B, T, C] bf16; differentiable w.r.t. the trainable PEFT tensors only."""

    @staticmethod
    def forward(ctx, model: "GPT", idx: Tensor, varlen: Optional[Dict], names: Tuple[str, ...], *trainables: Tensor):
        eng = model._engine_ready()
        packed = eng.pack_trainable(dict(zip(names, trainables)))
        ctx.model, ctx.names, ctx.unpad = model, names, None
        B, T = idx.shape
        if varlen is not None and "row_map" in varlen:
            # a right-padded [B, T] batch executed without its pad rows (GPT.skip_pad): gather the real tokens into one packed
            # row, run the decoder over it, scatter the hidden states back to [B, T] (pad rows = 0; they carry no loss)
            rows, count = varlen["row_map"], varlen["count"]
            ids = idx.reshape(-1).index_select(0, varlen["row_map_long"]).view(1, -1)
            hidden_p = eng.forward_hidden(ids, packed, save=True, varlen=varlen, dropout_seed=model._dropout_seed())
            hidden = eng._cap_buf("hidden_unpacked", (B * T, hidden_p.shape[1]), torch.bfloat16)
            ops.scatter_rows(hidden_p, rows, count, hidden)
            ctx.unpad = (rows, count, hidden_p.shape[0])
        else:
            hidden = eng.forward_hidden(idx, packed, save=True, varlen=varlen, dropout_seed=model._dropout_seed())
        ctx.step_id = eng.step_id
        return hidden.view(B, T, -1)

    @staticmethod
    def backward(ctx, d_hidden: Tensor):
        model = ctx.model
        eng = model._engine
        d2 = d_hidden.reshape(-1, d_hidden.shape[-1]).to(torch.bfloat16).contiguous()
        if ctx.unpad is not None:
            rows, count, Mp = ctx.unpad
            d2p = eng._cap_buf("d_hidden_packed", (Mp, d2.shape[1]), torch.bfloat16)
            ops.gather_rows(d2, rows, count, d2p)
            d2 = d2p
        grads = eng.backward(d2, ctx.step_id)
        scale = model._pending_grad_scale
        model._pending_grad_scale = None
        if scale is not None:
            # the engine hands out views of a few batched buffers (all gates, all prefixes, one tensor per LoRA wgrad): scale
            # each underlying buffer once, in place, instead of one elementwise launch per parameter
            bases = {}
            for g in grads.values():
                b = g._base if g._base is not None else g
                bases[id(b)] = b
            torch._foreach_mul_(list(bases.values()), scale)
        return (None, None, None, None, *[grads.get(n) for n in ctx.names])


class _LMHeadFn(torch.autograd.Function):
    """hidden [..., C] -> logits [..., V] (materialised)."""

    @staticmethod
    def forward(ctx, model: "GPT", hidden: Tensor, out_dtype: torch.dtype):
        eng = model._engine_ready()
        h2 = hidden.reshape(-1, hidden.shape[-1]).to(torch.bfloat16).contiguous()
        logits = torch.empty((h2.shape[0], eng.spec.vocab), device=h2.device, dtype=torch.bfloat16)
        eng.lm_head(h2, out=logits)
        ctx.model = model
        return logits.view(*hidden.shape[:-1], -1).to(out_dtype)

    @staticmethod
    def backward(ctx, d_logits: Tensor):
        eng = ctx.model._engine
        d2 = d_logits.reshape(-1, d_logits.shape[-1]).to(torch.bfloat16).contiguous()
        d_hidden = eng.lm_head_backward(d2)
        return None, d_hidden.view(*d_logits.shape[:-1], -1).clone(), None


class _LMHeadCEFn(torch.autograd.Function):
    """Fused vocab projection + cross entropy: hidden [M, C], targets [M] (-1 = ignore) -> scalar loss.
    denom > 0: loss = sum(row losses) / denom (the train reduction divides by ALL positions, quirk Q2);
    denom == 0: mean over valid targets (ger/utils.py:451-454)."""

    @staticmethod
    def forward(ctx, model: "GPT", hidden: Tensor, targets: Tensor, denom: float):
        eng = model._engine_ready()
        h2 = hidden.reshape(-1, hidden.shape[-1]).to(torch.bfloat16).contiguous()
        M, C = h2.shape
        need_grad = bool(ctx.needs_input_grad[1])
        tg = targets.reshape(-1).contiguous()
        # Only rows with a target != -1 carry loss or gradient (the response span: a few percent of the prompt), so only
        # they go through the vocabulary projection: ordered compaction on the device, GEMMs bounded by the device-side
        # count -- the host never learns it, there is no sync.
        count = eng._buf("ce_count", (1,), torch.int32)
        rows = eng._cap_buf("ce_row_idx", (M,), torch.int32)
        tg_c = eng._cap_buf("ce_targets_c", (M,), torch.int64)
        h_c = eng._cap_buf("ce_hidden_c", (M, C), torch.bfloat16)
        ops.compact_targets(tg, count, rows, tg_c)
        ops.gather_rows(h2, rows, count, h_c)
        logits = eng.lm_head(h_c, m_limit=count)
        row = eng._cap_buf("ce_rows", (M,), torch.float32)
        out2 = torch.empty(2, device=h2.device, dtype=torch.float32)
        ops.cross_entropy(logits, tg_c, row, (1.0 / denom) if denom > 0 else 1.0, 2 if need_grad else 0)
        ops.loss_reduce(row, tg_c, out2)
        ctx.model, ctx.shape, ctx.denom = model, hidden.shape, denom
        ctx.save_for_backward(out2)
        ctx.logits = logits if need_grad else None
        ctx.compact = (rows, count)
        return out2[0] / denom if denom > 0 else out2[0] / out2[1]

    @staticmethod
    def backward(ctx, d_loss: Tensor):
        model = ctx.model
        eng = model._engine
        rows, count = ctx.compact
        d_hidden_c = eng.lm_head_backward(ctx.logits, m_limit=count, name="d_hidden_c")
        d_hidden = eng._cap_buf("d_hidden", tuple(d_hidden_c.shape), torch.bfloat16)
        ops.scatter_rows(d_hidden_c, rows, count, d_hidden)
        (out2,) = ctx.saved_tensors
        scale = d_loss.float() if ctx.denom > 0 else d_loss.float() / out2[1]
        # the scalar is applied to the (small) final gradients by _DecoderFn.backward, in fp32
        model._pending_grad_scale = scale
        return None, d_hidden.view(ctx.shape), None, None


class LazyLogits:
    """Logits that have not been computed: a view [.., V] over ln_f hidden states.  Slicing the leading
    dims stays lazy (`logits[-1][..., :-1, :]`, `logits[0, -1]`); anything else materialises through the
    lm_head GEMM.  `ger.utils.chunked_cross_entropy` consumes these without ever building [B, T, V]."""

    def __init__(self, model: "GPT", hidden: Tensor, base: Optional[Tensor] = None, t0: int = 0, t1: Optional[int] = None,
                 packed_meta: Optional[Tuple[int, int]] = None):
        self.model, self.hidden = model, hidden
        self.packed_meta = packed_meta          # (n sequences, longest) of a packed batch: the loss denominator follows it
        self.base = base if base is not None else (hidden if hidden.dim() == 3 else None)   # [B, T, C] it was cut from
        self.t0 = t0
        self.t1 = t1 if t1 is not None else (hidden.shape[1] if hidden.dim() == 3 else None)

    # -- tensor-like surface -------------------------------------------------------------------
    @property
    def shape(self):
        return torch.Size((*self.hidden.shape[:-1], self.model.config.padded_vocab_size))

    def size(self, dim: Optional[int] = None):
        return self.shape if dim is None else self.shape[dim]

    def dim(self):
        return self.hidden.dim()

    @property
    def device(self):
        return self.hidden.device

    @property
    def dtype(self):
        return self.model._logits_dtype()

    def __len__(self):
        return self.shape[0]

    def __getitem__(self, item):
        if not isinstance(item, tuple):
            item = (item,)
        nd = self.hidden.dim()
        if Ellipsis in item:
            k = item.index(Ellipsis)
            item = item[:k] + (slice(None),) * (nd - (len(item) - 1)) + item[k + 1:]
        item = item + (slice(None),) * (nd - len(item))
        if len(item) == nd and item[-1] == slice(None):
            lead = item[:-1]
            if (self.base is not None and nd == 3 and lead[0] == slice(None) and isinstance(lead[1], slice)
                    and lead[1].step in (None, 1)):
                a, b, _ = lead[1].indices(self.t1 - self.t0)
                return LazyLogits(self.model, self.hidden[:, a:b], self.base, self.t0 + a, self.t0 + b, packed_meta=self.packed_meta)
            return LazyLogits(self.model, self.hidden[lead], None, packed_meta=self.packed_meta)
        return self.materialize()[item]

    def materialize(self) -> Tensor:
        return _LMHeadFn.apply(self.model, self.hidden, self.model._logits_dtype())

    def __truediv__(self, other):
        return self.materialize() / other

    def __mul__(self, other):
        return self.materialize() * other

    def float(self):
        return self.materialize().float()

    def reshape(self, *a):
        return self.materialize().reshape(*a)

    def view(self, *a):
        return self.materialize().view(*a)

    def argmax(self, *a, **k):
        return self.materialize().argmax(*a, **k)

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        def conv(x):
            if isinstance(x, LazyLogits):
                return x.materialize()
            if isinstance(x, (list, tuple)):
                return type(x)(conv(y) for y in x)
            return x
        return func(*conv(args), **{k: conv(v) for k, v in (kwargs or {}).items()})


# ---------------------------------------------------------------------------------------------------
# modules
# ---------------------------------------------------------------------------------------------------
class GPT(nn.Module):
    """Reference ger/model.py:23-151.  Sub-module and parameter names match the reference's state dict."""

    _variant = "base"

    def __init__(self, config: Config) -> None:
        super().__init__()
        assert config.padded_vocab_size is not None
        self.config = config
        self.lm_head = nn.Linear(config.n_embd, config.padded_vocab_size, bias=False)
        self.transformer = nn.ModuleDict(dict(
            wte=nn.Embedding(config.padded_vocab_size, config.n_embd),
            h=nn.ModuleList(Block(config) for _ in range(config.n_layer)),
            ln_f=config.norm_class(config.n_embd, eps=config.norm_eps)))
        self._init_runtime()

    # -- runtime state shared by all variants ----------------------------------------------------
    def _init_runtime(self) -> None:
        self.rope_cache: Optional[RoPECache] = None
        self.mask_cache: Optional[Tensor] = None
        self.kv_caches: List[KVCache] = []
        self.max_seq_length = self.config.block_size
        self._engine: Optional[DecoderEngine] = None
        self._frozen_sig = None
        self._pending_grad_scale: Optional[Tensor] = None
        self._cache_batch = 0
        # opt-in: run right-padded [B, T] batches without their pad rows (packed cu_seqlens attention); logits at pad
        # positions are then those of a zero hidden state instead of the reference's (ignored, label -1) values
        self.skip_pad = False

    def _engine_spec(self) -> EngineSpec:
        c = self.config
        return EngineSpec(n_layer=c.n_layer, n_head=c.n_head, n_query_groups=c.n_query_groups, n_embd=c.n_embd,
                          intermediate_size=c.intermediate_size, vocab=c.padded_vocab_size, block_size=c.block_size,
                          norm_eps=c.norm_eps, variant="base")

    def _trainable_filter(self, name: str) -> bool:
        """Names the engine can produce gradients for."""
        return False

    def _dropout_seed(self) -> Optional[int]:
        """Seed of this forward's LoRA dropout masks, drawn from torch's CPU generator (so `torch.manual_seed` / Fabric's
        `seed_everything` make a run repeatable, finetune/ger.py:105); None in eval mode or when the model has no dropout."""
        p = float(getattr(self.config, "dropout", 0.0) or 0.0)
        if not self.training or p <= 0.0:
            return None
        return int(torch.randint(0, 2 ** 62, (1,), dtype=torch.int64).item())

    def _frozen_items(self):
        return [(n, p) for n, p in self.named_parameters() if not self._trainable_filter(n) and "noise_classifier" not in n]

    def _engine_ready(self) -> DecoderEngine:
        """Create the engine on first use and (re)pack the frozen weights whenever they changed."""
        dev = self.lm_head_weight().device
        if dev.type != "cuda":
            raise RuntimeError("dualhyp_b200 runs on a B200 GPU only: move the model to cuda (there is no CPU path)")
        if self._engine is None or self._engine.device != dev:
            self.config.check_engine_supported()
            self._engine = DecoderEngine(self._engine_spec(), dev)
            self._frozen_sig = None
        frozen = self._frozen_items()
        sig = tuple((p.data_ptr(), p._version) for _, p in frozen)
        if sig != self._frozen_sig:
            self._engine.load_frozen({n: p for n, p in frozen})
            self._frozen_sig = sig
        return self._engine

    def lm_head_weight(self) -> Tensor:
        return self.lm_head.weight

    def _logits_dtype(self) -> torch.dtype:
        # fp32 masters => the reference's bf16-mixed run yields fp32 logits; bf16 weights => bf16-true inference
        return torch.float32 if self.lm_head_weight().dtype == torch.float32 else torch.bfloat16

    def _trainables(self) -> Tuple[Tuple[str, ...], Tuple[Tensor, ...]]:
        names, tensors = [], []
        # a merged LoRA module computes linear(x) only (reference ger/lora.py:159-161): its factors are already inside W
        merged = tuple(mn + ".lora_" for mn, m in self.named_modules() if getattr(m, "merged", False))
        for n, p in self.named_parameters():
            if "noise_classifier" in n:
                continue
            if merged and n.startswith(merged):
                continue
            if self._trainable_filter(n):
                names.append(n)
                tensors.append(p)
            elif p.requires_grad and torch.is_grad_enabled():
                raise NotImplementedError(
                    f"parameter {n!r} requires grad, but the B200 path differentiates only the PEFT tensors of this variant; "
                    "call mark_only_*_as_trainable(model) first (full fine-tuning is outside the DualHyp hot path)")
        return tuple(names), tuple(tensors)

    # -- forward ---------------------------------------------------------------------------------
    def _hidden(self, idx: Tensor, varlen: Optional[Dict] = None) -> Tensor:
        names, tensors = self._trainables()
        if varlen is None and self.skip_pad and idx.size(0) > 1:
            varlen = self._pad_free_layout(idx)
        if torch.is_grad_enabled() and any(t.requires_grad for t in tensors):
            return _DecoderFn.apply(self, idx, varlen, names, *tensors)
        eng = self._engine_ready()
        packed = eng.pack_trainable(dict(zip(names, tensors)))
        B, T = idx.shape
        if varlen is not None and "row_map" in varlen:
            ids = idx.reshape(-1).index_select(0, varlen["row_map_long"]).view(1, -1)
            hidden_p = eng.forward_hidden(ids, packed, save=False, varlen=varlen)
            hidden = eng._cap_buf("hidden_unpacked", (B * T, hidden_p.shape[1]), torch.bfloat16)
            ops.scatter_rows(hidden_p, varlen["row_map"], varlen["count"], hidden)
            return hidden.view(B, T, -1)
        return eng.forward_hidden(idx, packed, save=False, varlen=varlen).view(B, T, -1)

    def _pad_free_layout(self, idx: Tensor, lengths: Optional[Tensor] = None) -> Optional[Dict]:
        """Layout of a right-padded [B, T] batch without its pad rows.  The reference's collate pads ids with 0
        (data/av_dataset.py:260-292) and every sample ends with the eos id, so the length of a row is the position of its last
        non-zero id + 1 (one small device -> host read per batch, unless the caller supplies `lengths`)."""
        B, T = idx.shape
        if lengths is None:
            pos = torch.arange(1, T + 1, device=idx.device, dtype=torch.int32)
            lengths = ((idx != 0).to(torch.int32) * pos).amax(dim=1).clamp_(min=1).cpu()
        lens = [int(x) for x in lengths.tolist()]
        if sum(lens) == B * T:
            return None
        dev = idx.device
        cu = torch.zeros(B + 1, dtype=torch.int32)
        cu[1:] = torch.tensor(lens, dtype=torch.int32).cumsum(0)
        Mp = int(cu[-1])
        positions = torch.cat([torch.arange(n, dtype=torch.int32) for n in lens])
        row_map = torch.cat([torch.arange(b * T, b * T + n, dtype=torch.int32) for b, n in enumerate(lens)])
        meta = torch.cat((cu, positions, row_map, torch.tensor([Mp], dtype=torch.int32)))
        meta = meta.pin_memory().to(dev, non_blocking=True) if dev.type == "cuda" else meta
        cu_d, pos_d, map_d, cnt_d = meta[:B + 1], meta[B + 1:B + 1 + Mp], meta[B + 1 + Mp:B + 1 + 2 * Mp], meta[B + 1 + 2 * Mp:]
        return dict(cu_seqlens=cu_d, positions=pos_d, max_len=max(lens), n_seq=B, row_map=map_d, row_map_long=map_d.long(), count=cnt_d)

    def forward_packed(self, input_ids: Tensor, cu_seqlens: Tensor, positions: Tensor, max_len: int, lm_head_chunk_size: int = 0):
        """Causal forward over sequences packed into ONE row without pad tokens (`dualhyp_b200.collate.pack_varlen`; SURVEY 8f
        row 2): input_ids int64 [M], cu_seqlens int32 [n + 1], positions int32 [M] (position of every token inside its
        sequence), max_len = the longest sequence.  Returns logits [1, M, V] (or the chunk list) like `forward`; for training,
        `chunked_cross_entropy(logits, batch["targets"])` keeps the reference's train reduction: the sum over valid targets
        divided by n * (max_len - 1), the number of positions of the right-padded batch the reference would have built
        (ger/utils.py:442-448, quirk Q2)."""
        n = cu_seqlens.numel() - 1
        assert input_ids.dim() == 1 and positions.numel() == input_ids.numel() and max_len <= self.config.block_size
        varlen = dict(cu_seqlens=cu_seqlens.to(torch.int32), positions=positions.to(torch.int32), max_len=int(max_len), n_seq=n)
        hidden = self._hidden(input_ids.view(1, -1), varlen)
        meta = (n, int(max_len))
        M = input_ids.numel()
        if lm_head_chunk_size > 0:
            return [LazyLogits(self, hidden[:, a:a + lm_head_chunk_size], hidden, a, min(M, a + lm_head_chunk_size), packed_meta=meta)
                    for a in range(0, M, lm_head_chunk_size)]
        return LazyLogits(self, hidden, packed_meta=meta)

    def _forward_impl(self, idx: Tensor, input_pos: Optional[Tensor], lm_head_chunk_size: int):
        T = idx.size(1)
        block_size = self.config.block_size
        use_kv_cache = input_pos is not None
        if use_kv_cache:
            assert self.max_seq_length >= T, f"Cannot forward sequence of length {T}, max seq length is only {self.max_seq_length}"
        assert self.max_seq_length <= block_size, f"Cannot attend to {self.max_seq_length}, block size is only {block_size}"
        assert block_size >= T, f"Cannot forward sequence of length {T}, block size is only {block_size}"
        if not use_kv_cache:
            hidden = self._hidden(idx)
            if lm_head_chunk_size > 0:
                return [LazyLogits(self, hidden[:, a:a + lm_head_chunk_size], hidden, a, min(T, a + lm_head_chunk_size))
                        for a in range(0, T, lm_head_chunk_size)]
            return _LMHeadFn.apply(self, hidden, self._logits_dtype())
        return LazyLogits(self, self._cached_forward(idx, input_pos))

    @torch.no_grad()
    def _cached_forward(self, idx: Tensor, input_pos: Tensor) -> Tensor:
        """KV-cache inference as generate/base.py drives it: a prefill with input_pos = arange(T), then one
        token per call.  The cache spans max_seq_length slots like the reference's (ger/model.py:132-151) but
        holds the n_query_groups kv heads un-expanded, and attention reads only the filled prefix."""
        eng = self._engine_ready()
        names, tensors = self._trainables()
        packed = eng.pack_trainable(dict(zip(names, tensors)))
        B, T = idx.shape
        if not eng.cache_k or self._cache_batch != B or eng.cache_k[0].shape[0] != B or eng.cache_max_len < self.max_seq_length:
            eng.alloc_cache(B, self.max_seq_length)      # generate_batch / DecodeService may have left a shorter cache
            self._cache_batch = B
            self.kv_caches = list(zip(eng.cache_k, eng.cache_v))
        if T > 1:
            if input_pos.numel() != T or int(input_pos[0]) != 0:
                raise NotImplementedError("multi-token cached forward must be the prefill (input_pos = arange(T))")
            return eng.prefill(idx, packed).view(B, T, -1)
        pos = input_pos.to(torch.int32).reshape(-1)
        if pos.numel() == 1 and B > 1:
            pos = pos.expand(B)
        return eng.decode_step(idx.reshape(-1), pos.contiguous(), packed).view(B, 1, -1)

    def forward(self, idx: Tensor, max_seq_length: Optional[int] = None, input_pos: Optional[Tensor] = None):
        """Base signature, reference ger/model.py:65-67."""
        if max_seq_length is not None:
            assert max_seq_length <= self.config.block_size
        out = self._forward_impl(idx, input_pos, 0)
        return out.materialize() if isinstance(out, LazyLogits) else out

    def reset_cache(self) -> None:
        """ger/model.py:58-63."""
        self.kv_caches.clear()
        if self._engine is not None:
            self._engine.drop_cache()

    @classmethod
    def from_name(cls, name: str, **kwargs: Any):
        return cls(Config.from_name(name, **kwargs))

    def build_rope_cache(self, idx: Tensor) -> RoPECache:
        return build_rope_cache(seq_len=self.config.block_size, n_elem=int(self.config.rotary_percentage * self.config.head_size),
                                dtype=torch.bfloat16, device=idx.device)

    def build_mask_cache(self, idx: Tensor) -> Tensor:
        ones = torch.ones((self.config.block_size, self.config.block_size), device=idx.device, dtype=torch.bool)
        return torch.tril(ones).unsqueeze(0).unsqueeze(0)

    # fused loss entry used by ger.utils.chunked_cross_entropy
    def fused_loss(self, hidden: Tensor, targets: Tensor, denom: float) -> Tensor:
        return _LMHeadCEFn.apply(self, hidden, targets, float(denom))


class Block(nn.Module):
    """Parameter container with the reference's names (ger/model.py:154-188); the math runs inside the engine."""

    def __init__(self, config: Config) -> None:
        super().__init__()
        self.norm_1 = config.norm_class(config.n_embd, eps=config.norm_eps)
        self.attn = CausalSelfAttention(config)
        if not config.shared_attention_norm:
            self.norm_2 = config.norm_class(config.n_embd, eps=config.norm_eps)
        self.mlp = config.mlp_class(config)
        self.config = config

    def forward(self, *args, **kwargs):
        raise NotImplementedError("Block is executed as part of GPT.forward on the B200 engine; it has no standalone forward")


class CausalSelfAttention(nn.Module):
    def __init__(self, config: Config) -> None:
        super().__init__()
        shape = (config.n_head + 2 * config.n_query_groups) * config.head_size
        self.attn = nn.Linear(config.n_embd, shape, bias=config.bias)
        self.proj = nn.Linear(config.n_embd, config.n_embd, bias=config.bias)
        self.config = config

    def forward(self, *args, **kwargs):
        raise NotImplementedError("CausalSelfAttention is executed as part of GPT.forward on the B200 engine")


class LLaMAMLP(nn.Module):
    def __init__(self, config: Config) -> None:
        super().__init__()
        self.fc_1 = nn.Linear(config.n_embd, config.intermediate_size, bias=config.bias)
        self.fc_2 = nn.Linear(config.n_embd, config.intermediate_size, bias=config.bias)
        self.proj = nn.Linear(config.intermediate_size, config.n_embd, bias=config.bias)

    def forward(self, *args, **kwargs):
        raise NotImplementedError("LLaMAMLP is executed as part of GPT.forward on the B200 engine")


class GptNeoxMLP(nn.Module):
    def __init__(self, config: Config) -> None:
        super().__init__()
        raise NotImplementedError("GptNeoxMLP models are outside the DualHyp hot path (TinyLlama uses LLaMAMLP)")


def build_rope_cache(seq_len: int, n_elem: int, dtype: torch.dtype, device: torch.device, base: int = 10000,
                     condense_ratio: int = 1) -> RoPECache:
    """ger/model.py:319-346 (host-side table construction; fp32 math, bf16/fp16 rounding per `dtype`)."""
    theta = 1.0 / (base ** (torch.arange(0, n_elem, 2, device=device) / n_elem))
    seq_idx = torch.arange(seq_len, device=device) / condense_ratio
    idx_theta = torch.outer(seq_idx, theta).repeat(1, 2)
    cos, sin = torch.cos(idx_theta), torch.sin(idx_theta)
    if dtype == torch.bfloat16:
        return cos.bfloat16(), sin.bfloat16()
    if dtype in (torch.float16, torch.bfloat16, torch.int8):
        return cos.half(), sin.half()
    return cos, sin


def apply_rope(x: Tensor, cos: Tensor, sin: Tensor) -> Tensor:
    """ger/model.py:349-355 for x (B, nh, T, hs) bf16 on the GPU, through the rotary kernel (NeoX halves)."""
    from .fused_rotary_embedding import apply_rotary_emb_func
    hs = x.size(-1)
    xt = x.transpose(1, 2).contiguous()                 # kernel layout (B, T, nh, hs)
    out = apply_rotary_emb_func(xt, cos[..., : hs // 2].float().contiguous(), sin[..., : hs // 2].float().contiguous(), False, False)
    return out.transpose(1, 2)
