"""DecoderEngine: the B200 execution of the lit-gpt-style decoder the DualHyp modules run
(reference ger/model.py:65-114,163-268,312-316; ger/lora.py:504-549; ger/adapter.py:116-146).

The engine owns bf16 copies of the frozen weights in the layouts the kernels want, the RoPE tables, the
activation buffers that the backward reads, and the KV caches; it sequences hand-written sm_100a kernels
through the C ABI (dualhyp_b200.ops).  PyTorch supplies device memory and the stream, nothing else.

Per layer, forward (M = B*T token rows):
    rmsnorm -> [LoRA a = n A^T] -> QKV GEMM (+LoRA K-extension, RoPE epilogue, interleaved row kept as is)
    -> [prefix K/V GEMM] -> flash attention (+gated prefix) -> proj GEMM (+LoRA, residual epilogue, fp32 stream)
    -> rmsnorm -> fc_1|fc_2 GEMM (SwiGLU epilogue) -> proj GEMM (residual epilogue)
Backward is dgrad only for the frozen weights, plus the small LoRA / adapter wgrads.  Every frozen weight is
kept twice in HBM, as stored ([out, in]: the K-major B operand of the forward GEMM) and transposed ([in, out]:
the K-major B operand of its dgrad), so that both directions run on the CTA-pair tcgen05 kernel; the extra
2.1 GB is nothing against 180 GB and the copies are made once at load time.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass
from typing import Dict, List, Optional

import torch

from . import ops

Tensor = torch.Tensor
LORA_PAD = 64          # LoRA rank (x3 for QKV) padded to one 64-wide K block


@dataclass
class EngineSpec:
    n_layer: int
    n_head: int
    n_query_groups: int
    n_embd: int
    intermediate_size: int
    vocab: int
    block_size: int
    norm_eps: float
    variant: str = "base"              # base | lora | adapter
    r: int = 0
    alpha: float = 1.0
    lora_qkv: bool = False
    lora_proj: bool = False
    lora_start_layer: int = 0
    lora_dropout: float = 0.0          # nn.Dropout(p) on the input of every LoRA branch in training (ger/lora.py:94-98,165)
    adapter_prompt_length: int = 0
    adapter_start_layer: int = 0

    @property
    def qpk(self) -> int:
        return self.n_head // self.n_query_groups

    @property
    def qkv_out(self) -> int:
        return (self.n_head + 2 * self.n_query_groups) * 64

    @property
    def kv_embd(self) -> int:
        return self.n_embd // self.qpk

    def has_lora(self, layer: int) -> bool:
        return self.variant == "lora" and self.r > 0 and layer >= self.lora_start_layer and (self.lora_qkv or self.lora_proj)

    def has_adapter(self, layer: int) -> bool:
        return self.variant == "adapter" and layer >= self.adapter_start_layer

    @property
    def first_trainable_layer(self) -> int:
        if self.variant == "adapter":
            return self.adapter_start_layer
        if self.variant == "lora" and self.r > 0 and (self.lora_qkv or self.lora_proj):
            return self.lora_start_layer
        return self.n_layer


def build_rope_tables(block_size: int, device) -> tuple:
    """ger/model.py:319-346 with the call site's dtype=bfloat16 (:120-126): fp32 math, tables rounded to bf16,
    halves repeated (`repeat(1, 2)`).  Built on the host so they are bit-identical to the CPU oracle's."""
    theta = 1.0 / (10000 ** (torch.arange(0, 64, 2) / 64))
    idx_theta = torch.outer(torch.arange(block_size) / 1, theta).repeat(1, 2)
    return torch.cos(idx_theta).bfloat16().to(device).contiguous(), torch.sin(idx_theta).bfloat16().to(device).contiguous()


def interleave_fc(w1: Tensor, w2: Tensor) -> Tensor:
    """[fc_1 rows 0..31 | fc_2 rows 0..31 | fc_1 rows 32..63 | ...]: the SwiGLU epilogue needs both halves of a
    column pair inside one 64-wide accumulator chunk."""
    inter = w1.shape[0]
    return torch.stack((w1.view(inter // 32, 32, -1), w2.view(inter // 32, 32, -1)), dim=1).reshape(2 * inter, -1).contiguous()


class DecoderEngine:
    def __init__(self, spec: EngineSpec, device: torch.device):
        assert spec.n_embd % 256 == 0 and spec.intermediate_size % 64 == 0 and spec.vocab % 64 == 0
        assert spec.adapter_prompt_length <= 16
        if spec.variant == "lora" and spec.r > 0:
            assert 3 * spec.r <= LORA_PAD, "LoRA rank too large for the single-block K extension (3r <= 64)"
        self.spec = spec
        self.device = device
        self.cos, self.sin = build_rope_tables(spec.block_size, device)
        self.layers: List[Dict[str, Tensor]] = []
        self.wte: Optional[Tensor] = None
        self.ln_f: Optional[Tensor] = None
        self.lm_w: Optional[Tensor] = None
        self.lm_w_t: Optional[Tensor] = None
        self.lm_scale: Optional[Tensor] = None
        self.lm_bias: Optional[Tensor] = None
        self._bufs: Dict = {}
        self._saved: Optional[Dict] = None
        self.step_id = 0
        # KV cache
        self.cache_k: List[Tensor] = []
        self.cache_v: List[Tensor] = []
        self.cache_max_len = 0

    # ------------------------------------------------------------------------------------------
    # weights
    # ------------------------------------------------------------------------------------------
    def load_frozen(self, sd: Dict[str, Tensor]) -> None:
        """Pack the frozen decoder weights from a (reference-keyed) state dict: bf16, fc_1/fc_2 interleaved."""
        s, dev = self.spec, self.device
        lk = ".linear.weight" if any(k.endswith("attn.attn.linear.weight") for k in sd) else ".weight"

        def bf(k):
            return sd[k].detach().to(dev, torch.bfloat16).contiguous()

        def f32(k):
            return sd[k].detach().to(dev, torch.float32).contiguous()

        self.layers = []
        for i in range(s.n_layer):
            p = f"transformer.h.{i}."
            lw = dict(
                norm1=f32(p + "norm_1.weight"), norm2=f32(p + "norm_2.weight"),
                wqkv=bf(p + "attn.attn" + lk), wproj=bf(p + "attn.proj" + lk),
                w12=interleave_fc(bf(p + "mlp.fc_1" + lk), bf(p + "mlp.fc_2" + lk)), wmlp=bf(p + "mlp.proj" + lk))
            if i >= s.first_trainable_layer:      # dgrad operands ([in, out] row-major)
                for k in ("wqkv", "wproj", "w12", "wmlp"):
                    lw[k + "_t"] = lw[k].t().contiguous()
            self.layers.append(lw)
        self.wkv_all = None
        if s.variant == "adapter" and s.adapter_start_layer < s.n_layer:
            # k rows then v rows of every group, for all adapter layers: the operand of the batched prefix projection
            G, qpk = s.n_query_groups, s.qpk
            rows = torch.cat([torch.arange((g * (qpk + 2) + qpk) * 64, (g * (qpk + 2) + qpk + 2) * 64) for g in range(G)]).to(dev)
            self.wkv_all = torch.stack([self.layers[i]["wqkv"].index_select(0, rows) for i in range(s.adapter_start_layer, s.n_layer)]).contiguous()
        self.wte = sd["transformer.wte.weight"].detach().to(dev).contiguous()
        if self.wte.dtype not in (torch.float32, torch.bfloat16):
            self.wte = self.wte.float()
        self.ln_f = f32("transformer.ln_f.weight")
        if "lm_head.linear.weight" in sd:
            self.lm_w = bf("lm_head.linear.weight")
            scale, bias = f32("lm_head.adapter_scale"), f32("lm_head.adapter_bias")
            # AdapterV2Linear is frozen at scale=1, bias=0 (ger/lora.py:67-68); skip the epilogue work when trivial
            self.lm_scale = None if bool((scale == 1).all()) else scale
            self.lm_bias = None if bool((bias == 0).all()) else bias
        else:
            self.lm_w = bf("lm_head.weight")
            self.lm_scale = self.lm_bias = None
        self.lm_w_t = self.lm_w.t().contiguous() if s.first_trainable_layer < s.n_layer else None

    def _persistent(self, name: str, shape, dtype) -> Tensor:
        """Zero-initialised buffer that lives as long as the engine (created outside inference mode, so that both the training
        step and `generate` may update it in place)."""
        key = ("persistent", name, tuple(shape), dtype)
        t = self._bufs.get(key)
        if t is None:
            with torch.inference_mode(False):
                t = torch.zeros(shape, device=self.device, dtype=dtype)
            self._bufs[key] = t
        return t

    def _pack_lora_batched(self, tr: Dict[str, Tensor], packed: Dict[int, Dict[str, Tensor]]) -> bool:
        """The bf16 operand forms of ALL LoRA layers in ~17 launches (per-layer packing costs ~20 ATen launches per layer, 440 of
        the 1179 launches of a LoRA step: profiles/r02_lora_launch_summary_before.txt).  The forms live in persistent buffers
        [L, ...] whose zero padding is written once; a layer's entry in `packed` is a view of its slice.  Returns False when a
        LoRA module is absent from `tr` (merged into W): the caller then packs layer by layer."""
        s = self.spec
        lay = [i for i in range(s.n_layer) if s.has_lora(i)]
        if not lay:
            return True
        L, C, r, kv = len(lay), s.n_embd, s.r, s.kv_embd
        scaling = s.alpha / s.r
        pre = "transformer.h.{}.attn.{}"
        bf = torch.bfloat16

        def stack(fmt):
            return torch.stack([tr[fmt.format(i)].detach() for i in lay]).to(self.device)

        for kind, on, n_a, out_rows in (("attn", s.lora_qkv, 3 * r, s.qkv_out), ("proj", s.lora_proj, r, C)):
            if not on:
                continue
            fa, fb = pre.format("{}", kind + ".lora_A"), pre.format("{}", kind + ".lora_B")
            if any(fa.format(i) not in tr or fb.format(i) not in tr for i in lay):
                return False
            key = "qkv" if kind == "attn" else "proj"
            a_pad = self._persistent(key + "_a_all", (L, LORA_PAD, C), bf)
            a_pad_t = self._persistent(key + "_a_t_all", (L, C, LORA_PAD), bf)
            b_ext = self._persistent(key + "_b_all", (L, out_rows, LORA_PAD), bf)
            b_ext_t = self._persistent(key + "_b_t_all", (L, LORA_PAD, out_rows), bf)
            A, Bm = stack(fa), stack(fb)                                   # [L, n_a, C] / [L, out_rows, r] fp32
            a_pad[:, :n_a].copy_(A.float() * scaling)
            if kind == "attn":                                              # quirk Q3: q rows use after_A[:, 0:r], k rows [r:2r], v rows [2r:3r]
                b_ext[:, :C, 0:r].copy_(Bm[:, :C])
                b_ext[:, C:C + kv, r:2 * r].copy_(Bm[:, C:C + kv])
                b_ext[:, C + kv:, 2 * r:3 * r].copy_(Bm[:, C + kv:])
            else:
                b_ext[:, :, :r].copy_(Bm)
            a_pad_t.copy_(a_pad.transpose(1, 2))
            b_ext_t.copy_(b_ext.transpose(1, 2))
            for j, i in enumerate(lay):
                d = packed.setdefault(i, {})
                d[key + "_a"], d[key + "_b"], d[key + "_a_t"], d[key + "_b_t"] = a_pad[j], b_ext[j], a_pad_t[j], b_ext_t[j]
        return True

    def pack_trainable(self, tr: Dict[str, Tensor]) -> Dict[int, Dict[str, Tensor]]:
        """Per-step bf16 forms of the small trainable tensors (reference-keyed, fp32 masters)."""
        s, dev = self.spec, self.device
        packed: Dict[int, Dict[str, Tensor]] = {}
        scaling = (s.alpha / s.r) if s.r > 0 else 0.0
        lora_done = s.variant == "lora" and s.r > 0 and self._pack_lora_batched(tr, packed)
        if s.variant == "lora" and s.r > 0 and not lora_done:
            packed.clear()
        for i in range(s.n_layer):
            p = f"transformer.h.{i}.attn."
            d: Dict[str, Tensor] = {}
            if s.has_adapter(i):
                d["prefix"] = tr[p + "adapter_wte.weight"].detach().to(dev, torch.bfloat16).contiguous()
                d["gate"] = tr[p + "gating_factor"].detach().to(dev, torch.float32).reshape(-1).contiguous()
            if s.has_lora(i) and not lora_done:
                C, r = s.n_embd, s.r
                if s.lora_qkv and (p + "attn.lora_A") in tr:       # absent: the module was merged into W (LoRALinear.merge)
                    A, Bm = tr[p + "attn.lora_A"].detach().to(dev), tr[p + "attn.lora_B"].detach().to(dev)
                    a_pad = torch.zeros((LORA_PAD, C), device=dev, dtype=torch.bfloat16)
                    a_pad[: 3 * r] = (A.float() * scaling).to(torch.bfloat16)
                    b_ext = torch.zeros((s.qkv_out, LORA_PAD), device=dev, dtype=torch.bfloat16)
                    kv = s.kv_embd
                    b_ext[:C, 0:r] = Bm[:C].to(torch.bfloat16)                      # q rows use after_A[:, 0:r]
                    b_ext[C:C + kv, r:2 * r] = Bm[C:C + kv].to(torch.bfloat16)      # k rows use [:, r:2r]
                    b_ext[C + kv:, 2 * r:3 * r] = Bm[C + kv:].to(torch.bfloat16)    # v rows use [:, 2r:3r]
                    d["qkv_a"], d["qkv_b"] = a_pad, b_ext
                    d["qkv_a_t"], d["qkv_b_t"] = a_pad.t().contiguous(), b_ext.t().contiguous()
                if s.lora_proj and (p + "proj.lora_A") in tr:
                    A, Bm = tr[p + "proj.lora_A"].detach().to(dev), tr[p + "proj.lora_B"].detach().to(dev)
                    a_pad = torch.zeros((LORA_PAD, C), device=dev, dtype=torch.bfloat16)
                    a_pad[:r] = (A.float() * scaling).to(torch.bfloat16)
                    b_ext = torch.zeros((C, LORA_PAD), device=dev, dtype=torch.bfloat16)
                    b_ext[:, :r] = Bm.to(torch.bfloat16)
                    d["proj_a"], d["proj_b"] = a_pad, b_ext
                    d["proj_a_t"], d["proj_b_t"] = a_pad.t().contiguous(), b_ext.t().contiguous()
            if d:
                packed.setdefault(i, {}).update(d)
        if s.variant == "adapter" and s.adapter_start_layer < s.n_layer:
            packed[-1] = {"prefix_all": torch.stack([packed[i]["prefix"] for i in range(s.adapter_start_layer, s.n_layer)]).contiguous()}
        return packed

    # ------------------------------------------------------------------------------------------
    # buffers
    # ------------------------------------------------------------------------------------------
    def _buf(self, name: str, shape, dtype) -> Tensor:
        key = (name, tuple(shape), dtype)
        t = self._bufs.get(key)
        if t is None:
            t = torch.empty(shape, device=self.device, dtype=dtype)
            self._bufs[key] = t
        return t

    def _cap_buf(self, name: str, shape, dtype) -> Tensor:
        """Grow-only flat buffer viewed as `shape`: the inference paths see a different row count for every batch (prompt
        lengths vary), and an exact-shape cache would keep one full activation set per distinct count."""
        numel = 1
        for d in shape:
            numel *= int(d)
        key = ("cap", name, dtype)
        t = self._bufs.get(key)
        if t is None or t.numel() < numel:
            g = 22 if numel >= (1 << 20) else 12                 # 4 Mi-element granules (4 Ki for the small per-row vectors)
            cap = ((numel + (1 << g) - 1) >> g) << g
            t = self._bufs[key] = torch.empty(cap, device=self.device, dtype=dtype)
        return t[:numel].view(shape)

    def release_buffers(self) -> None:
        self._bufs.clear()
        self._saved = None

    # ------------------------------------------------------------------------------------------
    # forward over full sequences (training / evaluation / prefill)
    # ------------------------------------------------------------------------------------------
    def forward_hidden(self, idx: Tensor, packed: Dict[int, Dict[str, Tensor]], save: bool = False, write_cache: bool = False,
                       varlen: Optional[Dict] = None, dropout_seed: Optional[int] = None) -> Tensor:
        """idx int64 [B, T] -> ln_f output bf16 [B*T, C].  save=True keeps what backward() needs.
        varlen: idx is ONE row of packed sequences [1, M] without pad tokens; `varlen` holds cu_seqlens int32 [n + 1], positions
        int32 [M] (RoPE position of every row inside its sequence), max_len, and for the cache-writing prefill seq_ids int32 [M].
        Training over packed rows (save=True) keeps the same activations; only the attention kernels see sequence boundaries."""
        s = self.spec
        B, T = idx.shape
        if varlen is not None:
            assert B == 1, "a packed variable-length batch is ONE row of concatenated sequences"
            assert varlen["max_len"] <= s.block_size
        else:
            assert T <= s.block_size, f"Cannot forward sequence of length {T}, block size is only {s.block_size}"
        M, C, I, G, qpk = B * T, s.n_embd, s.intermediate_size, s.n_query_groups, s.qpk
        vpos = varlen["positions"] if varlen is not None else None
        L = s.n_layer
        bf, f32 = torch.bfloat16, torch.float32
        idx = idx.contiguous()
        per = (lambda name, i: f"{name}{i}") if save else (lambda name, i: name)
        _buf = self._cap_buf          # grow-only flat buffers viewed at the step's row count (packed batches change it every step)
        # the residual stream alternates between two buffers unless the backward needs every stage
        x = _buf(per("x", 0), (M, C), f32)
        ops.embedding_fwd(idx.view(-1), self.wte, x)
        n = _buf("n", (M, C), bf)
        act = _buf("act", (M, I), bf)
        # LoRA dropout (training only): the LoRA branch multiplies dropout(x), the frozen path the plain x (ger/lora.py:159-166).
        # The mask is Philox(seed, offset = 2 * layer [+ 1 for proj]); backward() regenerates it from the same pair for d(x).
        use_do = save and dropout_seed is not None and s.variant == "lora" and s.lora_dropout > 0
        saved = dict(B=B, T=T, layers=[], packed=packed, varlen=varlen, dropout_seed=dropout_seed if use_do else None) if save else None
        pkv_all = self._project_prefixes(packed, "pkv_all")
        for i, lw in enumerate(self.layers):
            pk = packed.get(i, {})
            nd_i = yd_i = None
            rstd1 = _buf(per("rstd1_", i), (M,), f32)
            n_i = _buf(per("n1_", i), (M, C), bf) if (save and "qkv_a" in pk) else n
            ops.rmsnorm_fwd(x, lw["norm1"], n_i, rstd1, s.norm_eps)
            qkv = _buf(per("qkv", i), (M, s.qkv_out), bf)
            a_s = None
            if "qkv_a" in pk:
                a_s = _buf(per("qkv_as", i), (M, LORA_PAD), bf)
                src = n_i
                if use_do:       # dropout(n) is kept for the factor gradient (27 MB per layer at 6.6 k tokens; regenerating it costs a pass)
                    src = nd_i = _buf(per("lora_do_qkv", i), (M, C), bf)
                    ops.dropout_bf16(n_i, src, s.lora_dropout, dropout_seed, 2 * i)
                ops.gemm(src, pk["qkv_a"], a_s)
                ops.gemm(n_i, lw["wqkv"], qkv, A2=a_s, B2=pk["qkv_b"], epilogue=ops.EPI_QKV_ROPE, rope_cos=self.cos, rope_sin=self.sin,
                         positions=vpos, seq_len=T, q_per_kv=qpk)
            else:
                ops.gemm(n_i, lw["wqkv"], qkv, epilogue=ops.EPI_QKV_ROPE, rope_cos=self.cos, rope_sin=self.sin, positions=vpos,
                         seq_len=T, q_per_kv=qpk)
            prefix_kv = gate = None
            if "prefix" in pk:
                prefix_kv = pkv_all[i - s.adapter_start_layer]
                gate = pk["gate"]
            if write_cache and varlen is not None:
                ops.kv_cache_write(qkv, self.cache_k[i], self.cache_v[i], varlen["seq_ids"], vpos, G, qpk)
            elif write_cache:
                self._write_prefill_cache(i, qkv, B, T)
            y = _buf(per("y", i), (M, C), bf)
            lse = _buf(per("lse", i), (B, s.n_head, T), f32)
            if varlen is not None:
                ops.attention_fwd_varlen(qkv, y, lse, varlen["cu_seqlens"], varlen["max_len"], G, qpk, prefix_kv, gate,
                                         s.adapter_prompt_length if prefix_kv is not None else 0)
            else:
                ops.attention_fwd(qkv, y, lse, B, T, G, qpk, prefix_kv, gate, s.adapter_prompt_length if prefix_kv is not None else 0)
            xmid = _buf(per("xmid", i), (M, C), f32)
            ap_s = None
            if "proj_a" in pk:
                ap_s = _buf(per("proj_as", i), (M, LORA_PAD), bf)
                src = y
                if use_do:
                    src = yd_i = _buf(per("lora_do_proj", i), (M, C), bf)
                    ops.dropout_bf16(y, src, s.lora_dropout, dropout_seed, 2 * i + 1)
                ops.gemm(src, pk["proj_a"], ap_s)
                ops.gemm(y, lw["wproj"], xmid, A2=ap_s, B2=pk["proj_b"], epilogue=ops.EPI_RESID_F32, aux=x)
            else:
                ops.gemm(y, lw["wproj"], xmid, epilogue=ops.EPI_RESID_F32, aux=x)
            rstd2 = _buf(per("rstd2_", i), (M,), f32)
            ops.rmsnorm_fwd(xmid, lw["norm2"], n, rstd2, s.norm_eps)
            h12 = _buf(per("h12_", i), (M, 2 * I), bf) if save else None
            ops.gemm(n, lw["w12"], h12, epilogue=ops.EPI_SWIGLU, out2=act)
            x_next = _buf(per("x", i + 1), (M, C), f32) if save else _buf("x_alt" if (i % 2 == 0) else "x", (M, C), f32)
            ops.gemm(act, lw["wmlp"], x_next, epilogue=ops.EPI_RESID_F32, aux=xmid)
            if save:
                saved["layers"].append(dict(x=x, xmid=xmid, rstd1=rstd1, rstd2=rstd2, qkv=qkv, y=y, lse=lse, h12=h12, n1=n_i,
                                            a_s=a_s, ap_s=ap_s, prefix_kv=prefix_kv, gate=gate, n1_do=nd_i, y_do=yd_i))
            x = x_next
        hidden = _buf("hidden", (M, C), bf)
        rstd_f = _buf("rstd_f", (M,), f32)
        ops.rmsnorm_fwd(x, self.ln_f, hidden, rstd_f, s.norm_eps)
        if save:
            saved["x_final"], saved["rstd_f"] = x, rstd_f
            self.step_id += 1
            saved["step_id"] = self.step_id
            self._saved = saved
        return hidden

    def _project_prefixes(self, packed: Dict[int, Dict[str, Tensor]], name: str) -> Optional[Tensor]:
        """k / v slots of `attn(adapter_wte)` for every adapter layer, one launch (ger/adapter.py:129-137)."""
        s = self.spec
        if -1 not in packed:
            return None
        La = s.n_layer - s.adapter_start_layer
        key = (name, (La, s.adapter_prompt_length, s.qkv_out), torch.bfloat16)
        fresh = key not in self._bufs
        pkv_all = self._buf(name, (La, s.adapter_prompt_length, s.qkv_out), torch.bfloat16)
        if fresh:
            pkv_all.zero_()          # the q slots are never written nor read
        ops.prefix_project_fwd(packed[-1]["prefix_all"], self.wkv_all, pkv_all, s.n_query_groups, s.qpk)
        return pkv_all

    def _wgrad_plan(self, rows: int, cols: int, K: int) -> int:
        """Split-K factor of a LoRA factor gradient (see _wgrad)."""
        tiles = ((rows + 127) // 128) * ((cols + 63) // 64)
        return max(1, min(ops.sm_count() // tiles, (K // 64) // 4, 16))

    def _wgrad_into(self, A: Tensor, B: Tensor, ws: Tensor, alpha: float = 1.0) -> None:
        """ws [S, rows, cols] fp32 = the S split-K slices of alpha * A^T B (the caller sums them, for all layers at once)."""
        S = ws.shape[0]
        if S == 1:
            ops.gemm(A, B, ws[0], a_layout=1, b_layout=1, epilogue=ops.EPI_F32, alpha=alpha)
        else:
            ops.gemm(A, B, ws, a_layout=1, b_layout=1, epilogue=ops.EPI_F32, alpha=alpha, split_k=S)

    def lm_head(self, hidden: Tensor, out: Optional[Tensor] = None, m_limit: Optional[Tensor] = None) -> Tensor:
        """bf16 logits [M, V] = adapter_scale * (hidden W^T + adapter_bias)  (ger/lora.py:70-71).  m_limit (device int32):
        only the first *m_limit rows are needed (valid-row compaction of the fused loss)."""
        M = hidden.shape[0]
        logits = out if out is not None else self._cap_buf("logits", (M, self.spec.vocab), torch.bfloat16)
        ops.gemm(hidden, self.lm_w, logits, col_scale=self.lm_scale, col_bias=self.lm_bias, m_limit=m_limit, b_frozen=True)
        return logits

    # ------------------------------------------------------------------------------------------
    # backward: d(hidden) -> gradients of the trainable tensors (reference-keyed, fp32)
    # ------------------------------------------------------------------------------------------
    def lm_head_backward(self, dlogits: Tensor, m_limit: Optional[Tensor] = None, name: str = "d_hidden") -> Tensor:
        """d(hidden) bf16 [M, C] = dlogits [M, V] @ W [V, C]   (W^T [C, V] is the K-major B operand)."""
        M = dlogits.shape[0]
        d_hidden = self._cap_buf(name, (M, self.spec.n_embd), torch.bfloat16)
        if self.lm_scale is not None:
            raise NotImplementedError("training through a non-trivial adapter_scale is not supported (it is frozen at 1)")
        ops.gemm(dlogits, self.lm_w_t, d_hidden, m_limit=m_limit)
        return d_hidden

    def backward(self, d_hidden: Tensor, step_id: Optional[int] = None) -> Dict[str, Tensor]:
        s = self.spec
        sv = self._saved
        if sv is None:
            raise RuntimeError("backward() without a saved forward")
        if step_id is not None and step_id != sv["step_id"]:
            raise RuntimeError("the engine's saved activations belong to a later forward; call backward before the next forward")
        B, T = sv["B"], sv["T"]
        vl = sv.get("varlen")          # packed sequences: B = 1, T = total rows
        M, C, I, G, qpk = B * T, s.n_embd, s.intermediate_size, s.n_query_groups, s.qpk
        bf, f32 = torch.bfloat16, torch.float32
        packed = sv["packed"]
        do_seed = sv.get("dropout_seed")
        grads: Dict[str, Tensor] = {}
        first = s.first_trainable_layer
        if first >= s.n_layer:
            return grads
        dx = self._cap_buf("dx", (M, C), f32)
        dxb = self._cap_buf("dxb", (M, C), bf)
        ops.rmsnorm_bwd(sv["x_final"], self.ln_f, sv["rstd_f"], d_hidden.contiguous(), None, dx, dxb)
        dh12 = self._cap_buf("dh12", (M, 2 * I), bf)
        dn = self._cap_buf("dn", (M, C), bf)
        dy = self._cap_buf("dy", (M, C), bf)
        dqkv = self._cap_buf("dqkv", (M, s.qkv_out), bf)
        ws = self._cap_buf("attn_ws", (ops.attention_bwd_workspace(B, T, G, qpk),), torch.uint8)
        scaling = (s.alpha / s.r) if s.r > 0 else 0.0
        d_pkv_all = d_gate_all = None
        if s.variant == "adapter":
            La = s.n_layer - s.adapter_start_layer
            d_pkv_all = torch.zeros((La, s.adapter_prompt_length, s.qkv_out), device=self.device, dtype=f32)
            d_gate_all = torch.zeros((La, s.n_head), device=self.device, dtype=f32)
        # LoRA factor gradients: every layer's split-K slices go to ONE workspace per kind, summed for all layers at the end
        # (4 reductions + 3 slice copies per step instead of ~12 ATen launches per layer)
        lora_lay = [i for i in range(first, s.n_layer) if s.has_lora(i)] if s.variant == "lora" else []
        lora_pos = {i: j for j, i in enumerate(lora_lay)}
        lws: Dict[str, Tensor] = {}
        if lora_lay:
            Ll = len(lora_lay)
            for key, rows, cols in (("qkv_dB", s.qkv_out, LORA_PAD), ("qkv_dA", LORA_PAD, C), ("proj_dB", C, LORA_PAD), ("proj_dA", LORA_PAD, C)):
                lws[key] = self._cap_buf("lora_ws_" + key, (Ll, self._wgrad_plan(rows, cols, M), rows, cols), f32)
        for i in range(s.n_layer - 1, first - 1, -1):
            lw, a, pk = self.layers[i], sv["layers"][i], packed.get(i, {})
            p = f"transformer.h.{i}.attn."
            # ---- MLP: d(act) = dx Wp ; SwiGLU backward in the epilogue ; d(n2) = d(h12) W12 ----
            ops.gemm(dxb, lw["wmlp_t"], dh12, epilogue=ops.EPI_SWIGLU_BWD, aux=a["h12"])
            ops.gemm(dh12, lw["w12_t"], dn)
            ops.rmsnorm_bwd(a["xmid"], lw["norm2"], a["rstd2"], dn, dx, dx, dxb)
            # ---- attention output projection ----
            if "proj_a" in pk:
                dap = self._cap_buf("dap", (M, LORA_PAD), bf)
                ops.gemm(dxb, pk["proj_b_t"], dap)
                self._wgrad_into(dxb, a["ap_s"], lws["proj_dB"][lora_pos[i]])
                if do_seed is None:
                    self._wgrad_into(dap, a["y"], lws["proj_dA"][lora_pos[i]], alpha=scaling)
                    ops.gemm(dxb, lw["wproj_t"], dy, A2=dap, B2=pk["proj_a_t"])
                else:
                    # the LoRA branch saw dropout(y): its factor gradient multiplies the saved dropout(y), and its share of d(y) passes
                    # through the same mask (regenerated from (seed, offset)), so it cannot ride the dgrad's K-extension
                    tmp = self._cap_buf("lora_do_g", (M, C), bf)
                    self._wgrad_into(dap, a["y_do"], lws["proj_dA"][lora_pos[i]], alpha=scaling)
                    ops.gemm(dxb, lw["wproj_t"], dy)
                    ops.gemm(dap, pk["proj_a_t"], tmp)
                    ops.dropout_bf16(tmp, dy, s.lora_dropout, do_seed, 2 * i + 1, accumulate=True)
            else:
                ops.gemm(dxb, lw["wproj_t"], dy)
            # ---- attention ----
            need_dqkv = not (s.variant == "adapter" and i == first)
            d_pkv = d_gate = None
            if a["prefix_kv"] is not None:
                d_pkv, d_gate = d_pkv_all[i - s.adapter_start_layer], d_gate_all[i - s.adapter_start_layer]
            plen = s.adapter_prompt_length if a["prefix_kv"] is not None else 0
            if vl is not None:
                ops.attention_bwd_varlen(a["qkv"], a["y"], dy, a["lse"], dqkv if need_dqkv else None, ws, vl["cu_seqlens"], vl["max_len"],
                                         G, qpk, a["prefix_kv"], a["gate"], d_pkv, d_gate, plen, need_dqkv, rope_cos=self.cos, rope_sin=self.sin)
            else:
                ops.attention_bwd(a["qkv"], a["y"], dy, a["lse"], dqkv if need_dqkv else None, ws, B, T, G, qpk, a["prefix_kv"], a["gate"],
                                  d_pkv, d_gate, plen, need_dqkv, rope_cos=self.cos, rope_sin=self.sin)
            if d_pkv is not None:
                grads[p + "gating_factor"] = d_gate.view(1, 1, s.n_head, 1)
            if not need_dqkv:
                continue
            # ---- QKV projection ----
            if "qkv_a" in pk:
                da_s = self._cap_buf("da_s", (M, LORA_PAD), bf)
                ops.gemm(dqkv, pk["qkv_b_t"], da_s)
                self._wgrad_into(dqkv, a["a_s"], lws["qkv_dB"][lora_pos[i]])
                if do_seed is None:
                    self._wgrad_into(da_s, a["n1"], lws["qkv_dA"][lora_pos[i]], alpha=scaling)
                    if i > first:
                        ops.gemm(dqkv, lw["wqkv_t"], dn, A2=da_s, B2=pk["qkv_a_t"])
                else:
                    tmp = self._cap_buf("lora_do_g", (M, C), bf)
                    self._wgrad_into(da_s, a["n1_do"], lws["qkv_dA"][lora_pos[i]], alpha=scaling)
                    if i > first:
                        ops.gemm(dqkv, lw["wqkv_t"], dn)
                        ops.gemm(da_s, pk["qkv_a_t"], tmp)
                        ops.dropout_bf16(tmp, dn, s.lora_dropout, do_seed, 2 * i, accumulate=True)
            elif i > first:
                ops.gemm(dqkv, lw["wqkv_t"], dn)
            if i > first:
                ops.rmsnorm_bwd(a["x"], lw["norm1"], a["rstd1"], dn, dx, dx, dxb)
        if lora_lay:
            r, kv = s.r, s.kv_embd
            pre = "transformer.h.{}.attn."
            if s.lora_qkv and any("qkv_a" in packed.get(i, {}) for i in lora_lay):
                dA = lws["qkv_dA"].sum(1)                                           # [L, 64, C]
                dBf = lws["qkv_dB"].sum(1)                                          # [L, qkv_out, 64]
                dB = torch.empty((len(lora_lay), s.qkv_out, r), device=self.device, dtype=f32)
                dB[:, :C].copy_(dBf[:, :C, 0:r])                                    # quirk Q3 scatter, inverted
                dB[:, C:C + kv].copy_(dBf[:, C:C + kv, r:2 * r])
                dB[:, C + kv:].copy_(dBf[:, C + kv:, 2 * r:3 * r])
                for j, i in enumerate(lora_lay):
                    if "qkv_a" in packed.get(i, {}):
                        grads[pre.format(i) + "attn.lora_A"] = dA[j, : 3 * r]
                        grads[pre.format(i) + "attn.lora_B"] = dB[j]
            if s.lora_proj and any("proj_a" in packed.get(i, {}) for i in lora_lay):
                dA = lws["proj_dA"].sum(1)
                dB = lws["proj_dB"].sum(1)
                for j, i in enumerate(lora_lay):
                    if "proj_a" in packed.get(i, {}):
                        grads[pre.format(i) + "proj.lora_A"] = dA[j, :r]
                        grads[pre.format(i) + "proj.lora_B"] = dB[j, :, :r]
        if d_pkv_all is not None:      # d(adapter_wte) of all layers in one launch: d(prefix k/v) through the k / v weight rows
            d_wte_all = torch.empty((d_pkv_all.shape[0], s.adapter_prompt_length, C), device=self.device, dtype=f32)
            ops.prefix_project_bwd(d_pkv_all, self.wkv_all, d_wte_all, G, qpk)
            for i in range(s.adapter_start_layer, s.n_layer):
                grads[f"transformer.h.{i}.attn.adapter_wte.weight"] = d_wte_all[i - s.adapter_start_layer]
        return grads

    # ------------------------------------------------------------------------------------------
    # KV-cache decode (ger/model.py:132-151,248-261; generate/base.py:19-82)
    # ------------------------------------------------------------------------------------------
    def alloc_cache(self, nseq: int, max_len: int) -> None:
        s = self.spec
        if self.cache_k and self.cache_k[0].shape[0] == nseq and self.cache_max_len == max_len:
            return
        shape = (nseq, s.n_query_groups, max_len, 64)       # native GQA: 4 kv heads, not 32 expanded ones (quirk Q4)
        self.cache_k = [torch.zeros(shape, device=self.device, dtype=torch.bfloat16) for _ in range(s.n_layer)]
        self.cache_v = [torch.zeros(shape, device=self.device, dtype=torch.bfloat16) for _ in range(s.n_layer)]
        self.cache_max_len = max_len

    def alloc_cache_for(self, nseq: int, need: int) -> None:
        """Cache for a decode that will reach `need` positions: rows of the next multiple of 256 slots, not the reference's
        block_size (quirk Q5) -- the decode attention's score buffer (shared memory) follows the row length: 1.97 -> 1.89 ms per
        token step of 128 sequences with 768-slot rows instead of 2048.  An existing cache is kept while it is long enough and at most twice too
        long (a re-allocation also costs a re-capture of the token-step graph)."""
        want = -(-int(need) // 256) * 256
        if self.cache_k and self.cache_k[0].shape[0] == nseq and want <= self.cache_max_len <= 2 * want:
            return
        self.alloc_cache(nseq, want)

    def drop_cache(self) -> None:
        self.cache_k, self.cache_v, self.cache_max_len = [], [], 0

    def _write_prefill_cache(self, layer: int, qkv: Tensor, B: int, T: int) -> None:
        seq = self._buf("pf_seq", (B * T,), torch.int32)
        pos = self._buf("pf_pos", (B * T,), torch.int32)
        if layer == 0:
            seq.copy_(torch.arange(B, device=self.device, dtype=torch.int32).repeat_interleave(T))
            pos.copy_(torch.arange(T, device=self.device, dtype=torch.int32).repeat(B))
        ops.kv_cache_write(qkv, self.cache_k[layer], self.cache_v[layer], seq, pos, self.spec.n_query_groups, self.spec.qpk)

    def prefill_packed(self, ids: Tensor, cu_seqlens: Tensor, positions: Tensor, seq_ids: Tensor, max_len: int,
                       packed: Dict[int, Dict[str, Tensor]]) -> Tensor:
        """Causal forward over prompts packed into one row without pad tokens (ids int64 [M]); K/V of every token go to the
        cache slot (seq_ids[m], positions[m]).  Returns ln_f output bf16 [M, C]."""
        assert self.cache_k and self.cache_max_len >= max_len
        return self.forward_hidden(ids.view(1, -1), packed, save=False, write_cache=True,
                                   varlen=dict(cu_seqlens=cu_seqlens, positions=positions, seq_ids=seq_ids, max_len=max_len))

    def prefill(self, idx: Tensor, packed: Dict[int, Dict[str, Tensor]]) -> Tensor:
        """Full causal forward over right-padded prompts [B, T]; K/V of every position go to the cache."""
        assert self.cache_k and self.cache_k[0].shape[0] >= idx.shape[0] and self.cache_max_len >= idx.shape[1]
        return self.forward_hidden(idx, packed, save=False, write_cache=True)

    def decode_plan_supported(self, n_seq: int, packed: Dict[int, Dict[str, Tensor]]) -> bool:
        """The persistent decode kernel (csrc/decode.cu) covers the base and adapter decoders (a LoRA model decodes through it once
        merged, generate/lora.py:112 merges before generating; un-merged LoRA factors take the per-GEMM path).  OPT-IN
        (DH_DECODE_MEGA=1): measured on B200 it is slower than the per-GEMM graph (2.39 vs 1.80 ms per token step of 128 sequences,
        profiles/r02_decode_mega_trace.txt: 8 grid barriers per layer cost 28 us of waiting)."""
        import os
        s = self.spec
        if os.environ.get("DH_DECODE_MEGA", "0") != "1" or not self.cache_k:
            return False
        if any(("qkv_a" in d or "proj_a" in d) for d in packed.values()):
            return False
        return (1 <= n_seq <= 128 and s.n_embd % 128 == 0 and s.qkv_out % 128 == 0 and (2 * s.intermediate_size) % 128 == 0
                and s.qpk <= 8 and s.n_layer <= 64 and self.cache_max_len % 64 == 0)

    def build_decode_plan(self, n_seq: int, positions: Tensor, packed: Dict[int, Dict[str, Tensor]], pkv_all: Optional[Tensor]):
        """Stage table + TMA descriptors of `dh_decode_layers` for the current cache, the `dec_*` activation buffers of this
        batch size, `positions` (a device tensor whose ADDRESS is baked in) and the packed adapter tensors."""
        import ctypes as C
        import os
        from . import _lib
        s = self.spec
        Cd, I = s.n_embd, s.intermediate_size
        bf, f32 = torch.bfloat16, torch.float32
        assert positions.dtype == torch.int32 and positions.is_contiguous() and positions.numel() == n_seq
        bufs = dict(x=self._buf("dec_x", (n_seq, Cd), f32), x2=self._buf("dec_x2", (n_seq, Cd), f32), n=self._buf("dec_n", (n_seq, Cd), bf),
                    qkv=self._buf("dec_qkv", (n_seq, s.qkv_out), bf), y=self._buf("dec_y", (n_seq, Cd), bf), act=self._buf("dec_act", (n_seq, I), bf),
                    hidden=self._buf("dec_hidden", (n_seq, Cd), bf))
        if "dec_barrier" not in self._bufs:
            self._bufs["dec_barrier"] = torch.zeros(2, device=self.device, dtype=torch.int64)     # arrive counter, launch generation
        barrier = self._bufs["dec_barrier"]
        layers = (_lib.DecodeLayer * s.n_layer)()
        keep = [positions, barrier, self.cos, self.sin, self.ln_f, pkv_all] + list(bufs.values())
        for i, lw in enumerate(self.layers):
            L, pk = layers[i], packed.get(i, {})
            L.wqkv, L.wproj, L.w12, L.wmlp = lw["wqkv"].data_ptr(), lw["wproj"].data_ptr(), lw["w12"].data_ptr(), lw["wmlp"].data_ptr()
            L.norm2 = lw["norm2"].data_ptr()
            L.norm_next = (self.layers[i + 1]["norm1"] if i + 1 < s.n_layer else self.ln_f).data_ptr()
            L.cache_k, L.cache_v = self.cache_k[i].data_ptr(), self.cache_v[i].data_ptr()
            if "prefix" in pk:
                L.prefix_kv, L.gate = pkv_all[i - s.adapter_start_layer].data_ptr(), pk["gate"].data_ptr()
                keep.append(pk["gate"])
        d = _lib.DecodeDesc()
        d.n_layer, d.n_seq, d.C, d.I, d.n_groups, d.q_per_kv = s.n_layer, n_seq, Cd, I, s.n_query_groups, s.qpk
        d.max_len, d.prefix_len, d.eps = self.cache_max_len, (s.adapter_prompt_length if s.variant == "adapter" else 0), s.norm_eps
        for k, t in bufs.items():
            setattr(d, k, t.data_ptr())
        d.positions, d.rope_cos, d.rope_sin, d.barrier = positions.data_ptr(), self.cos.data_ptr(), self.sin.data_ptr(), barrier.data_ptr()
        d.layers = C.cast(layers, C.POINTER(_lib.DecodeLayer))
        d.ws, d.ws_bytes = 0, 0
        if os.environ.get("DH_DECODE_TRACE", "0") == "1":            # developer knob: per-stage globaltimer stamps (scripts/decode_mega_trace.py)
            tr = self._bufs["dec_trace"] = torch.zeros(3 * s.n_layer * 24 + 8 * ops.sm_count() * 2, device=self.device, dtype=torch.int64)
            d.trace = tr.data_ptr()
            keep.append(tr)
        ws_bytes = ops.decode_workspace_bytes(d)
        ws = self._cap_buf("dec_mega_ws", (ws_bytes // 4,), f32)
        d.ws, d.ws_bytes = ws.data_ptr(), ws.numel() * 4
        keep += [ws, self.cache_k, self.cache_v]
        plan = ops.DecodePlan(d, keep)
        plan.n_seq = n_seq
        return plan

    def decode_step(self, tokens: Tensor, positions: Tensor, packed: Dict[int, Dict[str, Tensor]],
                    pkv_all: Optional[Tensor] = None, plan=None) -> Tensor:
        """One new token per sequence: tokens int64 [n], positions int32 [n] (0-based position of that token).
        Returns ln_f output bf16 [n, C].  Allocation-free after the first call of a given n (all buffers come from
        `_buf`), so a whole step can be captured into a CUDA graph (generate.DecodeLoop); `pkv_all` = the adapter
        prefix K/V of `project_prefixes()` when the caller hoisted it out of the token loop."""
        s = self.spec
        n_seq = tokens.shape[0]
        C, I, G, qpk = s.n_embd, s.intermediate_size, s.n_query_groups, s.qpk
        bf, f32 = torch.bfloat16, torch.float32
        x = self._buf("dec_x", (n_seq, C), f32)
        x2 = self._buf("dec_x2", (n_seq, C), f32)
        n = self._buf("dec_n", (n_seq, C), bf)
        qkv = self._buf("dec_qkv", (n_seq, s.qkv_out), bf)
        y = self._buf("dec_y", (n_seq, C), bf)
        act = self._buf("dec_act", (n_seq, I), bf)
        seq_ids = self._buf("dec_seq", (n_seq,), torch.int32)
        lens = self._buf("dec_lens", (n_seq,), torch.int32)
        if self._bufs.get("dec_seq_n") != n_seq:
            seq_ids.copy_(torch.arange(n_seq, device=self.device, dtype=torch.int32))
            self._bufs["dec_seq_n"] = n_seq
        torch.add(positions, 1, out=lens)
        ops.embedding_fwd(tokens.contiguous(), self.wte, x)
        if pkv_all is None:
            pkv_all = self._project_prefixes(packed, "dec_pkv_all")
        hidden = self._buf("dec_hidden", (n_seq, C), bf)
        if plan is not None:
            # every layer in ONE persistent kernel (csrc/decode.cu); `plan` was built for these buffers, this cache and `positions`
            assert plan.n_seq == n_seq
            ops.rmsnorm_fwd(x, self.layers[0]["norm1"], n, None, s.norm_eps)
            ops.decode_layers(plan)
            return hidden
        # The two residual GEMMs (attention proj, mlp proj; N = C gives only C / 128 tiles) run split-K over all SMs and
        # leave fp32 partial products; the norm that follows sums them, adds the residual and normalises in one launch.
        sk_ok = n_seq <= 256 and C % 128 == 0
        s_proj, s_mlp = (ops.decode_split_k(C, C), ops.decode_split_k(C, I)) if sk_ok else (1, 1)
        ws = self._buf("dec_splitk", (max(s_proj, s_mlp), n_seq, C), f32) if (s_proj > 1 or s_mlp > 1) else None
        s_qkv = ops.decode_split_k(s.qkv_out, C) if (n_seq <= 256 and s.qkv_out % 128 == 0) else 1
        ws_q = self._buf("dec_splitk_qkv", (s_qkv, n_seq, s.qkv_out), f32) if s_qkv > 1 else None
        # fc_1|fc_2 (2 I / 128 = 88 tiles on 148 SMs) as stream-K over all SMs, the CTA that completes a tile applying SwiGLU
        # (csrc/gemm.cu gemm_streamk_kernel; the workspace of partial tiles + arrival counters is zeroed once and belongs to this GEMM
        # shape).  OPT-IN (DH_DECODE_STREAMK=1): measured SLOWER at every batch -- token step 1.41 vs 1.16 ms at 16 sequences, 1.58 vs 1.27
        # at 48, 2.18 vs 1.64 at 128, 3.48 vs 2.54 at 256: an fp32 partial tile is twice the bytes of the weight tile it came from, and
        # parking + re-reading it through L2 costs more than the 60 idle SMs of the 88-tile grid.
        ws12 = None
        if os.environ.get("DH_DECODE_STREAMK", "0") == "1" and n_seq <= 256 and (2 * I) % 128 == 0 and C % 64 == 0:
            nb = ops.gemm_streamk_workspace_bytes(n_seq, 2 * I, C)
            key = ("persistent", "dec_streamk_ws12", n_seq)
            ws12 = self._bufs.get(key)
            if ws12 is None or ws12.numel() < nb:
                with torch.inference_mode(False):
                    ws12 = self._bufs[key] = torch.zeros(nb, device=self.device, dtype=torch.uint8)
        ops.rmsnorm_fwd(x, self.layers[0]["norm1"], n, None, s.norm_eps)
        for i, lw in enumerate(self.layers):
            pk = packed.get(i, {})
            a2q = b2q = None
            if "qkv_a" in pk:
                a2q = self._buf("dec_as", (n_seq, LORA_PAD), bf)
                ops.gemm(n, pk["qkv_a"], a2q)
                b2q = pk["qkv_b"]
            if s_qkv > 1:
                # split-K partial products; one consumer sums them, applies RoPE and writes the qkv row and the KV cache
                ops.gemm(n, lw["wqkv"], ws_q, A2=a2q, B2=b2q, epilogue=ops.EPI_F32, split_k=s_qkv, b_frozen=True)
                ops.splitk_qkv_rope_cache(ws_q, qkv, self.cache_k[i], self.cache_v[i], seq_ids, positions, self.cos, self.sin, G, qpk)
            else:
                ops.gemm(n, lw["wqkv"], qkv, A2=a2q, B2=b2q, epilogue=ops.EPI_QKV_ROPE, rope_cos=self.cos, rope_sin=self.sin,
                         positions=positions, seq_len=1, q_per_kv=qpk, b_frozen=True)
                ops.kv_cache_write(qkv, self.cache_k[i], self.cache_v[i], seq_ids, positions, G, qpk)
            prefix_kv = gate = None
            if "prefix" in pk:
                prefix_kv = pkv_all[i - s.adapter_start_layer]
                gate = pk["gate"]
            ops.attention_decode(qkv, self.cache_k[i], self.cache_v[i], lens, y, G, qpk, prefix_kv, gate,
                                 s.adapter_prompt_length if prefix_kv is not None else 0)
            a2 = b2 = None
            if "proj_a" in pk:
                a2 = self._buf("dec_aps", (n_seq, LORA_PAD), bf)
                ops.gemm(y, pk["proj_a"], a2)
                b2 = pk["proj_b"]
            if s_proj > 1:
                ops.gemm(y, lw["wproj"], ws[:s_proj], A2=a2, B2=b2, epilogue=ops.EPI_F32, split_k=s_proj, b_frozen=True)
                ops.splitk_resid_rmsnorm(ws[:s_proj], x, lw["norm2"], x2, n, s.norm_eps)
            else:
                ops.gemm(y, lw["wproj"], x2, A2=a2, B2=b2, epilogue=ops.EPI_RESID_F32, aux=x, b_frozen=True)
                ops.rmsnorm_fwd(x2, lw["norm2"], n, None, s.norm_eps)
            ops.gemm(n, lw["w12"], None, epilogue=ops.EPI_SWIGLU, out2=act, b_frozen=True, workspace=ws12)
            last = i + 1 == len(self.layers)
            w_next, n_next = (self.ln_f, hidden) if last else (self.layers[i + 1]["norm1"], n)
            if s_mlp > 1:
                ops.gemm(act, lw["wmlp"], ws[:s_mlp], epilogue=ops.EPI_F32, split_k=s_mlp, b_frozen=True)
                ops.splitk_resid_rmsnorm(ws[:s_mlp], x2, w_next, x, n_next, s.norm_eps)
            else:
                ops.gemm(act, lw["wmlp"], x, epilogue=ops.EPI_RESID_F32, aux=x2, b_frozen=True)
                ops.rmsnorm_fwd(x, w_next, n_next, None, s.norm_eps)
        return hidden
