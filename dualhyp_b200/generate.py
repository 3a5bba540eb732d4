"""Greedy decoding on the B200 engine.

`generate()` mirrors generate/base.py:19-82 (one sequence, KV cache, temperature/top_k accepted): with the
reference's `top_k=1` setting the sampling step keeps a single candidate, so the next token is the arg-max
(exact ties resolve to the lowest index here instead of randomly) -- computed on the device
by `dh_argmax` with no per-token host sync unless an `eos_id` has to be polled.

`generate_batch()` is the batched form: right-padded prompts prefilled together,
one decode step per new token for all live sequences, per-sequence lengths and the EOS flags on the device.
The per-token work (lm_head -> arg-max -> record -> 22 decoder layers over the KV cache -> position bump) is
captured ONCE into a CUDA graph (`DecodeLoop`) and replayed per token: the reference's loop costs ~400 kernel
launches and a host sync per token (generate/base.py:57-80); here a token is one `cudaGraphLaunch`."""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch

from . import ops

Tensor = torch.Tensor


class DecodeLoop:
    """Device-resident state of a batched greedy decode plus the captured graph of one token step.

    State (all on the device, fixed addresses so that the graph can be replayed and re-used across calls):
    `last` = ln_f hidden state of each sequence's newest token, `pos` = position the next token will take,
    `out_ids[:, step]` = generated ids, `done` = EOS seen, `step` = tokens generated so far, `eos` = the id to
    stop at (-1: none).  One `body()` = one new token for every sequence."""

    def __init__(self, eng, packed: Dict[int, Dict[str, Tensor]], n: int, max_new: int):
        dev = eng.device
        self.eng, self.n, self.max_new = eng, n, max_new
        self.packed = packed
        self.pkv_all = eng._project_prefixes(packed, "dec_pkv_all")
        self.last = eng._buf("dec_hidden", (n, eng.spec.n_embd), torch.bfloat16)      # decode_step writes its result here
        self.logits = torch.empty((n, eng.spec.vocab), dtype=torch.bfloat16, device=dev)
        self.nxt = torch.zeros(n, dtype=torch.int64, device=dev)
        self.pos = torch.zeros(n, dtype=torch.int32, device=dev)
        self.out_ids = torch.zeros((n, max_new), dtype=torch.int64, device=dev)
        self.done = torch.zeros(n, dtype=torch.bool, device=dev)
        self.step = torch.zeros(1, dtype=torch.int64, device=dev)
        self.eos = torch.full((1,), -1, dtype=torch.int64, device=dev)
        self.eos_id: Optional[int] = None
        self.kernels_per_token = 0
        self.trainable_sig = None
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.cache_sig = None
        self.plan = None              # persistent-kernel plan of the token step (engine.build_decode_plan), tied to the cache
        self.plan_sig = None
        self.sample = None            # (temperature, top_k) of generate/base.py:62-70 when it is not the top-1 setting; None = arg-max

    def refresh(self, packed: Dict[int, Dict[str, Tensor]]) -> None:
        """New values of the trainable tensors into the buffers the captured graph reads."""
        for i, d in packed.items():
            for k, t in d.items():
                self.packed[i][k].copy_(t)
        self.pkv_all = self.eng._project_prefixes(self.packed, "dec_pkv_all")

    def reset(self, last: Tensor, lens: Tensor, eos_id: Optional[int]) -> None:
        self.last.copy_(last)
        self.pos.copy_(lens)
        self.out_ids.zero_()
        self.done.zero_()
        self.step.zero_()
        self.eos_id = eos_id
        self.eos.fill_(-1 if eos_id is None else int(eos_id))

    def _cache_signature(self):
        # everything the captured launches / the plan bake in about the cache: row count, row length (the kernels' max_len stride)
        # and the K and V base pointers -- the caching allocator can hand a re-allocated cache of another shape the same pointers
        eng = self.eng
        return (self.n, eng.cache_max_len, tuple(t.data_ptr() for t in eng.cache_k), tuple(t.data_ptr() for t in eng.cache_v))

    def _ensure_plan(self) -> None:
        eng = self.eng
        sig = self._cache_signature()
        if sig == self.plan_sig:
            return
        self.plan = eng.build_decode_plan(self.n, self.pos, self.packed, self.pkv_all) if eng.decode_plan_supported(self.n, self.packed) else None
        self.plan_sig = sig

    def body(self) -> None:
        eng = self.eng
        self._ensure_plan()
        l0 = ops.launches()
        eng.lm_head(self.last, out=self.logits)
        if self.sample is None:
            ops.argmax(self.logits, self.nxt)
        else:
            # generate/base.py:62-70 for every sequence: temperature, keep the top_k logits, softmax, multinomial
            temperature, top_k = self.sample
            lg = self.logits.float() / temperature
            if top_k is not None:
                v, _ = torch.topk(lg, min(int(top_k), lg.size(-1)))
                lg = torch.where(lg < v[:, [-1]], torch.full_like(lg, float("-inf")), lg)
            self.nxt.copy_(torch.multinomial(torch.softmax(lg, dim=-1), num_samples=1).squeeze(1))
        self.out_ids.index_copy_(1, self.step, self.nxt.unsqueeze(1))
        self.done.logical_or_(self.nxt == self.eos)
        hidden = eng.decode_step(self.nxt, self.pos, self.packed, self.pkv_all, plan=self.plan)
        assert hidden.data_ptr() == self.last.data_ptr()
        self.pos.add_(1)
        self.step.add_(1)
        self.kernels_per_token = ops.launches() - l0          # this library's kernels in one token step

    def capture(self) -> None:
        """Record one body() into a graph.  Must follow at least one eager body() of the same shape (first-use
        initialisation inside the C ABI -- function attributes, the device query -- is not capturable)."""
        sig = self._cache_signature()
        if self.graph is not None and sig == self.cache_sig:
            return
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        l0 = ops.launches()
        with torch.cuda.graph(g):
            self.body()
        ops._count(l0 - ops.launches())                        # capturing launches nothing
        self.graph, self.cache_sig = g, sig

    def run(self, n_tokens: int, use_graph: bool, poll_every: int) -> int:
        """Generate up to n_tokens; returns how many steps ran.  The first token of a fresh loop runs eagerly."""
        polling = self.eos_id is not None and poll_every > 0
        steps = 0
        while steps < n_tokens:
            if use_graph and (self.graph is not None or steps > 0):
                self.capture()
                self.graph.replay()
                ops._count(self.kernels_per_token)             # keep ops.launches() truthful under replay
            else:
                self.body()
            steps += 1
            if polling and steps % poll_every == 0 and bool(self.done.all()):
                break
        return steps


def _decode_loop(model, eng, n: int, max_new: int) -> DecodeLoop:
    """One loop (and one captured graph) per batch size; `out_ids` has room for at least 256 new tokens so that
    calls with different `max_new_tokens` (generate/base.py is called once per utterance) share it.  The bf16 operand
    forms of the trainable tensors are re-packed only when a tensor changed (data pointer / version counter): an
    inference service decodes thousands of batches against the same adapter."""
    names, tensors = model._trainables()
    sig = tuple((t.data_ptr(), t._version) for t in tensors)
    loops = eng._bufs.setdefault("decode_loops", {})
    loop = loops.get(n)
    if loop is None or loop.max_new < max_new:
        loop = loops[n] = DecodeLoop(eng, eng.pack_trainable(dict(zip(names, tensors))), n, max(256, max_new))
    elif loop.trainable_sig != sig:
        loop.refresh(eng.pack_trainable(dict(zip(names, tensors))))
    loop.trainable_sig = sig
    return loop


@torch.inference_mode()
def generate(model, idx: Tensor, max_returned_tokens: int, *, temperature: float = 1.0, top_k: Optional[int] = None,
             eos_id: Optional[int] = None) -> Tensor:
    T = idx.size(0)
    assert max_returned_tokens > T
    if model.max_seq_length < max_returned_tokens - 1:
        raise NotImplementedError(f"max_seq_length {model.max_seq_length} needs to be >= {max_returned_tokens - 1}")
    if top_k == 1:
        # the DualHyp inference drivers' setting (inference/ger.py:78): one candidate survives the top-k cut, i.e. arg-max
        return generate_batch(model, [idx], max_returned_tokens - T, eos_id=eos_id)[0]
    if temperature <= 0:
        raise ValueError("temperature must be positive")
    # generate/base.py:62-70 proper: temperature + top-k (None = whole vocabulary) + multinomial, on the device with torch's CUDA
    # generator (so the draws differ from the reference's CPU stream); eager token loop, no graph
    return generate_batch(model, [idx], max_returned_tokens - T, eos_id=eos_id, use_graph=False, sample=(float(temperature), top_k))[0]


@torch.inference_mode()
def generate_batch(model, prompts: Sequence[Tensor], max_new_tokens: int, eos_id: Optional[int] = None,
                   poll_every: int = 8, use_graph: bool = True, return_device: bool = False, packed_prefill: bool = True,
                   sample: Optional[tuple] = None):
    """Greedy-decode every prompt (1-D int64 tensors, host or device).  Returns prompt + generated ids per sequence,
    cut after the first `eos_id` (inclusive) when given.  `return_device=True` returns the raw device tensor
    `[n, steps]` of generated ids instead (no device-to-host copy; used by bench.py's HBM-resident leg).
    `packed_prefill` (default): the prompts are prefilled as ONE packed row with cu_seqlens (no flops on pad tokens, SURVEY
    8f row 2); False right-pads them to the longest prompt like the reference's collate_fn -- both give the same bits.
    `sample = (temperature, top_k)`: draw every token like generate/base.py:62-70 instead of taking the arg-max (eager loop)."""
    eng = model._engine_ready()
    dev = eng.device
    n = len(prompts)
    lens = torch.tensor([int(p.numel()) for p in prompts], dtype=torch.int32)
    Tmax = int(lens.max())
    total = Tmax + max_new_tokens
    assert total - 1 <= model.max_seq_length, "prompt + new tokens exceed max_seq_length"
    lens_d = lens.to(dev)
    eng.alloc_cache_for(n, total)
    loop = _decode_loop(model, eng, n, max_new_tokens)
    stage = (lambda t: t.pin_memory().to(dev, non_blocking=True)) if dev.type == "cuda" else (lambda t: t)
    if packed_prefill:
        # prompts packed into one row (no pad tokens): cu_seqlens / per-token positions / sequence ids are built on the host
        cu = torch.zeros(n + 1, dtype=torch.int32)
        cu[1:] = lens.cumsum(0)
        M = int(cu[-1])
        pos_h = torch.cat([torch.arange(int(L), dtype=torch.int32) for L in lens.tolist()])
        seq_h = torch.repeat_interleave(torch.arange(n, dtype=torch.int32), lens.long())
        if all(p.is_cuda for p in prompts):
            ids = torch.cat(list(prompts))
        else:
            ids = stage(torch.cat([p.cpu() for p in prompts]))
        meta = stage(torch.cat((cu, pos_h, seq_h)))                                        # one small H2D copy
        cu_d, pos_d, seq_d = meta[: n + 1], meta[n + 1: n + 1 + M], meta[n + 1 + M:]
        hidden = eng.prefill_packed(ids, cu_d, pos_d, seq_d, Tmax, loop.packed)
        last = hidden[(cu_d[1:] - 1).long()]                                              # ln_f state of each prompt's last token
    else:
        if all(p.is_cuda for p in prompts):
            ids = torch.nn.utils.rnn.pad_sequence(list(prompts), batch_first=True, padding_value=0)
        else:
            prev = eng._bufs.get(("pinned_ids_event", n))
            if prev is not None:
                prev.synchronize()
            stage = eng._bufs.get(("pinned_ids", n))                # pinned staging rows, allocated once per batch size
            if stage is None or stage.shape[1] < Tmax:
                stage = eng._bufs[("pinned_ids", n)] = torch.zeros((n, max(Tmax, model.max_seq_length)), dtype=torch.int64).pin_memory()
            ids_h = stage[:, :Tmax]
            ids_h.zero_()
            for i, p in enumerate(prompts):
                ids_h[i, : p.numel()] = p.cpu()
            ids = ids_h.to(dev, non_blocking=True)
            copied = torch.cuda.Event()
            copied.record()
            eng._bufs[("pinned_ids_event", n)] = copied             # the next call waits for this copy before re-using the rows
        hidden = eng.prefill(ids, loop.packed).view(n, Tmax, -1)
        last = hidden[torch.arange(n, device=dev), (lens_d - 1).long()]                    # ln_f state of each prompt's last token
    loop.reset(last, lens_d, eos_id)
    loop.sample = sample
    try:
        steps = loop.run(max_new_tokens, use_graph and sample is None, poll_every)
    finally:
        loop.sample = None
    if return_device:
        return loop.out_ids[:, :steps]
    out_ids = loop.out_ids[:, :steps].cpu()
    result = []
    for i, p in enumerate(prompts):
        gen = out_ids[i]
        if eos_id is not None:
            hit = (gen == eos_id).nonzero()
            if hit.numel():
                gen = gen[: int(hit[0]) + 1]
        result.append(torch.cat((p.cpu().to(torch.int64), gen)))
    return result
