"""Continuous batching for the greedy decode.

The reference decodes one utterance at a time and syncs with the host after every token (inference/ger.py:60-83,
generate/base.py:57-80).  `generate_batch()` already decodes a batch under one CUDA graph, but a batch ends when its
LONGEST answer ends.  `DecodeService` keeps `n_slots` sequences in flight instead: every slot owns one row of the KV cache
and of the decode state; when a sequence stops (EOS or its own token budget) its slot is handed to the next waiting prompt,
which is prefilled (packed, no pad rows) straight into that cache row while the other slots keep their state.  The token
step is the same captured graph for the whole run -- per-slot counters, budgets and the stop flags live on the device, the
host only polls them every `poll_every` tokens.

Per-sequence results are the ones `generate_batch` gives (rows of a decode step do not interact: GEMM rows, per-sequence
attention), which tests/test_model_gpu.py checks."""
from __future__ import annotations

from collections import deque
from typing import Dict, List, Optional, Sequence, Union

import torch

from . import ops
from .generate import DecodeLoop

Tensor = torch.Tensor


class SlotLoop(DecodeLoop):
    """DecodeLoop whose sequences start and stop independently: `cnt[s]` tokens generated so far by slot s, `limit[s]`
    its budget, `done[s]` set by EOS or the budget.  A stopped (or empty) slot still rides through the graph -- its row is
    recomputed at a fixed position and its output lands in the spare last column -- which costs nothing extra: the step is
    bound by streaming the weights, not by the rows."""

    def __init__(self, eng, packed: Dict[int, Dict[str, Tensor]], n: int, max_new: int):
        super().__init__(eng, packed, n, max_new)
        dev = eng.device
        self.out_ids = torch.zeros((n, max_new + 1), dtype=torch.int64, device=dev)
        self.cnt = torch.zeros(n, dtype=torch.int64, device=dev)
        self.limit = torch.zeros(n, dtype=torch.int64, device=dev)
        self.done.fill_(True)

    def body(self) -> None:
        eng = self.eng
        self._ensure_plan()
        l0 = ops.launches()
        eng.lm_head(self.last, out=self.logits)
        ops.argmax(self.logits, self.nxt)
        self.out_ids.scatter_(1, self.cnt.unsqueeze(1), self.nxt.unsqueeze(1))
        self.cnt.add_(self.done.logical_not())                       # the token counts only for a live slot
        self.done.logical_or_(self.nxt == self.eos)
        self.done.logical_or_(self.cnt >= self.limit)
        hidden = eng.decode_step(self.nxt, self.pos, self.packed, self.pkv_all, plan=self.plan)
        assert hidden.data_ptr() == self.last.data_ptr()
        self.pos.add_(self.done.logical_not())                       # a stopped slot stays where it is
        self.kernels_per_token = ops.launches() - l0

    def admit(self, slots: Tensor, last: Tensor, lens: Tensor, limits: Tensor) -> None:
        """New sequences into `slots` (int64 [k]): ln_f state of their last prompt token, prompt lengths, token budgets."""
        self.last.index_copy_(0, slots, last)
        self.pos.index_copy_(0, slots, lens.to(torch.int32))
        self.cnt.index_fill_(0, slots, 0)
        self.limit.index_copy_(0, slots, limits)
        self.done.index_fill_(0, slots, False)


class DecodeService:
    """`n_slots` sequences decoded concurrently with slots refilled as sequences finish.

        svc = DecodeService(model, n_slots=64, max_new_tokens=150, eos_id=tokenizer.eos_id)
        outputs = svc.run(prompts)          # prompt + generated ids per prompt, in the order given

    `max_new_tokens` is the largest budget a request may ask for (inference/ger.py:71 uses 150); `run()` takes one budget
    for all prompts or one per prompt."""

    def __init__(self, model, n_slots: int, max_new_tokens: int, eos_id: Optional[int] = None, poll_every: int = 8,
                 use_graph: bool = True):
        assert n_slots >= 1 and max_new_tokens >= 1 and poll_every >= 1
        self.model, self.n, self.max_new = model, n_slots, max_new_tokens
        self.eos_id, self.poll_every, self.use_graph = eos_id, poll_every, use_graph
        self.stats = {"token_steps": 0, "prefills": 0, "prefill_tokens": 0, "generated": 0}

    def _loop(self, eng) -> SlotLoop:
        names, tensors = self.model._trainables()
        sig = tuple((t.data_ptr(), t._version) for t in tensors)
        loops = eng._bufs.setdefault("serve_loops", {})
        loop = loops.get(self.n)
        if loop is None or loop.max_new < self.max_new:
            loop = loops[self.n] = SlotLoop(eng, eng.pack_trainable(dict(zip(names, tensors))), self.n, self.max_new)
        elif loop.trainable_sig != sig:
            loop.refresh(eng.pack_trainable(dict(zip(names, tensors))))
        loop.trainable_sig = sig
        loop.eos_id = self.eos_id
        loop.eos.fill_(-1 if self.eos_id is None else int(self.eos_id))
        loop.done.fill_(True)                                        # every slot empty: a finite state that decodes harmlessly
        loop.pos.zero_()
        loop.cnt.zero_()
        loop.last.zero_()
        return loop

    def _admit(self, eng, loop: SlotLoop, slots: List[int], prompts: List[Tensor], limits: List[int]) -> None:
        dev = eng.device
        lens = torch.tensor([int(p.numel()) for p in prompts], dtype=torch.int32)
        k = len(prompts)
        cu = torch.zeros(k + 1, dtype=torch.int32)
        cu[1:] = lens.cumsum(0)
        M = int(cu[-1])
        pos_h = torch.cat([torch.arange(int(L), dtype=torch.int32) for L in lens.tolist()])
        seq_h = torch.repeat_interleave(torch.tensor(slots, dtype=torch.int32), lens.long())     # cache row = slot
        stage = (lambda t: t.pin_memory().to(dev, non_blocking=True)) if dev.type == "cuda" else (lambda t: t)
        if all(p.is_cuda for p in prompts):
            ids = torch.cat(list(prompts))
        else:
            ids = stage(torch.cat([p.cpu() for p in prompts]))
        meta = stage(torch.cat((cu, pos_h, seq_h, lens, torch.tensor(slots, dtype=torch.int32), torch.tensor(limits, dtype=torch.int32))))
        o = 0
        cu_d, o = meta[o: o + k + 1], o + k + 1
        pos_d, o = meta[o: o + M], o + M
        seq_d, o = meta[o: o + M], o + M
        lens_d, o = meta[o: o + k], o + k
        slots_d, o = meta[o: o + k].long(), o + k
        limits_d = meta[o: o + k].long()
        hidden = eng.prefill_packed(ids, cu_d, pos_d, seq_d, int(lens.max()), loop.packed)
        loop.admit(slots_d, hidden[(cu_d[1:] - 1).long()], lens_d, limits_d)
        self.stats["prefills"] += 1
        self.stats["prefill_tokens"] += M

    @torch.inference_mode()
    def run(self, prompts: Sequence[Tensor], max_new_tokens: Union[None, int, Sequence[int]] = None) -> List[Tensor]:
        model = self.model
        eng = model._engine_ready()
        if max_new_tokens is None:
            budgets = [self.max_new] * len(prompts)
        elif isinstance(max_new_tokens, int):
            budgets = [max_new_tokens] * len(prompts)
        else:
            budgets = [int(b) for b in max_new_tokens]
        assert len(budgets) == len(prompts) and all(1 <= b <= self.max_new for b in budgets), "token budget outside [1, max_new_tokens]"
        for p, b in zip(prompts, budgets):
            if p.numel() + b - 1 > model.max_seq_length:
                raise ValueError(f"prompt of {p.numel()} tokens + {b} new tokens exceeds max_seq_length {model.max_seq_length}")
        eng.alloc_cache_for(self.n, max(p.numel() + b for p, b in zip(prompts, budgets)))
        loop = self._loop(eng)
        waiting = deque(range(len(prompts)))
        owner = [-1] * self.n                                        # request index held by each slot
        results: List[Optional[Tensor]] = [None] * len(prompts)
        while waiting or any(o >= 0 for o in owner):
            free = [s for s in range(self.n) if owner[s] < 0]
            if free and waiting:
                take = [waiting.popleft() for _ in range(min(len(free), len(waiting)))]
                slots = free[: len(take)]
                self._admit(eng, loop, slots, [prompts[i] for i in take], [budgets[i] for i in take])
                for s, i in zip(slots, take):
                    owner[s] = i
            self.stats["token_steps"] += loop.run(self.poll_every, self.use_graph, 0)
            state = torch.stack((loop.done.long(), loop.cnt)).cpu()  # the one host sync per `poll_every` tokens
            fin = [s for s in range(self.n) if owner[s] >= 0 and bool(state[0, s])]
            if fin:
                rows = loop.out_ids[torch.tensor(fin, device=eng.device)].cpu()
                for r, s in zip(rows, fin):
                    i = owner[s]
                    gen = r[: int(state[1, s])]
                    results[i] = torch.cat((prompts[i].cpu().to(torch.int64), gen))
                    self.stats["generated"] += int(gen.numel())
                    owner[s] = -1
        return results  # type: ignore[return-value]
