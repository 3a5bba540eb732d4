"""Greedy decoding on the B200 engine.

`generate()` mirrors generate/base.py:19-82 (one sequence, KV cache, temperature/top_k accepted): with the
reference's `top_k=1` setting the sampling step keeps a single candidate, so the next token is the arg-max
(exact ties resolve to the lowest index here instead of randomly) -- computed on the device
by `dh_argmax` with no per-token host sync unless an `eos_id` has to be polled.

`generate_batch()` is the batched form: right-padded prompts prefilled together,
one decode step per new token for all live sequences, per-sequence lengths on the device."""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from . import ops

Tensor = torch.Tensor


@torch.inference_mode()
def generate(model, idx: Tensor, max_returned_tokens: int, *, temperature: float = 1.0, top_k: Optional[int] = None,
             eos_id: Optional[int] = None) -> Tensor:
    T = idx.size(0)
    assert max_returned_tokens > T
    if model.max_seq_length < max_returned_tokens - 1:
        raise NotImplementedError(f"max_seq_length {model.max_seq_length} needs to be >= {max_returned_tokens - 1}")
    if top_k is not None and top_k != 1:
        raise NotImplementedError("the DualHyp inference drivers decode with top_k=1 (inference/ger.py:78); sampling is not on this path")
    out = generate_batch(model, [idx], max_returned_tokens - T, eos_id=eos_id)[0]
    return out


@torch.inference_mode()
def generate_batch(model, prompts: Sequence[Tensor], max_new_tokens: int, eos_id: Optional[int] = None,
                   poll_every: int = 8) -> List[Tensor]:
    """Greedy-decode every prompt (1-D int64 tensors).  Returns prompt + generated ids per sequence, cut after
    the first `eos_id` (inclusive) when given."""
    eng = model._engine_ready()
    dev = eng.device
    names, tensors = model._trainables()
    packed = eng.pack_trainable(dict(zip(names, tensors)))
    n = len(prompts)
    lens = torch.tensor([int(p.numel()) for p in prompts], dtype=torch.int32)
    Tmax = int(lens.max())
    total = Tmax + max_new_tokens
    assert total - 1 <= model.max_seq_length, "prompt + new tokens exceed max_seq_length"
    ids = torch.zeros((n, Tmax), dtype=torch.int64)
    for i, p in enumerate(prompts):
        ids[i, : p.numel()] = p.cpu()
    ids = ids.to(dev)
    lens_d = lens.to(dev)
    eng.alloc_cache(n, max(model.max_seq_length, total))
    hidden = eng.prefill(ids, packed).view(n, Tmax, -1)
    last = hidden[torch.arange(n, device=dev), (lens_d - 1).long()].contiguous()       # ln_f state of each prompt's last token
    out_ids = torch.zeros((n, max_new_tokens), dtype=torch.int64, device=dev)
    nxt = torch.empty(n, dtype=torch.int64, device=dev)
    logits = torch.empty((n, eng.spec.vocab), dtype=torch.bfloat16, device=dev)
    pos = lens_d.clone()                                                               # position of the token being generated
    done = torch.zeros(n, dtype=torch.bool, device=dev)
    steps = 0
    for step in range(max_new_tokens):
        eng.lm_head(last, out=logits)
        ops.argmax(logits, nxt)
        out_ids[:, step] = nxt
        steps = step + 1
        if eos_id is not None:
            done |= nxt == eos_id
            if (step + 1) % poll_every == 0 and bool(done.all()):
                break
        if step + 1 < max_new_tokens:
            last = eng.decode_step(nxt, pos, packed)
            pos = pos + 1
    out_ids = out_ids[:, :steps].cpu()
    result = []
    for i, p in enumerate(prompts):
        gen = out_ids[i]
        if eos_id is not None:
            hit = (gen == eos_id).nonzero()
            if hit.numel():
                gen = gen[: int(hit[0]) + 1]
        result.append(torch.cat((p.cpu().to(torch.int64), gen)))
    return result
