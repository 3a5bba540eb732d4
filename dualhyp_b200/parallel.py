"""Data parallelism for the PEFT fine-tune: every rank holds the full frozen bf16 model and the
full trainable set; per optimizer step ONE all-reduce over a single flat fp32 bucket of the trainable gradients
(1.6 MB adapter ... 27 MB RelPrompt), then a fused AdamW over the same flat bucket.  The reference reaches for
FSDP + activation checkpointing when --d > 1 (finetune/ger.py:67-73); for a 1.1 B frozen model on 180 GB
GPUs that is pure overhead and is deliberately not reproduced.  Inference shards utterances by index, no comm."""
from __future__ import annotations

from typing import Iterable, List, Optional

import torch
import torch.distributed as dist

from . import ops

Tensor = torch.Tensor


class FlatTrainables:
    """Re-homes the trainable parameters (and their .grad) as views of two flat fp32 buffers."""

    def __init__(self, params: Iterable[torch.nn.Parameter]):
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        assert self.params, "no trainable parameters"
        dev = self.params[0].device
        n = sum(p.numel() for p in self.params)
        self.flat = torch.empty(n, device=dev, dtype=torch.float32)
        self.grad = torch.zeros(n, device=dev, dtype=torch.float32)
        o = 0
        for p in self.params:
            k = p.numel()
            self.flat[o:o + k].copy_(p.data.reshape(-1).float())
            p.data = self.flat[o:o + k].view(p.shape)
            p.grad = self.grad[o:o + k].view(p.shape)
            o += k
        self.numel = n
        self.exp_avg: Optional[Tensor] = None
        self.exp_avg_sq: Optional[Tensor] = None
        self.step_count = 0

    def zero_grad(self) -> None:
        self.grad.zero_()
        o = 0
        for p in self.params:          # autograd may have replaced .grad if it was set to None by a caller
            k = p.numel()
            if p.grad is None or p.grad.data_ptr() != self.grad[o:o + k].data_ptr():
                p.grad = self.grad[o:o + k].view(p.shape)
            o += k

    def allreduce(self, async_op: bool = False):
        """Sum over ranks (callers pre-scale the loss by 1/world or pass grad_scale=1/world to adamw_step)."""
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            return dist.all_reduce(self.grad, op=dist.ReduceOp.SUM, async_op=async_op)
        return None

    def adamw_step(self, lr: float, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.01, grad_scale: float = 1.0) -> None:
        if self.exp_avg is None:
            self.exp_avg = torch.zeros_like(self.flat)
            self.exp_avg_sq = torch.zeros_like(self.flat)
        self.step_count += 1
        ops.adamw_step(self.flat, self.grad, self.exp_avg, self.exp_avg_sq, lr, betas[0], betas[1], eps, weight_decay, self.step_count,
                       grad_scale)
        # the kernel wrote the parameters through the flat buffer: tell autograd / every (data_ptr, _version) cache key
        # (generate._decode_loop, serve.DecodeService, GPT._frozen_sig) that they changed
        for p in self.params:
            torch.autograd.graph.increment_version(p)


def shard_indices(n_items: int, rank: int, world: int) -> List[int]:
    """Utterance sharding for inference: item i goes to rank i % world."""
    return list(range(rank, n_items, world))
