"""`ger.utils` mirror: the pieces of reference ger/utils.py the DualHyp drivers import
(finetune/ger.py:28-35, inference/ger.py:25): chunked_cross_entropy (hot path), checkpoint helpers,
precision default, parameter counting, CSV logger factory, quantization context."""
from __future__ import annotations

import sys
from contextlib import contextmanager
from pathlib import Path
from typing import Any, Dict, List, Mapping, Optional, Union

import torch

from .. import ops
from .config import find_multiple  # noqa: F401  (re-exported like the reference)

Tensor = torch.Tensor


class _CrossEntropyFn(torch.autograd.Function):
    """Cross entropy over materialised logits [M, V] with ignore_index=-1, on the `dh_cross_entropy` kernel.
    denom > 0: sum/denom; denom == 0: mean over valid targets."""

    @staticmethod
    def forward(ctx, logits: Tensor, targets: Tensor, denom: float):
        M, V = logits.shape
        work = logits.detach().to(torch.bfloat16).contiguous()
        if work.data_ptr() == logits.data_ptr():
            work = work.clone()
        row = torch.empty(M, device=logits.device, dtype=torch.float32)
        out2 = torch.empty(2, device=logits.device, dtype=torch.float32)
        tg = targets.reshape(-1).contiguous()
        need_grad = bool(ctx.needs_input_grad[0])
        ops.cross_entropy(work, tg, row, (1.0 / denom) if denom > 0 else 1.0, need_grad)
        ops.loss_reduce(row, tg, out2)
        ctx.denom, ctx.in_dtype = denom, logits.dtype
        ctx.save_for_backward(work if need_grad else out2, out2)
        return out2[0] / denom if denom > 0 else out2[0] / out2[1]

    @staticmethod
    def backward(ctx, d_loss: Tensor):
        grad, out2 = ctx.saved_tensors
        scale = d_loss.float() if ctx.denom > 0 else d_loss.float() / out2[1]
        return (grad.float() * scale).to(ctx.in_dtype), None, None


def _ce(logits: Tensor, targets: Tensor, denom: float) -> Tensor:
    if not logits.is_cuda:
        raise RuntimeError("chunked_cross_entropy runs on the B200 only (no CPU path)")
    return _CrossEntropyFn.apply(logits.reshape(-1, logits.size(-1)), targets, float(denom))


def chunked_cross_entropy(logits, targets: Tensor, chunk_size: int = 128) -> Tensor:
    """Reference ger/utils.py:424-463, same call signature and the same two reductions:

    * list of lm_head chunks and chunk_size > 0 (the training call, finetune/ger.py:281): per-position CE with
      ignore_index=-1, `cat(...).mean()`  => sum over valid targets / ALL positions;
    * chunk_size == 0 (validation, finetune/ger.py:346): mean over valid targets;
    * flat logits with chunk_size > 0: again per-position losses averaged over all positions.

    When the chunks are `LazyLogits` from this package's GPT.forward, the vocabulary projection and the loss
    run fused (lm_head GEMM -> in-place softmax-gradient) and no [B, T, V] tensor is ever built."""
    from .model import LazyLogits
    if isinstance(logits, LazyLogits):
        logits = [logits]
        lazy_single = True
    else:
        lazy_single = False
    if isinstance(logits, list) and len(logits) > 0 and all(isinstance(l, LazyLogits) for l in logits):
        model = logits[0].model
        same_base = all(l.base is not None and l.base is logits[0].base for l in logits)
        contiguous_cover = same_base and all(logits[i].t1 == logits[i + 1].t0 for i in range(len(logits) - 1))
        n_pos = sum(int(l.shape[0]) * int(l.shape[1]) if l.dim() == 3 else int(l.shape[0]) for l in logits)
        if logits[0].packed_meta is not None:
            # packed batch (GPT.forward_packed): the reference would have right-padded these n sequences to the longest and
            # averaged over ALL n * (T - 1) shifted positions (quirk Q2); the targets are already shifted per sequence
            n_seq, longest = logits[0].packed_meta
            n_pos = n_seq * (longest - 1)
        if contiguous_cover and targets.dim() == 1 and logits[0].base.shape[0] == 1:
            targets = targets.unsqueeze(0)                  # packed batch: one row of concatenated sequences
        if contiguous_cover and targets.dim() == 2:
            base = logits[0].base
            Bsz, T, _ = base.shape
            t0, t1 = logits[0].t0, logits[-1].t1
            assert targets.shape == (Bsz, t1 - t0), f"targets {tuple(targets.shape)} do not match logits [{Bsz}, {t1 - t0}, V]"
            if t0 == 0 and t1 == T:
                hidden, tg = base, targets
            else:   # project every row, ignore the ones outside the covered span (no strided copy of the hidden states)
                hidden = base
                tg = torch.full((Bsz, T), -1, dtype=targets.dtype, device=targets.device)
                tg[:, t0:t1] = targets
            denom = 0.0 if (chunk_size == 0) else float(n_pos)
            return model.fused_loss(hidden, tg, denom)
        hidden = torch.cat([l.hidden.reshape(-1, l.hidden.shape[-1]) for l in logits], dim=0)
        denom = 0.0 if chunk_size == 0 else float(n_pos)
        return model.fused_loss(hidden, targets.reshape(-1), denom)
    if isinstance(logits, list):
        logits = [l.materialize() if isinstance(l, LazyLogits) else l for l in logits]
        flat = torch.cat([l.reshape(-1, l.size(-1)) for l in logits], dim=0)
        return _ce(flat, targets, 0.0 if chunk_size == 0 else float(flat.shape[0]))
    flat = logits.reshape(-1, logits.size(-1))
    return _ce(flat, targets, 0.0 if chunk_size == 0 else float(flat.shape[0]))


def map_old_state_dict_weights(state_dict: Dict, mapping: Mapping, prefix: str) -> Dict:
    for checkpoint_name, attribute_name in mapping.items():
        full_checkpoint_name = prefix + checkpoint_name
        if full_checkpoint_name in state_dict:
            state_dict[prefix + attribute_name] = state_dict.pop(full_checkpoint_name)
    return state_dict


def num_parameters(module: torch.nn.Module, requires_grad: Optional[bool] = None) -> int:
    return sum(p.numel() for p in module.parameters() if requires_grad is None or p.requires_grad == requires_grad)


def get_default_supported_precision(training: bool, tpu: bool = False) -> str:
    """ger/utils.py:475-492."""
    if tpu:
        return "32-true"
    if not torch.cuda.is_available() or torch.cuda.is_bf16_supported():
        return "bf16-mixed" if training else "bf16-true"
    return "16-mixed" if training else "16-true"


def check_valid_checkpoint_dir(checkpoint_dir: Path) -> None:
    """ger/utils.py:239-270: the four files a lit-gpt checkpoint directory must hold."""
    files = {
        "lit_model.pth": (checkpoint_dir / "lit_model.pth").is_file(),
        "lit_config.json": (checkpoint_dir / "lit_config.json").is_file(),
        "tokenizer.json OR tokenizer.model": (checkpoint_dir / "tokenizer.json").is_file()
        or (checkpoint_dir / "tokenizer.model").is_file(),
        "tokenizer_config.json": (checkpoint_dir / "tokenizer_config.json").is_file(),
    }
    if checkpoint_dir.is_dir():
        if all(files.values()):
            return
        problem = f" is missing the files: {[f for f, exists in files.items() if not exists]!r}"
    else:
        problem = " is not a checkpoint directory"
    print(f"--checkpoint_dir {str(checkpoint_dir.absolute())!r}{problem}.", file=sys.stderr)
    raise SystemExit(1)


class lazy_load:
    """Context manager yielding a checkpoint dict (ger/utils.py:98-236 defers tensor reads; torch's mmap does
    the same job here): `with lazy_load(path) as ckpt: model.load_state_dict(ckpt.get("model", ckpt))`."""

    def __init__(self, fn: Union[str, Path]):
        self.sd = torch.load(str(fn), map_location="cpu", mmap=True, weights_only=False)

    def __enter__(self):
        return self.sd

    def __exit__(self, exc_type, exc_val, exc_tb):
        del self.sd


def load_checkpoint(fabric, model, checkpoint_path: Path, strict: bool = True) -> None:
    with lazy_load(checkpoint_path) as sd:
        model.load_state_dict(sd.get("model", sd), strict=strict)


@contextmanager
def quantization(mode: Optional[str] = None):
    """ger/utils.py:40-92 needs the absent `quantize.*` package in the reference too; only mode=None works."""
    if mode is not None:
        raise NotImplementedError("quantized weights are outside the DualHyp hot path")
    yield


class _CSVLogger:
    """Tiny stand-in for lightning's CSVLogger (the drivers create one and never log to it)."""

    def __init__(self, root_dir=".", name="logs", flush_logs_every_n_steps: int = 100, **_: Any):
        self.root_dir, self.name, self.metrics = str(root_dir), name, []
        self.log_dir = str(Path(root_dir) / name)

    def log_metrics(self, metrics: Dict[str, float], step: Optional[int] = None) -> None:
        self.metrics.append({**metrics, "step": step})

    def save(self) -> None:
        import csv
        if not self.metrics:
            return
        merged: Dict[Any, Dict] = {}
        for m in self.metrics:
            merged.setdefault(m.get("step"), {}).update(m)
        rows = [v for _, v in sorted(merged.items(), key=lambda kv: (kv[0] is None, kv[0]))]
        keys = sorted({k for r in rows for k in r})
        Path(self.log_dir).mkdir(parents=True, exist_ok=True)
        with open(Path(self.log_dir) / "metrics.csv", "w", newline="") as f:
            w = csv.DictWriter(f, fieldnames=keys)
            w.writeheader()
            w.writerows(rows)


def step_csv_logger(*args: Any, cls=_CSVLogger, **kwargs: Any):
    return cls(*args, **kwargs)
