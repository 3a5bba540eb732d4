"""Trainable-only checkpoints.

The reference's `save_lora_checkpoint` / `save_adapter_checkpoint` write the WHOLE 1.1 B-parameter model every
time (finetune/ger.py:356-358, finetune/relprompt.py:598-600: `fabric.save(path, {"model": model})`, ~4.4 GB in
fp32) although only the PEFT tensors change.  `save_trainable` writes just the tensors selected by the variant's
filter (`lora_filter` ger/lora.py:442-443, `adapter_filter` ger/adapter.py:164-165, plus the RelPrompt
`noise_classifier` heads, ger/relprompt.py:79-119): 1.6 MB (adapter) ... 27 MB (RelPrompt r=16).

The file stays loadable by the reference's own tooling: it is a plain `{"model": {name: tensor}}` dict with the
reference's parameter names, i.e. exactly what `lazy_load` + `model.load_state_dict(ckpt["model"], strict=False)`
(inference/ger.py:98-101) expects on top of the frozen base checkpoint.  `load_trainable` is that call plus a
check that every stored tensor found its parameter and that no trainable parameter was left without a value."""
from __future__ import annotations

from pathlib import Path
from typing import Callable, Dict, Iterable, Optional, Union

import torch

Tensor = torch.Tensor
FORMAT = "dualhyp-trainable-v1"


def default_filter(name: str, _value=None) -> bool:
    """Union of the reference's PEFT filters: LoRA factors, adapter prefix / gate, AdapterV2 scale / bias, and the
    RelPrompt reliability predictors."""
    return ("lora_" in name or "adapter_wte" in name or "gating_factor" in name or "adapter_scale" in name
            or "adapter_bias" in name or "noise_classifier" in name)


def trainable_state(model: torch.nn.Module, key_filter: Optional[Callable[[str, Tensor], bool]] = None) -> Dict[str, Tensor]:
    """CPU copies of the selected tensors, reference-keyed.  Without a filter: every parameter with requires_grad
    (what `mark_only_*_as_trainable` left on), else the filter decides."""
    out: Dict[str, Tensor] = {}
    params = dict(model.named_parameters())
    for name, value in model.state_dict().items():
        p = params.get(name)
        keep = key_filter(name, value) if key_filter is not None else (p is not None and p.requires_grad)
        if keep:
            out[name] = value.detach().to("cpu").clone()
    return out


def save_trainable(model: torch.nn.Module, file_path: Union[str, Path], key_filter: Optional[Callable] = None,
                   extra: Optional[dict] = None) -> int:
    """Write `{"model": {...}, "format": ..., **extra}`; returns the number of bytes of tensor data stored."""
    sd = trainable_state(model, key_filter)
    if not sd:
        raise ValueError("no trainable tensors selected: call mark_only_*_as_trainable(model) or pass a key_filter")
    file_path = Path(file_path)
    file_path.parent.mkdir(parents=True, exist_ok=True)
    payload = {"model": sd, "format": FORMAT}
    if extra:
        payload.update(extra)
    tmp = file_path.with_suffix(file_path.suffix + ".tmp")
    torch.save(payload, str(tmp))
    tmp.replace(file_path)                       # atomic: a killed job never leaves a half-written checkpoint
    return sum(t.numel() * t.element_size() for t in sd.values())


def load_trainable(model: torch.nn.Module, file_path: Union[str, Path], require_all: bool = True) -> Iterable[str]:
    """Load a trainable-only (or a full reference) checkpoint on top of the model's current weights.  Raises on
    names the model does not have or shape mismatches; with require_all, also when a parameter that requires grad
    got no value from the file.  Returns the names loaded."""
    ckpt = torch.load(str(file_path), map_location="cpu", weights_only=False)
    sd = ckpt.get("model", ckpt)
    own = model.state_dict()
    unknown = [k for k in sd if k not in own]
    if unknown:
        raise KeyError(f"checkpoint tensors without a parameter in the model: {unknown[:5]}{' ...' if len(unknown) > 5 else ''}")
    bad = [k for k, v in sd.items() if tuple(v.shape) != tuple(own[k].shape)]
    if bad:
        raise ValueError(f"shape mismatch for {bad[:5]}: checkpoint {[tuple(sd[k].shape) for k in bad[:5]]} vs model "
                         f"{[tuple(own[k].shape) for k in bad[:5]]}")
    if require_all:
        missing = [n for n, p in model.named_parameters() if p.requires_grad and n not in sd]
        if missing:
            raise KeyError(f"trainable parameters missing from the checkpoint: {missing[:5]}{' ...' if len(missing) > 5 else ''}")
    model.load_state_dict(sd, strict=False)
    return list(sd)
