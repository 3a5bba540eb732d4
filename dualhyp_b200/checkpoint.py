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
# RelPrompt: `resize_token_embeddings(3)` appends randomly initialised wte rows for <<C>>/<<M>>/<<N>> (ger/relprompt.py:215-230)
# and `mark_only_lora_as_trainable` then freezes wte (finetune/relprompt.py:168-169).  The LoRA factors are trained AGAINST those
# rows, so a trainable-only checkpoint has to carry them: they are stored under this extra key (ignored by the reference's
# `load_state_dict(strict=False)`), rows [padded_vocab_size:] of transformer.wte.weight.
EXTRA_WTE_KEY = "transformer.wte.weight::extra_rows"


def _extra_wte_rows(model: torch.nn.Module) -> Optional[Tensor]:
    cfg = getattr(model, "config", None)
    wte = getattr(getattr(model, "transformer", {}), "wte", None) if hasattr(model, "transformer") else None
    if cfg is None or wte is None or getattr(cfg, "padded_vocab_size", None) is None:
        return None
    w = wte.weight
    return w.detach()[cfg.padded_vocab_size:] if w.shape[0] > cfg.padded_vocab_size else None


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
    extra_rows = _extra_wte_rows(model)
    if extra_rows is not None:
        sd[EXTRA_WTE_KEY] = extra_rows.to("cpu").clone()
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
    sd = dict(ckpt.get("model", ckpt))
    extra_rows = sd.pop(EXTRA_WTE_KEY, None)
    own_extra = _extra_wte_rows(model)
    if own_extra is not None and extra_rows is None and "transformer.wte.weight" not in sd:
        raise KeyError(f"the model has {own_extra.shape[0]} resized wte rows (resize_token_embeddings) but the checkpoint holds none: "
                       "the LoRA factors were trained against embeddings this file does not provide")
    if extra_rows is not None:
        if own_extra is None or tuple(own_extra.shape) != tuple(extra_rows.shape):
            raise ValueError(f"checkpoint carries {tuple(extra_rows.shape)} extra wte rows, the model has "
                             f"{None if own_extra is None else tuple(own_extra.shape)}: call resize_token_embeddings first")
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
    loaded = list(sd)
    if extra_rows is not None:
        with torch.no_grad():
            w = model.transformer.wte.weight
            w[model.config.padded_vocab_size:].copy_(extra_rows.to(w.device, w.dtype))     # in place: bumps the version, the engine re-packs
        loaded.append(EXTRA_WTE_KEY)
    return loaded
