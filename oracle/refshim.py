"""TEST INFRASTRUCTURE ONLY - import shims that let the *unmodified* reference import on CPU.

The reference (`/root/reference`) depends on packages that are absent from this image
(`lightning`, `lightning_utilities`, `xformers`).  None of them takes part in the arithmetic of the
hot path; they are only touched at import time:

* `lightning_utilities.core.imports.RequirementCache`   -- ger/model.py:11,20 ; ger/__init__.py:5-16
* `xformers.ops.SwiGLU`                                  -- ger/model.py:15 (isinstance only, in dead code)
* a hollow `lightning` tree                              -- ger/utils.py:15-17 ; ger/speed_monitor.py:7-10 ;
                                                            generate/base.py:6-8

`install()` registers hollow stand-ins in `sys.modules` and puts the reference on `sys.path`.
Nothing under `/root/reference` is modified or copied.  Only `oracle/make_golden.py` and the
`tests/` that pin the oracle against the live reference call this; product code never does, and
nothing that runs on the GPU box does (the reference does not exist there).
"""
from __future__ import annotations

import os
import sys
import types

REFERENCE_ROOT = os.environ.get("DUALHYP_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "ger", "lora.py"))


def _mod(name: str, **attrs):
    m = sys.modules.get(name)
    if m is None:
        m = types.ModuleType(name)
        sys.modules[name] = m
    m.__dict__.update(attrs)
    if "." in name:
        parent, child = name.rsplit(".", 1)
        _mod(parent)
        setattr(sys.modules[parent], child, m)
    if not hasattr(m, "__path__"):
        m.__path__ = []  # mark as package so sub-imports resolve
    return m


class _RequirementCache:
    """flash-attn is reported absent so the reference takes its torch SDPA branch (ger/model.py:287-289)."""

    def __init__(self, req: str, *a, **k):
        self.req = req

    def __bool__(self) -> bool:
        return "flash-attn" not in self.req

    def __str__(self) -> str:
        return self.req


class _Any:
    def __init__(self, *a, **k):
        pass


def install() -> None:
    """Idempotent. After this: `import ger.lora, ger.adapter, ger.relprompt; from generate.base import generate`."""
    import torch

    if not available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")
    if "ger" in sys.modules and not getattr(sys.modules["ger"], "__file__", "").startswith(REFERENCE_ROOT):
        raise RuntimeError("a different `ger` package is already imported in this process")
    ident = lambda fn: fn
    _mod("lightning_utilities.core.imports", RequirementCache=_RequirementCache)
    _mod("xformers.ops", SwiGLU=type("SwiGLU", (torch.nn.Module,), {}))
    _mod("lightning", Fabric=_Any, Callback=_Any, LightningModule=_Any, Trainer=_Any,
         seed_everything=lambda s: torch.manual_seed(s))
    _mod("lightning.fabric.strategies", FSDPStrategy=_Any, XLAStrategy=_Any, DDPStrategy=_Any)
    _mod("lightning.fabric.utilities.load", _lazy_load=None)
    _mod("lightning.fabric.utilities.rank_zero", rank_zero_only=ident)
    _mod("lightning.pytorch.utilities.rank_zero", rank_zero_only=ident)
    _mod("lightning.fabric.accelerators.xla", _XLA_GREATER_EQUAL_2_1=False)
    _mod("lightning.fabric.loggers", CSVLogger=_Any)
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
