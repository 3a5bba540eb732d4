"""dualhyp_b200: a B200-native (sm_100a) implementation of the DualHyp decoder hot path behind the
reference's own `ger` module API.  See DESIGN.md / INTEGRATION.md."""
import sys

__version__ = "0.1.0"


def install_as_ger() -> None:
    """Make `import ger`, `import ger.lora`, ... resolve to this package's mirror (dualhyp_b200.ger), so the
    reference's drivers run unchanged on top of the B200 kernels.  Refuses to shadow an already-imported `ger`."""
    import importlib
    existing = sys.modules.get("ger")
    mirror = importlib.import_module("dualhyp_b200.ger")
    if existing is not None and existing is not mirror:
        raise RuntimeError("another `ger` package is already imported; call install_as_ger() before importing the drivers")
    sys.modules["ger"] = mirror
    for sub in ("config", "model", "lora", "adapter", "relprompt", "rmsnorm", "fused_rotary_embedding", "utils", "speed_monitor",
                "tokenizer"):
        sys.modules[f"ger.{sub}"] = importlib.import_module(f"dualhyp_b200.ger.{sub}")
