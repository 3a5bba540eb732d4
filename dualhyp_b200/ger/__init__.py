"""Host-side mirror of the reference's `ger` package for the decoder hot path: same module names
(`ger.model`, `ger.lora`, `ger.adapter`, `ger.relprompt`, `ger.rmsnorm`, `ger.fused_rotary_embedding`,
`ger.utils`, `ger.config`, `ger.speed_monitor`), same classes, constructor arguments, forward()
signatures and state-dict keys -- executed by hand-written sm_100a kernels (libdualhyp.so).

`dualhyp_b200.install_as_ger()` registers this package under the name `ger`, so the reference's
drivers (`finetune/ger.py`, `finetune/relprompt.py`, `generate/*.py`, `inference/*.py`) import it in place
of their own."""
from .config import Config
from .model import GPT

__all__ = ["GPT", "Config", "Tokenizer"]


def __getattr__(name):
    if name == "Tokenizer":
        from .tokenizer import Tokenizer
        return Tokenizer
    raise AttributeError(name)
