"""`ger.tokenizer` name kept for `from ger import Tokenizer` (generate/base.py:14).  Host-side text handling
is outside the hot path; the live drivers use transformers.AutoTokenizer.  This thin class wraps
that with the reference wrapper's encode/decode surface."""
from pathlib import Path
from typing import Optional

import torch


class Tokenizer:
    def __init__(self, checkpoint_dir: Path) -> None:
        from transformers import AutoTokenizer
        self.processor = AutoTokenizer.from_pretrained(str(checkpoint_dir))
        self.bos_id = self.processor.bos_token_id
        self.eos_id = self.processor.eos_token_id

    @property
    def vocab_size(self) -> int:
        return self.processor.vocab_size

    def encode(self, string: str, device: Optional[torch.device] = None, bos: bool = False, eos: bool = False,
               max_length: int = -1) -> torch.Tensor:
        tokens = self.processor.encode(string, add_special_tokens=False)
        if bos and self.bos_id is not None:
            tokens = [self.bos_id] + tokens
        if eos and self.eos_id is not None:
            tokens = tokens + [self.eos_id]
        if max_length > 0:
            tokens = tokens[:max_length]
        return torch.tensor(tokens, dtype=torch.int, device=device)

    def decode(self, tensor: torch.Tensor) -> str:
        tokens = [tensor.item()] if tensor.ndim == 0 else tensor.tolist()
        return self.processor.decode(tokens)
