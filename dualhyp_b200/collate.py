"""Prompt assembly and batch collation for the decoder hot path (reference
data/av_dataset.py:210-292 GER, :373-429 DualHyp, :546-647 RelPrompt).

Host-side only.  The template strings themselves are data the caller passes in (the reference keeps them in
data/prompts.py as `{"prompt_1", "prompt_2", "prompt_3"}` dicts); this module restates how they are filled, how a
sample becomes (input_ids, labels) and how samples become a batch:

* `ger_prompt`, `dualhyp_prompt`, `relprompt_prompt` -- placeholder substitution of the three dataset classes;
* `encode_sample` -- tokenise prompt and prompt + response + eos, labels = -1 over the prompt span (:246-252);
* `collate_right_pad` -- the reference's `collate_fn` for the text fields: ids padded with 0, labels with -1;
* `pack_varlen` -- the same samples as ONE row of concatenated tokens with `cu_seqlens` and per-token positions
  (no pad tokens).  The B200 attention kernels take one length per batch today, so training consumes
  `collate_right_pad`; `pack_varlen` is the layout the packed (cu_seqlens) attention will consume."""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence

import torch

Tensor = torch.Tensor
IGNORE = -1


def ger_prompt(fmt: Dict[str, str], hyps: Sequence[str], max_nhyps: Optional[int] = None, language: Optional[str] = None) -> str:
    """data/av_dataset.py:210-222: best hypothesis, then the other hypotheses one per line."""
    p1 = fmt["prompt_1"]
    if language:
        p1 = p1.replace("speech recognition system", f"{language} speech recognition system")      # :112
    other = list(hyps[1:max_nhyps] if max_nhyps is not None else hyps[1:])
    return p1 + hyps[0] + fmt["prompt_2"] + "\n" + "\n".join(other) + fmt["prompt_3"]


def dualhyp_prompt(fmt: Dict[str, str], asr_hyps: Sequence[str], vsr_hyps: Sequence[str], max_nhyps: Optional[int] = None) -> str:
    """data/av_dataset.py:373-395: ASR / VSR best hypotheses into prompt_1, the others into prompt_2."""
    a_other = list(asr_hyps[1:max_nhyps] if max_nhyps is not None else asr_hyps[1:])
    v_other = list(vsr_hyps[1:max_nhyps] if max_nhyps is not None else vsr_hyps[1:])
    return (fmt["prompt_1"].replace("<<<ASR_NHYPS>>>", asr_hyps[0]).replace("<<<VSR_NHYPS>>>", vsr_hyps[0])
            + fmt["prompt_2"].replace("<<<ASR_NHYPS>>>", "\n".join(a_other)).replace("<<<VSR_NHYPS>>>", "\n".join(v_other))
            + fmt["prompt_3"])


def relprompt_prompt(fmt: Dict[str, str], asr_hyps: Sequence[str], vsr_hyps: Sequence[str], audio_masks: Optional[Sequence[str]] = None,
                     video_masks: Optional[Sequence[str]] = None, max_nhyps: Optional[int] = None) -> str:
    """data/av_dataset.py:546-570.  With masks given (training: the chunk labels) the `<<<ASR_MASKS>>>` / `<<<VSR_MASKS>>>`
    placeholders are filled; without (inference) they stay for `relprompt_infer.fill_prompt_text`."""
    a_other = list(asr_hyps[1:max_nhyps] if max_nhyps is not None else asr_hyps[1:])
    v_other = list(vsr_hyps[1:max_nhyps] if max_nhyps is not None else vsr_hyps[1:])
    p = (fmt["prompt_1"].replace("<<<ASR_BEST_NHYPS>>>", asr_hyps[0]).replace("<<<VSR_BEST_NHYPS>>>", vsr_hyps[0])
         .replace("<<<ASR_NHYPS>>>", "\n".join(a_other)).replace("<<<VSR_NHYPS>>>", "\n".join(v_other)))
    if audio_masks is not None:
        p = p.replace("<<<ASR_MASKS>>>", "".join(audio_masks))
    if video_masks is not None:
        p = p.replace("<<<VSR_MASKS>>>", "".join(video_masks))
    return p + fmt["prompt_3"]


def encode_sample(encode: Callable[[str], List[int]], prompt_no_response: str, response: str, eos_token: str) -> Dict[str, Tensor]:
    """data/av_dataset.py:223-258 (the non-chat-template branch): the prompt and prompt + response + eos are tokenised
    separately; labels copy the ids of the response span and are -1 over the prompt span."""
    ids_no_resp = list(encode(prompt_no_response))
    ids = list(encode(prompt_no_response + response + eos_token))
    labels = [IGNORE] * len(ids_no_resp) + ids[len(ids_no_resp):]
    return {"input_ids_no_response": torch.tensor(ids_no_resp, dtype=torch.int64), "input_ids": torch.tensor(ids, dtype=torch.int64),
            "labels": torch.tensor(labels, dtype=torch.int64), "input": prompt_no_response + response + eos_token}


def collate_right_pad(samples: Sequence[Dict[str, Tensor]]) -> Dict[str, object]:
    """data/av_dataset.py:260-292, text fields: right-pad to the longest sample, ids with 0 and labels with -1."""
    max_len = max(int(s["input_ids"].numel()) for s in samples)
    ids = torch.zeros((len(samples), max_len), dtype=torch.int64)
    labels = torch.full((len(samples), max_len), IGNORE, dtype=torch.int64)
    for i, s in enumerate(samples):
        n = int(s["input_ids"].numel())
        ids[i, :n] = s["input_ids"]
        labels[i, :n] = s["labels"]
    return {"input_ids": ids, "labels": labels, "input_ids_no_response": [s["input_ids_no_response"] for s in samples]}


def pack_varlen(samples: Sequence[Dict[str, Tensor]]) -> Dict[str, Tensor]:
    """The same batch without pad tokens: ids / labels concatenated, `cu_seqlens` int32 [n + 1], `positions` int32 (RoPE
    position of each token inside its sample).  The label of the LAST token of every sample is -1 in the shifted
    targets, so that no sample predicts the first token of the next one."""
    lens = [int(s["input_ids"].numel()) for s in samples]
    cu = torch.zeros(len(samples) + 1, dtype=torch.int32)
    cu[1:] = torch.tensor(lens, dtype=torch.int32).cumsum(0)
    ids = torch.cat([s["input_ids"] for s in samples])
    labels = torch.cat([s["labels"] for s in samples])
    positions = torch.cat([torch.arange(n, dtype=torch.int32) for n in lens])
    targets = torch.full_like(labels, IGNORE)                       # next-token targets, per sample (finetune/ger.py:280-281)
    for a, b in zip(cu[:-1].tolist(), cu[1:].tolist()):
        targets[a:b - 1] = labels[a + 1:b]
    return {"input_ids": ids, "labels": labels, "targets": targets, "cu_seqlens": cu, "positions": positions,
            "pad_fraction_saved": torch.tensor(1.0 - sum(lens) / (len(lens) * max(lens)))}
