"""Text side of the reference's dataset classes, under the names the drivers import (SURVEY.md section 8(b) "shims", 8f row 2):

    from data.av_dataset import AVDataset, DualHypothesesAVDataset, DualHypothesesAlignAVDataset, DualHypothesesAlign2AVDataset
                                                                       (finetune/ger.py:36, inference/ger.py:27)

`install_as_data()` registers this module as `data.av_dataset`.  What is kept of data/av_dataset.py: the JSON index by `Uid`
(:39-78), the random choice among a uid's records (:123, :344), the prompt assembly and label masking of `get_prompt` (:210-258,
:373-429; through `dualhyp_b200.collate`), truncation to `max_input_length` (:138-140), `collate_fn` for the text fields
(:260-292) and `get_max_seq_length`.  What is NOT here: audio / video loading and corruption (whisper, cv2, h5py, albumentations
are outside the decoder hot path, SURVEY section 2 OUT-OF-SCOPE): `audio` and `video` are None in every item and batch, the
constructor accepts and ignores the media arguments.  The chat-template branch (:225-241) needs a Hugging Face tokenizer and is
supported when the tokenizer has `apply_chat_template`.

`DualHypothesesAlignAVDataset` / `DualHypothesesAlign2AVDataset` do not exist in the reference (its drivers import names its
data module never defines, and `get_prompts_format` raises on "DualAlign*").  They are defined here as the ALIGNED variants of
`DualHypothesesAVDataset`: the ASR and the VSR hypothesis lists come from the SAME record of the uid instead of two independent
draws (`Align`), and `Align2` additionally keeps the n-best order (no shuffling of the other-hypotheses).  Both use the DualHyp
templates.  This is a stated choice, not a restatement.

`pack=True` in `collate_fn` adds the pad-free layout (`collate.pack_varlen`) the packed attention kernels consume."""
from __future__ import annotations

import json
import random
import sys
from collections import OrderedDict
from typing import Dict, List, Optional, Sequence, Union

import torch

from . import collate as C

GER_PROMPTS = {
    "prompt_1": ("Below is the best-hypotheses transcribed from speech recognition system. Please try to revise it using the words which "
                 "are only included into other-hypothesis, and write the response for the true transcription.\n\n### Best-hypothesis:\n"),
    "prompt_2": "\n\n### Other-hypothesis:",
    "prompt_3": "\n\n### Response:\n",
}
DUALHYP_PROMPTS = {
    "prompt_1": ("Below are the best-hypothesis transcribed from speech recognition systems, ASR and VSR, respectively. Please try to "
                 "revise it using the words which are only included into other-hypotheses, and write the response for the true "
                 "transcription.\n\n### ASR Best-hypothesis:\n<<<ASR_NHYPS>>>\n\n### VSR Best-hypothesis:\n<<<VSR_NHYPS>>>"),
    "prompt_2": "\n\n### ASR Other-hypotheses:\n<<<ASR_NHYPS>>>\n\n### VSR Other-hypotheses:\n<<<VSR_NHYPS>>>",
    "prompt_3": "\n\n### Response:\n",
}


def get_prompts_format(name: str) -> Dict[str, str]:
    """data/prompts.py:24-35 for the two text-only formats; "DualAlign" / "DualAlign2" (unknown to the reference) map to DualHyp."""
    if name == "GER":
        return dict(GER_PROMPTS)
    if name in ("DualHyp", "DualAlign", "DualAlign2"):
        return dict(DUALHYP_PROMPTS)
    raise ValueError(f"Unknown prompt name: {name}")


def _shuffled(seq: Sequence[str]) -> List[str]:
    """data/utils.py random_sample_sequence(seq, len(seq)): a random permutation of the other-hypotheses."""
    return random.sample(list(seq), len(seq))


class AVDataset(torch.utils.data.Dataset):
    """GER: one n-best list per sample (data/av_dataset.py:21-323, text fields only)."""

    default_format = "GER"

    def __init__(self, split: str, json_path: Union[str, Sequence[str]], max_input_length: int = -1, max_nhyps: Optional[int] = None,
                 nhyps_key: str = "nhyps_asr", random_sample_nhyps: bool = True, tokenizer=None, prompts_format: Optional[str] = None,
                 apply_chat_template: bool = False, language: Optional[str] = None, **media_kwargs):
        assert split in ("train", "val", "test")
        self.split, self.max_input_length, self.max_nhyps, self.nhyps_key = split, max_input_length, max_nhyps, nhyps_key
        self.random_sample_nhyps, self.tokenizer, self.apply_chat_template, self.language = random_sample_nhyps, tokenizer, apply_chat_template, language
        self.media_kwargs = media_kwargs                    # occlusion_patch_dir, audio_*, visual_*, transform_*, maximum_*: not used here
        paths = [json_path] if isinstance(json_path, str) else list(json_path)
        self.data: List[dict] = []
        self.uid2sample: "OrderedDict[str, List[dict]]" = OrderedDict()
        for path in paths:
            with open(path, "r") as f:
                records = json.load(f)
            suffix = "_pretrain" if (not isinstance(json_path, str) and "_pretrain" in path) else ""      # :58-66, :74-78: own uids
            for rec in records:
                self.data.append(rec)
                self.uid2sample.setdefault(rec["Uid"] + suffix, []).append(rec)
        self.idx2uid = list(self.uid2sample)
        self.fmt = get_prompts_format(prompts_format or self.default_format)
        if tokenizer is not None and not hasattr(tokenizer, "eos_token"):
            tokenizer.eos_token = "</s>"                                                   # :116-118

    def __len__(self) -> int:
        return len(self.uid2sample)

    # ---- prompt assembly -------------------------------------------------------------------------
    def _others(self, hyps: Sequence[str]) -> List[str]:
        other = list(hyps[1:self.max_nhyps] if self.max_nhyps is not None else hyps[1:])
        return _shuffled(other) if self.random_sample_nhyps else other

    def _encode(self, prompt_no_response: str, caption: str) -> Dict[str, object]:
        assert self.tokenizer is not None, "a tokenizer is needed to build prompts"
        if self.apply_chat_template:                                                       # :225-241
            messages = [{"role": "system", "content": "You are a helpful AI assistant."}, {"role": "user", "content": prompt_no_response}]
            prompt_ids = list(self.tokenizer.apply_chat_template(messages, tokenize=True, add_generation_prompt=True))
            answer_ids = list(self.tokenizer(caption, add_special_tokens=False)["input_ids"]) + [self.tokenizer.eos_token_id]
            out = {"input_ids_no_response": torch.tensor(prompt_ids, dtype=torch.int64),
                   "input_ids": torch.tensor(prompt_ids + answer_ids, dtype=torch.int64),
                   "labels": torch.tensor([C.IGNORE] * len(prompt_ids) + answer_ids, dtype=torch.int64),
                   "input": prompt_no_response + caption + self.tokenizer.eos_token}
        else:
            out = C.encode_sample(self.tokenizer.encode, prompt_no_response, caption, self.tokenizer.eos_token)
        if self.max_input_length > 0:                                                      # :138-140
            out["input_ids"] = out["input_ids"][:self.max_input_length]
            out["labels"] = out["labels"][:self.max_input_length]
        return out

    def get_prompt(self, sample: dict) -> Dict[str, object]:
        hyps = sample[self.nhyps_key]["hyps"]
        text = C.ger_prompt(self.fmt, [hyps[0]] + self._others(hyps), None, self.language)
        return self._encode(text, sample["Caption"])

    def __getitem__(self, idx: int) -> Dict[str, object]:
        sample = random.choice(self.uid2sample[self.idx2uid[idx]])
        return {"audio": None, "video": None, "uid": sample.get("Uid", ""), "ground_truth": sample.get("Caption", ""), **self.get_prompt(sample)}

    # ---- batching ----------------------------------------------------------------------------------
    def collate_fn(self, samples: Sequence[Dict[str, object]], pack: bool = False) -> Dict[str, object]:
        batch = {"input": [s["input"] for s in samples], "uid": [s["uid"] for s in samples], "ground_truth": [s["ground_truth"] for s in samples],
                 **C.collate_right_pad(samples), "audio": None, "video": None}
        if pack:
            batch["packed"] = C.pack_varlen(samples)
        return batch

    def get_max_seq_length(self):
        """Longest prompt + response over one draw of every uid, and the mean (the reference scans a pre-tokenised `data`)."""
        lengths = [int(self[i]["input_ids"].numel()) for i in range(len(self))]
        longest = max(range(len(lengths)), key=lengths.__getitem__)
        return max(lengths), lengths[longest], longest, sum(lengths) / len(lengths)


class DualHypothesesAVDataset(AVDataset):
    """DualHyp: an ASR and a VSR n-best list, drawn independently among the uid's records (data/av_dataset.py:326-429)."""

    default_format = "DualHyp"
    aligned = False
    keep_order = False

    def __init__(self, split, json_path, prompts_format: str = "DualHyp", **kwargs):
        super().__init__(split, json_path, prompts_format=prompts_format, **kwargs)
        self.nhyps_key_asr, self.nhyps_key_vsr = "nhyps_asr", "nhyps_vsr"
        if self.language is not None:
            self.fmt["prompt_1"] = self.fmt["prompt_1"].replace("speech recognition system", f"{self.language} speech recognition system")

    def _others(self, hyps):
        other = list(hyps[1:self.max_nhyps] if self.max_nhyps is not None else hyps[1:])
        return _shuffled(other) if (self.random_sample_nhyps and not self.keep_order) else other

    def get_prompt(self, sample_1: dict, sample_2: Optional[dict] = None) -> Dict[str, object]:
        sample_2 = sample_1 if sample_2 is None else sample_2
        asr, vsr = sample_1[self.nhyps_key_asr]["hyps"], sample_2[self.nhyps_key_vsr]["hyps"]
        text = C.dualhyp_prompt(self.fmt, [asr[0]] + self._others(asr), [vsr[0]] + self._others(vsr), None)
        return self._encode(text, sample_1["Caption"])

    def __getitem__(self, idx: int) -> Dict[str, object]:
        records = self.uid2sample[self.idx2uid[idx]]
        if self.aligned:
            sample_1 = sample_2 = random.choice(records)
        else:
            sample_1, sample_2 = random.choices(records, k=2)                               # :344
        return {"audio": None, "video": None, "uid": sample_1.get("Uid", ""), "ground_truth": sample_1.get("Caption", ""),
                **self.get_prompt(sample_1, sample_2)}


class DualHypothesesAlignAVDataset(DualHypothesesAVDataset):
    """Not in the reference (see the module docstring): both hypothesis lists from the same record."""

    default_format = "DualAlign"
    aligned = True

    def __init__(self, split, json_path, prompts_format: str = "DualAlign", **kwargs):
        super().__init__(split, json_path, prompts_format=prompts_format, **kwargs)


class DualHypothesesAlign2AVDataset(DualHypothesesAlignAVDataset):
    """Not in the reference: the aligned variant that also keeps the n-best order of the other-hypotheses."""

    default_format = "DualAlign2"
    keep_order = True

    def __init__(self, split, json_path, prompts_format: str = "DualAlign2", **kwargs):
        super().__init__(split, json_path, prompts_format=prompts_format, **kwargs)


def install_as_data() -> None:
    """Make `from data.av_dataset import ...` resolve to this module.  Refuses to shadow an already-imported `data` package."""
    import types
    this = sys.modules[__name__]
    existing = sys.modules.get("data.av_dataset")
    if existing is not None and existing is not this:
        raise RuntimeError("another `data.av_dataset` is already imported; call install_as_data() before importing the drivers")
    pkg = sys.modules.get("data")
    if pkg is None:
        pkg = types.ModuleType("data")
        pkg.__path__ = []                      # a namespace-like package holding only the shim
        sys.modules["data"] = pkg
    pkg.av_dataset = this
    sys.modules["data.av_dataset"] = this
