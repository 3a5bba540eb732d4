"""RelPrompt end-to-end inference glue (reference inference/relprompt.py:102-173).

The reference, per utterance: encoder features -> `model.{audio,visual}_noise_classifier` -> arg-max over the three
reliability classes per 0.4 s chunk -> the strings `<<C>>` / `<<M>>` / `<<N>>` joined and substituted for the
`<<<ASR_MASKS>>>` / `<<<VSR_MASKS>>>` placeholders of the prompt text -> re-tokenise -> greedy `generate()` with the
KV cache.  The Whisper / BRAVEn encoders are out of scope: the encoder OUTPUTS are passed in as
tensors.  Two entry points:

* `predict_masks` / `mask_strings` -- the classifier + arg-max + string step, batched;
* `fill_mask_tokens` -- the same substitution done on token ids (each mask token is ONE id after
  `resize_token_embeddings(3)`, inference/relprompt.py:341, so splicing ids is equivalent to re-tokenising the text
  and needs no tokenizer on the device path);
* `relprompt_generate` -- the whole step for a batch of utterances on `generate_batch`."""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from .generate import generate_batch

Tensor = torch.Tensor
IDX_TO_MASK_TOKEN = {0: "<<C>>", 1: "<<M>>", 2: "<<N>>"}          # inference/relprompt.py:92
ASR_PLACEHOLDER, VSR_PLACEHOLDER = "<<<ASR_MASKS>>>", "<<<VSR_MASKS>>>"   # data/prompts.py:16


@torch.no_grad()
def predict_masks(model, audio_feats: Tensor, visual_feats: Tensor, audio_frames: Optional[Sequence[int]] = None,
                  visual_frames: Optional[Sequence[int]] = None) -> Tuple[List[Tensor], List[Tensor]]:
    """Reliability class per chunk (inference/relprompt.py:112-118).  audio_feats [B, Ta, whisper_dim] (50 fps),
    visual_feats [B, Tv, raven_dim] (25 fps); optional true frame counts of right-padded batches (the number of
    chunks of utterance b is ceil(frames / pool), AvgPool1d ceil_mode).  Returns int64 class ids per utterance."""
    a_logits = model.audio_noise_classifier(audio_feats)            # [B, ceil(Ta / 20), 3]
    v_logits = model.visual_noise_classifier(visual_feats)          # [B, ceil(Tv / 10), 3]
    a_pred, v_pred = a_logits.argmax(-1), v_logits.argmax(-1)
    a_pool, v_pool = model.audio_noise_classifier.pool_size, model.visual_noise_classifier.pool_size
    out_a, out_v = [], []
    for b in range(a_pred.shape[0]):
        na = a_pred.shape[1] if audio_frames is None else -(-int(audio_frames[b]) // a_pool)
        nv = v_pred.shape[1] if visual_frames is None else -(-int(visual_frames[b]) // v_pool)
        out_a.append(a_pred[b, :na])
        out_v.append(v_pred[b, :nv])
    return out_a, out_v


def mask_strings(pred: Tensor) -> str:
    """inference/relprompt.py:145-150: class ids -> '<<C>><<M>>...'."""
    return "".join(IDX_TO_MASK_TOKEN[int(i)] for i in pred.tolist())


def fill_prompt_text(prompt: str, audio_pred: Tensor, visual_pred: Tensor) -> str:
    """inference/relprompt.py:153 on the prompt text."""
    return prompt.replace(ASR_PLACEHOLDER, mask_strings(audio_pred)).replace(VSR_PLACEHOLDER, mask_strings(visual_pred))


def fill_mask_tokens(prompt_ids: Tensor, audio_pred: Tensor, visual_pred: Tensor, asr_slot: int, vsr_slot: int,
                     first_mask_id: int) -> Tensor:
    """The substitution on token ids: `prompt_ids` holds ONE placeholder id at positions `asr_slot` < `vsr_slot`;
    each is replaced by the run of mask-token ids `first_mask_id + class` (the three ids `resize_token_embeddings(3)`
    appended: 32000..32002 for TinyLlama)."""
    assert 0 <= asr_slot < vsr_slot < prompt_ids.numel()
    dev = prompt_ids.device
    a = audio_pred.to(dev, torch.int64) + first_mask_id
    v = visual_pred.to(dev, torch.int64) + first_mask_id
    return torch.cat((prompt_ids[:asr_slot], a, prompt_ids[asr_slot + 1:vsr_slot], v, prompt_ids[vsr_slot + 1:]))


@torch.no_grad()
def relprompt_generate(model, prompt_ids: Sequence[Tensor], slots: Sequence[Tuple[int, int]], audio_feats: Tensor,
                       visual_feats: Tensor, max_new_tokens: int = 150, eos_id: Optional[int] = None,
                       audio_frames: Optional[Sequence[int]] = None, visual_frames: Optional[Sequence[int]] = None,
                       first_mask_id: Optional[int] = None):
    """One batch of RelPrompt inference: predicted reliability masks spliced into each prompt, then batched greedy
    decoding (150 new tokens at most in the reference, inference/relprompt.py:158).  Returns (ids per utterance,
    audio class ids, visual class ids)."""
    if first_mask_id is None:
        first_mask_id = model.config.padded_vocab_size          # resize_token_embeddings appended the three ids here
    a_pred, v_pred = predict_masks(model, audio_feats, visual_feats, audio_frames, visual_frames)
    filled = [fill_mask_tokens(p, a_pred[b], v_pred[b], slots[b][0], slots[b][1], first_mask_id)
              for b, p in enumerate(prompt_ids)]
    outs = generate_batch(model, filled, max_new_tokens, eos_id=eos_id)
    return outs, a_pred, v_pred
