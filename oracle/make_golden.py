"""TEST INFRASTRUCTURE ONLY - generate golden vectors by running the UNMODIFIED reference on CPU.

Run in the build container (where `/root/reference` is mounted):

    python -m oracle.make_golden            # writes tests/golden/*.npz

The reference ships no tests or known-answer vectors, so these fixtures are the pin
for `oracle/oracle.py`: weights and inputs come from `dualhyp_b200.synthetic` (a counter hash, bit
reproducible everywhere), outputs come from the reference's own `ger.lora.GPT`, `ger.adapter`
(patched composition), `ger.relprompt.GPT`, `ger.utils.chunked_cross_entropy` and
`generate.base.generate`, imported through `oracle/refshim.py`.  The fixtures travel to the GPU box;
the reference does not.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import refshim  # noqa: E402
from dualhyp_b200 import synthetic  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

# A TinyLlama-shaped decoder small enough for fixtures: head_size 64, GQA with 2 q-heads per kv head.
TINY = dict(block_size=256, vocab_size=512, padded_vocab_size=512, n_layer=3, n_head=4, n_embd=256,
            rotary_percentage=1.0, parallel_residual=False, bias=False, _norm_class="RMSNorm", norm_eps=1e-5,
            _mlp_class="LLaMAMLP", intermediate_size=512, n_query_groups=2)
LORA = dict(r=8, alpha=16, dropout=0.0, to_query=True, to_key=True, to_value=True, to_projection=True,
            to_mlp=False, to_head=False)
B, T = 2, 160            # two lm_head chunks (128 + 32) so the chunk-boundary shift is exercised
SEED = 1337


def batch(vocab=512, extra=None):
    return synthetic.synth_batch(B, T, vocab, SEED, prompt_range=(90, 130), response_range=(8, 24),
                                 extra_token_ids=extra)


def np32(t):
    return t.detach().float().cpu().numpy()


def driver_losses(model, idx, labels, fwd):
    from ger.utils import chunked_cross_entropy
    logits = fwd(idx, lm_head_chunk_size=128)                       # finetune/ger.py:278
    logits[-1] = logits[-1][..., :-1, :]                            # :280
    train = chunked_cross_entropy(logits, labels[..., 1:])          # :281
    full = fwd(idx)
    val = chunked_cross_entropy(full[..., :-1, :], labels[..., 1:], chunk_size=0)   # :345-346
    return train, val, full


def golden_lora():
    import ger.lora as L
    cfg = L.Config(name="tiny", **TINY, **LORA)
    model = L.GPT(cfg)
    sd = synthetic.synth_state_dict(cfg, "lora", SEED)
    model.load_state_dict(sd, strict=True)
    L.mark_only_lora_as_trainable(model)
    idx, labels = batch()
    train, val, full = driver_losses(model, idx, labels, lambda i, **k: model(i, **k))
    train.backward()
    out = {"idx": idx.numpy(), "labels": labels.numpy(), "loss_train": np32(train), "loss_val": np32(val),
           "logits_sub": np32(full[:, ::8, :])}
    for n, p in model.named_parameters():
        if p.requires_grad:
            out["grad." + n] = np32(p.grad)
    # greedy decode through the reference KV cache (inference/ger.py:74-83)
    from generate.base import generate
    prompt = idx[0, :40].clone()
    with torch.no_grad():
        ids = generate(model, prompt, 40 + 16, temperature=0.2, top_k=1, eos_id=None)
    model.reset_cache()
    out["gen_prompt"] = prompt.numpy()
    out["gen_ids"] = ids.numpy()
    np.savez_compressed(os.path.join(GOLDEN_DIR, "tiny_lora.npz"), **out)
    print("lora  : train", float(train), "val", float(val), "gen", ids[40:].tolist())


def golden_adapter():
    import ger.adapter as A
    import ger.model as M
    cfg = A.Config(name="tiny", **TINY, adapter_prompt_length=10, adapter_start_layer=1)
    model = A.GPT(cfg)
    sd = synthetic.synth_state_dict(cfg, "adapter", SEED)
    model.load_state_dict(sd, strict=True)
    A.mark_only_adapter_as_trainable(model)
    model.rope_cache = None
    model.kv_caches = []

    def fwd(i, lm_head_chunk_size=0):
        # ger/adapter.py GPT.forward is broken as shipped (SURVEY headline 2); compose the unmodified
        # pieces: base forward (ger/model.py:65-114) over adapter Blocks, chunking as ger/adapter.py:69-72.
        for b in model.transformer.h:
            if hasattr(b.attn, "adapter_kv_cache"):
                b.attn.adapter_kv_cache = None                      # quirk Q8
        if lm_head_chunk_size == 0:
            return M.GPT.forward(model, i)
        head = model.lm_head
        model.lm_head = torch.nn.Identity()
        try:
            x = M.GPT.forward(model, i)
        finally:
            model.lm_head = head
        return [head(x_i) for x_i in x.split(lm_head_chunk_size, dim=1)]

    idx, labels = batch()
    train, val, full = driver_losses(model, idx, labels, fwd)
    train.backward()
    out = {"idx": idx.numpy(), "labels": labels.numpy(), "loss_train": np32(train), "loss_val": np32(val),
           "logits_sub": np32(full[:, ::8, :])}
    for n, p in model.named_parameters():
        if p.requires_grad:
            out["grad." + n] = np32(p.grad)
    np.savez_compressed(os.path.join(GOLDEN_DIR, "tiny_adapter.npz"), **out)
    print("adapt : train", float(train), "val", float(val))


def golden_relprompt():
    import ger.relprompt as R
    from ger.utils import chunked_cross_entropy
    import torch.nn.functional as F
    cfg = R.Config(name="tiny", **TINY, **{**LORA, "r": 8}, whisper_dim=48, raven_dim=32, pool_size=5)
    model = R.GPT(cfg)
    model.resize_token_embeddings(3)                                 # finetune/relprompt.py:168
    sd = synthetic.synth_state_dict(cfg, "relprompt", SEED, extra_wte_rows=3)
    model.load_state_dict(sd, strict=True)
    R.mark_only_lora_as_trainable(model)
    model.eval()                                                     # classifier dropout -> identity
    idx, labels = batch(extra=(512, 513, 514))
    af = synthetic.synth_features(B, 103, 48, SEED, 301)             # ragged vs pool 10: ceil(103/10) = 11
    vf = synthetic.synth_features(B, 52, 32, SEED, 302)              # pool 5: ceil(52/5) = 11
    at = synthetic.hash_ints((B, 10), SEED, 303, 0, 3)               # targets shorter than predictions
    vt = synthetic.hash_ints((B, 12), SEED, 304, 0, 3)               # targets longer than predictions
    al = model.audio_noise_classifier(af)
    vl = model.visual_noise_classifier(vf)
    n_a, n_v = min(al.size(1), at.size(1)), min(vl.size(1), vt.size(1))
    ml = F.cross_entropy(al[:, :n_a].reshape(-1, 3), at[:, :n_a].reshape(-1)) + \
        F.cross_entropy(vl[:, :n_v].reshape(-1, 3), vt[:, :n_v].reshape(-1))   # finetune/relprompt.py:385-387
    logits = model(idx, audio_query=None, lip_query=None, max_seq_length=T, lm_head_chunk_size=128)
    logits[-1] = logits[-1][..., :-1, :]
    llm = chunked_cross_entropy(logits, labels[..., 1:])
    loss = llm + 0.02 * ml                                           # :401-403
    loss.backward()
    out = {"idx": idx.numpy(), "labels": labels.numpy(), "audio_targets": at.numpy(), "visual_targets": vt.numpy(),
           "audio_logits": np32(al), "visual_logits": np32(vl), "mask_loss": np32(ml), "llm_loss": np32(llm),
           "loss": np32(loss)}
    for n, p in model.named_parameters():
        if p.requires_grad:
            g = p.grad
            if "conv2.weight" in n:
                g = g[::4, ::4, :]
            out["grad." + n] = np32(g)
    from generate.relprompt import generate as generate_rp
    np.savez_compressed(os.path.join(GOLDEN_DIR, "tiny_relprompt.npz"), **out)
    print("relpr : loss", float(loss), "llm", float(llm), "mask", float(ml))


def golden_leaf():
    from ger.rmsnorm import RMSNorm
    from ger.model import build_rope_cache, apply_rope
    from ger.utils import chunked_cross_entropy
    x = synthetic.hash_uniform((3, 7, 256), SEED, 401, 2.0, bf16_exact=False)
    w = synthetic.hash_uniform((256,), SEED, 402, 1.5, bf16_exact=False)
    n = RMSNorm(256, eps=1e-5)
    n.weight.data.copy_(w)
    cos, sin = build_rope_cache(seq_len=64, n_elem=64, dtype=torch.bfloat16, device=torch.device("cpu"))
    q = synthetic.hash_uniform((2, 4, 64, 64), SEED, 403, 1.0, bf16_exact=False)
    logits = synthetic.hash_uniform((2, 40, 97), SEED, 404, 4.0, bf16_exact=False)
    tg = synthetic.hash_ints((2, 40), SEED, 405, -1, 97)
    tg[:, :11] = -1
    out = {"rms_y": np32(n(x)), "rope_cos": np32(cos), "rope_sin": np32(sin), "rope_q": np32(apply_rope(q, cos, sin)),
           "ce_targets": tg.numpy(),
           "ce_list_train": np32(chunked_cross_entropy(list(logits.split(16, dim=1)), tg)),
           "ce_list_val": np32(chunked_cross_entropy(list(logits.split(16, dim=1)), tg, chunk_size=0)),
           "ce_flat_chunked": np32(chunked_cross_entropy(logits, tg, chunk_size=16)),
           "ce_flat_val": np32(chunked_cross_entropy(logits, tg, chunk_size=0))}
    np.savez_compressed(os.path.join(GOLDEN_DIR, "leaf_ops.npz"), **out)
    print("leaf  : ok")


def main():
    torch.manual_seed(0)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    refshim.install()
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    golden_leaf()
    golden_lora()
    golden_adapter()
    golden_relprompt()


if __name__ == "__main__":
    main()
