"""CPU: pin the oracle (oracle/oracle.py) against golden vectors produced by the unmodified reference
(oracle/make_golden.py), and against the live reference where it is mounted."""
import numpy as np
import pytest
import torch

from dualhyp_b200 import synthetic
from oracle import oracle as O
from oracle import refshim
from tests.helpers import load_golden, rel_err, t, tiny_cfg

SEED = 1337


def test_leaf_ops():
    g = load_golden("leaf_ops.npz")
    x = synthetic.hash_uniform((3, 7, 256), SEED, 401, 2.0, bf16_exact=False)
    w = synthetic.hash_uniform((256,), SEED, 402, 1.5, bf16_exact=False)
    assert torch.allclose(O.rmsnorm(x, w, 1e-5), t(g["rms_y"]), rtol=1e-6, atol=1e-6)
    cos, sin = O.build_rope_cache(64, 64)
    assert cos.dtype == torch.bfloat16
    assert torch.equal(cos.float(), t(g["rope_cos"])) and torch.equal(sin.float(), t(g["rope_sin"]))
    q = synthetic.hash_uniform((2, 4, 64, 64), SEED, 403, 1.0, bf16_exact=False)
    assert torch.allclose(O.apply_rope(q, cos, sin), t(g["rope_q"]), rtol=1e-6, atol=1e-6)
    logits = synthetic.hash_uniform((2, 40, 97), SEED, 404, 4.0, bf16_exact=False)
    tg = t(g["ce_targets"])
    chunks = list(logits.split(16, dim=1))
    assert torch.allclose(O.chunked_cross_entropy(chunks, tg), t(g["ce_list_train"]), rtol=1e-6)
    assert torch.allclose(O.chunked_cross_entropy(chunks, tg, chunk_size=0), t(g["ce_list_val"]), rtol=1e-6)
    assert torch.allclose(O.chunked_cross_entropy(logits, tg, chunk_size=16), t(g["ce_flat_chunked"]), rtol=1e-6)
    assert torch.allclose(O.chunked_cross_entropy(logits, tg, chunk_size=0), t(g["ce_flat_val"]), rtol=1e-6)
    # quirk Q2: the train reduction divides by ALL positions, the val reduction by the valid ones
    n_valid = int((tg != -1).sum())
    assert abs(float(g["ce_list_train"]) * tg.numel() / n_valid - float(g["ce_list_val"])) < 1e-5


@pytest.mark.parametrize("variant", ["lora", "adapter"])
def test_decoder_against_golden(variant):
    g = load_golden(f"tiny_{variant}.npz")
    cfg = tiny_cfg(variant)
    sd = synthetic.synth_state_dict(cfg, variant, SEED)
    idx, labels = t(g["idx"]), t(g["labels"])
    full = O.gpt_forward(sd, cfg, idx)
    assert rel_err(full[:, ::8, :], t(g["logits_sub"])) < 2e-6
    loss, grads = O.loss_and_grads(sd, cfg, idx, labels, "train")
    assert abs(float(loss) - float(g["loss_train"])) < 1e-6
    assert abs(float(O.val_loss(sd, cfg, idx, labels)) - float(g["loss_val"])) < 1e-5
    gk = [k for k in g if k.startswith("grad.")]
    assert len(gk) == len(grads) and len(gk) > 0
    for k in gk:
        assert rel_err(grads[k[5:]], t(g[k])) < 1e-5, k


def test_greedy_decode_against_golden():
    g = load_golden("tiny_lora.npz")
    cfg = tiny_cfg("lora")
    sd = synthetic.synth_state_dict(cfg, "lora", SEED)
    prompt = t(g["gen_prompt"])
    ids = O.generate_greedy(sd, cfg, prompt, prompt.numel() + 16)
    assert ids.tolist() == g["gen_ids"].tolist()


def test_relprompt_against_golden():
    g = load_golden("tiny_relprompt.npz")
    cfg = tiny_cfg("relprompt")
    sd = synthetic.synth_state_dict(cfg, "relprompt", SEED, extra_wte_rows=3)
    B = g["idx"].shape[0]
    af = synthetic.synth_features(B, 103, 48, SEED, 301)
    vf = synthetic.synth_features(B, 52, 32, SEED, 302)
    al = O.noise_mask_classifier(af, sd, "audio_noise_classifier.", 2 * cfg.pool_size)
    vl = O.noise_mask_classifier(vf, sd, "visual_noise_classifier.", cfg.pool_size)
    assert al.shape == g["audio_logits"].shape and vl.shape == g["visual_logits"].shape   # ceil_mode
    assert rel_err(al, t(g["audio_logits"])) < 2e-6 and rel_err(vl, t(g["visual_logits"])) < 2e-6
    keys = O.trainable_keys(sd, cfg)
    work = {k: (v.clone().requires_grad_(True) if k in keys else v) for k, v in sd.items()}
    loss, llm, ml = O.relprompt_loss(work, cfg, t(g["idx"]), t(g["labels"]), af, vf,
                                     t(g["audio_targets"]), t(g["visual_targets"]))
    assert abs(float(ml) - float(g["mask_loss"])) < 1e-5
    assert abs(float(llm) - float(g["llm_loss"])) < 1e-6
    assert abs(float(loss) - float(g["loss"])) < 1e-6
    loss.backward()
    for k in g:
        if not k.startswith("grad."):
            continue
        gr = work[k[5:]].grad
        if "conv2.weight" in k:
            gr = gr[::4, ::4, :]
        assert rel_err(gr, t(g[k])) < 1e-5, k


@pytest.mark.reference
@pytest.mark.skipif(not refshim.available(), reason="reference checkout not mounted")
def test_oracle_against_live_reference_random_shapes():
    """Where the reference is mounted: a second config/seed, straight against the live modules."""
    refshim.install()
    import ger.lora as L
    tiny = dict(block_size=128, vocab_size=256, padded_vocab_size=256, n_layer=2, n_head=8, n_embd=512,
                rotary_percentage=1.0, parallel_residual=False, bias=False, _norm_class="RMSNorm", norm_eps=1e-5,
                _mlp_class="LLaMAMLP", intermediate_size=768, n_query_groups=2)
    rcfg = L.Config(name="t2", **tiny, r=4, alpha=8, dropout=0.0, to_query=True, to_key=True, to_value=True,
                    to_projection=True)
    model = L.GPT(rcfg)
    ocfg = O.OracleConfig(n_layer=2, n_head=8, n_query_groups=2, head_size=64, n_embd=512, intermediate_size=768,
                          padded_vocab_size=256, block_size=128, variant="lora", r=4, alpha=8, to_query=True,
                          to_key=True, to_value=True, to_projection=True)
    sd = synthetic.synth_state_dict(ocfg, "lora", 7)
    model.load_state_dict(sd, strict=True)
    idx, labels = synthetic.synth_batch(3, 100, 256, 7, prompt_range=(50, 80))
    with torch.no_grad():
        ref = model(idx)
    assert rel_err(O.gpt_forward(sd, ocfg, idx), ref) < 2e-6


@pytest.mark.skipif(not refshim.available(), reason="reference not mounted")
def test_oracle_lora_dropout_against_live_reference():
    """LoRA dropout (ger/lora.py:94-98,165,373) in the LIVE reference's train mode against the oracle with the masks installed in
    `oracle.LORA_DROPOUT_MASKS`.  On the CPU an `nn.Dropout` call consumes the generator exactly like `F.dropout` on a tensor of ones
    of the same shape, so replaying the seed and drawing the masks in the reference's call order (layer by layer: the QKV branch,
    then the projection branch) reproduces them; logits and LoRA gradients must then agree to fp32 round-off.  This is what the
    GPU test `test_lora_dropout_matches_oracle_on_the_same_masks` leans on."""
    refshim.install()
    import ger.lora as L
    import torch.nn.functional as F
    tiny = dict(block_size=128, vocab_size=256, padded_vocab_size=256, n_layer=2, n_head=8, n_embd=512,
                rotary_percentage=1.0, parallel_residual=False, bias=False, _norm_class="RMSNorm", norm_eps=1e-5,
                _mlp_class="LLaMAMLP", intermediate_size=768, n_query_groups=2)
    p = 0.25
    rcfg = L.Config(name="t3", **tiny, r=4, alpha=8, dropout=p, to_query=True, to_key=True, to_value=True, to_projection=True)
    model = L.GPT(rcfg)
    ocfg = O.OracleConfig(n_layer=2, n_head=8, n_query_groups=2, head_size=64, n_embd=512, intermediate_size=768,
                          padded_vocab_size=256, block_size=128, variant="lora", r=4, alpha=8, to_query=True,
                          to_key=True, to_value=True, to_projection=True)
    sd = synthetic.synth_state_dict(ocfg, "lora", 9)
    model.load_state_dict(sd, strict=True)
    L.mark_only_lora_as_trainable(model)
    model.train()
    B, T = 2, 40
    idx, _ = synthetic.synth_batch(B, T, 256, 9, prompt_range=(20, 30))
    torch.manual_seed(31)
    ref = model(idx)
    ref.square().mean().backward()
    torch.manual_seed(31)
    masks = {}
    for i in range(2):
        for which in ("attn", "proj"):
            masks[(i, which)] = F.dropout(torch.ones(B, T, 512), p, training=True)
    work = {k: (v.clone().requires_grad_(True) if "lora_" in k else v) for k, v in sd.items()}
    O.LORA_DROPOUT_MASKS = masks
    try:
        got = O.gpt_forward(work, ocfg, idx)
    finally:
        O.LORA_DROPOUT_MASKS = None
    assert rel_err(got.detach(), ref.detach()) < 2e-6
    assert rel_err(O.gpt_forward(sd, ocfg, idx), ref.detach()) > 1e-4          # without the masks the outputs differ
    got.square().mean().backward()
    for name, prm in model.named_parameters():
        if prm.requires_grad:
            assert rel_err(work[name].grad, prm.grad) < 1e-4, name


def test_peaky_fixture_walks_the_permutation():
    """`synth_state_dict(peaky=True)` (the fixture of the 64-token exact greedy test): the head is a blend of the random head and a
    permutation of the embedding, so the oracle's greedy decode walks the permutation's pre-images with a margin far above bf16
    noise -- checked here on the toy config, the GPU test checks the margin again at TinyLlama-1.1B."""
    import math
    cfg = tiny_cfg("adapter")
    assert math.gcd(synthetic.PEAKY_A, cfg.padded_vocab_size) == 1
    old = synthetic.PEAKY_MIX
    synthetic.PEAKY_MIX = 0.5            # 3 toy layers perturb the residual stream relatively more than 22 real ones: stronger blend
    try:
        sd = synthetic.synth_state_dict(cfg, "adapter", SEED, peaky=True)
    finally:
        synthetic.PEAKY_MIX = old
    prompt = synthetic.synth_prompts(1, cfg.padded_vocab_size, SEED + 1, prompt_range=(20, 30))[0]
    out, logits = O.generate_greedy(sd, cfg, prompt, prompt.numel() + 12, return_logits=True)
    V = cfg.padded_vocab_size
    inv = pow(synthetic.PEAKY_A, -1, V)
    tkn = int(prompt[-1])
    for i, lg in enumerate(logits):
        tkn = ((tkn - synthetic.PEAKY_B) * inv) % V
        top2 = lg.float().topk(2).values
        assert int(out[prompt.numel() + i]) == tkn
        assert float(top2[0] - top2[1]) > 0.05 * float(lg.abs().max())


@pytest.mark.skipif(not refshim.available(), reason="reference not mounted")
def test_oracle_classifier_dropout_against_live_reference():
    """The Dropout(0.1) between the two convolutions of `NoiseMaskClassifier` (ger/relprompt.py:135,141) in the live reference's
    train mode against `oracle.noise_mask_classifier(..., dropout_mask=)`: the mask is drawn on the (B, 256, T) activation, replayed
    here from the same seed."""
    refshim.install()
    import ger.relprompt as R
    import torch.nn.functional as F
    B, T, C, pool = 2, 53, 48, 5
    head = R.NoiseMaskClassifier(C, pool_size=pool)
    head.train()
    sd = {"h." + k: v.detach().clone() for k, v in head.state_dict().items()}
    x = synthetic.synth_features(B, T, C, 5, 77)
    torch.manual_seed(5)
    ref = head(x)
    torch.manual_seed(5)
    mask = F.dropout(torch.ones(B, 256, T), head.dropout.p, training=True).transpose(1, 2)
    got = O.noise_mask_classifier(x, sd, "h.", pool, dropout_mask=mask)
    assert rel_err(got, ref.detach()) < 2e-6
    assert rel_err(O.noise_mask_classifier(x, sd, "h.", pool), ref.detach()) > 1e-4
