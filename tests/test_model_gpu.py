"""GPU parity of the whole hot path through the reference-facing API (dualhyp_b200.ger.* -> engine -> C ABI):
logits, both loss reductions, trainable-tensor gradients and greedy ids, against (a) the golden vectors the
unmodified reference produced (tests/golden, fp32 CPU) and (b) the oracle on the same synthetic inputs.

Tolerances (bf16 compute vs fp32 reference): loss |d| <= 2e-2 abs and 5e-3 rel; logits
Frobenius rel error <= 2e-2 (a few bf16 ulps at the logit scale); trainable-grad cosine >= 0.999 per tensor
(>= 0.995 for the tiny toy model whose gradients are noisier), and relative norm error <= 5e-2."""
import numpy as np
import pytest
import torch

from dualhyp_b200 import synthetic
from dualhyp_b200.generate import generate, generate_batch
from dualhyp_b200.ger import adapter, lora, relprompt
from dualhyp_b200.ger.utils import chunked_cross_entropy
from oracle import oracle as O
from tests.helpers import full_cfg, load_golden, rel_err, t, tiny_cfg

pytestmark = pytest.mark.gpu
DEV = "cuda"
SEED = 1337
LORA = dict(r=8, alpha=16, to_query=True, to_key=True, to_value=True, to_projection=True)


def cos_sim(a, b):
    a, b = a.double().flatten().cpu(), b.double().flatten().cpu()
    return float((a @ b) / (a.norm() * b.norm()).clamp_min(1e-30))


def build(variant, name="tiny-llama-test", **over):
    if variant == "lora":
        cfg = lora.Config.from_name(name, **{**LORA, **over})
        m = lora.GPT(cfg)
        sd = synthetic.synth_state_dict(cfg, "lora", SEED)
        mark = lora.mark_only_lora_as_trainable
    elif variant == "adapter":
        cfg = adapter.Config.from_name(name, **over)
        m = adapter.GPT(cfg)
        sd = synthetic.synth_state_dict(cfg, "adapter", SEED)
        mark = adapter.mark_only_adapter_as_trainable
    else:
        cfg = relprompt.Config.from_name(name, **{**LORA, **over})
        m = relprompt.GPT(cfg)
        m.resize_token_embeddings(3)
        sd = synthetic.synth_state_dict(cfg, "relprompt", SEED, extra_wte_rows=3)
        mark = relprompt.mark_only_lora_as_trainable
    m.load_state_dict(sd, strict=True)
    mark(m)
    return m.to(DEV), sd


def train_step(m, idx, labels):
    """The driver's micro-step verbatim (finetune/ger.py:278-285)."""
    logits = m(idx, lm_head_chunk_size=128)
    logits[-1] = logits[-1][..., :-1, :]
    loss = chunked_cross_entropy(logits, labels[..., 1:])
    loss.backward()
    return loss


def check_grads(m, ref_grads, min_cos, what):
    n = 0
    for name, p in m.named_parameters():
        if not p.requires_grad or name not in ref_grads:
            continue
        assert p.grad is not None, f"{what}: no gradient for {name}"
        ref = ref_grads[name]
        c = cos_sim(p.grad, ref)
        nr = float(p.grad.double().norm().cpu() / ref.double().norm().clamp_min(1e-30))
        assert c >= min_cos and abs(nr - 1) <= 5e-2, f"{what}: grad {name} cosine {c:.5f} norm ratio {nr:.4f}"
        n += 1
    assert n > 0


@pytest.mark.parametrize("variant", ["lora", "adapter"])
def test_tiny_against_reference_golden(variant):
    g = load_golden(f"tiny_{variant}.npz")
    m, _ = build(variant, **({"adapter_start_layer": 1} if variant == "adapter" else {}))
    idx, labels = t(g["idx"]).to(DEV), t(g["labels"]).to(DEV)
    with torch.no_grad():
        full = m(idx)
        val = chunked_cross_entropy(full[..., :-1, :], labels[..., 1:], chunk_size=0)
    assert full.shape == (2, 160, 512) and full.dtype == torch.float32
    e = rel_err(full[:, ::8, :].cpu(), t(g["logits_sub"]))
    assert e < 2e-2, f"logits rel err {e:.3e}"
    assert abs(float(val) - float(g["loss_val"])) < 2e-2, (float(val), float(g["loss_val"]))
    loss = train_step(m, idx, labels)
    ref = float(g["loss_train"])
    assert abs(float(loss) - ref) < max(2e-2 * 0.25, 5e-3 * ref), (float(loss), ref)
    check_grads(m, {k[5:]: t(g[k]) for k in g if k.startswith("grad.")}, 0.995, variant)


def test_tiny_relprompt_against_reference_golden(monkeypatch):
    g = load_golden("tiny_relprompt.npz")
    m, _ = build("relprompt", whisper_dim=48, raven_dim=32, pool_size=5)
    m.eval()
    B = g["idx"].shape[0]
    af = synthetic.synth_features(B, 103, 48, SEED, 301, DEV)
    vf = synthetic.synth_features(B, 52, 32, SEED, 302, DEV)
    # the fp32 CUDA-core head reproduces the reference's fp32 logits to 1e-4; the default tensor-core head (bf16 operands, what the
    # rest of this test trains through) to bf16 tolerance
    monkeypatch.setenv("DH_RELPRED", "fp32")
    al, vl = m.audio_noise_classifier(af), m.visual_noise_classifier(vf)
    assert rel_err(al.detach().cpu(), t(g["audio_logits"])) < 1e-4 and rel_err(vl.detach().cpu(), t(g["visual_logits"])) < 1e-4
    monkeypatch.delenv("DH_RELPRED")
    al, vl = m.audio_noise_classifier(af), m.visual_noise_classifier(vf)
    assert rel_err(al.detach().cpu(), t(g["audio_logits"])) < 1.5e-2 and rel_err(vl.detach().cpu(), t(g["visual_logits"])) < 1.5e-2
    at, vt = t(g["audio_targets"]).to(DEV), t(g["visual_targets"]).to(DEV)
    # finetune/relprompt.py:360-387 (trim to the common length), through this package's CE kernel
    na, nv = min(al.size(1), at.size(1)), min(vl.size(1), vt.size(1))
    import torch.nn.functional as F
    ml = F.cross_entropy(al[:, :na].reshape(-1, 3), at[:, :na].reshape(-1)) + \
        F.cross_entropy(vl[:, :nv].reshape(-1, 3), vt[:, :nv].reshape(-1))     # 3-class host-side glue of the driver
    idx, labels = t(g["idx"]).to(DEV), t(g["labels"]).to(DEV)
    logits = m(idx, audio_query=None, lip_query=None, max_seq_length=160, lm_head_chunk_size=128)
    logits[-1] = logits[-1][..., :-1, :]
    llm = chunked_cross_entropy(logits, labels[..., 1:])
    loss = llm + 0.02 * ml
    loss.backward()
    assert abs(float(ml) - float(g["mask_loss"])) < 1e-2
    assert abs(float(llm) - float(g["llm_loss"])) < 5e-3
    ref = {}
    for k in g:
        if k.startswith("grad."):
            ref[k[5:]] = t(g[k])
    n = 0
    for name, p in m.named_parameters():
        if not p.requires_grad:
            continue
        got = p.grad[::4, ::4, :] if "conv2.weight" in name else p.grad       # the fixture stores conv2 subsampled
        c = cos_sim(got, ref[name])
        nr = float(got.double().norm().cpu() / ref[name].double().norm().clamp_min(1e-30))
        assert c >= 0.995 and abs(nr - 1) <= 5e-2, f"relprompt grad {name}: cosine {c:.5f} norm ratio {nr:.4f}"
        n += 1
    assert n == len(ref)


def test_greedy_ids_match_reference_golden():
    g = load_golden("tiny_lora.npz")
    m, sd = build("lora")
    prompt = t(g["gen_prompt"]).to(DEV)
    ids = generate(m, prompt, prompt.numel() + 16, temperature=0.2, top_k=1, eos_id=None)
    m.reset_cache()
    ref = g["gen_ids"].tolist()
    # identical ids wherever the fp32 reference's top-1/top-2 logit margin exceeds the bf16 logit tolerance
    cfg = tiny_cfg("lora")
    _, step_logits = O.generate_greedy(sd, cfg, t(g["gen_prompt"]), prompt.numel() + 16, return_logits=True)
    got = ids.tolist()
    for i, lg in enumerate(step_logits):
        top2 = lg.topk(2).values
        margin = float(top2[0] - top2[1])
        pos = prompt.numel() + i
        if got[pos] != ref[pos]:
            assert margin < 0.05, f"token {i}: got {got[pos]} ref {ref[pos]} with a clear margin {margin:.3f}"
            break
    else:
        assert got == ref


def test_batched_decode_equals_single_and_uncached():
    """Batched KV-cache decode == single-sequence decode (same kernels, same bits), and == uncached arg-max over
    the full sequence wherever the top-1/top-2 margin is clear of bf16 noise (a random-init toy model has
    near-uniform logits, so the comparison stops at the first position whose margin is inside the noise)."""
    m, sd = build("adapter", adapter_start_layer=1)
    prompts = synthetic.synth_prompts(3, 512, SEED, prompt_range=(20, 60))
    outs = generate_batch(m, [p.to(DEV) for p in prompts], 8)
    checked = 0
    for p, o in zip(prompts, outs):
        single = generate(m, p.to(DEV), p.numel() + 8, temperature=0.2, top_k=1)
        assert single.tolist() == o.tolist()
        with torch.no_grad():
            full = m(o[:-1].view(1, -1).to(DEV))[0, p.numel() - 1:].float().cpu()
        for i in range(8):
            top2 = full[i].topk(2).values
            if float(top2[0] - top2[1]) < 0.02 * float(full[i].abs().max()):     # inside bf16 noise of the logit scale
                break
            assert int(full[i].argmax()) == int(o[p.numel() + i]), f"token {i} differs with margin {float(top2[0] - top2[1]):.3f}"
            checked += 1
    assert checked >= 1
    m.reset_cache()


def test_graph_decode_equals_eager_and_is_reusable():
    """The captured per-token CUDA graph (generate.DecodeLoop) produces the same ids as the eager launch sequence,
    is re-used by later calls with other prompts / token budgets / trainable values, and honours `eos_id` (the cut
    is inclusive, generate/base.py:72-74)."""
    m, _ = build("adapter", adapter_start_layer=1)
    pa = [p.to(DEV) for p in synthetic.synth_prompts(4, 512, SEED + 1, prompt_range=(20, 60))]
    pb = [p.to(DEV) for p in synthetic.synth_prompts(4, 512, SEED + 2, prompt_range=(10, 90))]
    eager_a = generate_batch(m, pa, 12, use_graph=False)
    graph_a = generate_batch(m, pa, 12, use_graph=True)           # first token eager, then capture + replay
    loop = m._engine._bufs["decode_loops"][4]
    g0 = loop.graph
    assert g0 is not None
    graph_b = generate_batch(m, pb, 9, use_graph=True)            # pure replay of the same graph
    eager_b = generate_batch(m, pb, 9, use_graph=False)
    assert m._engine._bufs["decode_loops"][4].graph is g0
    for e, g in zip(eager_a + eager_b, graph_a + graph_b):
        assert e.tolist() == g.tolist()
    # a changed trainable tensor must reach the replayed graph (DecodeLoop.refresh)
    with torch.no_grad():
        for n_, p_ in m.named_parameters():
            if "gating_factor" in n_:
                p_.mul_(-3.0)
    eager_c = generate_batch(m, pa, 12, use_graph=False)
    graph_c = generate_batch(m, pa, 12, use_graph=True)
    assert [x.tolist() for x in eager_c] == [x.tolist() for x in graph_c]
    assert [x.tolist() for x in eager_c] != [x.tolist() for x in eager_a], "the gate change did not alter the decode"
    # EOS: stop id = the 3rd generated token of sequence 0
    eos = int(graph_c[0][pa[0].numel() + 2])
    cut = generate_batch(m, pa, 12, eos_id=eos, poll_every=4)
    for full, c in zip(graph_c, cut):
        gen = full[len(full) - 12:].tolist()
        want = full.tolist() if eos not in gen else full[: len(full) - 12 + gen.index(eos) + 1].tolist()
        assert c.tolist() == want
    m.reset_cache()


@pytest.mark.parametrize("n", [100, 200])
def test_wide_batch_decode_equals_single(n):
    """100 / 200 sequences per token step (one / two 128-row activation boxes of the decode GEMM, n x G attention blocks): every
    sequence decodes to the ids it gets alone -- bench.py decodes 256 utterances per GPU and step."""
    m, _ = build("adapter", adapter_start_layer=1)
    prompts = [p.to(DEV) for p in synthetic.synth_prompts(n, 512, SEED + 9, prompt_range=(8, 70))]
    wide = generate_batch(m, prompts, 10)
    for i in (0, 1, 37, 63, 64, 65, 98, n - 1, n - 2, n // 2 + 29):
        assert wide[i].tolist() == generate_batch(m, [prompts[i]], 10)[0].tolist(), i
    m.reset_cache()


def test_continuous_batching_service_equals_per_utterance_decode():
    """serve.DecodeService: 3 cache slots serve 7 prompts with different token budgets, slots are refilled
    as sequences stop (budget or EOS); every utterance gets the ids of decoding it on its own (generate/base.py:19-82)."""
    from dualhyp_b200.serve import DecodeService
    m, _ = build("adapter", adapter_start_layer=1)
    prompts = [p.to(DEV) for p in synthetic.synth_prompts(7, 512, SEED + 5, prompt_range=(12, 80))]
    budgets = [5, 9, 3, 12, 6, 4, 8]
    alone = [generate_batch(m, [p], b)[0] for p, b in zip(prompts, budgets)]
    svc = DecodeService(m, n_slots=3, max_new_tokens=12, poll_every=2)
    got = svc.run(prompts, budgets)
    for a, g, p, b in zip(alone, got, prompts, budgets):
        assert g.numel() == p.numel() + b
        assert a.tolist() == g.tolist()
    assert svc.stats["generated"] == sum(budgets) and svc.stats["prefills"] >= 3
    # fewer token steps than decoding [0:3], [3:6], [6:7] as three batches that each wait for their longest answer
    assert svc.stats["token_steps"] <= 9 + 12 + 8 + 3 * 2
    # EOS: an id the answers contain stops those sequences early (cut inclusive); the freed slot is re-used
    eos = int(alone[3][prompts[3].numel() + 4])
    alone_e = [generate_batch(m, [p], b, eos_id=eos)[0] for p, b in zip(prompts, budgets)]
    svc_e = DecodeService(m, n_slots=3, max_new_tokens=12, eos_id=eos, poll_every=3)
    got_e = svc_e.run(prompts, budgets)
    assert [x.tolist() for x in alone_e] == [x.tolist() for x in got_e]
    assert alone_e[3].numel() < alone[3].numel()
    # the batch decode of the same width still works afterwards (shared per-width buffers), and so does a second run
    again = generate_batch(m, prompts[:3], 6)
    assert [x.tolist()[: p.numel() + 3] for x, p in zip(again, prompts[:3])] == [x.tolist()[: p.numel() + 3] for x, p in zip(alone[:3], prompts[:3])]
    assert [x.tolist() for x in svc.run(prompts[::-1], budgets[::-1])] == [x.tolist() for x in alone[::-1]]
    m.reset_cache()


def test_relprompt_inference_glue():
    """inference/relprompt.py:102-173 on tensors: classifier arg-max (checked against the oracle's classifier), mask
    tokens spliced into the prompts, batched decode == the same prompts decoded one by one through generate()."""
    from dualhyp_b200 import relprompt_infer as R
    m, sd = build("relprompt", whisper_dim=48, raven_dim=32, pool_size=5)
    m.eval()
    cfg = tiny_cfg("relprompt")
    Bn = 3
    af = synthetic.synth_features(Bn, 93, cfg.whisper_dim, SEED, 71, DEV)
    vf = synthetic.synth_features(Bn, 47, cfg.raven_dim, SEED, 72, DEV)
    a_pred, v_pred = R.predict_masks(m, af, vf, audio_frames=[93, 80, 41], visual_frames=[47, 40, 21])
    ref_a = O.noise_mask_classifier(af.cpu(), sd, "audio_noise_classifier.", 2 * cfg.pool_size)
    ref_v = O.noise_mask_classifier(vf.cpu(), sd, "visual_noise_classifier.", cfg.pool_size)
    assert [p.numel() for p in a_pred] == [10, 8, 5] and [p.numel() for p in v_pred] == [10, 8, 5]
    for b in range(Bn):
        ra, rv = ref_a[b, : a_pred[b].numel()], ref_v[b, : v_pred[b].numel()]
        clear_a = (ra.topk(2).values[:, 0] - ra.topk(2).values[:, 1]) > 2e-2 * ra.abs().max()      # bf16 head: skip near-ties
        clear_v = (rv.topk(2).values[:, 0] - rv.topk(2).values[:, 1]) > 2e-2 * rv.abs().max()
        assert torch.equal(a_pred[b].cpu()[clear_a], ra.argmax(-1)[clear_a])
        assert torch.equal(v_pred[b].cpu()[clear_v], rv.argmax(-1)[clear_v])
    prompts = [p.to(DEV) for p in synthetic.synth_prompts(Bn, 500, SEED + 5, prompt_range=(24, 40))]
    slots = [(5, 17), (3, 20), (8, 9)]
    first = m.transformer.wte.weight.shape[0] - 3
    outs, _, _ = R.relprompt_generate(m, prompts, slots, af, vf, max_new_tokens=6, audio_frames=[93, 80, 41],
                                      visual_frames=[47, 40, 21], first_mask_id=first)
    filled = [R.fill_mask_tokens(prompts[b], a_pred[b], v_pred[b], slots[b][0], slots[b][1], first) for b in range(Bn)]
    direct = generate_batch(m, filled, 6)                                     # the same decode, prompts spliced by hand
    for b in range(Bn):
        assert filled[b].numel() == prompts[b].numel() - 2 + a_pred[b].numel() + v_pred[b].numel()
        assert int(filled[b].max()) >= first                                  # mask-token rows of the resized wte are used
        assert outs[b][: filled[b].numel()].tolist() == filled[b].tolist()
        assert outs[b].tolist() == direct[b].tolist()
    m.reset_cache()


def test_packed_prefill_equals_padded_prefill():
    """generate_batch prefills the prompts as one packed row (cu_seqlens, no pad tokens) by default; right-padding them
    like the reference's collate_fn must give the same ids (the per-row GEMM results and the per-sequence attention tiles
    do not depend on where a sequence sits in the batch)."""
    for variant in ("adapter", "lora"):
        m, _ = build(variant, **({"adapter_start_layer": 1} if variant == "adapter" else {}))
        prompts = [p.to(DEV) for p in synthetic.synth_prompts(5, 512, SEED + 7, prompt_range=(3, 150))]
        a = generate_batch(m, prompts, 10, packed_prefill=True)
        b = generate_batch(m, prompts, 10, packed_prefill=False)
        c = generate_batch(m, [p.cpu() for p in prompts], 10, packed_prefill=True)        # host prompts
        assert [x.tolist() for x in a] == [x.tolist() for x in b] == [x.tolist() for x in c]
        m.reset_cache()


@pytest.mark.parametrize("variant,name,n,plen", [("adapter", "tiny-llama-test", 5, (3, 60)), ("adapter", "tiny-llama-test", 128, (3, 100)),
                                                 ("lora", "tiny-llama-test", 7, (3, 60)), ("adapter", "tiny-llama-1.1b", 37, (100, 300))])
def test_persistent_decode_kernel_matches_per_gemm_path(variant, name, n, plen, monkeypatch):
    """(The persistent kernel is opt-in, DH_DECODE_MEGA=1: measured slower than the per-GEMM graph)
    One decode token step of all layers as ONE persistent kernel (csrc/decode.cu: stream-K weight streaming over all SMs,
    grid barriers, partial tiles summed by the stage that follows) against the per-GEMM kernel sequence it replaces: same
    ln_f state (bf16 tolerance: the fp32 partial sums are associated differently), same new K / V rows in the cache, and
    bit-identical results from launch to launch (the stages sum their partial tiles in a fixed order)."""
    monkeypatch.setenv("DH_DECODE_MEGA", "1")
    m, _ = build(variant, name=name, **({"adapter_start_layer": 1} if (variant == "adapter" and name == "tiny-llama-test") else {}))
    if variant == "lora":
        lora.merge_lora_weights(m)               # generate/lora.py:112 merges before decoding; un-merged factors take the per-GEMM path
    vocab = m.config.padded_vocab_size
    prompts = [p.to(DEV) for p in synthetic.synth_prompts(n, vocab, SEED + 21, prompt_range=plen)]
    with torch.inference_mode():
        generate_batch(m, prompts, 2, use_graph=False)                 # fills the cache rows, leaves the engine's decode buffers in place
        eng = m._engine
        names, tensors = m._trainables()
        packed = eng.pack_trainable(dict(zip(names, tensors)))
        pkv = eng._project_prefixes(packed, "dec_pkv_all")
        assert eng.decode_plan_supported(n, packed)
        pos = torch.tensor([int(p.numel()) + 2 for p in prompts], dtype=torch.int32, device=DEV)
        tok = synthetic.hash_ints((n,), SEED, 77, 3, vocab).to(DEV)
        rows = torch.arange(n, device=DEV)
        h_old = eng.decode_step(tok, pos, packed, pkv).clone()
        k_old = [ck[rows, :, pos.long()].clone() for ck in eng.cache_k]
        v_old = [cv[rows, :, pos.long()].clone() for cv in eng.cache_v]
        for ck, cv in zip(eng.cache_k, eng.cache_v):
            ck[rows, :, pos.long()] = 0
            cv[rows, :, pos.long()] = 0
        plan = eng.build_decode_plan(n, pos, packed, pkv)
        h_new = eng.decode_step(tok, pos, packed, pkv, plan=plan).clone()
        torch.cuda.synchronize()
        assert torch.isfinite(h_new.float()).all()
        e = rel_err(h_new.float().cpu(), h_old.float().cpu())
        assert e < 2e-2, f"ln_f state rel err {e:.3e}"
        for i in range(len(eng.cache_k)):
            ek = rel_err(eng.cache_k[i][rows, :, pos.long()].float().cpu(), k_old[i].float().cpu())
            ev = rel_err(eng.cache_v[i][rows, :, pos.long()].float().cpu(), v_old[i].float().cpu())
            assert ek < 2e-2 and ev < 2e-2, f"layer {i}: new K / V rows rel err {ek:.3e} / {ev:.3e}"
        h_again = eng.decode_step(tok, pos, packed, pkv, plan=plan).clone()       # second launch: generation counter, determinism
        assert torch.equal(h_new, h_again)
    m.reset_cache()


def _unpad(idx, labels):
    """Samples of a synthetic right-padded batch (the response span is the last labelled stretch of every row)."""
    out = []
    for b in range(idx.shape[0]):
        n = int((labels[b] != -1).nonzero().max()) + 1
        out.append({"input_ids": idx[b, :n].clone(), "labels": labels[b, :n].clone()})
    return out


def packed_train_step(m, samples):
    """The driver's micro-step on a packed batch (collate.pack_varlen -> GPT.forward_packed -> chunked_cross_entropy)."""
    from dualhyp_b200.collate import pack_varlen
    pk = {k: v.to(DEV) for k, v in pack_varlen(samples).items()}
    longest = max(int(s["input_ids"].numel()) for s in samples)
    logits = m.forward_packed(pk["input_ids"], pk["cu_seqlens"], pk["positions"], longest, lm_head_chunk_size=128)
    loss = chunked_cross_entropy(logits, pk["targets"])
    loss.backward()
    return loss


@pytest.mark.parametrize("variant", ["adapter", "lora"])
def test_packed_training_step_equals_padded_step(variant):
    """the same samples as one packed row (no pad tokens, cu_seqlens attention forward AND backward) give the
    loss and PEFT gradients of the reference's right-padded batch -- including the train reduction's denominator, the
    B * (T - 1) positions of the padded batch (quirk Q2).  Also the opt-in `skip_pad` mode of the unchanged [B, T] call."""
    m, _ = build(variant, **({"adapter_start_layer": 1} if variant == "adapter" else {}))
    idx, labels = synthetic.synth_batch(5, 160, 512, SEED + 3, prompt_range=(20, 130))
    samples = _unpad(idx, labels)
    longest = max(int(s["input_ids"].numel()) for s in samples)
    assert longest < 160 and min(int(s["input_ids"].numel()) for s in samples) < longest
    idx_p, labels_p = idx[:, :longest].to(DEV), labels[:, :longest].to(DEV)       # what the reference's collate_fn builds
    loss_pad = train_step(m, idx_p, labels_p)
    g_pad = {n: p.grad.clone() for n, p in m.named_parameters() if p.requires_grad}
    m.zero_grad(set_to_none=True)
    loss_pk = packed_train_step(m, samples)
    assert abs(float(loss_pk) - float(loss_pad)) <= 2e-6 * max(1.0, abs(float(loss_pad))), (float(loss_pk), float(loss_pad))
    for n, p in m.named_parameters():
        if p.requires_grad:
            assert cos_sim(p.grad, g_pad[n]) > 0.99999, n
            assert abs(float(p.grad.double().norm() / g_pad[n].double().norm().clamp_min(1e-30)) - 1) < 1e-3, n
    m.zero_grad(set_to_none=True)
    m.skip_pad = True
    loss_sp = train_step(m, idx_p, labels_p)
    m.skip_pad = False
    assert abs(float(loss_sp) - float(loss_pad)) <= 2e-6 * max(1.0, abs(float(loss_pad)))
    for n, p in m.named_parameters():
        if p.requires_grad:
            assert cos_sim(p.grad, g_pad[n]) > 0.99999, n
    # validation reduction (mean over valid targets) on the packed row
    with torch.no_grad():
        from dualhyp_b200.collate import pack_varlen
        pk = {k: v.to(DEV) for k, v in pack_varlen(samples).items()}
        v_pk = chunked_cross_entropy(m.forward_packed(pk["input_ids"], pk["cu_seqlens"], pk["positions"], longest), pk["targets"], chunk_size=0)
        v_pad = chunked_cross_entropy(m(idx_p)[..., :-1, :], labels_p[..., 1:], chunk_size=0)
    assert abs(float(v_pk) - float(v_pad)) < 2e-3, (float(v_pk), float(v_pad))


def test_reference_generate_loop_runs_on_the_model():
    """generate/base.py's loop restated verbatim: model(x, input_pos) -> logits[0, -1] / temperature -> topk(1) ->
    where(< v) -> softmax -> multinomial, one token at a time through the KV cache.  bf16 logits can tie exactly at
    the maximum; the reference then samples among the tied ids (quirk Q7) while `dh_argmax` takes the lowest, so ids
    are compared up to the first exact tie and every sampled id must be a maximiser."""
    m, _ = build("lora")
    prompt = synthetic.synth_prompts(1, 512, 3, prompt_range=(30, 40))[0].to(DEV)
    T = prompt.numel()
    fast = generate(m, prompt, T + 6, temperature=0.2, top_k=1)
    m.reset_cache()
    idx = torch.empty(T + 6, dtype=prompt.dtype, device=DEV)
    idx[:T] = prompt
    input_pos = torch.arange(0, T, device=DEV)
    compared = 0
    with torch.inference_mode():
        for step in range(6):
            x = idx.index_select(0, input_pos).view(1, -1)
            logits = m(x, input_pos)
            logits = logits[0, -1] / 0.2
            n_max = int((logits == logits.max()).sum())
            v, _ = torch.topk(logits, 1)
            logits = torch.where(logits < v[[-1]], -float("Inf"), logits)
            probs = torch.nn.functional.softmax(logits, dim=-1)
            nxt = torch.multinomial(probs, num_samples=1).to(dtype=idx.dtype)
            assert float(probs[nxt]) > 0
            if n_max > 1:
                break
            assert int(nxt) == int(fast[T + step]), f"step {step}"
            compared += 1
            input_pos = input_pos[-1:] + 1
            idx = idx.index_copy(0, input_pos, nxt)
    m.reset_cache()
    assert compared >= 1


@pytest.mark.parametrize("variant,B,T", [("adapter", 2, 512), ("lora", 1, 384)])
def test_full_tinyllama_step_against_oracle(variant, B, T):
    """TinyLlama-1.1B shapes, random-init weights: loss and PEFT gradients vs the fp32 CPU oracle."""
    over = {} if variant == "lora" else {}
    m, sd = build(variant, name="tiny-llama-1.1b", **over)
    cfg = full_cfg(variant)
    idx, labels = synthetic.synth_batch(B, T, 32000, SEED, prompt_range=(int(T * 0.6), int(T * 0.9)))
    loss = train_step(m, idx.to(DEV), labels.to(DEV))
    torch.set_num_threads(max(1, torch.get_num_threads()))
    ref_loss, ref_grads = O.loss_and_grads(sd, cfg, idx, labels, "train")
    assert abs(float(loss) - float(ref_loss)) < max(2e-2 * float(ref_loss), 5e-3), (float(loss), float(ref_loss))
    check_grads(m, ref_grads, 0.999, variant)


def test_full_tinyllama_greedy_decode_against_oracle():
    """TinyLlama-1.1B shapes (the decode kernels at their real sizes: split-K skinny GEMMs over K = 2048 / 5632, mma.sync
    decode attention over a few hundred cached keys, CUDA-graph replay): batched greedy ids vs the fp32 CPU oracle's
    `generate_greedy` (generate/base.py restated).  bf16 logits of a random-init model are close to ties, so ids must be
    identical unless the oracle's top-1 / top-2 margin is inside the bf16 noise of the logit scale (3 %); a sequence is
    compared up to its first such divergence (later tokens then follow a different prefix)."""
    m, sd = build("adapter", name="tiny-llama-1.1b")
    cfg = full_cfg("adapter")
    prompts = synthetic.synth_prompts(2, 32000, SEED + 9, prompt_range=(150, 210))
    n_new = 5
    outs = generate_batch(m, [p.to(DEV) for p in prompts], n_new)
    clear = 0
    for p, o in zip(prompts, outs):
        ref, step_logits = O.generate_greedy(sd, cfg, p, p.numel() + n_new, return_logits=True)
        assert o[: p.numel()].tolist() == p.tolist()
        for i, lg in enumerate(step_logits):
            top2 = lg.float().topk(2).values
            margin, scale = float(top2[0] - top2[1]), float(lg.float().abs().max())
            pos = p.numel() + i
            if int(o[pos]) != int(ref[pos]):
                assert margin < 0.03 * scale, f"token {i}: got {int(o[pos])} oracle {int(ref[pos])} margin {margin:.4f} scale {scale:.3f}"
                break                                   # a near-tie resolved the other way: the sequences diverge from here
            clear += 1
    assert clear >= 4, f"only {clear} decode positions agreed with the oracle"
    m.reset_cache()


def test_lora_dropout_matches_oracle_on_the_same_masks():
    """LoRA dropout in training (ger/lora.py:94-98,165,373; finetune/ger.py:401 trains with p = 0.05).  The engine draws Philox
    masks from (seed, 2 * layer [+ 1]); the test reads the same masks back through dh_dropout_bf16 on a tensor of ones, installs
    them in the oracle (oracle.LORA_DROPOUT_MASKS) and compares loss and LoRA gradients; the same torch seed repeats the step
    bit for bit, another seed changes it, eval() ignores dropout."""
    from dualhyp_b200 import ops
    p_drop = 0.25
    m, sd = build("lora", dropout=p_drop)
    cfg = tiny_cfg("lora")
    B, T, C = 3, 48, cfg.n_embd
    idx, labels = synthetic.synth_batch(B, T, 512, SEED + 5, prompt_range=(20, 40))
    idx_d, labels_d = idx.to(DEV), labels.to(DEV)
    m.train()
    torch.manual_seed(777)
    seed = int(torch.randint(0, 2 ** 62, (1,), dtype=torch.int64).item())       # what GPT._dropout_seed() will draw next
    torch.manual_seed(777)
    loss = train_step(m, idx_d, labels_d)
    g1 = {n: p.grad.clone() for n, p in m.named_parameters() if p.grad is not None}
    ones = torch.ones((B * T, C), device=DEV, dtype=torch.bfloat16)
    masks = {}
    for i in range(cfg.n_layer):
        for off, which in ((0, "attn"), (1, "proj")):
            mk = torch.empty_like(ones)
            ops.dropout_bf16(ones, mk, p_drop, seed, 2 * i + off)
            masks[(i, which)] = mk.float().cpu().view(B, T, C)
    assert 0.6 < float((masks[(0, "attn")] != 0).float().mean()) < 0.9
    O.LORA_DROPOUT_MASKS = masks
    try:
        ref_loss, ref_grads = O.loss_and_grads(sd, cfg, idx, labels, "train")
    finally:
        O.LORA_DROPOUT_MASKS = None
    assert abs(float(loss) - float(ref_loss)) < max(2e-2 * 0.25, 5e-3 * float(ref_loss)), (float(loss), float(ref_loss))
    check_grads(m, ref_grads, 0.995, "lora dropout")
    ref0, _ = O.loss_and_grads(sd, cfg, idx, labels, "train")                    # the masks matter: p = 0 gives another loss
    assert abs(float(ref0) - float(ref_loss)) > 1e-4
    m.zero_grad(set_to_none=True)
    torch.manual_seed(777)
    loss2 = train_step(m, idx_d, labels_d)
    assert float(loss2) == float(loss)
    for n, p in m.named_parameters():
        if p.grad is not None:
            assert torch.equal(p.grad, g1[n]), n
    m.zero_grad(set_to_none=True)
    torch.manual_seed(778)
    assert float(train_step(m, idx_d, labels_d)) != float(loss)
    m.eval()
    m.zero_grad(set_to_none=True)
    le = train_step(m, idx_d, labels_d)
    assert abs(float(le) - float(ref0)) < max(2e-2 * 0.25, 5e-3 * float(ref0))


@pytest.mark.parametrize("variant", ["adapter", "lora"])
def test_full_tinyllama_greedy_64_tokens_identical_to_oracle(variant):
    """TinyLlama-1.1B shapes, ALL 64 greedy ids of every sequence identical to the fp32 CPU oracle's `generate_greedy`
    (generate/base.py:19-82 restated).  A random-init decoder has near-tied logits, so the fixture is the `peaky` one of
    dualhyp_b200.synthetic (head tied to a permutation of the embedding, blended with the random head): the test first checks that
    the ORACLE's top-1 / top-2 margin is at least 10 % of the logit scale at every step -- several times the bf16 noise -- and then
    demands exact equality.  Adapter: prefix attention + gates in every layer; LoRA: the un-merged factors ride the decode
    GEMMs' K-extension.  Batched decode over the CUDA graph, prompts of different lengths."""
    if variant == "adapter":
        cfg_m = adapter.Config.from_name("tiny-llama-1.1b")
        m = adapter.GPT(cfg_m)
    else:
        cfg_m = lora.Config.from_name("tiny-llama-1.1b", **LORA)
        m = lora.GPT(cfg_m)
    sd = synthetic.synth_state_dict(cfg_m, variant, SEED, peaky=True)
    m.load_state_dict(sd, strict=True)
    m = m.to(DEV)
    cfg = full_cfg(variant)
    prompts = synthetic.synth_prompts(2, 32000, SEED + 9, prompt_range=(150, 210))
    n_new = 64
    outs = generate_batch(m, [p.to(DEV) for p in prompts], n_new)
    for p, o in zip(prompts, outs):
        ref, step_logits = O.generate_greedy(sd, cfg, p, p.numel() + n_new, return_logits=True)
        for i, lg in enumerate(step_logits):
            top2 = lg.float().topk(2).values
            assert float(top2[0] - top2[1]) >= 0.1 * float(lg.float().abs().max()), f"fixture not peaky at step {i}"
        assert o.tolist() == ref.tolist()
        assert len(set(o[p.numel():].tolist())) >= 16, "the decode should walk distinct ids, not repeat one"
    m.reset_cache()


@pytest.mark.parametrize("variant", ["adapter", "lora"])
def test_bench_shape_step_against_golden(variant):
    """The fine-tune step at bench.py's shapes (adapter 16 x 512 = the CTA-pair GEMMs' persistent schedule over 32 x N/256 tiles,
    the fused loss's valid-row compaction at ~3 % of the rows; LoRA 8 x 512) against the fp32 CPU oracle's loss and PEFT
    gradients, computed once on the build box (oracle/make_golden_bench_shape.py -> tests/golden/bench_shape_*.npz).  Then the
    same samples packed into one row without their pad tokens (what bench.py times) must give the same loss and gradients."""
    from oracle.make_golden_bench_shape import bench_batch, model_config
    g = load_golden(f"bench_shape_{variant}.npz")
    cfg_m = model_config(variant)
    m = (adapter.GPT if variant == "adapter" else lora.GPT)(cfg_m)
    sd = synthetic.synth_state_dict(cfg_m, variant, SEED)
    m.load_state_dict(sd, strict=True)
    (adapter.mark_only_adapter_as_trainable if variant == "adapter" else lora.mark_only_lora_as_trainable)(m)
    m = m.to(DEV)
    idx, labels = bench_batch(variant)
    assert list(idx.shape) == g["shape"].tolist() and int(idx.sum()) == int(g["ids_checksum"]) and int(labels.sum()) == int(g["labels_checksum"])
    loss = train_step(m, idx.to(DEV), labels.to(DEV))
    ref_loss = float(g["loss_train"])
    assert abs(float(loss) - ref_loss) < max(2e-2 * ref_loss, 5e-3), (float(loss), ref_loss)
    got = {n: p.grad.detach().clone() for n, p in m.named_parameters() if p.requires_grad}
    n_full = 0
    for key in g:
        if key.startswith("norm/"):
            name = key[5:]
            nr = float(got[name].double().norm().cpu()) / max(float(g[key]), 1e-30)
            assert abs(nr - 1) <= 5e-2, f"{variant}: grad {name} norm ratio {nr:.4f}"
        elif key.startswith("grad/"):
            name = key[5:]
            c = cos_sim(got[name], t(g[key]))
            assert c >= MIN_COS_BENCH[variant], f"{variant}: grad {name} cosine {c:.5f}"
            n_full += 1
    assert n_full >= 6
    # the packed form of the same batch (no pad rows): same loss, same gradients up to bf16 re-association
    m.zero_grad(set_to_none=True)
    samples = _unpad(idx, labels)
    loss_p = packed_train_step(m, samples)
    # the train reduction divides by B * (T - 1) of the PADDED batch (quirk Q2, ger/utils.py:442-448); the packed call states the
    # collate's T = the longest sample, the golden batch is padded to 512
    longest = max(int(s_["input_ids"].numel()) for s_ in samples)
    scale = (longest - 1) / (idx.shape[1] - 1)
    assert abs(float(loss_p) * scale - float(loss)) < 2e-3 * max(1.0, abs(float(loss))), (float(loss_p), scale, float(loss))
    for name, p in m.named_parameters():
        if p.requires_grad:
            assert cos_sim(p.grad, got[name]) >= 0.999, name


MIN_COS_BENCH = {"adapter": 0.999, "lora": 0.999}


@pytest.mark.parametrize("B,T,C,pool", [(4, 500, 1280, 20), (4, 250, 1024, 10), (3, 103, 96, 10)])
def test_reliability_head_tensor_core_path_at_cfg4_shapes(B, T, C, pool):
    """`NoiseMaskClassifier` (ger/relprompt.py:126-147) at the shapes RelPrompt really runs (whisper_dim 1280 x 500 frames pooled
    by 20, raven_dim 1024 x 250 frames pooled by 10, ger/relprompt.py:212-213; 25 mask targets each) plus a ragged toy shape,
    on the default tensor-core path: logits and the gradients of all six parameter tensors against the fp32 CPU oracle.
    Tolerance: bf16 operands with fp32 accumulation over K = 3 C up to 3840 -- logits 1.5e-2 relative (Frobenius), gradients
    cosine >= 0.995 and norm within 5 % (see grads_close); the fp32 CUDA-core path (DH_RELPRED=fp32) is held to 1e-4 by
    tests/test_kernels_gpu.py::test_relpred_fwd_bwd.
    Then training mode: the Dropout(0.1) between the convolutions (:141) on the mask the kernel drew, read back through
    dh_dropout_bf16 on ones and replayed by the oracle."""
    from dualhyp_b200 import ops
    head = relprompt.NoiseMaskClassifier(C, pool_size=pool).to(DEV)
    p = "h."
    sd = {p + n: synthetic.hash_uniform(tuple(t.shape), 5, 80 + i, 1.0 / (3 * t.shape[1]) ** 0.5 if t.dim() == 3 else 0.06, DEV, bf16_exact=False)
          for i, (n, t) in enumerate(head.named_parameters())}
    with torch.no_grad():
        for n, t in head.named_parameters():
            t.copy_(sd[p + n])
    x = synthetic.synth_features(B, T, C, 3, 70, DEV)
    n_out = -(-T // pool)
    dl = synthetic.hash_uniform((B, n_out, 3), 11, 71, 1.0, DEV)

    def run(mask):
        work = {k: v.cpu().clone().requires_grad_(True) for k, v in sd.items()}
        ref = O.noise_mask_classifier(x.cpu(), work, p, pool, dropout_mask=mask)
        ref.backward(dl.cpu())
        return ref.detach(), {k: v.grad for k, v in work.items()}

    head.eval()
    logits = head(x)
    logits.backward(dl)
    ref, ref_g = run(None)
    assert tuple(logits.shape) == (B, n_out, 3)
    assert rel_err(logits.detach().cpu(), ref) < 1.5e-2, rel_err(logits.detach().cpu(), ref)
    def grads_close(ref_grads, what):
        # ReLU makes the gradient discontinuous in the pre-activations: a bf16-operand pre-activation within ~1 % of zero lands on
        # the other side for ~1 % of the elements, each a full-size error -- so the bound is the cosine / norm pair the decoder's
        # gradient checks use, not a Frobenius 1e-2 (measured here: 2-5 % Frobenius, growing from conv2.bias back to conv1.weight)
        for n, t_ in head.named_parameters():
            c = cos_sim(t_.grad, ref_grads[p + n])
            nr = float(t_.grad.double().norm().cpu() / ref_grads[p + n].double().norm().clamp_min(1e-30))
            assert c >= 0.995 and abs(nr - 1) <= 5e-2, f"{what}: grad {n} cosine {c:.5f} norm ratio {nr:.4f}"

    grads_close(ref_g, "eval")
    # ---- training mode: dropout between the convolutions ----
    head.train()
    head.zero_grad(set_to_none=True)
    torch.manual_seed(4242)
    seed = int(torch.randint(0, 2 ** 62, (1,), dtype=torch.int64).item())
    torch.manual_seed(4242)
    logits_t = head(x)
    logits_t.backward(dl)
    ones = torch.ones((B * T, 256), device=DEV, dtype=torch.bfloat16)
    mk = torch.empty_like(ones)
    ops.dropout_bf16(ones, mk, head.dropout.p, seed, 0)
    mask = mk.float().cpu().view(B, T, 256)
    assert 0.85 < float((mask != 0).float().mean()) < 0.95
    ref_t, ref_gt = run(mask)
    assert rel_err(ref_t, ref) > 1e-3                       # the mask matters
    assert rel_err(logits_t.detach().cpu(), ref_t) < 1.5e-2
    grads_close(ref_gt, "train")


def test_generate_temperature_top_k_sampling():
    """generate/base.py:62-70 beyond the drivers' top_k = 1: temperature + top-k + multinomial.  (a) every sampled id lies inside the
    top-k set of the model's own logits for that position (recomputed without the cache; 3 extra ranks of slack for bf16 near-ties),
    (b) a fixed torch seed repeats the draw, another seed changes it, (c) temperature -> 0 over the whole vocabulary picks a maximum of every row."""
    m, _ = build("adapter", adapter_start_layer=1)
    prompt = synthetic.synth_prompts(1, 512, SEED + 3, prompt_range=(20, 40))[0].to(DEV)
    n_new, k = 12, 4
    torch.manual_seed(11)
    a = generate(m, prompt, prompt.numel() + n_new, temperature=0.8, top_k=k)
    torch.manual_seed(11)
    b = generate(m, prompt, prompt.numel() + n_new, temperature=0.8, top_k=k)
    torch.manual_seed(12)
    c = generate(m, prompt, prompt.numel() + n_new, temperature=0.8, top_k=k)
    assert a.tolist() == b.tolist() and a.tolist() != c.tolist()
    assert a[: prompt.numel()].tolist() == prompt.tolist() and a.numel() == prompt.numel() + n_new
    with torch.no_grad():
        full = m(a[:-1].view(1, -1).to(DEV))[0, prompt.numel() - 1:].float().cpu()
    for i in range(n_new):
        assert int(a[prompt.numel() + i]) in full[i].topk(k + 3).indices.tolist(), f"token {i} outside the top-{k} set"
    cold = generate(m, prompt, prompt.numel() + n_new, temperature=1e-4, top_k=None)
    with torch.no_grad():
        full_c = m(cold[:-1].view(1, -1).to(DEV))[0, prompt.numel() - 1:].float().cpu()
    for i in range(n_new):                              # every cold draw is an arg-max of its own row, up to bf16 ties
        row = full_c[i]
        assert float(row[int(cold[prompt.numel() + i])]) >= float(row.max()) - 0.03 * float(row.abs().max()), f"cold token {i} is not a maximum"
    m.reset_cache()
