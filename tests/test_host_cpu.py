"""CPU: host-side logic of the `ger` mirror -- config registry, module trees / state-dict keys against the
reference's (live where mounted, and by the golden key lists), checkpoint key remapping, LazyLogits slicing,
synthetic-input determinism, flop accounting."""
import pytest
import torch

from dualhyp_b200 import flops, synthetic
from dualhyp_b200.ger import adapter, lora, relprompt
from dualhyp_b200.ger.config import Config
from dualhyp_b200.ger.model import LazyLogits
from oracle import refshim
from tests.helpers import load_golden

LORA = dict(r=8, alpha=16, to_query=True, to_key=True, to_value=True, to_projection=True)


def test_tinyllama_config_resolves_by_hf_name():
    c = Config.from_name("TinyLlama-1.1B-Chat-v1.0")
    assert (c.n_layer, c.n_head, c.n_query_groups, c.n_embd, c.intermediate_size, c.padded_vocab_size, c.head_size) == \
        (22, 32, 4, 2048, 5632, 32000, 64)
    c.check_engine_supported()
    with pytest.raises(ValueError):
        Config.from_name("no-such-model")


def test_parameter_counts_match_survey():
    m = torch.device("meta")
    with m:
        lo = lora.GPT(lora.Config.from_name("tiny-llama-1.1b", **{**LORA, "r": 16}))
        ad = adapter.GPT(adapter.Config.from_name("tiny-llama-1.1b"))
    lora.mark_only_lora_as_trainable(lo)
    adapter.mark_only_adapter_as_trainable(ad)
    assert sum(p.numel() for p in lo.parameters() if p.requires_grad) == 4_505_600
    assert sum(p.numel() for p in ad.parameters() if p.requires_grad) == 410_240
    assert sum(p.numel() for p in lo.parameters()) == 1_104_617_984 + 0


def test_state_dict_keys_match_golden_gradient_keys():
    for variant, mod, extra in (("lora", lora, LORA), ("adapter", adapter, {"adapter_start_layer": 1})):
        g = load_golden(f"tiny_{variant}.npz")
        model = mod.GPT(mod.Config.from_name("tiny-llama-test", **extra))
        names = {n for n, _ in model.named_parameters()}
        for k in g:
            if k.startswith("grad."):
                assert k[5:] in names, k


@pytest.mark.skipif(not refshim.available(), reason="reference checkout not mounted")
def test_state_dict_keys_equal_live_reference():
    refshim.install()
    import ger.lora as RL
    rcfg = RL.Config(name="t", block_size=256, vocab_size=512, padded_vocab_size=512, n_layer=3, n_head=4, n_embd=256,
                     rotary_percentage=1.0, parallel_residual=False, bias=False, _norm_class="RMSNorm", norm_eps=1e-5,
                     _mlp_class="LLaMAMLP", intermediate_size=512, n_query_groups=2, **LORA)
    ref = RL.GPT(rcfg).state_dict()
    mine = lora.GPT(lora.Config.from_name("tiny-llama-test", **LORA)).state_dict()
    assert list(ref.keys()) == list(mine.keys())
    assert all(ref[k].shape == mine[k].shape for k in ref)


def test_base_checkpoint_keys_are_remapped():
    cfg = lora.Config.from_name("tiny-llama-test", **LORA)
    m = lora.GPT(cfg)
    res = m.load_state_dict(synthetic.synth_state_dict(cfg, "base", 3), strict=False)
    assert not res.unexpected_keys
    assert all("lora_" in k or "adapter_" in k for k in res.missing_keys)


def test_lazy_logits_slicing_stays_lazy():
    class Fake:
        class config:
            padded_vocab_size = 11
    h = torch.zeros(2, 10, 4)
    chunks = [LazyLogits(Fake, h[:, a:a + 4], h, a, min(10, a + 4)) for a in range(0, 10, 4)]
    assert [tuple(c.shape) for c in chunks] == [(2, 4, 11), (2, 4, 11), (2, 2, 11)]
    last = chunks[-1][..., :-1, :]
    assert isinstance(last, LazyLogits) and (last.t0, last.t1) == (8, 9) and last.base is h
    row = LazyLogits(Fake, h)[0, -1]
    assert isinstance(row, LazyLogits) and tuple(row.shape) == (11,)


def test_synthetic_inputs_are_deterministic_and_structured():
    a = synthetic.synth_batch(4, 512, 32000, 1337)
    b = synthetic.synth_batch(4, 512, 32000, 1337)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    ids, labels = a
    for r in range(4):
        valid = (labels[r] >= 0).nonzero().flatten()
        assert 8 <= valid.numel() <= 24 and 330 <= int(valid[0]) <= 480
        assert torch.equal(labels[r][valid], ids[r][valid])
        assert (ids[r][int(valid[-1]) + 1:] == 0).all()          # right padding with id 0, label -1
    w = synthetic.hash_uniform((64, 64), 1, 2, 0.1)
    assert torch.equal(w, w.bfloat16().float())                   # bf16-exact weights


def test_flop_accounting_matches_survey_figures():
    c = Config.from_name("tiny-llama-1.1b")
    assert abs(flops.forward_flops_per_token(c, 512) / 1e9 - 2.115) < 0.01
    c.lora_start_layer = 0
    assert abs((flops.forward_flops_per_token(c, 512) + flops.backward_flops_per_token(c, "lora", 512)) / 1e9 - 4.29) < 0.02
    c.adapter_start_layer = 2
    assert abs((flops.forward_flops_per_token(c, 512) + flops.backward_flops_per_token(c, "adapter", 512)) / 1e9 - 4.10) < 0.03


def test_trainable_only_checkpoint_roundtrip(tmp_path):
    """only the PEFT tensors are written (reference-keyed, loadable with strict=False on top of the
    frozen base), the load checks names / shapes / completeness, and the frozen weights are left untouched."""
    from dualhyp_b200 import checkpoint
    cfg = lora.Config.from_name("tiny-llama-test", **LORA)
    a, b = lora.GPT(cfg), lora.GPT(cfg)
    lora.mark_only_lora_as_trainable(a)
    lora.mark_only_lora_as_trainable(b)
    with torch.no_grad():
        for n, p in a.named_parameters():
            if p.requires_grad:
                p.normal_(0, 0.1)
    path = tmp_path / "iter-000010.pth"
    nbytes = checkpoint.save_trainable(a, path)
    n_train = sum(p.numel() for p in a.parameters() if p.requires_grad)
    assert nbytes == 4 * n_train and path.stat().st_size < nbytes + 64 * 1024
    stored = torch.load(str(path), weights_only=False)
    assert stored["format"] == checkpoint.FORMAT and all("lora_" in k for k in stored["model"])
    assert set(stored["model"]) == {n for n, p in a.named_parameters() if p.requires_grad}
    frozen_before = {n: p.clone() for n, p in b.named_parameters() if not p.requires_grad}
    loaded = checkpoint.load_trainable(b, path)
    assert len(loaded) == len(stored["model"])
    for (n, pa), (_, pb) in zip(a.named_parameters(), b.named_parameters()):
        if pa.requires_grad:
            assert torch.equal(pa, pb), n
        else:
            assert torch.equal(pb, frozen_before[n]), n
    # the reference's own filter selects the same tensors
    assert checkpoint.trainable_state(a, lora.lora_filter).keys() == stored["model"].keys()
    # a checkpoint of another variant is rejected, an incomplete one too
    ad = adapter.GPT(adapter.Config.from_name("tiny-llama-test", adapter_start_layer=1))
    adapter.mark_only_adapter_as_trainable(ad)
    with pytest.raises(KeyError):
        checkpoint.load_trainable(ad, path)
    partial = {"model": {k: v for i, (k, v) in enumerate(stored["model"].items()) if i % 2 == 0}}
    torch.save(partial, str(tmp_path / "partial.pth"))
    with pytest.raises(KeyError):
        checkpoint.load_trainable(b, tmp_path / "partial.pth")
    checkpoint.load_trainable(b, tmp_path / "partial.pth", require_all=False)


def test_relprompt_checkpoint_carries_resized_wte_rows(tmp_path):
    """resize_token_embeddings(3) appends random wte rows that mark_only_lora_as_trainable freezes (finetune/relprompt.py:168-169):
    a trainable-only checkpoint must carry them, restore them, and refuse a file without them."""
    from dualhyp_b200 import checkpoint
    from dualhyp_b200.ger import relprompt
    cfg = relprompt.Config.from_name("tiny-llama-test", r=4, alpha=8, whisper_dim=48, raven_dim=32, pool_size=5)
    torch.manual_seed(0)
    a = relprompt.GPT(cfg)
    a.resize_token_embeddings(3)
    relprompt.mark_only_lora_as_trainable(a)
    torch.manual_seed(1)
    b = relprompt.GPT(cfg)
    b.resize_token_embeddings(3)
    relprompt.mark_only_lora_as_trainable(b)
    V = cfg.padded_vocab_size
    with torch.no_grad():
        b.transformer.wte.weight[:V].copy_(a.transformer.wte.weight[:V])     # the shared frozen base checkpoint
    assert not torch.equal(a.transformer.wte.weight[V:], b.transformer.wte.weight[V:])
    path = tmp_path / "rel.pth"
    checkpoint.save_trainable(a, path)
    stored = torch.load(str(path), weights_only=False)["model"]
    assert checkpoint.EXTRA_WTE_KEY in stored and stored[checkpoint.EXTRA_WTE_KEY].shape == (3, cfg.n_embd)
    assert any("noise_classifier" in k for k in stored) and "transformer.wte.weight" not in stored
    v0 = b.transformer.wte.weight._version
    loaded = checkpoint.load_trainable(b, path)
    assert checkpoint.EXTRA_WTE_KEY in loaded and b.transformer.wte.weight._version > v0      # the engine's frozen signature sees it
    assert torch.equal(a.transformer.wte.weight, b.transformer.wte.weight)
    for (n, pa), (_, pb) in zip(a.named_parameters(), b.named_parameters()):
        if pa.requires_grad:
            assert torch.equal(pa, pb), n
    # a file without the rows is refused by a resized model; a file with rows is refused by a model that was not resized
    del stored[checkpoint.EXTRA_WTE_KEY]
    torch.save({"model": stored}, str(tmp_path / "norows.pth"))
    with pytest.raises(KeyError):
        checkpoint.load_trainable(b, tmp_path / "norows.pth")
    c = relprompt.GPT(cfg)
    relprompt.mark_only_lora_as_trainable(c)
    with pytest.raises(ValueError):
        checkpoint.load_trainable(c, path)


def test_merged_lora_modules_leave_the_trainable_set_and_bump_the_weight_version():
    """after merge_lora_weights() the factors live inside W; the engine must not add them a second time
    (GPT._trainables drops them) and must re-pack W (the in-place add bumps the parameter version the frozen signature reads)."""
    cfg = lora.Config.from_name("tiny-llama-test", **LORA)
    m = lora.GPT(cfg)
    lora.mark_only_lora_as_trainable(m)
    with torch.no_grad():
        for n, p in m.named_parameters():
            if "lora_B" in n:
                p.normal_(0, 0.05)
    names, _ = m._trainables()
    assert len(names) == 4 * cfg.n_layer
    w = m.transformer.h[0].attn.attn.linear.weight
    v0, before = w._version, w.detach().clone()
    lora.merge_lora_weights(m)
    assert w._version > v0 and not torch.equal(before, w)
    names2, _ = m._trainables()
    assert names2 == ()
    at = m.transformer.h[0].attn.attn
    r = cfg.r
    A, Bm = at.lora_A.float(), at.lora_B.float()
    q = cfg.n_embd
    kv = at.kv_embd_size
    delta = torch.cat((Bm[:q] @ A[:r], Bm[q:q + kv] @ A[r:2 * r], Bm[q + kv:] @ A[2 * r:])) * (cfg.alpha / cfg.r)
    assert torch.allclose(w.detach() - before, delta, atol=1e-6)
    lora.merge_lora_weights(m)                  # idempotent
    assert torch.allclose(w.detach() - before, delta, atol=1e-6)


def test_flat_bucket_adamw_bumps_parameter_versions(monkeypatch):
    """the fused AdamW writes the parameters through the flat buffer; the (data_ptr, _version) cache keys of
    the decode loops must see it."""
    from dualhyp_b200 import parallel
    ps = [torch.nn.Parameter(torch.randn(4, 3)), torch.nn.Parameter(torch.randn(5))]
    b = parallel.FlatTrainables(ps)
    monkeypatch.setattr(parallel.ops, "adamw_step", lambda *a, **k: None)
    v = [p._version for p in ps]
    b.adamw_step(lr=1e-3)
    assert all(p._version > v0 for p, v0 in zip(ps, v))


def test_relprompt_mask_token_splice_equals_text_substitution():
    """splicing mask-token ids == the reference's text replace + re-tokenise (inference/relprompt.py:145-157)
    for a tokenizer in which every mask token is one id (what add_special_tokens + resize_token_embeddings(3) gives)."""
    from dualhyp_b200 import relprompt_infer as R
    vocab = {"<<C>>": 500, "<<M>>": 501, "<<N>>": 502, R.ASR_PLACEHOLDER: 7, R.VSR_PLACEHOLDER: 7}
    words = "revise asr hyps : x y z ### Audio Mask: {} ### VSR : u v ### Video Mask: {} end".split()

    def encode(text):       # whitespace tokenizer with single-id special tokens
        out = []
        for w in text.split():
            while w:
                for tok, i in vocab.items():
                    if w.startswith(tok):
                        out.append(i)
                        w = w[len(tok):]
                        break
                else:
                    out.append(10 + (sum(map(ord, w)) % 400))
                    w = ""
        return torch.tensor(out)

    template = " ".join(words).format(R.ASR_PLACEHOLDER, R.VSR_PLACEHOLDER)
    a_pred, v_pred = torch.tensor([0, 2, 2, 1, 0]), torch.tensor([1, 1, 0])
    assert R.mask_strings(a_pred) == "<<C>><<N>><<N>><<M>><<C>>"
    via_text = encode(R.fill_prompt_text(template, a_pred, v_pred))
    ids = encode(template)
    slots = [i for i, t in enumerate(ids.tolist()) if t == 7]
    via_ids = R.fill_mask_tokens(ids, a_pred, v_pred, slots[0], slots[1], first_mask_id=500)
    assert via_ids.tolist() == via_text.tolist()
    assert via_ids.numel() == ids.numel() - 2 + 5 + 3


def test_prompt_assembly_label_masking_and_collation():
    """SURVEY 8f row 2 (host side): the three dataset classes' placeholder substitution, the -1 label span over the
    prompt, right-padded collation as the reference's collate_fn, and the pad-free packed layout."""
    import os
    from dualhyp_b200 import collate as Cc
    if os.path.isfile("/root/reference/data/prompts.py"):           # the real templates where the checkout is mounted
        ns = {}
        exec(open("/root/reference/data/prompts.py").read(), ns)
        ger, dual, relp = ns["GER_PROMPTS"], ns["DualHyp_PROMPTS"], ns["RelPrompt_PROMPTS"]
    else:
        ger = {"prompt_1": "revise:\n", "prompt_2": "\n\nothers:", "prompt_3": "\n\nresponse:\n"}
        dual = {"prompt_1": "asr <<<ASR_NHYPS>>> vsr <<<VSR_NHYPS>>>", "prompt_2": "\nA:\n<<<ASR_NHYPS>>>\nV:\n<<<VSR_NHYPS>>>", "prompt_3": "\nresponse:\n"}
        relp = {"prompt_1": "a <<<ASR_BEST_NHYPS>>> <<<ASR_NHYPS>>> m <<<ASR_MASKS>>> v <<<VSR_BEST_NHYPS>>> <<<VSR_NHYPS>>> m <<<VSR_MASKS>>>",
                "prompt_2": "", "prompt_3": "\nresponse:\n"}
    asr = ["the cat sat", "the cat sad", "a cat sat", "the cap sat", "the cat set"]
    vsr = ["the bat sat", "the cat sat", "the mat sat", "the cat fat", "the cat hat"]
    g = Cc.ger_prompt(ger, asr)
    assert g.startswith(ger["prompt_1"] + asr[0]) and g.endswith(ger["prompt_3"]) and "\n".join(asr[1:]) in g
    assert asr[4] not in Cc.ger_prompt(ger, asr, max_nhyps=3)
    d = Cc.dualhyp_prompt(dual, asr, vsr)
    assert "<<<" not in d and asr[0] in d and vsr[0] in d and "\n".join(vsr[1:]) in d and d.endswith(dual["prompt_3"])
    r_inf = Cc.relprompt_prompt(relp, asr, vsr)
    assert "<<<ASR_MASKS>>>" in r_inf and "<<<VSR_MASKS>>>" in r_inf and "<<<ASR_NHYPS>>>" not in r_inf
    r_tr = Cc.relprompt_prompt(relp, asr, vsr, ["<<C>>", "<<N>>"], ["<<M>>"])
    assert "<<C>><<N>>" in r_tr and "<<<" not in r_tr.replace("<<C>>", "").replace("<<N>>", "").replace("<<M>>", "")

    enc = lambda t: [1] + [3 + (ord(c) % 200) for c in t]          # char tokenizer with a BOS id, like the llama one
    s1 = Cc.encode_sample(enc, d, "the cat sat", "</s>")
    s2 = Cc.encode_sample(enc, g, "a cat", "</s>")
    n1 = s1["input_ids_no_response"].numel()
    assert torch.equal(s1["input_ids"][:n1], s1["input_ids_no_response"])
    assert (s1["labels"][:n1] == -1).all() and torch.equal(s1["labels"][n1:], s1["input_ids"][n1:])
    b = Cc.collate_right_pad([s1, s2])
    T = max(s1["input_ids"].numel(), s2["input_ids"].numel())
    assert b["input_ids"].shape == (2, T) and b["labels"].shape == (2, T)
    short = 1 if s2["input_ids"].numel() < s1["input_ids"].numel() else 0
    ns_ = [s1, s2][short]["input_ids"].numel()
    assert (b["input_ids"][short, ns_:] == 0).all() and (b["labels"][short, ns_:] == -1).all()
    pk = Cc.pack_varlen([s1, s2])
    assert pk["cu_seqlens"].tolist() == [0, s1["input_ids"].numel(), s1["input_ids"].numel() + s2["input_ids"].numel()]
    assert pk["positions"][pk["cu_seqlens"][1]] == 0 and pk["targets"][pk["cu_seqlens"][1] - 1] == -1
    # packed next-token targets == the padded batch's shifted labels on the valid positions
    for i, s in enumerate([s1, s2]):
        a, e = pk["cu_seqlens"][i].item(), pk["cu_seqlens"][i + 1].item()
        assert torch.equal(pk["targets"][a:e - 1], b["labels"][i, 1:e - a])


class _CharTokenizer:
    """Stand-in for the llama tokenizer: BOS id 1, one id per character."""
    eos_token = "</s>"

    def encode(self, text):
        return [1] + [3 + (ord(c) % 200) for c in text]


def _write_nbest_json(path, n_uids=3, variants=2):
    import json
    recs = []
    for u in range(n_uids):
        for v in range(variants):
            recs.append({"Uid": f"utt{u}", "Caption": f"the true words {u}",
                         "nhyps_asr": {"hyps": [f"asr{u}.{v} best"] + [f"asr{u}.{v} other {k}" for k in range(4)]},
                         "nhyps_vsr": {"hyps": [f"vsr{u}.{v} best"] + [f"vsr{u}.{v} other {k}" for k in range(4)]}})
    with open(path, "w") as f:
        json.dump(recs, f)
    return recs


def test_av_dataset_shim_exports_the_names_the_drivers_import(tmp_path):
    """SURVEY 8(b) shims / 8f row 2: `from data.av_dataset import AVDataset, DualHypothesesAVDataset, DualHypothesesAlignAVDataset,
    DualHypothesesAlign2AVDataset` (finetune/ger.py:36) resolves after install_as_data(); the classes take the keyword arguments the
    driver passes (finetune/ger.py:149-160), index the JSON by Uid, build the reference's prompt text and mask the prompt span."""
    import random
    import sys
    from dualhyp_b200 import av_dataset as A
    A.install_as_data()
    from data.av_dataset import AVDataset, DualHypothesesAVDataset, DualHypothesesAlignAVDataset, DualHypothesesAlign2AVDataset
    assert sys.modules["data.av_dataset"] is A
    path = str(tmp_path / "train.json")
    recs = _write_nbest_json(path)
    tok = _CharTokenizer()
    driver_kwargs = dict(nhyps_key="nhyps_asr", max_input_length=-1, max_nhyps=3, tokenizer=tok, audio_corruption_enabled=False,
                         visual_corruption_enabled=False, apply_chat_template=False)

    ger = AVDataset(split="train", json_path=path, prompts_format="GER", random_sample_nhyps=False, **driver_kwargs)
    assert len(ger) == 3 and ger.idx2uid == ["utt0", "utt1", "utt2"]
    random.seed(5)
    item = ger[1]
    rec = next(r for r in recs if r["Uid"] == "utt1" and item["input"].find(r["nhyps_asr"]["hyps"][0]) >= 0)
    h = rec["nhyps_asr"]["hyps"]
    expect = A.GER_PROMPTS["prompt_1"] + h[0] + A.GER_PROMPTS["prompt_2"] + "\n" + "\n".join(h[1:3]) + A.GER_PROMPTS["prompt_3"]
    assert item["input"] == expect + rec["Caption"] + "</s>" and item["uid"] == "utt1" and item["ground_truth"] == rec["Caption"]
    n = item["input_ids_no_response"].numel()
    assert item["input_ids_no_response"].tolist() == tok.encode(expect)
    assert (item["labels"][:n] == -1).all() and torch.equal(item["labels"][n:], item["input_ids"][n:])
    assert item["audio"] is None and item["video"] is None

    dual = DualHypothesesAVDataset(split="train", json_path=path, prompts_format="DualHyp", random_sample_nhyps=False, **driver_kwargs)
    seen_mixed = False
    for s in range(20):                                               # independent draws: ASR and VSR lists of different records do occur
        random.seed(s)
        t = dual[0]["input"]
        seen_mixed |= ("asr0.0" in t and "vsr0.1" in t) or ("asr0.1" in t and "vsr0.0" in t)
        assert "<<<" not in t and t.endswith("the true words 0</s>")
    assert seen_mixed

    al = DualHypothesesAlignAVDataset(split="train", json_path=path, prompts_format="DualAlign", **driver_kwargs)
    al2 = DualHypothesesAlign2AVDataset(split="train", json_path=path, prompts_format="DualAlign2", **driver_kwargs)
    for s in range(20):                                               # aligned: both lists always from the same record
        random.seed(s)
        t = al[2]["input"]
        assert ("asr2.0" in t) != ("asr2.1" in t) and ("asr2.0" in t) == ("vsr2.0" in t)
        t2 = al2[2]["input"]
        v = "0" if "asr2.0" in t2 else "1"
        assert f"asr2.{v} other 0\nasr2.{v} other 1" in t2 and f"vsr2.{v} other 0\nvsr2.{v} other 1" in t2       # n-best order kept
    short = DualHypothesesAlignAVDataset(split="val", json_path=path, **{**driver_kwargs, "max_input_length": 40})
    assert short[0]["input_ids"].numel() == 40 and short[0]["labels"].numel() == 40

    batch = al.collate_fn([al[0], al[1], al[2]], pack=True)
    assert batch["input_ids"].shape == batch["labels"].shape and batch["input_ids"].shape[0] == 3 and len(batch["uid"]) == 3
    assert batch["packed"]["cu_seqlens"].numel() == 4 and batch["packed"]["input_ids"].numel() == int(batch["packed"]["cu_seqlens"][-1])
    assert al.get_max_seq_length()[0] >= 40
    with pytest.raises(ValueError):
        AVDataset(split="train", json_path=path, prompts_format="NoSuchFormat", tokenizer=tok)
    both = AVDataset(split="train", json_path=[path, str(tmp_path / "x_pretrain.json")], tokenizer=tok) if _write_nbest_json(str(tmp_path / "x_pretrain.json"), 2, 1) else None
    assert len(both) == 5 and "utt0_pretrain" in both.uid2sample


@pytest.mark.skipif(not __import__("os").path.isfile("/root/reference/data/prompts.py"), reason="reference checkout not mounted")
def test_av_dataset_shim_templates_equal_the_live_reference():
    from dualhyp_b200 import av_dataset as A
    ns = {}
    exec(open("/root/reference/data/prompts.py").read(), ns)
    assert A.GER_PROMPTS == ns["GER_PROMPTS"] and A.DUALHYP_PROMPTS == ns["DualHyp_PROMPTS"]
    assert A.get_prompts_format("GER") == ns["get_prompts_format"]("GER") and A.get_prompts_format("DualHyp") == ns["get_prompts_format"]("DualHyp")
