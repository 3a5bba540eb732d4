"""Host logic of serve.DecodeService (slot admission, per-slot counters / budgets / EOS flags, polling, result assembly) on CPU,
against a stand-in engine whose "language model" is a checkable function of each sequence's own cache row:
    next token = (sum of the sequence's tokens so far + position of the newest one) mod V.
A slot that reads another sequence's cache row, a stale position, or a wrong budget gives a wrong id.  The real engine is
exercised by tests/test_model_gpu.py::test_continuous_batching_service_equals_per_utterance_decode."""
import torch

from dualhyp_b200 import ops
from dualhyp_b200.serve import DecodeService

V, C, ROW = 53, 8, 256


class _Spec:
    n_embd, vocab = C, V


class _Engine:
    def __init__(self):
        self.device, self.spec, self._bufs = torch.device("cpu"), _Spec(), {}
        self.cache = None
        self.prefills = []
        self.cache_k, self.cache_v, self.cache_max_len = [], [], ROW

    def decode_plan_supported(self, n_seq, packed):
        return False                   # the persistent decode kernel needs a GPU: the stand-in takes the per-step path

    def _buf(self, name, shape, dtype):
        return self._bufs.setdefault((name, tuple(shape), dtype), torch.zeros(shape, dtype=dtype))

    def _project_prefixes(self, packed, name):
        return None

    def pack_trainable(self, named):
        return {}

    def alloc_cache_for(self, nseq, need):
        assert need <= ROW
        if self.cache is None or self.cache.shape[0] != nseq:
            self.cache = torch.full((nseq, ROW), -10 ** 6, dtype=torch.int64)        # poison: reading an unwritten slot shows

    def _state(self, slot, pos):
        return int(self.cache[slot, : pos + 1].sum() + pos) % V

    def prefill_packed(self, ids, cu, pos, seq, max_len, packed):
        self.prefills.append(sorted(set(seq.tolist())))
        hidden = torch.zeros((ids.numel(), C), dtype=torch.bfloat16)
        for m in range(ids.numel()):
            self.cache[int(seq[m]), int(pos[m])] = int(ids[m])
            hidden[m, 0] = self._state(int(seq[m]), int(pos[m]))
        return hidden

    def lm_head(self, last, out=None):
        out.zero_()
        out.scatter_(1, last[:, :1].long(), 1.0)
        return out

    def decode_step(self, tokens, positions, packed, pkv_all=None, plan=None):
        assert plan is None
        hidden = self._buf("dec_hidden", (tokens.numel(), C), torch.bfloat16)
        for s in range(tokens.numel()):
            self.cache[s, int(positions[s])] = int(tokens[s])
            hidden[s, 0] = self._state(s, int(positions[s])) if self.cache[s, 0] >= 0 else 0     # empty slot: any finite state
        return hidden


class _Model:
    max_seq_length = ROW

    def __init__(self):
        self._engine = _Engine()

    def _engine_ready(self):
        return self._engine

    def _trainables(self):
        return (), ()


def _alone(prompt, budget, eos=None):
    toks = prompt.tolist()
    for _ in range(budget):
        nxt = (sum(toks) + len(toks) - 1) % V
        toks.append(nxt)
        if eos is not None and nxt == eos:
            break
    return toks


def _prompts(n, seed):
    g = torch.Generator().manual_seed(seed)
    return [torch.randint(1, V, (int(torch.randint(3, 40, (1,), generator=g)),), generator=g) for _ in range(n)]


def test_slots_are_refilled_and_every_utterance_decodes_as_if_alone(monkeypatch):
    monkeypatch.setattr(ops, "argmax", lambda logits, out: out.copy_(logits.float().argmax(-1)))
    prompts = _prompts(11, 3)
    budgets = [4, 9, 1, 12, 6, 3, 8, 12, 2, 5, 7]
    m = _Model()
    svc = DecodeService(m, n_slots=3, max_new_tokens=12, poll_every=2, use_graph=False)
    got = svc.run(prompts, budgets)
    assert [g.tolist() for g in got] == [_alone(p, b) for p, b in zip(prompts, budgets)]
    assert svc.stats["generated"] == sum(budgets)
    assert m._engine.prefills[0] == [0, 1, 2] and len(m._engine.prefills) > 3, "freed slots must be refilled while others decode"
    assert svc.stats["token_steps"] < sum(budgets), "sequences must share token steps"
    # one budget for all, a second run on the same service object and loop
    assert [g.tolist() for g in svc.run(prompts[:4], 5)] == [_alone(p, 5) for p in prompts[:4]]


def test_eos_stops_a_sequence_inclusively_and_frees_its_slot(monkeypatch):
    monkeypatch.setattr(ops, "argmax", lambda logits, out: out.copy_(logits.float().argmax(-1)))
    prompts = _prompts(9, 4)
    free_run = [_alone(p, 10) for p in prompts]
    eos = free_run[2][len(prompts[2]) + 3]                       # the 4th generated token of utterance 2
    want = [_alone(p, 10, eos) for p in prompts]
    assert len(want[2]) == len(prompts[2]) + 4
    svc = DecodeService(_Model(), n_slots=4, max_new_tokens=10, eos_id=eos, poll_every=3, use_graph=False)
    assert [g.tolist() for g in svc.run(prompts)] == want


def test_budget_and_length_checks():
    svc = DecodeService(_Model(), n_slots=2, max_new_tokens=4, use_graph=False)
    p = [torch.arange(1, 6)]
    for bad in (0, 5):
        try:
            svc.run(p, bad)
        except AssertionError:
            continue
        raise AssertionError(f"budget {bad} accepted")
    try:
        svc.run([torch.ones(ROW, dtype=torch.int64)], 4)
    except ValueError:
        return
    raise AssertionError("a prompt that cannot fit max_seq_length was accepted")


def test_generate_batch_host_logic_on_the_stand_in_engine(monkeypatch):
    """generate.generate_batch on the same stand-in: packed prompt staging (cu_seqlens, per-token positions, cache rows), the
    device-side step counter, the inclusive EOS cut and the result assembly."""
    from dualhyp_b200.generate import generate_batch
    monkeypatch.setattr(ops, "argmax", lambda logits, out: out.copy_(logits.float().argmax(-1)))
    prompts = _prompts(5, 8)
    m = _Model()
    got = generate_batch(m, prompts, 7, use_graph=False)
    assert [g.tolist() for g in got] == [_alone(p, 7) for p in prompts]
    assert m._engine.prefills == [[0, 1, 2, 3, 4]]
    eos = got[1][len(prompts[1]) + 2].item()
    cut = generate_batch(m, prompts, 7, eos_id=eos, poll_every=2, use_graph=False)
    for c, p in zip(cut, prompts):
        full = _alone(p, 7)
        gen = full[len(p):]
        assert c.tolist() == (full if eos not in gen else full[: len(p) + gen.index(eos) + 1])
    raw = generate_batch(m, prompts, 3, use_graph=False, return_device=True)
    assert raw.shape == (5, 3) and raw.tolist() == [_alone(p, 3)[len(p):] for p in prompts]
