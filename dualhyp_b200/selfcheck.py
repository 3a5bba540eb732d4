"""smoke(): one small invocation of the hot path on cuda:0, checked against the oracle (the oracle is imported
here as the checker only, as the task's rules allow for __graft_entry__.smoke())."""
from __future__ import annotations

import torch


def smoke_check() -> None:
    from . import synthetic
    from .ger import adapter
    from .ger.utils import chunked_cross_entropy
    from oracle import oracle as O

    dev = "cuda:0"
    cfg = adapter.Config.from_name("tiny-llama-test", adapter_start_layer=1)
    model = adapter.GPT(cfg)
    sd = synthetic.synth_state_dict(cfg, "adapter", 1337)
    model.load_state_dict(sd, strict=True)
    adapter.mark_only_adapter_as_trainable(model)
    model = model.to(dev)
    idx, labels = synthetic.synth_batch(2, 160, 512, 1337, prompt_range=(90, 130))
    logits = model(idx.to(dev), lm_head_chunk_size=128)
    logits[-1] = logits[-1][..., :-1, :]
    loss = chunked_cross_entropy(logits, labels.to(dev)[..., 1:])
    loss.backward()
    torch.cuda.synchronize()
    ocfg = O.OracleConfig(n_layer=3, n_head=4, n_query_groups=2, head_size=64, n_embd=256, intermediate_size=512,
                          padded_vocab_size=512, block_size=256, variant="adapter", adapter_prompt_length=10, adapter_start_layer=1)
    ref_loss, ref_grads = O.loss_and_grads(sd, ocfg, idx, labels, "train")
    assert abs(float(loss) - float(ref_loss)) < 5e-3, (float(loss), float(ref_loss))
    for n, p in model.named_parameters():
        if p.requires_grad:
            a, b = p.grad.double().flatten().cpu(), ref_grads[n].double().flatten()
            cos = float(a @ b / (a.norm() * b.norm()))
            assert cos > 0.995, (n, cos)
    print(f"smoke ok: loss {float(loss):.5f} (oracle {float(ref_loss):.5f})")
