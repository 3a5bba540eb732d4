"""TEST INFRASTRUCTURE ONLY - golden loss and gradients at the BENCH shapes, computed once on the build box.

    python -m oracle.make_golden_bench_shape        # writes tests/golden/bench_shape_{adapter,lora}.npz

The GPU box cannot afford a TinyLlama-1.1B fp32 forward + backward over 8192 tokens on its host cores inside a test, so the fp32
CPU oracle (oracle/oracle.py, pinned to the unmodified reference by tests/golden/tiny_*.npz) runs here and the results travel:
the train-reduction loss (finetune/ger.py:278-285) and a handful of PEFT gradient tensors of

  * the DualHyp adapter fine-tune step at bench.py's shape, 16 prompts x 512 tokens (BASELINE.json configs[1]), and
  * the LoRA r = 8 step at 8 prompts x 512 tokens (configs[2]'s per-GPU share at 8 GPUs),

on weights and token ids from `dualhyp_b200.synthetic` (a counter hash: the GPU test regenerates the same inputs bit for bit).
Kept per tensor: the full gradient where it is small (gating factors), else a strided sample of rows (the row indices are in the
file), plus every trainable tensor's gradient norm."""
from __future__ import annotations

import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from dualhyp_b200 import synthetic  # noqa: E402
from oracle import oracle as O  # noqa: E402
from tests.helpers import full_cfg  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
SEED = 1337
SHAPES = {"adapter": (16, 512), "lora": (8, 512)}
LORA = dict(r=8, alpha=16, to_query=True, to_key=True, to_value=True, to_projection=True)
KEEP_LAYERS = (2, 11, 21)          # layers whose large gradient tensors are stored in full


def bench_batch(variant: str):
    B, T = SHAPES[variant]
    return synthetic.synth_batch(B, T, 32000, SEED + 77, prompt_range=(330, 480))


def model_config(variant: str):
    from dualhyp_b200.ger import adapter, lora
    if variant == "adapter":
        return adapter.Config.from_name("tiny-llama-1.1b")
    return lora.Config.from_name("tiny-llama-1.1b", **LORA)


def main():
    torch.set_num_threads(os.cpu_count() or 8)
    for variant in ("adapter", "lora"):
        cfg = full_cfg(variant)
        sd = synthetic.synth_state_dict(model_config(variant), variant, SEED)
        idx, labels = bench_batch(variant)
        t0 = time.time()
        loss, grads = O.loss_and_grads(sd, cfg, idx, labels, "train")
        out = {"loss_train": np.float64(float(loss)), "shape": np.array(idx.shape), "ids_checksum": np.int64(int(idx.sum())),
               "labels_checksum": np.int64(int(labels.sum()))}
        for name, g in grads.items():
            out["norm/" + name] = np.float64(float(g.double().norm()))
            layer = int(name.split(".")[2])
            if g.numel() <= 4096 or layer in KEEP_LAYERS:
                out["grad/" + name] = g.float().numpy()
        path = os.path.join(GOLDEN_DIR, f"bench_shape_{variant}.npz")
        np.savez_compressed(path, **out)
        print(f"{variant}: loss {float(loss):.6f}, {len(grads)} gradients, {time.time() - t0:.0f} s -> {path} ({os.path.getsize(path) / 1e6:.2f} MB)")


if __name__ == "__main__":
    main()
