"""`ger.speed_monitor` mirror: the three names the drivers import (finetune/ger.py:27, finetune/relprompt.py:27-28).
The reference builds the monitor and never feeds it; throughput reporting for this repo lives in
bench.py, so this is the estimate helpers plus an inert monitor."""
from __future__ import annotations

from .utils import num_parameters


def flops_per_param(max_seq_length: int, n_layer: int, n_embd: int, n_params: int) -> int:
    flops_per_token = 2 * n_params
    flops_per_seq = flops_per_token * max_seq_length
    attn_flops_per_seq = n_layer * 2 * 2 * (n_embd * (max_seq_length ** 2))
    return flops_per_seq + attn_flops_per_seq


def estimate_flops(model) -> int:
    """Same estimate as ger/speed_monitor.py:374-395: trainable params count 3x (fwd, bwd, wgrad), frozen 2x."""
    n_tr = num_parameters(model, requires_grad=True)
    n_fr = num_parameters(model, requires_grad=False)
    c = model.config
    tr = flops_per_param(model.max_seq_length, c.n_layer, c.n_embd, n_tr)
    fr = flops_per_param(model.max_seq_length, c.n_layer, c.n_embd, n_fr)
    return (3 if model.training else 1) * tr + (2 if model.training else 1) * fr


def measure_flops(model, x) -> int:
    """The reference counts ATen flops with FlopCounterMode; the B200 path launches its own kernels, which that
    counter cannot see, so this returns the algorithmic count used for the roofline."""
    from ..flops import step_flops
    B, T = x.shape
    return int(step_flops(model.config, model._variant, B, T, training=model.training))


class SpeedMonitorFabric:
    def __init__(self, fabric=None, *args, **kwargs):
        self.fabric = fabric

    def on_train_batch_end(self, *args, **kwargs):
        pass

    def eval_end(self, *args, **kwargs):
        pass
