"""bench.py -- the driver's measurement contract for the DualHyp hot path on B200.

    python bench.py --gpus 1 --steps 10 --warmup 3                 # this repo's CUDA path
    python bench.py --impl reference --steps 3 --warmup 1          # the CPU restatement of the reference path
    torchrun ... bench.py --gpus N ...                             # one rank per GPU, weak scaling

Workload at every N (BASELINE.json configs[1]): one DualHyp *adapter fine-tune step* on TinyLlama-1.1B
(random init), 5 ASR + 5 VSR synthetic hypotheses per prompt, T = 512, bf16, 16 prompts per GPU:
forward + train-reduction loss + backward (PEFT gradients) + gradient all-reduce (N > 1) + fused AdamW.
`value` = tokens/s over all ranks with inputs resident in HBM; `e2e` = the same step through the public
`ger` API starting from pinned HOST token ids / labels and ending with the loss read back on the host.
`--workload lora` switches to configs[2] (LoRA r=8; global batch 64 split over the ranks).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "DualHyp train tokens/s (adapter fine-tune step, TinyLlama-1.1B, 5+5 N-best)"
UNIT = "tokens/s"
SEQ_LEN = 512
PER_GPU_BATCH = 16
LORA = dict(r=8, alpha=16, dropout=0.0, to_query=True, to_key=True, to_value=True, to_projection=True)
# DRAM bytes per launch of the dominant GEMMs from one `ncu --set full` capture of this command (profiles/r01g_ncu_full_summary.txt);
# key = (M, N, K, epilogue).  Algorithmic bytes of the SwiGLU GEMM: A 33.6 MB + B 46.1 MB + h12 184.5 MB + act 92.3 MB = 356.5 MB.
NCU_DRAM_BYTES = {(8192, 11264, 2048, 5): 372.5e6, (8192, 2048, 5632, 2): 523.8e6, (8192, 2560, 2048, 4): 58.6e6, (8192, 2048, 2048, 2): 179.8e6}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="adapter", choices=["adapter", "lora"])
    ap.add_argument("--batch", type=int, default=0, help="prompts per GPU (default 16; lora: 64 / gpus)")
    ap.add_argument("--seq", type=int, default=SEQ_LEN)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    return ap.parse_args()


def workload_name(args) -> str:
    if args.workload == "adapter":
        return "DualHyp adapter fine-tune step, TinyLlama-1.1B random-init, 5 ASR + 5 VSR synthetic hyps per prompt, seq 512, bf16"
    return "DualHyp LoRA (r=8) fine-tune step, TinyLlama-1.1B random-init, global batch 64 x seq 512, bf16, data parallel"


def per_gpu_batch(args, world: int) -> int:
    if args.batch > 0:
        return args.batch
    return PER_GPU_BATCH if args.workload == "adapter" else max(1, 64 // world)


# ------------------------------------------------------------------------------------------------
# clocks: sample nvidia-smi DURING the timed region
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.thread = [], None, None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={gpu_index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle (CPU restatement of the reference path) on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_step_tokens_per_s(workload: str, seq: int, steps: int, warmup: int):
    """One bounded sample per step: the same fine-tune step (forward + train loss + PEFT backward) for ONE prompt of
    `seq` tokens, fp32, all host threads.  Returns (tokens/s, ms per step, cores, sample description)."""
    import torch
    from dualhyp_b200 import synthetic
    from oracle import oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    if workload == "adapter":
        cfg = O.OracleConfig(variant="adapter")
    else:
        cfg = O.OracleConfig(variant="lora", r=8, alpha=16, to_query=True, to_key=True, to_value=True, to_projection=True)
    sd = synthetic.synth_state_dict(cfg, cfg.variant, 1337)
    idx, labels = synthetic.synth_batch(1, seq, 32000, 1337, prompt_range=(int(seq * 0.65), int(seq * 0.93)))
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        O.loss_and_grads(sd, cfg, idx, labels, "train")
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    total = sum(times)
    sample = f"oracle (CPU restatement of ger.{workload} path) fp32, 1 prompt x {seq} tokens fwd+loss+bwd per step, {steps} steps"
    return seq * len(times) / total, 1e3 * total / len(times), cores, sample


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    tps, ms, cores, sample = cpu_step_tokens_per_s(args.workload, args.seq, args.steps, args.warmup)
    line = {"impl": "reference", "metric": METRIC, "value": round(tps, 2), "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms, 2), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": {"workload": workload_name(args), "seq_len": args.seq, "prompts_per_step": 1},
            "cpu_baseline": {"value": round(tps, 2), "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": round(tps, 2), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# native arm
# ------------------------------------------------------------------------------------------------
def build_model(workload: str, device):
    import torch
    from dualhyp_b200 import synthetic
    from dualhyp_b200.ger import adapter, lora
    if workload == "adapter":
        cfg = adapter.Config.from_name("tiny-llama-1.1b")
        model = adapter.GPT(cfg)
        model.load_state_dict(synthetic.synth_state_dict(cfg, "adapter", 1337), strict=True)
        adapter.mark_only_adapter_as_trainable(model)
    else:
        cfg = lora.Config.from_name("tiny-llama-1.1b", **LORA)
        model = lora.GPT(cfg)
        model.load_state_dict(synthetic.synth_state_dict(cfg, "lora", 1337), strict=True)
        lora.mark_only_lora_as_trainable(model)
    return model.to(device)


def gemm_epi_name(e: int) -> str:
    return ["bf16", "f32", "resid_f32", "resid_bf16", "qkv_rope", "swiglu", "swiglu_bwd"][e]


def run_native(args):
    import torch
    import torch.distributed as dist
    from dualhyp_b200 import flops, ops, synthetic
    from dualhyp_b200.ger.utils import chunked_cross_entropy
    from dualhyp_b200.parallel import FlatTrainables

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, T = per_gpu_batch(args, world), args.seq
    model = build_model(args.workload, dev)
    bucket = FlatTrainables(p for p in model.parameters() if p.requires_grad)
    idx_h, labels_h = synthetic.synth_batch(B, T, 32000, 1337 + rank)
    idx_h, labels_h = idx_h.pin_memory(), labels_h.pin_memory()
    idx_d, labels_d = idx_h.to(dev), labels_h.to(dev)

    def step(idx, labels):
        """The driver's micro-step (finetune/ger.py:278-292) with accumulation 1."""
        logits = model(idx, lm_head_chunk_size=128)
        logits[-1] = logits[-1][..., :-1, :]
        loss = chunked_cross_entropy(logits, labels[..., 1:])
        loss.backward()
        bucket.allreduce()
        bucket.adamw_step(lr=1e-4, weight_decay=0.01, grad_scale=1.0 / world)
        bucket.zero_grad()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    for _ in range(max(args.warmup, 3)):
        step(idx_d, labels_d)
    sampler = ClockSampler(local) if rank == 0 else None
    l0 = ops.launches()
    ms_total = timed(lambda: step(idx_d, labels_d), args.steps)
    n_launch = ops.launches() - l0
    clocks = sampler.stop() if sampler else None
    tokens = world * B * T * args.steps
    value = tokens / (ms_total / 1e3)

    # ---- end to end: pinned host inputs -> H2D -> step -> loss read back on the host, every step ----
    losses = []

    def e2e_step():
        i_d = idx_h.to(dev, non_blocking=True)
        l_d = labels_h.to(dev, non_blocking=True)
        losses.append(step(i_d, l_d).item())
    for _ in range(2):
        e2e_step()
    ms_e2e = timed(e2e_step, args.steps)
    e2e_value = tokens / (ms_e2e / 1e3)

    line = {"metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": round(ms_total / args.steps, 3), "higher_is_better": True, "scaling": "weak" if args.workload == "adapter" else "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(args), "model": "TinyLlama-1.1B (random init)", "prompts_per_gpu": B, "seq_len": T,
                       "global_batch": world * B, "parallelism": f"dp{world}", "optimizer": "fused AdamW on the flat PEFT bucket",
                       "l2": "per-step working set (~9 GB of activations + 2.2 GB weights) >> 126 MB L2; no explicit flush",
                       "loss": round(losses[-1], 4)},
            "e2e": {"value": round(e2e_value, 1), "unit": UNIT, "h2d_bytes_per_step": 2 * B * T * 8, "d2h_bytes_per_step": 4,
                    "ms_per_step": round(ms_e2e / args.steps, 3)},
            "gpu_launches": n_launch, "clocks": clocks}
    # the vocabulary projection only runs on rows with a target (valid-row compaction): credit only those flops
    rho = float((labels_h[..., 1:] != -1).sum()) / float(B * T)
    line["config"]["lm_head_row_fraction"] = round(rho, 4)
    step_fl = flops.step_flops(model.config, args.workload, B, T, training=True, rho=rho)
    line["step_tflops_per_gpu"] = round(step_fl / (ms_total / args.steps / 1e3) / 1e12, 1)

    # ---- roofline of the dominant kernel: per-launch CUDA events in a dedicated instrumented pass ----
    if rank == 0 and not args.no_roofline:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:  # noqa: BLE001
            pass
        peak_tf = float(peaks.get("bf16_tflops_sustained", 1400.0))
        peak_src = "measured sustained (MEASURED_PEAKS.json)" if peaks else "fallback 1.4 PF/s sustained"
        with ops.profile() as recs:
            for _ in range(2):
                step(idx_d, labels_d)
            torch.cuda.synchronize()
        agg, total_ms = {}, 0.0
        for name, meta, e0, e1 in recs:
            ms = e0.elapsed_time(e1)
            total_ms += ms
            key = (name, meta)
            a = agg.setdefault(key, [0, 0.0])
            a[0] += 1
            a[1] += ms
        gemm_ms = sum(v[1] for k, v in agg.items() if k[0] == "gemm")
        gemm_fl = sum(v[0] * 2.0 * k[1][0] * k[1][1] * (k[1][2] + k[1][3]) for k, v in agg.items() if k[0] == "gemm")
        (dname, dmeta), (dcount, dms) = max(((k, v) for k, v in agg.items() if k[0] == "gemm"), key=lambda kv: kv[1][1])
        M_, N_, K_, K2_, epi, al, bl = dmeta
        fl = 2.0 * M_ * N_ * (K_ + K2_)
        achieved = fl / (dms / dcount / 1e3) / 1e12
        line["roofline"] = {"bound": "tensor", "achieved": round(achieved, 1), "peak": peak_tf, "unit": "TFLOP/s",
                            "frac": round(achieved / peak_tf, 4), "traffic": NCU_DRAM_BYTES.get((M_, N_, K_ + K2_, epi)), "peak_source": peak_src,
                            "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full (profiles/r01g_ncu_full_summary.txt)",
                            "kernel": f"gemm_kernel<256,{gemm_epi_name(epi)}> M={M_} N={N_} K={K_}", "launches": dcount,
                            "avg_ms": round(dms / dcount, 4), "share_of_step": round(dms / total_ms, 4),
                            "all_gemms": {"tflops": round(gemm_fl / (gemm_ms / 1e3) / 1e12, 1), "share_of_step": round(gemm_ms / total_ms, 4)},
                            "by_op_ms_per_step": {n: round(sum(v[1] for k, v in agg.items() if k[0] == n) / 2, 3)
                                                  for n in sorted({k[0] for k in agg})},
                            "gemm_by_shape": [{"M": k[1][0], "N": k[1][1], "K": k[1][2] + k[1][3], "epi": gemm_epi_name(k[1][4]),
                                               "n": v[0] // 2, "avg_ms": round(v[1] / v[0], 4),
                                               "tflops": round(2.0 * k[1][0] * k[1][1] * (k[1][2] + k[1][3]) / (v[1] / v[0] / 1e3) / 1e12, 1)}
                                              for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]) if k[0] == "gemm"][:12]}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        tps, ms, cores, sample = cpu_step_tokens_per_s(args.workload, T, 2, 1)
        line["cpu_baseline"] = {"value": round(tps, 2), "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
