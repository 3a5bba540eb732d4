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
`--workload decode` measures the second half of BASELINE.json's metric (configs[4]): greedy KV-cache generation,
64 synthetic DualHyp prompts per GPU (the 512-utterance test set sharded 8 ways), 64 new tokens each, in utt/s;
the default run appends a short sample of it as `"infer"` to the training line.
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
DECODE_METRIC = "DualHyp infer utterances/s (greedy KV-cache decode, adapter TinyLlama-1.1B, 5+5 N-best prompts, 64 new tokens)"
DECODE_UTTS, DECODE_NEW = 256, 64     # 256 sequences per token step: two 128-row activation boxes stream each weight tile once
# LoRA as the shipped scripts train it (finetune/ger.py:397-407): r = 8, alpha = 16, dropout 0.05 on the LoRA branch, q/k/v/proj
LORA = dict(r=8, alpha=16, dropout=0.05, to_query=True, to_key=True, to_value=True, to_projection=True)



def measured_traffic():
    """profiles/traffic.json (scripts/capture_traffic.py: ncu --set full of a build) -- valid only for the kernel sources it was
    captured from (`_lib.source_id()`: comment- and space-insensitive hash of csrc/ + the header; the library file is not hashed,
    nvcc does not reproduce it byte for byte); returns ({} , why) otherwise."""
    try:
        from dualhyp_b200 import _lib as _l
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        sha = _l.source_id()
    except Exception as e:  # noqa: BLE001
        return {}, f"no capture ({type(e).__name__})"
    if t.get("src_sha16") != sha:
        return {}, f"profiles/traffic.json was captured from sources {t.get('src_sha16')}, this run uses {sha}: not reported"
    return t, f"{t.get('source')}, kernel sources {sha}"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="adapter", choices=["adapter", "lora", "relprompt", "decode"])
    ap.add_argument("--batch", type=int, default=0, help="prompts per GPU (default 16; lora: 64 / gpus)")
    ap.add_argument("--seq", type=int, default=SEQ_LEN)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--no-infer", action="store_true", help="skip the decode sample appended to the training line")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs only (scripts/profile_step.sh): skip the end-to-end leg, `e2e` is then null")
    ap.add_argument("--padded", action="store_true",
                    help="train on right-padded [B, T] batches like the reference's collate_fn (tokens/s then counts pad tokens); "
                         "default: the same samples packed into one row without pad tokens, tokens/s counts REAL tokens")
    ap.add_argument("--lora-dropout", type=float, default=LORA["dropout"], help="LoRA branch dropout of the lora / relprompt workloads (reference default 0.05)")
    ap.add_argument("--no-lora-sample", action="store_true", help="skip the LoRA (configs[2]) sample appended to the default training line")
    ap.add_argument("--utts", type=int, default=DECODE_UTTS, help="decode workload: utterances per GPU per step")
    ap.add_argument("--new-tokens", type=int, default=DECODE_NEW)
    return ap.parse_args()


def metric_name(args) -> str:
    kind = {"adapter": "adapter fine-tune step", "lora": "LoRA r=8 fine-tune step", "relprompt": "RelPrompt fine-tune step"}.get(args.workload)
    return METRIC if args.workload == "adapter" else f"DualHyp train tokens/s ({kind}, TinyLlama-1.1B, 5+5 N-best)"


def workload_name(args) -> str:
    if args.workload == "decode":
        return ("DualHyp greedy generation with KV cache, TinyLlama-1.1B adapter random-init, synthetic prompts U[330,480] tokens, "
                f"{args.new_tokens} new tokens, bf16, utterances sharded across GPUs")
    if args.workload == "relprompt":
        return ("RelPrompt fine-tune step: reliability predictors on synthetic audio [B,500,1280] / visual [B,250,1024] feature frames + "
                "LoRA (r=16) TinyLlama-1.1B forward+backward over prompts with mask tokens, seq 512, bf16")
    if args.workload == "adapter":
        return "DualHyp adapter fine-tune step, TinyLlama-1.1B random-init, 5 ASR + 5 VSR synthetic hyps per prompt, seq 512, bf16"
    return "DualHyp LoRA (r=8) fine-tune step, TinyLlama-1.1B random-init, global batch 64 x seq 512, bf16, data parallel"


def per_gpu_batch(args, world: int) -> int:
    if args.batch > 0:
        return args.batch
    if args.workload == "relprompt":
        return 4                                   # B = 4 per GPU
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
    rel = None
    if workload == "adapter":
        cfg = O.OracleConfig(variant="adapter")
    elif workload == "relprompt":
        cfg = O.OracleConfig(variant="relprompt", r=16, alpha=32, to_query=True, to_key=True, to_value=True, to_projection=True)
    else:
        cfg = O.OracleConfig(variant="lora", r=8, alpha=16, to_query=True, to_key=True, to_value=True, to_projection=True)
    sd = synthetic.synth_state_dict(cfg, cfg.variant, 1337, extra_wte_rows=3 if workload == "relprompt" else 0)
    idx, labels = synthetic.synth_batch(1, seq, 32000, 1337, prompt_range=(int(seq * 0.65), int(seq * 0.93)))
    if workload == "relprompt":
        rel = (synthetic.synth_features(1, 500, 1280, 1337, 301), synthetic.synth_features(1, 250, 1024, 1337, 302),
               synthetic.hash_ints((1, 25), 1337, 303, 0, 3), synthetic.hash_ints((1, 25), 1337, 304, 0, 3))
        keys = O.trainable_keys(sd, cfg)

    def one_step():
        if rel is None:
            O.loss_and_grads(sd, cfg, idx, labels, "train")
            return
        work = {k: (v.detach().clone().requires_grad_(True) if k in keys else v.detach()) for k, v in sd.items()}
        loss = O.relprompt_loss(work, cfg, idx, labels, *rel)[0]
        torch.autograd.grad(loss, [work[k] for k in keys])

    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        one_step()
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    total = sum(times)
    sample = f"oracle (CPU restatement of ger.{workload} path) fp32, 1 prompt x {seq} tokens fwd+loss+bwd per step, {steps} steps"
    return seq * len(times) / total, 1e3 * total / len(times), cores, sample


def cpu_cfg1_forward_loss():
    """BASELINE.json configs[0] exactly: GER (ASR-only, 5-best) adapter forward + loss, TinyLlama-1.1B random-init,
    batch 4 x seq 256, fp32 on the host cores, best of 3.  Returns a dict for `cpu_baseline["cfg1"]`."""
    import torch
    from dualhyp_b200 import synthetic
    from oracle import oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = O.OracleConfig(variant="adapter")
    sd = synthetic.synth_state_dict(cfg, "adapter", 1337)
    idx, labels = synthetic.synth_batch(4, 256, 32000, 1337, prompt_range=(150, 230), response_range=(8, 24))
    best = None
    with torch.no_grad():
        for _ in range(3):
            t0 = time.perf_counter()
            loss = O.train_loss(sd, cfg, idx, labels)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
    return {"value": round(4 * 256 / best, 1), "unit": "tokens/s", "ms": round(1e3 * best, 1), "loss": round(float(loss), 5),
            "sample": "oracle fp32, GER adapter forward + train-reduction loss, batch 4 x seq 256, best of 3"}


def cpu_decode_utt_per_s(new_tokens: int):
    """Bounded sample of the reference decode loop (generate/base.py:19-82 as restated by the oracle) for ONE utterance
    of 405 prompt tokens: a run with 1 new token (prefill + first token) and a run with 5 give the per-token cost;
    utt time = prefill run + (new_tokens - 1) x per-token cost.  Returns (utt/s, ms per utterance, cores, sample)."""
    import torch
    from dualhyp_b200 import synthetic
    from oracle import oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = O.OracleConfig(variant="adapter")
    sd = synthetic.synth_state_dict(cfg, "adapter", 1337)
    prompt = synthetic.synth_prompts(1, 32000, 1337, prompt_range=(405, 405))[0]
    O.generate_greedy(sd, cfg, prompt[:64], 65)                     # warm the thread pool / allocator
    t0 = time.perf_counter()
    O.generate_greedy(sd, cfg, prompt, prompt.numel() + 1)
    t1 = time.perf_counter()
    O.generate_greedy(sd, cfg, prompt, prompt.numel() + 5)
    t2 = time.perf_counter()
    per_tok = max(((t2 - t1) - (t1 - t0)) / 4.0, 0.0)
    utt = (t1 - t0) + (new_tokens - 1) * per_tok
    sample = (f"oracle (CPU restatement of generate/base.py + ger.adapter) fp32, 1 utterance, 405-token prompt: prefill+1 token "
              f"{t1 - t0:.2f} s, {per_tok * 1e3:.0f} ms per further token (4 timed), extrapolated to {new_tokens} new tokens")
    return 1.0 / utt, 1e3 * utt, cores, sample


def decode_measure(model, dev, world: int, rank: int, utts: int, new_tokens: int, steps: int, warmup: int, timed_fn):
    """utt/s of generate_batch over `utts` prompts per GPU: HBM-resident prompts (`value`) and host prompts -> host ids
    (`e2e`); per-token graph replays timed separately for the HBM roofline of the decode step."""
    import torch
    from dualhyp_b200 import ops, synthetic
    from dualhyp_b200.generate import generate_batch
    prompts_h = synthetic.synth_prompts(utts, 32000, 4242 + rank)
    prompts_d = [p.to(dev) for p in prompts_h]
    for _ in range(max(warmup, 3)):
        generate_batch(model, prompts_d, new_tokens, return_device=True)
    l0 = ops.launches()
    ms_dev = timed_fn(lambda: generate_batch(model, prompts_d, new_tokens, return_device=True), steps)
    launches_eager_equiv = ops.launches() - l0
    ms_e2e = timed_fn(lambda: generate_batch(model, prompts_h, new_tokens), steps)
    # the token loop alone (graph replays): the HBM-bound part
    eng = model._engine
    loop = eng._bufs["decode_loops"][utts]
    lens = torch.tensor([p.numel() for p in prompts_h], dtype=torch.int32, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.inference_mode():
        loop.pos.copy_(lens)
        loop.step.zero_()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(new_tokens):
            loop.graph.replay()
        e1.record()
        torch.cuda.synchronize()
    tok_ms = e0.elapsed_time(e1) / new_tokens
    s = eng.spec
    w_bytes = 2.0 * (s.n_layer * (s.qkv_out * s.n_embd + s.n_embd * s.n_embd + 3 * s.n_embd * s.intermediate_size) + s.vocab * s.n_embd)
    kv_bytes = 2.0 * s.n_layer * 2 * s.n_query_groups * 64 * float(lens.sum() + utts * new_tokens / 2)
    return dict(ms_dev=ms_dev, ms_e2e=ms_e2e, tok_ms=tok_ms, step_bytes=w_bytes + kv_bytes, launches=launches_eager_equiv,
                h2d=int(sum(p.numel() for p in prompts_h) * 8), d2h=utts * new_tokens * 8,
                prompt_tokens=int(lens.sum()), kernels_per_token=loop.kernels_per_token)


def run_decode(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    model = build_model("adapter", dev)

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

    sampler = ClockSampler(local) if rank == 0 else None
    r = decode_measure(model, dev, world, rank, args.utts, args.new_tokens, args.steps, args.warmup, timed)
    clocks = sampler.stop() if sampler else None
    utts = world * args.utts * args.steps
    line = decode_line(args, r, world, utts, clocks)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ups, ms, cores, sample = cpu_decode_utt_per_s(args.new_tokens)
        line["cpu_baseline"] = {"value": round(ups, 5), "unit": "utt/s", "cores": cores, "kind": "port", "sample": sample}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def decode_line(args, r, world, utts, clocks, steps=None):
    steps = steps or args.steps
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:  # noqa: BLE001
        pass
    peak = float(peaks.get("hbm_gbs", 6500.0))
    ach = r["step_bytes"] / (r["tok_ms"] / 1e3) / 1e9
    return {"metric": DECODE_METRIC, "value": round(utts / (r["ms_dev"] / 1e3), 2), "unit": "utt/s", "n_gpus": world, "steps": steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(r["ms_dev"] / steps, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(args), "model": "TinyLlama-1.1B (random init)", "utterances_per_gpu": args.utts,
                       "new_tokens": args.new_tokens, "prompt_tokens_per_gpu": r["prompt_tokens"], "parallelism": f"dp{world} (utterance sharding, no collective)",
                       "l2": "each token step streams 2.2 GB of weights >> 126 MB L2; no explicit flush", "eos": "disabled for timing (fixed token count)"},
            "e2e": {"value": round(utts / (r["ms_e2e"] / 1e3), 2), "unit": "utt/s", "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
                    "ms_per_step": round(r["ms_e2e"] / steps, 3)},
            "gpu_launches": r["launches"], "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4),
                         "traffic": (measured_traffic()[0].get("decode", {}).get(str(args.utts), {}) or {}).get("dram_bytes_per_token_step"),
                         "traffic_source": measured_traffic()[1],
                         "peak_source": "measured copy bandwidth (MEASURED_PEAKS.json)" if peaks else "fallback 6.5 TB/s",
                         "kernel": "one decode token step (CUDA graph of %d kernels: lm_head, arg-max, 22 layers over the KV cache)" % r["kernels_per_token"],
                         "algorithmic_bytes_per_launch": round(r["step_bytes"]), "avg_ms": round(r["tok_ms"], 4),
                         "share_of_step": round(r["tok_ms"] * args.new_tokens * steps / r["ms_dev"], 4),
                         "prefill_ms_per_step": round(r["ms_dev"] / steps - r["tok_ms"] * args.new_tokens, 3)}}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.workload == "decode":
        ups, ms, cores, sample = cpu_decode_utt_per_s(args.new_tokens)
        line = {"impl": "reference", "metric": DECODE_METRIC, "value": round(ups, 5), "unit": "utt/s", "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": round(ms, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": {"workload": workload_name(args), "utterances_per_step": 1},
                "cpu_baseline": {"value": round(ups, 5), "unit": "utt/s", "cores": cores, "kind": "port", "sample": sample},
                "e2e": {"value": round(ups, 5), "unit": "utt/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
        print(json.dumps(line), flush=True)
        return
    tps, ms, cores, sample = cpu_step_tokens_per_s(args.workload, args.seq, args.steps, args.warmup)
    line = {"impl": "reference", "metric": metric_name(args), "value": round(tps, 2), "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms, 2), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": {"workload": workload_name(args), "seq_len": args.seq, "prompts_per_step": 1},
            "cpu_baseline": {"value": round(tps, 2), "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": round(tps, 2), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# native arm
# ------------------------------------------------------------------------------------------------
def build_model(workload: str, device, lora_dropout: float = LORA["dropout"]):
    import torch
    from dualhyp_b200 import synthetic
    from dualhyp_b200.ger import adapter, lora
    if workload == "adapter":
        cfg = adapter.Config.from_name("tiny-llama-1.1b")
        model = adapter.GPT(cfg)
        model.load_state_dict(synthetic.synth_state_dict(cfg, "adapter", 1337), strict=True)
        adapter.mark_only_adapter_as_trainable(model)
    elif workload == "relprompt":
        from dualhyp_b200.ger import relprompt
        cfg = relprompt.Config.from_name("tiny-llama-1.1b", r=16, alpha=32, dropout=lora_dropout, to_query=True, to_key=True, to_value=True,
                                         to_projection=True)
        model = relprompt.GPT(cfg)
        model.resize_token_embeddings(3)           # <<C>> <<M>> <<N>> (finetune/relprompt.py:168)
        model.load_state_dict(synthetic.synth_state_dict(cfg, "relprompt", 1337, extra_wte_rows=3), strict=True)
        relprompt.mark_only_lora_as_trainable(model)
        model.train()                              # LoRA dropout and the classifiers' Dropout(0.1) are live, as in finetune/relprompt.py
    else:
        cfg = lora.Config.from_name("tiny-llama-1.1b", **{**LORA, "dropout": lora_dropout})
        model = lora.GPT(cfg)
        model.load_state_dict(synthetic.synth_state_dict(cfg, "lora", 1337), strict=True)
        lora.mark_only_lora_as_trainable(model)
    return model.to(device)


def lora_sample(args, dev, world: int, rank: int, timed, steps: int = 5, warmup: int = 3):
    """BASELINE.json configs[2] next to the headline line: the LoRA r = 8 fine-tune step, GLOBAL batch 64 x seq 512 split over the
    ranks (strong scaling: 64 / N prompts per GPU), packed rows, LoRA dropout 0.05, one flat-bucket all-reduce + fused AdamW."""
    import torch
    import torch.distributed as dist
    from dualhyp_b200 import flops, synthetic
    from dualhyp_b200.collate import pack_varlen
    from dualhyp_b200.ger.utils import chunked_cross_entropy
    from dualhyp_b200.parallel import FlatTrainables
    B = max(1, 64 // world)
    model = build_model("lora", dev, args.lora_dropout)
    model.train()
    bucket = FlatTrainables(p for p in model.parameters() if p.requires_grad)
    idx_h, labels_h = synthetic.synth_batch(B, SEQ_LEN, 32000, 4337 + rank)
    samples = []
    for b in range(B):
        n = int((labels_h[b] != -1).nonzero().max()) + 1
        samples.append({"input_ids": idx_h[b, :n].clone(), "labels": labels_h[b, :n].clone()})
    longest = max(int(smp["input_ids"].numel()) for smp in samples)
    pk = {k: v.to(dev) for k, v in pack_varlen(samples).items()}

    def step():
        logits = model.forward_packed(pk["input_ids"], pk["cu_seqlens"], pk["positions"], longest, lm_head_chunk_size=128)
        loss = chunked_cross_entropy(logits, pk["targets"])
        loss.backward()
        bucket.allreduce()
        bucket.adamw_step(lr=1e-4, weight_decay=0.01, grad_scale=1.0 / world)
        bucket.zero_grad()
        return loss

    for _ in range(warmup):
        step()
    ms = timed(step, steps)
    tok = torch.tensor([float(pk["input_ids"].numel())], device=dev)
    if world > 1:
        dist.all_reduce(tok)
    lens = (pk["cu_seqlens"][1:] - pk["cu_seqlens"][:-1]).tolist()
    rho = float((pk["targets"] != -1).sum()) / float(pk["input_ids"].numel())
    fl = sum(flops.step_flops(model.config, "lora", 1, int(n), training=True, rho=rho) for n in lens)
    out = {"metric": "DualHyp train tokens/s (LoRA r=8 fine-tune step, TinyLlama-1.1B, 5+5 N-best)", "value": round(float(tok) * steps / (ms / 1e3), 1),
           "unit": UNIT, "ms_per_step": round(ms / steps, 3), "steps": steps, "scaling": "strong", "global_batch": B * world, "prompts_per_gpu": B,
           "tokens_per_step": round(float(tok)), "lora_dropout": args.lora_dropout, "step_tflops_per_gpu": round(fl / (ms / steps / 1e3) / 1e12, 1)}
    del model, bucket
    torch.cuda.empty_cache()
    return out


def gemm_epi_name(e: int) -> str:
    return ["bf16", "f32", "resid_f32", "resid_bf16", "qkv_rope", "swiglu", "swiglu_bwd"][e]


def run_native(args):
    if args.workload == "decode":
        return run_decode(args)
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
    model = build_model(args.workload, dev, args.lora_dropout)
    model.train()
    bucket = FlatTrainables(p for p in model.parameters() if p.requires_grad)
    idx_h, labels_h = synthetic.synth_batch(B, T, 32000, 1337 + rank)
    packed_mode = not args.padded and args.workload != "relprompt"
    pk_h = pk_d = None
    longest = T
    if packed_mode:
        # the same samples without their right pad (data/av_dataset.py:258-292 pads to the longest sample of the batch):
        # one packed row + cu_seqlens (dualhyp_b200.collate.pack_varlen), consumed by GPT.forward_packed
        from dualhyp_b200.collate import pack_varlen
        samples = []
        for b in range(B):
            n = int((labels_h[b] != -1).nonzero().max()) + 1
            samples.append({"input_ids": idx_h[b, :n].clone(), "labels": labels_h[b, :n].clone()})
        longest = max(int(smp["input_ids"].numel()) for smp in samples)
        pk = pack_varlen(samples)
        pk_h = {k: pk[k].pin_memory() for k in ("input_ids", "targets", "cu_seqlens", "positions")}
        pk_d = {k: v.to(dev) for k, v in pk_h.items()}
    idx_h, labels_h = idx_h.pin_memory(), labels_h.pin_memory()
    idx_d, labels_d = idx_h.to(dev), labels_h.to(dev)
    real_tokens = int(pk_h["input_ids"].numel()) if packed_mode else B * T

    rel = None
    if args.workload == "relprompt":
        import torch.nn.functional as F
        rel = dict(af=synthetic.synth_features(B, 500, 1280, 1337 + rank, 301, dev), vf=synthetic.synth_features(B, 250, 1024, 1337 + rank, 302, dev),
                   at=synthetic.hash_ints((B, 25), 1337 + rank, 303, 0, 3).to(dev), vt=synthetic.hash_ints((B, 25), 1337 + rank, 304, 0, 3).to(dev))
        # mask tokens inside the prompt span (ids 32000..32002), as the RelPrompt template places them
        idx_h[:, 40:65] = 32000 + synthetic.hash_ints((B, 25), 1337 + rank, 305, 0, 3)
        idx_h[:, 200:225] = 32000 + synthetic.hash_ints((B, 25), 1337 + rank, 306, 0, 3)
        idx_d = idx_h.to(dev)

    def mask_loss():
        """finetune/relprompt.py:356-387: CE(audio) + CE(visual) over the chunk labels, logits trimmed to the label length."""
        al, vl = model.audio_noise_classifier(rel["af"]), model.visual_noise_classifier(rel["vf"])
        na, nv = min(al.size(1), rel["at"].size(1)), min(vl.size(1), rel["vt"].size(1))
        return F.cross_entropy(al[:, :na].reshape(-1, 3), rel["at"][:, :na].reshape(-1)) + \
            F.cross_entropy(vl[:, :nv].reshape(-1, 3), rel["vt"][:, :nv].reshape(-1))

    ar_events = None

    def step(idx, labels, sync=True):
        """The driver's micro-step (finetune/ger.py:278-292) with accumulation 1.  sync=False skips the gradient
        all-reduce: the instrumented roofline pass runs on rank 0 alone and must not enter a collective.
        Packed mode: `idx` is the dict of packed tensors (input_ids, targets, cu_seqlens, positions)."""
        if packed_mode:
            logits = model.forward_packed(idx["input_ids"], idx["cu_seqlens"], idx["positions"], longest, lm_head_chunk_size=128)
            loss = chunked_cross_entropy(logits, idx["targets"])
        else:
            logits = model(idx, lm_head_chunk_size=128)
            logits[-1] = logits[-1][..., :-1, :]
            loss = chunked_cross_entropy(logits, labels[..., 1:])
        if rel is not None:
            loss = loss + 0.02 * mask_loss()
        loss.backward()
        if sync:
            if world > 1 and ar_events is not None:      # exposed time of the gradient exchange (collective + the wait for the slowest rank)
                ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                ea.record()
                bucket.allreduce()
                eb.record()
                ar_events.append((ea, eb))
            else:
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

    in_d = pk_d if packed_mode else idx_d
    for _ in range(max(args.warmup, 3)):
        step(in_d, labels_d)
    sampler = ClockSampler(local) if rank == 0 else None
    l0 = ops.launches()
    ar_events = [] if world > 1 else None
    ms_total = timed(lambda: step(in_d, labels_d), args.steps)
    n_launch = ops.launches() - l0
    ar_ms = None
    if ar_events:
        ar_t = torch.tensor([sum(a.elapsed_time(b) for a, b in ar_events) / len(ar_events)], device=dev)
        dist.all_reduce(ar_t, op=dist.ReduceOp.MAX)
        ar_ms = float(ar_t)
    ar_events = None
    clocks = sampler.stop() if sampler else None
    tok_t = torch.tensor([float(real_tokens)], device=dev)
    if world > 1:
        dist.all_reduce(tok_t)                      # every rank packs its own samples: sum the real token counts
    tokens_per_step = float(tok_t)
    tokens = tokens_per_step * args.steps
    value = tokens / (ms_total / 1e3)

    # ---- end to end: pinned host inputs -> H2D -> step -> loss read back on the host, every step ----
    losses = []

    def e2e_step():
        if packed_mode:
            losses.append(step({k: v.to(dev, non_blocking=True) for k, v in pk_h.items()}, None).item())
            return
        i_d = idx_h.to(dev, non_blocking=True)
        l_d = labels_h.to(dev, non_blocking=True)
        losses.append(step(i_d, l_d).item())
    if args.no_e2e:
        losses.append(float(step(in_d, labels_d)))
        ms_e2e = float("nan")
    else:
        for _ in range(2):
            e2e_step()
        ms_e2e = timed(e2e_step, args.steps)
    e2e_value = tokens / (ms_e2e / 1e3)

    line = {"metric": metric_name(args), "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": round(ms_total / args.steps, 3), "higher_is_better": True, "scaling": "strong" if args.workload == "lora" else "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(args), "model": "TinyLlama-1.1B (random init)", "prompts_per_gpu": B, "seq_len": T,
                       "global_batch": world * B, "parallelism": f"dp{world}", "optimizer": "fused AdamW on the flat PEFT bucket",
                       "l2": "per-step working set (~9 GB of activations + 2.2 GB weights) >> 126 MB L2; no explicit flush",
                       "loss": round(losses[-1], 4),
                       "tokens_counted": ("REAL tokens only: the samples are packed into one row without the right pad of the reference's "
                                          "collate_fn (cu_seqlens attention forward + backward)" if packed_mode else
                                          "all B x T positions of the right-padded batch (pad tokens included, as the reference computes them)"),
                       "tokens_per_step": round(tokens_per_step), "padded_positions_per_step": world * B * longest,
                       "padded_equivalent_tokens_per_s": round(world * B * longest * args.steps / (ms_total / 1e3), 1)},
            "e2e": None if args.no_e2e else {"value": round(e2e_value, 1), "unit": UNIT,
                    "h2d_bytes_per_step": (sum(v.numel() * v.element_size() for v in pk_h.values()) if packed_mode else 2 * B * T * 8),
                    "d2h_bytes_per_step": 4,
                    "ms_per_step": round(ms_e2e / args.steps, 3)},
            "gpu_launches": n_launch, "clocks": clocks}
    if ar_ms is not None:
        line["grad_allreduce"] = {"exposed_ms_per_step": round(ar_ms, 3), "bytes": int(bucket.numel * 4),
                                  "note": "CUDA events around the one flat-bucket ncclAllReduce on the compute stream, max over ranks: the collective plus "
                                          "the wait for the slowest rank's backward; every gradient materialises at the end of the backward (batched "
                                          "prefix / LoRA reductions), so there is nothing left to overlap it with"}
    # the vocabulary projection only runs on rows with a target (valid-row compaction): credit only those flops
    rho = float((labels_h[..., 1:] != -1).sum()) / float(real_tokens)
    line["config"]["lm_head_row_fraction"] = round(rho, 4)
    if packed_mode:        # per-token flops with each sequence's own causal length, summed over the real tokens of this rank
        lens = (pk_h["cu_seqlens"][1:] - pk_h["cu_seqlens"][:-1]).tolist()
        step_fl = sum(flops.step_flops(model.config, args.workload, 1, int(n), training=True, rho=rho) for n in lens)
    else:
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
                step(in_d, labels_d, sync=False)
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
        # the launch behind this shape: CTA-pair kernel for K-major operands, N % 256 == 0, M >= 1024 (csrc/gemm.cu dispatch_bn)
        kname = f"gemm_pair_kernel<{epi}>" if (al == 0 and bl == 0 and N_ % 256 == 0 and M_ >= 1024) else f"gemm_kernel<*,{epi}>"
        tr, traffic_src = measured_traffic()
        cfg_ok = tr.get("train_config") == {"workload": args.workload, "prompts_per_gpu": B, "packed": bool(packed_mode)}
        ent = tr.get("train", {}).get(kname) if cfg_ok else None
        traffic = ent["dram_bytes_per_launch"] if ent else None
        if tr and not cfg_ok:
            traffic_src = f"no capture for this workload (profiles/traffic.json holds {tr.get('train_config')})"
        # A + B read once; SwiGLU writes h12 (M x N) and act (M x N/2); the others write M x N (bf16) or read + write fp32 M x N
        alg_bytes = 2 * M_ * (K_ + K2_) + 2 * N_ * (K_ + K2_) + (2 * M_ * N_ + M_ * N_ if gemm_epi_name(epi) == "swiglu" else 2 * M_ * N_)
        line["roofline"] = {"bound": "tensor", "achieved": round(achieved, 1), "peak": peak_tf, "unit": "TFLOP/s",
                            "frac": round(achieved / peak_tf, 4), "traffic": traffic, "peak_source": peak_src,
                            "traffic_source": traffic_src,
                            "algorithmic_bytes": alg_bytes,
                            "kernel": f"{kname} ({gemm_epi_name(epi)} epilogue) M={M_} N={N_} K={K_ + K2_}", "launches": dcount,
                            "avg_ms": round(dms / dcount, 4), "share_of_step": round(dms / total_ms, 4),
                            "all_gemms": {"tflops": round(gemm_fl / (gemm_ms / 1e3) / 1e12, 1), "share_of_step": round(gemm_ms / total_ms, 4)},
                            "by_op_ms_per_step": {n: round(sum(v[1] for k, v in agg.items() if k[0] == n) / 2, 3)
                                                  for n in sorted({k[0] for k in agg})},
                            "gemm_by_shape": [{"M": k[1][0], "N": k[1][1], "K": k[1][2] + k[1][3], "epi": gemm_epi_name(k[1][4]),
                                               "n": v[0] // 2, "avg_ms": round(v[1] / v[0], 4),
                                               "tflops": round(2.0 * k[1][0] * k[1][1] * (k[1][2] + k[1][3]) / (v[1] / v[0] / 1e3) / 1e12, 1)}
                                              for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]) if k[0] == "gemm"][:12]}
    # ---- the metric's second half: a short sample of the greedy decode workload on the same model (full run: --workload decode) ----
    if args.workload == "adapter" and not args.no_infer:
        bucket.zero_grad()
        model._engine.release_buffers()
        r = decode_measure(model, dev, world, rank, args.utts, args.new_tokens, 2, 3, timed)
        if rank == 0:
            d = decode_line(args, r, world, world * args.utts * 2, None, steps=2)
            line["infer"] = {"metric": DECODE_METRIC, "value": d["value"], "unit": "utt/s", "e2e": d["e2e"], "ms_per_batch": d["ms_per_step"],
                             "utterances_per_gpu": args.utts, "new_tokens": args.new_tokens, "roofline": d["roofline"]}
    if args.workload == "adapter" and not args.no_lora_sample:
        model._engine.release_buffers()
        ls = lora_sample(args, dev, world, rank, timed)
        if rank == 0:
            line["lora"] = ls
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        tps, ms, cores, sample = cpu_step_tokens_per_s(args.workload, T, 4, 1)      # ~7 s of CPU work (1 warm-up + 4 timed steps)
        line["cpu_baseline"] = {"value": round(tps, 2), "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        if args.workload == "adapter":
            line["cpu_baseline"]["cfg1"] = cpu_cfg1_forward_loss()
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
