"""Turn `ncu --set full` captures of the CURRENT build into profiles/traffic.json, the file bench.py reads `roofline.traffic` from
(DRAM bytes per launch of the dominant training GEMM; DRAM bytes of one decode token step).  The file carries the sha256 of the
libdualhyp.so it was captured from; bench.py reports traffic = null when the library it runs is a different build.

On the GPU box (one GPU; each only after the plain command exited 0):
    ncu --set full --clock-control none -k regex:gemm_pair_kernel -s 700 -c 16 -o gpurun_out/tr_gemm python bench.py --steps 1 --warmup 3 \\
        --no-cpu-baseline --no-roofline --no-infer
    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/tr_decode.csv \\
        python scripts/decode_traffic.py 128
Here:
    ncu -i gpurun_out/tr_gemm.ncu-rep --page raw --csv > gpurun_out/tr_gemm_raw.csv
    python scripts/capture_traffic.py gpurun_out/tr_gemm_raw.csv [gpurun_out/tr_decode.csv [utterances]]"""
import csv
import hashlib
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def so_hash() -> str:
    sys.path.insert(0, ROOT)
    from dualhyp_b200 import _lib
    return _lib.source_id()


def main():
    out = {"src_sha16": so_hash(), "source": "ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch (scripts/capture_traffic.py)"}
    allrows = list(csv.DictReader(l for l in open(sys.argv[1]) if not l.startswith("==")))
    units = allrows[0]                       # the raw page's second line holds the unit of every column
    mult = {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}
    rows = [r for r in allrows if r.get("ID", "").strip().isdigit()]
    by = {}
    for r in rows:
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void dh::", "").replace("void ", "").replace("dh::", "")
        tot = 0.0
        for col in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            tot += float(r[col].replace(",", "")) * mult[units[col].strip().lower()]
        by.setdefault(name, []).append((tot, float(r["gpu__time_duration.sum"].replace(",", ""))))
    # one kernel name covers several GEMM shapes (the epilogue is the template argument): split launches of a name by duration class
    split = {}
    for name, v in by.items():
        v.sort(key=lambda x: x[1])
        groups = []
        for b, t in v:
            if groups and t < 1.25 * groups[-1][-1][1]:
                groups[-1].append((b, t))
            else:
                groups.append([(b, t)])
        for gi, g in enumerate(groups):
            split[name if len(groups) == 1 else f"{name}#{gi}"] = g
    by = {k: [b for b, _ in g] for k, g in split.items()}
    dur = {k: sum(t for _, t in g) / len(g) for k, g in split.items()}
    # the capture command of scripts/profile_step.sh: bench.py's default adapter step (16 packed prompts); bench.py reports these
    # bytes only for that workload
    out["train_config"] = {"workload": "adapter", "prompts_per_gpu": 16, "packed": True}
    out["train"] = {k: {"launches": len(v), "dram_bytes_per_launch": round(sum(v) / len(v)), "min": round(min(v)), "max": round(max(v)),
                        "ncu_duration_us": round(dur[k], 1)} for k, v in by.items()}
    if len(sys.argv) > 2:
        tot = 0.0
        steps = 0
        for r in csv.DictReader(l for l in open(sys.argv[2]) if not l.startswith("==")):
            if r.get("Metric Name", "").startswith("dram__bytes"):
                v = float(r["Metric Value"].replace(",", ""))
                u = r.get("Metric Unit", "byte").lower()
                tot += v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
            if "argmax" in r.get("Kernel Name", "") and r.get("Metric Name", "") == "dram__bytes_read.sum":
                steps += 1
        out["decode"] = {(sys.argv[3] if len(sys.argv) > 3 else "256"): {"token_steps": steps, "dram_bytes_per_token_step": round(tot / max(steps, 1))}}
    path = os.path.join(ROOT, "profiles", "traffic.json")
    json.dump(out, open(path, "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
