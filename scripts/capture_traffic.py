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
    python scripts/capture_traffic.py gpurun_out/tr_gemm_raw.csv [gpurun_out/tr_decode.csv]"""
import csv
import hashlib
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def so_hash() -> str:
    return hashlib.sha256(open(os.path.join(ROOT, "dualhyp_b200", "libdualhyp.so"), "rb").read()).hexdigest()[:16]


def main():
    out = {"so_sha16": so_hash(), "source": "ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch (scripts/capture_traffic.py)"}
    rows = list(csv.DictReader(l for l in open(sys.argv[1]) if not l.startswith("==")))
    rows = [r for r in rows if r.get("ID", "").strip().isdigit()]
    by = {}
    for r in rows:
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void dh::", "")
        rd, wr = float(r["dram__bytes_read.sum"].replace(",", "")), float(r["dram__bytes_write.sum"].replace(",", ""))
        unit = 1.0
        by.setdefault(name, []).append((rd + wr) * unit)
    out["train"] = {k: {"launches": len(v), "dram_bytes_per_launch": round(sum(v) / len(v)), "min": round(min(v)), "max": round(max(v))} for k, v in by.items()}
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
        out["decode"] = {"128": {"token_steps": steps, "dram_bytes_per_token_step": round(tot / max(steps, 1))}}
    path = os.path.join(ROOT, "profiles", "traffic.json")
    json.dump(out, open(path, "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
