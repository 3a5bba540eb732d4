"""Turn the captures of scripts/profile_step.sh (+ the decode launch list) into the tracked files under profiles/:
    python scripts/summarize_profiles.py <tag>          # reads gpurun_out/<tag>_{launches.csv,gemm.ncu-rep,attn.ncu-rep,decode.csv}
writes profiles/r02_launch_summary.txt, r02_ncu_full_summary.txt, r02_decode_launch_summary.txt and profiles/traffic.json
(through scripts/capture_traffic.py).  Needs `ncu` (to export the raw pages); no GPU."""
import collections
import csv
import hashlib
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
G = lambda name: os.path.join(ROOT, "gpurun_out", f"{tag}_{name}")  # noqa: E731
P = lambda name: os.path.join(ROOT, "profiles", name)  # noqa: E731
sys.path.insert(0, ROOT)
from dualhyp_b200 import _lib  # noqa: E402
sha = _lib.source_id()          # comment- and space-insensitive hash of the kernel sources (the .so is not reproducible byte for byte)

# ---- 1. launch list of one adapter step ----
rows = [(x["Kernel Name"], float(x["Metric Value"])) for x in csv.DictReader(l for l in open(G("launches.csv")) if not l.startswith("=="))]
idx = [i for i, r in enumerate(rows) if "adamw" in r[0]]
step = rows[idx[-2] + 1: idx[-1] + 1]
agg = collections.OrderedDict()
for n, v in step:
    d = agg.setdefault(re.sub(r"\(.*", "", n)[:72], [0, 0.0])
    d[0] += 1
    d[1] += v
tot = sum(v[1] for v in agg.values())
out = [f"ncu --metrics gpu__time_duration.sum --clock-control none launch list of ONE adapter fine-tune step (bench.py --steps 1 --warmup 3, 16 packed prompts), "
       f"sources {sha}", f"{len(step)} launches, {tot / 1e3:.1f} us sum of kernel durations (cold-cache, serialised: compare SHARES with bench.py's by_op_ms_per_step)"]
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    out.append(f"{k:72s} x{v[0]:5d} {v[1] / v[0] / 1e3:8.2f} us {v[1] / 1e3:9.1f} us/step {100 * v[1] / tot:5.1f}%")
dh = sum(v[1] for k, v in agg.items() if "dh::" in k)
out.append(f"dh:: kernels {dh / 1e3:.1f} us ({100 * dh / tot:.1f}%), ATen / other {(tot - dh) / 1e3:.1f} us in {sum(v[0] for k, v in agg.items() if 'dh::' not in k)} launches")
out.append(f"shares: GEMM kernels {100 * sum(v[1] for k, v in agg.items() if 'gemm' in k) / tot:.1f}%, attention kernels "
           f"{100 * sum(v[1] for k, v in agg.items() if 'attn' in k) / tot:.1f}%")
open(P("r02_launch_summary.txt"), "w").write("\n".join(out) + "\n")

# ---- 2. ncu --set full raw pages ----
REPS = ["gemm", "attn"] + (["dattn"] if os.path.exists(G("dattn.ncu-rep")) else [])      # dattn: the decode attention kernel (256 sequences)
for rep in REPS:
    with open(G(f"{rep}_raw.csv"), "w") as f:
        subprocess.run(["ncu", "-i", G(f"{rep}.ncu-rep"), "--page", "raw", "--csv"], stdout=f, stderr=subprocess.DEVNULL, check=True)
cols = ["Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "smsp__inst_executed.sum"]
out = [f"ncu --set full --clock-control none, sources {sha}, one adapter fine-tune step of 16 packed prompts",
       "columns: kernel, grid, duration us, DRAM read MB, DRAM write MB, SM throughput %, tensor pipe active % (sm__pipe_tensor_cycles_active), warps active %, "
       "regs/thread, warp instructions, issue-active %, L2 sector hit rate %", ""]
for rep in REPS:
    if rep == "dattn":
        out.append("decode attention (scripts/decode_traffic.py 256: eager token step, 256 sequences of 346..496 cached keys):")
    r = list(csv.reader(open(G(f"{rep}_raw.csv"))))
    ix = {h: i for i, h in enumerate(r[0])}
    units = r[1]
    mb = {"byte": 1e-6, "kbyte": 1e-3, "mbyte": 1.0, "gbyte": 1e3}
    for row in r[2:]:
        name = re.sub(r"\(.*", "", row[ix["Kernel Name"]]).replace("void ", "")
        v = [row[ix[c]] for c in cols]
        rd = float(v[2]) * mb[units[ix[cols[2]]].lower()]
        wr = float(v[3]) * mb[units[ix[cols[3]]].lower()]
        ex = [float(row[ix[c]]) if c in ix else float("nan") for c in ("smsp__issue_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct")]
        out.append(f"{name:28s} {v[0]:>13s} {float(v[1]):8.1f} {rd:8.1f} {wr:8.1f} {float(v[4]):6.1f} {float(v[5]):6.1f} {float(v[6]):6.1f} {v[7]:>4s} "
                   f"{float(v[8]) / 1e6:7.2f}M {ex[0]:6.1f} {ex[1]:6.1f}")
    out.append("")
open(P("r02_ncu_full_summary.txt"), "w").write("\n".join(out) + "\n")

# ---- 3. decode token step ----
per = collections.OrderedDict()
mult = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}
for r in csv.DictReader(l for l in open(G("decode.csv")) if not l.startswith("==")):
    k = (r["ID"], re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").replace("dh::", ""), r["Grid Size"])
    per.setdefault(k, {})[r["Metric Name"]] = float(r["Metric Value"].replace(",", "")) * mult.get(r["Metric Unit"].lower(), 1)
agg = collections.OrderedDict()
for (_, n, g), d in per.items():
    a = agg.setdefault((n, g), [0, 0.0, 0.0])
    a[0] += 1
    a[1] += d.get("gpu__time_duration.sum", 0)
    a[2] += d.get("dram__bytes_read.sum", 0) + d.get("dram__bytes_write.sum", 0)
tt, tb = sum(a[1] for a in agg.values()), sum(a[2] for a in agg.values())
out = [f"ncu launch list (gpu__time_duration.sum, dram__bytes_read.sum + dram__bytes_write.sum; --clock-control none) of TWO eager decode token steps, 256 sequences, "
       f"sources {sha} (scripts/decode_traffic.py 256)",
       f"{sum(a[0] for a in agg.values()) // 2} kernels per token step, {tt / 2:.0f} us sum of kernel durations per step (cold-cache, serialised), "
       f"{tb / 2 / 1e9:.2f} GB of DRAM traffic per step",
       f"{'kernel':44s} {'grid':>14s} {'x/step':>6s} {'us each':>8s} {'us/step':>8s} {'MB each':>8s} {'GB/s':>7s}"]
for (n, g), a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    out.append(f"{n[:44]:44s} {g:>14s} {a[0] / 2:6.0f} {a[1] / a[0]:8.1f} {a[1] / 2:8.0f} {a[2] / a[0] / 1e6:8.1f} {a[2] / a[1] / 1e3:7.0f}")
open(P("r02_decode_launch_summary.txt"), "w").write("\n".join(out) + "\n")

subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "capture_traffic.py"), G("gemm_raw.csv"), G("decode.csv"), "256"], check=True, stdout=subprocess.DEVNULL)
print(open(P("r02_launch_summary.txt")).read().splitlines()[1])
print("\n".join(open(P("r02_ncu_full_summary.txt")).read().splitlines()[3:32]))
print(open(P("r02_decode_launch_summary.txt")).read().splitlines()[1])
