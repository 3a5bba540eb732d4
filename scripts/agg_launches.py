"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by (kernel, grid): python scripts/agg_launches.py file.csv [last_n] [div]"""
import collections
import csv
import re
import sys

path = sys.argv[1]
last_n = int(sys.argv[2]) if len(sys.argv) > 2 else 0
div = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
rows = [(x["Kernel Name"], x.get("Grid Size"), float(x["Metric Value"])) for x in csv.DictReader(lines)]
if last_n:
    rows = rows[-last_n:]
agg = collections.OrderedDict()
for n, g, v in rows:
    a = agg.setdefault((re.sub(r"\(.*", "", n)[:60], g), [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(v[1] for v in agg.values())
print(f"{len(rows)} launches, {tot / 1e3 / div:.1f} us per step (sum of kernel durations / {div:g})")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[0]:60s} {k[1]:>16s} x{v[0] / div:6.1f} {v[1] / v[0] / 1e3:8.2f} us  {v[1] / div / 1e3:8.1f} us/step  {100 * v[1] / tot:5.1f}%")
