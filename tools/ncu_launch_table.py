#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel: launches, total and mean duration,
share of the listed time.  usage: ncu_launch_table.py launches.csv ["title"] [--alternate name1,name2 (kernels launched alternately for the
forward and the backward, e.g. the reference's two renderCUDA<3>)] [--passes k (divide totals by k)] > profiles/xxx.md"""
import csv
import re
import sys
from collections import OrderedDict

argv = sys.argv[1:]
alternate, passes = [], 1
if "--alternate" in argv:
    i = argv.index("--alternate"); alternate = argv[i + 1].split(","); del argv[i:i + 2]
if "--passes" in argv:
    i = argv.index("--passes"); passes = int(argv[i + 1]); del argv[i:i + 2]
sys.argv = [sys.argv[0]] + argv
seen = {}
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
H = rows[0]
ki, mi, vi, ui = H.index("Kernel Name"), H.index("Metric Name"), H.index("Metric Value"), H.index("Metric Unit")
agg = OrderedDict()
for r in rows[1:]:
    if r[mi] != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", r[ki]).replace("void ", "")
    name = re.sub(r"<.*", "", name) if len(name) > 70 else name
    if name in alternate:
        seen[name] = seen.get(name, 0) + 1
        name += " (forward)" if seen[name] % 2 == 1 else " (backward)"
    v = float(r[vi].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "nsecond": 1e-3, "ms": 1e3, "msecond": 1e3}.get(r[ui], 1.0)
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
total = sum(a[1] for a in agg.values())
if len(sys.argv) > 2:
    print(f"# {sys.argv[2]}\n")
print(f"{sum(a[0] for a in agg.values())} launches over {passes} pass(es), {total / 1e3 / passes:.3f} ms of kernel time per pass "
      "(ncu per-launch durations: cold caches, serialised launches -- shares, not absolute step times)\n")
print("| kernel | launches | total per pass [us] | mean [us] | share |")
print("|---|---|---|---|---|")
for name, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{name}` | {n} | {t / passes:.1f} | {t / n:.2f} | {100 * t / total:.1f}% |")
