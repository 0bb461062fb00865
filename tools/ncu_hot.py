#!/usr/bin/env python
"""Top stall-sample SASS instructions of one kernel launch in an ncu report.
usage: ncu_hot.py report.ncu-rep kernel_regex [launch_skip] [top_n]"""
import csv, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name",
                      "regex:" + rx, "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
H = rows[1]
si, ii, src = H.index("# Samples"), H.index("Instructions Executed"), H.index("Source")
body = [r for r in rows[2:] if len(r) > si and r[si].isdigit()]
tot = sum(int(r[si]) for r in body); toti = sum(int(r[ii]) for r in body)
print(rows[0][1][:90], "samples", tot, "warp-instr", toti)
stall_cols = [i for i, h in enumerate(H) if h.startswith("stall_")]
agg = {H[i]: sum(int(r[i] or 0) for r in body if len(r) > i and r[i].isdigit()) for i in stall_cols}
print("stalls:", sorted(((v, k) for k, v in agg.items() if v), reverse=True)[:8])
for n, r in sorted(((int(r[si]), r) for r in body), key=lambda x: -x[0])[:top]:
    print(f"{n:7d} {100*n/tot:5.1f}%  exec {int(r[ii]):>10d}  line#{body.index(r):4d} {r[src].strip()[:70]}")
