#!/usr/bin/env python
"""Executed warp instructions and stall samples per CUDA source line for one kernel launch of an ncu report:
joins the report's SASS page with nvdisasm's line info of the same kernel in the built library.
usage: ncu_lines.py report.ncu-rep kernel_regex launch_skip mangled_substring [lib.so]"""
import csv, re, subprocess, sys, collections, os, tempfile
rep, rx, skip, mangled = sys.argv[1:5]
lib = sys.argv[5] if len(sys.argv) > 5 else "animating-gaussian-splats_b200/libgsb200.so"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name",
                      "regex:" + rx, "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
H = rows[1]
si, ii = H.index("# Samples"), H.index("Instructions Executed")
body = [r for r in rows[2:] if len(r) > si and r[si].isdigit()]
# one row per instruction (the tool prints rows twice in some versions: dedupe by address)
seen, ins = set(), []
for r in body:
    if r[0] in seen:
        continue
    seen.add(r[0]); ins.append((int(r[si]), int(r[ii]), r[1].strip()))
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
lines = None
for f in sorted(os.listdir(tmp)):
    if not f.endswith(".cubin"):
        continue
    d = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
    if mangled not in d:
        continue
    # split by function
    cur, fn, per = None, None, {}
    for l in d.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", l)
        if m:
            fn = m.group(1); per[fn] = []; cur = None; continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
        if fn and re.match(r"\s*/\*[0-9a-f]{4,}\*/", l):
            per[fn].append(cur)
    for k, v in per.items():
        if mangled in k and len(v) == len(ins):
            lines = v
    if lines:
        break
if not lines:
    sys.exit(f"no function matching {mangled} with {len(ins)} instructions")
agg = collections.defaultdict(lambda: [0, 0])
for (smp, ex, _), ln in zip(ins, lines):
    agg[ln][0] += ex; agg[ln][1] += smp
tot_e = sum(v[0] for v in agg.values()); tot_s = sum(v[1] for v in agg.values())
print(f"total warp-instr {tot_e}  samples {tot_s}")
for ln, (e, s) in sorted(agg.items(), key=lambda kv: (kv[0] is None, kv[0])):
    if e * 200 > tot_e or s * 200 > tot_s:
        print(f"{str(ln):28s} instr {100*e/tot_e:5.1f}%  samples {100*s/tot_s:5.1f}%")
