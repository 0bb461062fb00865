#!/usr/bin/env python
"""Per-kernel DRAM traffic and duration of an ncu --set full report as JSON (profiles/r1_traffic.json).
usage: ncu_traffic.py report.ncu-rep > profiles/r1_traffic.json"""
import csv, json, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
H, U = rows[0], rows[1]
ki, ri, wi, ti = H.index("Kernel Name"), H.index("dram__bytes_read.sum"), H.index("dram__bytes_write.sum"), H.index("gpu__time_duration.sum")
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
tscale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}
k = {}
for r in rows[2:]:
    name = r[ki].split("(")[0].split("<")[0].replace("void ", "").replace("gsb::", "")
    full = r[ki]
    if name.startswith("seg_") and "<" in full:
        name = name + "<" + full.split("<", 1)[1].split(">")[0].replace("unsigned short", "u16").replace("unsigned int", "u32").replace("(int)", "").replace("(bool)", "") + ">"
    e = k.setdefault(name, {"launches": 0, "dram_bytes": 0.0, "time_ms": 0.0})
    e["launches"] += 1
    e["dram_bytes"] += (float(r[ri]) * scale[U[ri]] + float(r[wi]) * scale[U[wi]])
    e["time_ms"] += float(r[ti]) * tscale[U[ti]]
for e in k.values():
    e["dram_bytes_per_launch"] = e["dram_bytes"] / e["launches"]
    e["time_ms_per_launch"] = e["time_ms"] / e["launches"]
inst = [v for n, v in k.items() if n.startswith("seg_") and "u16" in n]
if inst:
    k["seg_sort_instances"] = {"launches": 1, "dram_bytes": sum(v["dram_bytes"] for v in inst),
                               "time_ms": sum(v["time_ms"] for v in inst),
                               "dram_bytes_per_launch": sum(v["dram_bytes"] for v in inst),
                               "time_ms_per_launch": sum(v["time_ms"] for v in inst),
                               "note": "sum of the three kernels of the one-pass instance sort (u16 keys)"}
print(json.dumps({"source": "ncu --set full --clock-control none, python bench.py --profile-only (C2 workload), one step; "
                            "units bytes / ms; tools/prof_r1.sh", "kernels": k}, indent=1))
