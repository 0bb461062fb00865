#!/usr/bin/env python
"""Summarise an ncu report (--set full) as a markdown table: one row per captured launch.
usage: ncu_summary.py report.ncu-rep > profiles/xxx.md"""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
H, U = rows[0], rows[1]
cols = [("Kernel Name", "kernel"), ("launch__grid_size", "grid"), ("launch__registers_per_thread", "regs"),
        ("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma pipe %"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu (SFU) pipe %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
        ("smsp__inst_executed.sum", "warp instr")]
idx = [(H.index(c), n) for c, n in cols if c in H]
print("| " + " | ".join(f"{n} [{U[i]}]" if U[i] else n for i, n in idx) + " |")
print("|" + "---|" * len(idx))
for r in rows[2:]:
    cells = []
    for i, n in idx:
        v = r[i]
        if n == "kernel":
            v = v.split("(")[0].replace("void ", "").replace("gsb::", "")
        else:
            try:
                f = float(v); v = f"{f:.4g}" if abs(f) < 1e6 else f"{f:.3e}"
            except ValueError:
                pass
        cells.append(v)
    print("| " + " | ".join(cells) + " |")
