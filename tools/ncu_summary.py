#!/usr/bin/env python
"""Markdown summary of an ncu --set full report (one row per captured launch). usage: tools/ncu_summary.py report.ncu-rep"""
import csv, io, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, data = rows[0], rows[2:]
cols = [("Kernel Name", "kernel"), ("launch__grid_size", "grid"), ("gpu__time_duration.sum", "us"),
        ("dram__bytes_read.sum", "DRAM rd MB"), ("dram__bytes_write.sum", "DRAM wr MB"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
        ("smsp__inst_executed.sum", "warp instr"), ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "ALU %"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "FMA %"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "LSU %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ %"), ("launch__registers_per_thread", "regs")]
idx = [hdr.index(c) for c, _ in cols]
print("| " + " | ".join(n for _, n in cols) + " |")
print("|" + "---|" * len(cols))
for r in data:
    v = []
    for i, (c, _) in zip(idx, cols):
        x = r[i]
        if c == "Kernel Name":
            x = x.split("(")[0].replace("void ", "")
        else:
            try:
                f = float(x.replace(",", ""))
                x = f"{f:.0f}" if f >= 100 else f"{f:.1f}"
            except ValueError:
                pass
        v.append(x)
    print("| " + " | ".join(v) + " |")
