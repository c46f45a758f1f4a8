#!/usr/bin/env python
"""Summarise an ncu report's source page per CUDA source line: instructions executed and stall samples.
usage: tools/ncu_src.py report.ncu-rep kernel_regex [top_n] [sort: s|i]"""
import csv, subprocess, sys, io
rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
key = sys.argv[4] if len(sys.argv) > 4 else "s"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k", f"regex:{kre}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None
data = {}
fname = ""
seen_fn = 0
for r in rows:
    if r and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r and r[0] == "Function Name":
        continue
    if r and r[0] == "Line No":
        hdr = r; iI, iN = hdr.index("Instructions Executed"), hdr.index("# Samples"); continue
    if hdr is None or not r or r[0] == "": continue
    try:
        k = (fname, int(r[0]))
        v = data.setdefault(k, [0, 0, r[1]])
        v[0] += int(r[iI] or 0); v[1] += int(r[iN] or 0)
    except (ValueError, IndexError):
        pass
tot_i = sum(d[0] for d in data.values()) or 1; tot_s = sum(d[1] for d in data.values()) or 1
print(f"total warp-instructions {tot_i}, samples {tot_s} (summed over matching launches)")
si = 1 if key == "s" else 0
for k, d in sorted(data.items(), key=lambda kv: -kv[1][si])[:top]:
    print(f"{100*d[0]/tot_i:5.1f}%i {100*d[1]/tot_s:5.1f}%s  {k[0]}:{k[1]:<4} {d[2].strip()[:140]}")
