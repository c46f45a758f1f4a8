#!/usr/bin/env python
"""Aggregate an ncu source page into named line ranges (phases) of one file.
usage: tools/ncu_phases.py report.ncu-rep kernel_regex file.cu name:lo-hi[,lo-hi] ..."""
import csv, subprocess, sys, io
rep, kre, fn = sys.argv[1], sys.argv[2], sys.argv[3]
phases = []
for a in sys.argv[4:]:
    name, rs = a.split(":")
    phases.append((name, [tuple(int(v) for v in r.split("-")) for r in rs.split(",")]))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k", f"regex:{kre}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None; fname = ""; data = {}
for r in rows:
    if r and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r and r[0] == "Function Name": continue
    if r and r[0] == "Line No":
        hdr = r; iI, iN = hdr.index("Instructions Executed"), hdr.index("# Samples"); continue
    if hdr is None or not r or r[0] == "": continue
    try:
        k = (fname, int(r[0])); v = data.setdefault(k, [0, 0]); v[0] += int(r[iI] or 0); v[1] += int(r[iN] or 0)
    except (ValueError, IndexError): pass
ti = sum(d[0] for d in data.values()) or 1; ts = sum(d[1] for d in data.values()) or 1
acc = {name: [0, 0] for name, _ in phases}; other = [0, 0]
for (f, ln), d in data.items():
    hit = False
    if f == fn:
        for name, rs in phases:
            if any(lo <= ln <= hi for lo, hi in rs):
                acc[name][0] += d[0]; acc[name][1] += d[1]; hit = True; break
    if not hit: other[0] += d[0]; other[1] += d[1]
print(f"total warp-instructions {ti}, samples {ts}")
for name, _ in phases: print(f"{name:>14}: {acc[name][0]:12d} {100*acc[name][0]/ti:5.1f}%i {100*acc[name][1]/ts:5.1f}%s")
print(f"{'other':>14}: {other[0]:12d} {100*other[0]/ti:5.1f}%i {100*other[1]/ts:5.1f}%s")
