#!/usr/bin/env python
"""profiles/ncu_traffic.json from an `ncu --set full` report of `bench.py --steps 1 --warmup 1 --batch 64 --no-cpu`:
per-stage DRAM bytes and warp instructions of one pass (one launch of each stage kernel; the pyramid launches summed).
usage: tools/ncu_traffic.py report.ncu-rep "source note" [frames per launch, default 64] > profiles/ncu_traffic.json"""
import csv, io, json, subprocess, sys
rep, note = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, data = rows[0], rows[1], rows[2:]
ik, ir, iw, ii = (hdr.index(c) for c in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum"))
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
stage_of = {"orb_pyr": "pyramid", "orb_fast": "fast_cells", "orb_blur": "blur", "orb_quadtree": "quadtree", "orb_describe": "describe"}
dram, instr = {}, {}
for r in data:
    name = r[ik].replace("void ", "")
    st = next((v for k, v in stage_of.items() if name.startswith(k)), None)
    if st is None:
        continue
    d = dram.setdefault(st, {"read": 0, "write": 0})
    d["read"] += int(float(r[ir].replace(",", "")) * scale[units[ir]])
    d["write"] += int(float(r[iw].replace(",", "")) * scale[units[iw]])
    instr[st] = instr.get(st, 0) + int(float(r[ii].replace(",", "")))
dram["pyramid"]["note"] = "level0 interior + border + 7 resize launches of one pass summed"
instr["note"] = "smsp__inst_executed.sum of the same capture; the extraction kernels are bound by instruction issue, not by DRAM"
print(json.dumps({"source": note, "frames_per_launch": int(sys.argv[3]) if len(sys.argv) > 3 else 64, "dram_bytes_per_launch": dram, "warp_instructions_per_launch": instr}, indent=2))
