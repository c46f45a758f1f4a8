#!/usr/bin/env python
"""Static SASS instruction count per CUDA source line of one kernel (needs -lineinfo).
usage: tools/sass_lines.py file.o kernel_substring [file.cu]   -> 'line count  [opcodes]' sorted by line."""
import subprocess, sys, re, tempfile, os, collections
obj, ksub = sys.argv[1], sys.argv[2]
only = sys.argv[3] if len(sys.argv) > 3 else None
d = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, stdout=subprocess.DEVNULL)
cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
out = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
cur_fn = None; cur = (None, 0); cnt = collections.Counter(); ops = collections.defaultdict(collections.Counter)
for ln in out.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+),", ln)
    if m: cur_fn = m.group(1); continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if cur_fn is None or ksub not in cur_fn: continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)", ln)
    if m:
        op = m.group(2).split(".")[0]
        if op in ("NOP",): continue
        cnt[cur] += 1; ops[cur][op] += 1
tot = sum(cnt.values())
print("total static instructions:", tot)
for (f, l) in sorted(cnt):
    if only and f != only: continue
    print(f"{f}:{l:<5d} {cnt[(f,l)]:4d}  " + " ".join(f"{o}x{n}" if n > 1 else o for o, n in ops[(f, l)].most_common()))
