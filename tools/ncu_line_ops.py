"""Executed instructions of ONE source line of a kernel in an .ncu-rep, by opcode and by SASS address range (what is the code the line table files under it?).

    python tools/ncu_line_ops.py report.ncu-rep kernel_name_substring object.o file.cuh line
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

rep, which, obj, fname, line = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4], int(sys.argv[5])
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
sections, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name": cur = {"name": r[1], "hdr": None, "data": []}; sections.append(cur)
    elif cur is not None and cur["hdr"] is None: cur["hdr"] = r
    elif cur is not None and len(r) == len(cur["hdr"]): cur["data"].append(r)
s = [x for x in sections if which in x["name"]][0]; hdr, data = s["hdr"], s["data"]
iI, iS, iSrc = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
secs = [i for i, l in enumerate(sass) if l.startswith("//--------------------- .text.")]
want = None
for i in secs:
    j = secs[secs.index(i) + 1] if secs.index(i) + 1 < len(secs) else len(sass)
    n = sum(1 for l in sass[i:j] if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s", l))
    if n == len(data): want = (i, j); break
assert want, "kernel section not found by instruction count"
loc, cur_loc = [], ("?", 0)
for l in sass[want[0]:want[1]]:
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m: cur_loc = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s", l): loc.append(cur_loc)
ops = collections.Counter(); tot = 0; runs = []; all_i = sum(int(r[iI]) for r in data)
for k, r in enumerate(data):
    if loc[k] == (fname, line):
        n = int(r[iI]); op = r[iSrc].split()[0] if not r[iSrc].startswith("@") else r[iSrc].split()[1]
        ops[op.split(".")[0]] += n; tot += n
        if runs and runs[-1][1] == k - 1: runs[-1][1] = k; runs[-1][2] += n
        else: runs.append([k, k, n])
print(f"{fname}:{line}: {tot} executed warp instructions = {100 * tot / all_i:.2f}% of the kernel")
print("by opcode:", dict(ops.most_common(14)))
print("contiguous SASS ranges (first index, last index, executed):")
for a, b, n in sorted(runs, key=lambda x: -x[2])[:8]:
    print(f"  [{a}, {b}] {n}   first: {data[a][iSrc][:70]}   max per-instruction count {max(int(data[k][iI]) for k in range(a, b + 1))}")
