"""Attribute the stall samples / executed instructions of one kernel launch in an .ncu-rep to source lines.

    python tools/ncu_by_source.py report.ncu-rep launch_index object.o [top_n]

ncu's CSV source page carries per-SASS-instruction metrics but no line numbers, so the line of every instruction is
taken from `nvdisasm -g` of the SAME build's object file (-lineinfo) and joined by instruction index.
Prints the totals per source file, per function-sized line bucket and the top lines.
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

rep, which, obj = sys.argv[1], sys.argv[2], sys.argv[3]          # `which`: launch index, or a substring of the kernel name (first match)
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
sections, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name": cur = {"name": r[1], "hdr": None, "data": []}; sections.append(cur)
    elif cur is not None and cur["hdr"] is None: cur["hdr"] = r
    elif cur is not None and len(r) == len(cur["hdr"]): cur["data"].append(r)
s = sections[int(which)] if which.isdigit() else [x for x in sections if which in x["name"]][0]; hdr, data = s["hdr"], s["data"]
iS, iI, iT = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
name = s["name"]
print(name[:120])

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
# find the section of this kernel: match on the demangled-name fragments through c++filt
secs = [i for i, l in enumerate(sass) if l.startswith("//--------------------- .text.")]
want = None
for i in secs:
    m = re.match(r"//-+ \.text\.(\S+) -+", sass[i])
    dem = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
    if dem.replace(" ", "").replace("(int)", "").replace("(bool)", "") .startswith(name.split("(rt::Ctrl")[0].replace(" ", "").replace("(int)", "").replace("(bool)", "").replace("void", "", 1)) or \
       dem.replace(" ", "") == name.replace(" ", ""):
        want = i; break
if want is None:
    # fall back: match by instruction count
    for i in secs:
        j = secs[secs.index(i) + 1] if secs.index(i) + 1 < len(secs) else len(sass)
        n = sum(1 for l in sass[i:j] if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s", l))
        if n == len(data): want = i; break
assert want is not None, "kernel section not found in " + obj
end = secs[secs.index(want) + 1] if secs.index(want) + 1 < len(secs) else len(sass)
loc, cur_loc = [], ("?", 0)
for l in sass[want:end]:
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m: cur_loc = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s", l): loc.append(cur_loc)
print("sass instructions: ncu", len(data), "nvdisasm", len(loc))
n = min(len(data), len(loc))
tot_s, tot_i = sum(int(r[iS]) for r in data), sum(int(r[iI]) for r in data)
by_file, by_line = collections.Counter(), collections.Counter()
inst_file, inst_line = collections.Counter(), collections.Counter()
thr_file = collections.Counter()
static_file = collections.Counter()
for k in range(n):
    f, ln = loc[k]
    by_file[f] += int(data[k][iS]); by_line[(f, ln)] += int(data[k][iS])
    inst_file[f] += int(data[k][iI]); inst_line[(f, ln)] += int(data[k][iI]); static_file[f] += 1; thr_file[f] += int(data[k][iT])
print("\nper file:            samples   warp-inst   static SASS   lanes/inst")
for f, v in by_file.most_common():
    print(f"  {f:18s} {100 * v / tot_s:6.1f}%   {100 * inst_file[f] / tot_i:6.1f}%   {static_file[f]:6d}   {thr_file[f] / max(inst_file[f], 1):6.1f}")
print("\ntop lines by samples:")
for (f, ln), v in by_line.most_common(topn):
    print(f"  {f}:{ln:<5d} {100 * v / tot_s:5.2f}% samples  {100 * inst_line[(f, ln)] / tot_i:5.2f}% inst")
