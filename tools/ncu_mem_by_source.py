"""Memory traffic of one kernel launch in an .ncu-rep by source line: which loads / stores are the uncoalesced ones.

    python tools/ncu_mem_by_source.py report.ncu-rep launch_index|name_substring object.o [top_n]

For every memory column of ncu's source page (requests, sectors, wavefronts: global / local / shared) the totals and the top source lines, with
sectors per request (32 = every lane its own sector, 16 = a fully coalesced 16-byte load, 1-2 = a broadcast).

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
mem_cols = [i for i, h in enumerate(hdr) if any(k in h for k in ("Sectors", "Requests", "Wavefronts")) and "Excessive" not in h]
def num(x):
    try: return float(x.replace(",", ""))
    except Exception: return 0.0
print("memory columns:", [hdr[i] for i in mem_cols])
tot = {i: sum(num(r[i]) for r in data[:n]) for i in mem_cols}
tot_i = sum(int(r[iI]) for r in data)
per = {i: collections.Counter() for i in mem_cols}
inst_line = collections.Counter()
for k in range(n):
    key = loc[k]
    inst_line[key] += int(data[k][iI])
    for i in mem_cols:
        v = num(data[k][i])
        if v: per[i][key] += v
print("\ntotals per launch:")
for i in mem_cols: print(f"  {hdr[i]:50s} {tot[i]:16.0f}")
def col(name):
    for i in mem_cols:
        if hdr[i] == name: return i
    return None
pairs = [("L1 Tag Requests Global", "L2 Theoretical Sectors Global"), ("L1 Tag Requests Local", "L2 Theoretical Sectors Local")]
for rq, sc in pairs:
    a, b = col(rq), col(sc)
    if a is None or b is None or not tot[b]: continue
    print(f"\ntop lines by {hdr[b]} (share of the launch, sectors per request, warp instructions executed on the line):")
    for key, v in per[b].most_common(topn):
        r = per[a].get(key, 0)
        print(f"  {key[0]}:{key[1]:<5d} {100 * v / tot[b]:5.2f}%   {v / r if r else 0:5.1f} sectors/request   {r:12.0f} requests   {100 * inst_line[key] / tot_i:5.2f}% inst")
lc = col("L2 Theoretical Sectors Local")
if lc is not None and tot[lc]:
    print(f"\ntop lines by {hdr[lc]} (share of the launch; sectors per warp instruction executed on the line):")
    for key, v in per[lc].most_common(topn):
        print(f"  {key[0]}:{key[1]:<5d} {100 * v / tot[lc]:5.2f}%   {v:12.0f} sectors   {100 * inst_line[key] / tot_i:5.2f}% inst")
