"""Which source lines execute the local-memory loads / stores (LDL / STL) of one kernel in an .ncu-rep.

    python tools/ncu_local_mem.py report.ncu-rep section_index object.o [top_n]
"""
import collections, csv, io, os, re, subprocess, sys, tempfile
rep, which, obj = sys.argv[1], int(sys.argv[2]), sys.argv[3]
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
sections, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name": cur = {"name": r[1], "hdr": None, "data": []}; sections.append(cur)
    elif cur is not None and cur["hdr"] is None: cur["hdr"] = r
    elif cur is not None and len(r) == len(cur["hdr"]): cur["data"].append(r)
s = sections[which]; hdr, data = s["hdr"], s["data"]
iSrc, iI = hdr.index("Source"), hdr.index("Instructions Executed")
print(s["name"][:110])
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
secs = [i for i, l in enumerate(sass) if l.startswith("//--------------------- .text.")]
want = None
for i in secs:
    j = secs[secs.index(i) + 1] if secs.index(i) + 1 < len(secs) else len(sass)
    if sum(1 for l in sass[i:j] if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s", l)) == len(data): want = (i, j); break
assert want, "kernel not found by instruction count"
loc, cur_loc = [], ("?", 0)
for l in sass[want[0]:want[1]]:
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m: cur_loc = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s", l): loc.append(cur_loc)
tot = sum(int(r[iI]) for r in data)
ld, st = collections.Counter(), collections.Counter()
for k, r in enumerate(data):
    op = r[iSrc].split()[0] if r[iSrc].split() else ""
    if op.startswith("@"): op = r[iSrc].split()[1]
    if op.startswith("LDL"): ld[loc[k]] += int(r[iI])
    elif op.startswith("STL"): st[loc[k]] += int(r[iI])
print(f"LDL {100 * sum(ld.values()) / tot:.1f}% of executed instructions, STL {100 * sum(st.values()) / tot:.1f}%")
allc = collections.Counter(); allc.update(ld); allc.update(st)
for (f, ln), v in allc.most_common(topn):
    print(f"  {f}:{ln:<5d} {100 * v / tot:5.2f}%  (ld {100 * ld[(f, ln)] / tot:.2f} st {100 * st[(f, ln)] / tot:.2f})")
