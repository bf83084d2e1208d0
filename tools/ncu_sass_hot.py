"""Hot SASS instructions of one launch in an .ncu-rep (source page): python tools/ncu_sass_hot.py rep [launch_index] [top_n]"""
import csv, io, subprocess, sys
rep, which, topn = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0, int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
sections, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name": cur = {"name": r[1], "hdr": None, "data": []}; sections.append(cur)
    elif cur is not None and cur["hdr"] is None: cur["hdr"] = r
    elif cur is not None and len(r) == len(cur["hdr"]): cur["data"].append(r)
s = sections[which]; hdr, data = s["hdr"], s["data"]
print(s["name"][:100], "| launches in report:", len(sections))
iS, iI, iT, isrc = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Avg. Threads Executed"), hdr.index("Source")
tot, totI = sum(int(r[iS]) for r in data), sum(int(r[iI]) for r in data)
print("samples", tot, "warp-insts", totI, "sass lines", len(data))
for c in ["stall_long_sb", "stall_wait", "stall_no_inst", "stall_short_sb", "stall_branch_resolving", "stall_math", "stall_lg", "stall_not_selected", "stall_selected", "stall_dispatch", "stall_mio", "stall_tex"]:
    if c in hdr:
        i = hdr.index(c); print(f"  {c:26s} {100 * sum(int(r[i] or 0) for r in data) / max(tot, 1):5.1f}%")
top = sorted(range(len(data)), key=lambda k: -int(data[k][iS]))[:topn]
for k in sorted(top):
    r = data[k]; print(f"{k:5d} smp {100 * int(r[iS]) / tot:5.2f}% inst {100 * int(r[iI]) / totI:5.2f}% thr {r[iT]:>5s}  {r[isrc][:100]}")
