"""HBM roofline of the ray-cast sweep at sizes whose tree does not fit L2 ("HBM bandwidth once they do not", VERDICT r01 missing #2).

    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:k_closest_persistent \
        --csv --page raw --log-file soup10.csv python tools/soup_sweep.py --sizes 10
    python tools/soup_roofline.py soup10.csv 10000000 profiles/soup_roofline_r02.json

tools/soup_sweep.py launches the kernel 4 x on the coherent batch (warm-up + 3 timed) and then 4 x on the incoherent one, 2^24 rays each; the last
launch of each group is used.  achieved = DRAM bytes of that launch / its duration (both from ncu, serialised and cold-clock: a lower bound of the live rate),
peak = MEASURED_PEAKS.json hbm_gbs."""
import csv
import json
import os
import sys

path, tris, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
hdr, units, data = rows[0], rows[1], rows[2:]
data = [r for r in data if "k_closest_persistent" in r[hdr.index("Kernel Name")]]
assert len(data) >= 8, len(data)


def val(r, m):
    i = hdr.index(m)
    return float(r[i].replace(",", "")) * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0}.get(units[i], 1.0)


root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pk = os.path.join(root, "MEASURED_PEAKS.json")
peak = json.load(open(pk))["hbm_gbs"] if os.path.exists(pk) else 6650.0
rays = 1 << 24
res = json.load(open(out)) if os.path.exists(out) else {"what": "ray-cast sweep (rt_closest_device), DRAM traffic per launch from ncu vs the measured HBM copy bandwidth", "rows": []}
for name, r in (("coherent", data[3]), ("incoherent", data[7])):
    b = val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")
    s = val(r, "gpu__time_duration.sum")
    res["rows"].append({"tris": tris, "batch": name, "rays": rays, "ms_under_ncu": s * 1e3, "Mrays_s_under_ncu": rays / s / 1e6, "dram_bytes_per_ray": b / rays,
                        "achieved_GBs": b / s / 1e9, "peak_GBs": peak, "frac": b / s / 1e9 / peak, "l2_hit_pct": val(r, "lts__t_sector_hit_rate.pct")})
json.dump(res, open(out, "w"), indent=1)
print(json.dumps(res["rows"][-2:]))
