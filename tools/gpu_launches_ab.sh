#!/bin/bash
# per-kernel launch durations (ncu, serialised, cold) of one 16-spp dragon render for every librtcore*.so variant present
for so in rustraytracer_b200/librtcore*.so; do
  tag=$(basename $so .so)
  RT_LIBRTCORE=$PWD/$so timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_$tag.csv python tools/profile_render.py plastic_dragon 0 1024 16 8 1 > gpurun_out/launches_$tag.log 2>&1
  python - "$tag" <<'PY'
import csv, sys, collections
tag = sys.argv[1]
rows = [r for r in csv.reader(l for l in open(f"gpurun_out/launches_{tag}.csv") if l.startswith('"'))]
h = rows[0]; ki = h.index("Kernel Name"); vi = h.index("Metric Value")
tot = collections.defaultdict(float); cnt = collections.Counter()
for r in rows[1:]:
    name = r[ki].split("(")[0][:70]; tot[name] += float(r[vi].replace(",", "")) / 1e6; cnt[name] += 1
print(tag)
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]): print(f"  {v:9.3f} ms  x{cnt[k]:4d}  {k}")
PY
done
