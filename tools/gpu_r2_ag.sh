#!/bin/bash
# round 2, GPU pass AG: L1 prefetch of the stack's new top entry (the next node / leaf record popped) at push time, A/B
set -x
for lib in librtcore.so librtcore_pf.so; do
  for sc in "plastic_dragon 0 1024 64 8 1" "cornell_box_statue 0 1024 64 8 1"; do
    for rep in 1 2; do echo "== $lib $sc"; RT_LIBRTCORE=$PWD/rustraytracer_b200/$lib timeout 300 python tools/profile_render.py $sc 2>&1 | tail -1; done
  done
  echo "== $lib soup"; RT_LIBRTCORE=$PWD/rustraytracer_b200/$lib timeout 600 python tools/soup_sweep.py --sizes 1,10 2>&1 | python -c "
import sys, json
for line in sys.stdin:
    try: d = json.loads(line); print(d['tris'], round(d['coherent_Mrays_s']), round(d['incoherent_Mrays_s']))
    except Exception: pass"
done
