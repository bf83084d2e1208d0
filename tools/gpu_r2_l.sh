#!/bin/bash
# round 2, GPU pass L: leaf size of the device-built tree (RT_LBVH_LEAF 1..4) x topology (radix / PLOC): dragon render + soup sweep
set -x
for t in radix ploc; do for l in 1 2 3; do
  echo "== $t leaf $l"
  RT_LBVH_LEAF=$l RT_LBVH=$t RT_BUILD=lbvh RT_TRACE_BUILD=1 timeout 300 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 > gpurun_out/ab_lbvh_${t}_$l.log 2>&1; grep "^bvh: device LBVH" gpurun_out/ab_lbvh_${t}_$l.log | tail -1; tail -n 1 gpurun_out/ab_lbvh_${t}_$l.log
  RT_LBVH_LEAF=$l RT_LBVH=$t RT_BUILD=lbvh timeout 600 python tools/soup_sweep.py --sizes 1,10 --out gpurun_out/soup_lbvh_${t}_$l.json 2>&1 | python -c "
import sys, json
for line in sys.stdin:
    try: d = json.loads(line); print(d['tris'], round(d['coherent_Mrays_s']), round(d['incoherent_Mrays_s']), d['bvh4_nodes'], d['max_stack'], d['tree_build_ms'])
    except Exception: pass"
done; done
(RT_LBVH_LEAF=2 timeout 600 python -m pytest tests -m gpu -x -q -k "lbvh or soup" 2>&1 | tail -4)
