#!/bin/bash
# device LBVH against the host SAH build: commit time and traversal speed (soup sweep + the dragon render)
set -x
RT_BUILD=lbvh timeout 900 python tools/soup_sweep.py --sizes 1,2,5,10,20 --out gpurun_out/soup_sweep_lbvh.json > gpurun_out/soup_lbvh.log 2>&1
for b in sah lbvh; do
  RT_BUILD=$b RT_TRACE_BUILD=1 timeout 200 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 > gpurun_out/dragon_$b.log 2>&1
done
tail -n 6 gpurun_out/soup_lbvh.log | cut -c1-900; tail -n 3 gpurun_out/dragon_sah.log gpurun_out/dragon_lbvh.log
