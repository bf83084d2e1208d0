#!/bin/bash
# round 2, GPU pass K: PLOC topology of the device build — parity (LBVH tests, soup parity on both trees), A/B against the radix tree and the host SAH tree
set -x
(RT_TRACE_BUILD=1 timeout 900 python -m pytest tests -m gpu -x -q -k "lbvh or soup or degenerate_uv" 2>&1 | tail -12) > gpurun_out/pytest_ploc.log 2>&1; cat gpurun_out/pytest_ploc.log
for t in ploc radix; do
  RT_LBVH=$t RT_BUILD=lbvh RT_TRACE_BUILD=1 timeout 300 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 > gpurun_out/ab_lbvh_$t.log 2>&1; grep "^bvh" gpurun_out/ab_lbvh_$t.log | tail -2; tail -n 1 gpurun_out/ab_lbvh_$t.log
  RT_LBVH=$t RT_BUILD=lbvh timeout 600 python tools/soup_sweep.py --sizes 1,10 --out gpurun_out/soup_lbvh_$t.json > gpurun_out/soup_lbvh_$t.log 2>&1; cut -c1-260 gpurun_out/soup_lbvh_$t.log
done
timeout 300 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 | tail -n 1
timeout 600 python tools/soup_sweep.py --sizes 1,10 --out gpurun_out/soup_sah.json > gpurun_out/soup_sah.log 2>&1; cut -c1-260 gpurun_out/soup_sah.log
