#!/bin/bash
# BASELINE.json configs[4] at its largest size: 50 M triangles, device LBVH (and host SAH with SAH=1); id checksums must agree between the two trees
set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/pytest_gpu.log 2>&1
cat gpurun_out/pytest_gpu.log
RT_BUILD=lbvh timeout 600 python tools/soup_sweep.py --sizes 1,20,50 --out gpurun_out/soup50_lbvh.json > gpurun_out/soup50_lbvh.log 2>&1
if [ -n "$SAH" ]; then timeout 900 python tools/soup_sweep.py --sizes 1,50 --out gpurun_out/soup50_sah.json > gpurun_out/soup50_sah.log 2>&1; fi
tail -n 4 gpurun_out/soup50_lbvh.log | cut -c1-1200
RT_BUILD=lbvh RT_TRACE_BUILD=1 timeout 200 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 2>&1 | tail -3
