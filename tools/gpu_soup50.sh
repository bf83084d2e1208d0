#!/bin/bash
# BASELINE.json configs[4] at its largest size: 50 M triangles, device LBVH and host SAH; id checksums must agree between the two trees
set -x
nvidia-smi --query-gpu=memory.total --format=csv; free -g | head -2; nproc
RT_BUILD=lbvh timeout 600 python tools/soup_sweep.py --sizes 1,50 --out gpurun_out/soup50_lbvh.json > gpurun_out/soup50_lbvh.log 2>&1
timeout 900 python tools/soup_sweep.py --sizes 1,50 --out gpurun_out/soup50_sah.json > gpurun_out/soup50_sah.log 2>&1
tail -n 4 gpurun_out/soup50_lbvh.log gpurun_out/soup50_sah.log | cut -c1-1200
