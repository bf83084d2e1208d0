#!/bin/bash
# round 2, GPU pass AD (final build): configs[4] parity at 1 / 10 / 50 M triangles on both trees (staged host -> device copies in use), soup sweep rates
set -x
timeout 2000 python tools/soup_parity.py --sizes 1,10,50 --out gpurun_out/soup_parity_r02.json 2>&1 | tail -2
timeout 900 python tools/soup_sweep.py --sizes 1,10,50 > gpurun_out/soup_sweep_final.log 2>&1; tail -5 gpurun_out/soup_sweep_final.log | cut -c1-400
