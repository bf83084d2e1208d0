#!/bin/bash
# Full ncu captures of the shade and shadow-connect kernels (one GPU; run only after the plain program exited 0).
set -x
timeout 200 python tools/profile_render.py plastic_dragon 0 1024 8 8 1 > gpurun_out/prof_plain2.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_shade --launch-skip 2 -c 4 -o gpurun_out/shade_full -f python tools/profile_render.py plastic_dragon 0 1024 8 8 1 > gpurun_out/ncu_s.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_connect_shadow --launch-skip 1 -c 2 -o gpurun_out/shadow_full -f python tools/profile_render.py plastic_dragon 0 1024 8 8 1 > gpurun_out/ncu_c.log 2>&1
RT_TRACE_ITER=1 timeout 200 python tools/profile_render.py plastic_dragon 0 1024 16 8 1 > gpurun_out/trace_iter.log 2>&1
tail -n 3 gpurun_out/ncu_s.log gpurun_out/ncu_c.log
