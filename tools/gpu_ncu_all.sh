#!/bin/bash
# ncu captures of the current build (one GPU).  gpurun brings back at most 64 MiB, a full-set report is ~17 MB per launch:
#   bash tools/gpu_ncu_all.sh trace   -> launch list + one full capture each of k_extend and k_connect
#   bash tools/gpu_ncu_all.sh shade   -> full captures of the two shade kernels of one wavefront
set -x
R="python tools/profile_render.py plastic_dragon 0 1024 8 8 1"
timeout 200 $R > gpurun_out/prof_plain3.log 2>&1 || exit 1
if [ "$1" = "trace" ]; then
  timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches2.csv $R > gpurun_out/ncu_l2.log 2>&1
  timeout 500 ncu --set full --clock-control none -k regex:k_extend --launch-skip 1 -c 1 -o gpurun_out/extend_full2 -f $R > gpurun_out/ncu_e2.log 2>&1
  timeout 500 ncu --set full --clock-control none -k regex:k_connect --launch-skip 1 -c 1 -o gpurun_out/connect_full2 -f $R > gpurun_out/ncu_c2.log 2>&1
else
  timeout 600 ncu --set full --clock-control none -k regex:k_shade --launch-skip 2 -c 2 -o gpurun_out/shade_full2 -f $R > gpurun_out/ncu_s2.log 2>&1
fi
cat gpurun_out/prof_plain3.log; du -sh gpurun_out
