#!/bin/bash
# round 2, GPU pass I: full-set ncu captures of the RIGHT launches (a render here is 16 iterations: the second render's iteration 1 = launch 17 of
# k_extend / k_connect, launches 34-35 of k_shade), A/B of k_connect's launch bound
set -x
(timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/pytest_gpu.log 2>&1; cat gpurun_out/pytest_gpu.log
R="python tools/profile_render.py plastic_dragon 0 1024 8 8 1"
for so in rustraytracer_b200/librtcore.so rustraytracer_b200/librtcore_c4.so rustraytracer_b200/librtcore_t4.so; do
  tag=$(basename $so .so)
  RT_LIBRTCORE=$PWD/$so timeout 200 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 > gpurun_out/ab_$tag.log 2>&1; echo $tag; tail -n 1 gpurun_out/ab_$tag.log
done
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_extend --launch-skip 17 -c 1 -o gpurun_out/extend_full_r02 -f $R > gpurun_out/ncu_e.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_connect --launch-skip 16 -c 1 -o gpurun_out/connect_full_r02 -f $R > gpurun_out/ncu_c.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_shade --launch-skip 34 -c 2 -o gpurun_out/shade_full_r02 -f $R > gpurun_out/ncu_s.log 2>&1
ls -la gpurun_out/*.ncu-rep
