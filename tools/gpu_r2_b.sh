#!/bin/bash
# round 2, GPU pass B: traversal variants A/B (64-spp dragon), ncu full capture of k_extend / k_connect of the default build, the fixed overflow test
set -x
(timeout 600 python -m pytest tests -m gpu -x -q -k "overflow or progressive or soup_parity" 2>&1 | tail -8) > gpurun_out/pytest_b.log 2>&1; cat gpurun_out/pytest_b.log
for so in rustraytracer_b200/librtcore_r01.so rustraytracer_b200/librtcore.so rustraytracer_b200/librtcore_od.so rustraytracer_b200/librtcore_odinl.so rustraytracer_b200/librtcore_pf.so rustraytracer_b200/librtcore_odpf.so; do
  tag=$(basename $so .so)
  RT_LIBRTCORE=$PWD/$so timeout 200 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 > gpurun_out/ab_$tag.log 2>&1
  tail -n 1 gpurun_out/ab_$tag.log
done
R="python tools/profile_render.py plastic_dragon 0 1024 8 8 1"
RT_TRACE_ITER=1 timeout 200 $R > gpurun_out/trace_iter.log 2>&1
timeout 500 ncu --set full --import-source on --clock-control none -k regex:k_extend --launch-skip 1 -c 1 -o gpurun_out/extend_full_r02a -f $R > gpurun_out/ncu_e.log 2>&1
timeout 500 ncu --set full --import-source on --clock-control none -k regex:k_connect --launch-skip 1 -c 1 -o gpurun_out/connect_full_r02a -f $R > gpurun_out/ncu_c.log 2>&1
ls -la gpurun_out/*.ncu-rep; head -12 gpurun_out/trace_iter.log
