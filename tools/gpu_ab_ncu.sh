#!/bin/bash
# A/B of every librtcore*.so variant (64-spp dragon render) + one full ncu capture of k_extend for the default build
set -x
(timeout 600 python -m pytest tests -m gpu -x -q -k "soup or dragon or kat or device_resident" 2>&1 | tail -5) > gpurun_out/ab_pytest.log 2>&1
for so in rustraytracer_b200/librtcore*.so; do
  tag=$(basename $so .so)
  RT_LIBRTCORE=$PWD/$so timeout 200 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 > gpurun_out/ab_$tag.log 2>&1
done
R="python tools/profile_render.py plastic_dragon 0 1024 8 8 1"
[ "$1" = "ncu" ] && timeout 500 ncu --set full --clock-control none -k regex:k_extend --launch-skip 1 -c 1 -o gpurun_out/extend_ab -f $R > gpurun_out/ncu_e_ab.log 2>&1
cat gpurun_out/ab_pytest.log; tail -n 1 gpurun_out/ab_librtcore*.log
