#!/bin/bash
# round 2, GPU pass X: shade inlining level 3 (record / texture / area-sample functions inlined) re-measured now that the BSDF is in registers; fresh per-line local-memory tables
set -x
for lib in librtcore.so librtcore_l3.so; do
  for sc in "plastic_dragon 0 1024 64 8 1" "cornell_box_statue 0 1024 64 8 1" "two_dragons 0 1024 32 16 1"; do
    for rep in 1 2; do echo "== $lib $sc"; RT_LIBRTCORE=$PWD/rustraytracer_b200/$lib timeout 300 python tools/profile_render.py $sc 2>&1 | tail -1; done
  done
done
R="python tools/profile_render.py plastic_dragon 0 1024 8 8 1"
O=rustraytracer_b200/csrc
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_shade --launch-skip 33 -c 3 -o /tmp/shade_full -f $R > gpurun_out/ncu_s.log 2>&1
python tools/ncu_mem_by_source.py /tmp/shade_full.ncu-rep "(int)0, (int)0" $O/shade_f32.o 30 > gpurun_out/shade_matte_mem_by_source_r02c.txt 2>&1
python tools/ncu_mem_by_source.py /tmp/shade_full.ncu-rep "(int)2, (int)1" $O/shade_f32.o 30 > gpurun_out/shade_plastic_mem_by_source_r02c.txt 2>&1
python tools/ncu_by_source.py /tmp/shade_full.ncu-rep 0 $O/shade_f32.o 40 > gpurun_out/shade_by_source_r02c.txt 2>&1
python tools/ncu_summary.py /tmp/shade_full.ncu-rep gpurun_out/shade_full_r02c_summary.csv
