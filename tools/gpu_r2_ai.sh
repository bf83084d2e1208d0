#!/bin/bash
# round 2, GPU pass AI: per-source-line samples / instructions of the two f32 shade kernels on the current build (where do the f64 divisions and square roots still run?)
set -x
R="python tools/profile_render.py plastic_dragon 0 1024 8 8 1"
O=rustraytracer_b200/csrc
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_shade --launch-skip 33 -c 3 -o /tmp/shade_full -f $R > gpurun_out/ncu_s.log 2>&1
python tools/ncu_by_source.py /tmp/shade_full.ncu-rep "(int)0, (int)0" $O/shade_f32.o 45 > gpurun_out/shade_matte_by_source_r02d.txt 2>&1
python tools/ncu_by_source.py /tmp/shade_full.ncu-rep "(int)2, (int)1" $O/shade_f32.o 45 > gpurun_out/shade_plastic_by_source_r02d.txt 2>&1
python tools/ncu_summary.py /tmp/shade_full.ncu-rep gpurun_out/shade_full_r02d_summary.csv
