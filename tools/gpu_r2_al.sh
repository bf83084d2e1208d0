#!/bin/bash
# round 2, GPU pass AL: per-source-line tables of the two f64 shade kernels (parity mode) on the final build
set -x
R="python tools/profile_render.py plastic_dragon 0 1024 8 8 0"
O=rustraytracer_b200/csrc
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_shade --launch-skip 33 -c 3 -o /tmp/shade64 -f $R > gpurun_out/ncu_s64.log 2>&1
python tools/ncu_by_source.py /tmp/shade64.ncu-rep "(int)0, (int)0" $O/render.o 45 > gpurun_out/shade64_matte_by_source_r02.txt 2>&1
python tools/ncu_by_source.py /tmp/shade64.ncu-rep "(int)2, (int)1" $O/render.o 45 > gpurun_out/shade64_plastic_by_source_r02.txt 2>&1
python tools/ncu_mem_by_source.py /tmp/shade64.ncu-rep "(int)0, (int)0" $O/render.o 20 > gpurun_out/shade64_matte_mem_by_source_r02.txt 2>&1
python tools/ncu_summary.py /tmp/shade64.ncu-rep gpurun_out/shade64_full_r02_summary.csv
