#!/bin/bash
# round 2, GPU pass U: local-memory sectors by source line after the __grid_constant__ / register-RNG change: shade (matte/shape and plastic/triangle bins), connect, extend
set -x
R="python tools/profile_render.py plastic_dragon 0 1024 8 8 1"
O=rustraytracer_b200/csrc
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_shade --launch-skip 33 -c 3 -o /tmp/shade_full -f $R > gpurun_out/ncu_s.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_connect --launch-skip 16 -c 1 -o /tmp/connect_full -f $R > gpurun_out/ncu_c.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_extend --launch-skip 17 -c 1 -o /tmp/extend_full -f $R > gpurun_out/ncu_e.log 2>&1
python tools/ncu_mem_by_source.py /tmp/shade_full.ncu-rep "(int)0, (int)0" $O/shade_f32.o 30 > gpurun_out/shade_matte_mem_by_source_r02.txt 2>&1
python tools/ncu_mem_by_source.py /tmp/shade_full.ncu-rep "(int)2, (int)1" $O/shade_f32.o 30 > gpurun_out/shade_plastic_mem_by_source_r02.txt 2>&1
python tools/ncu_mem_by_source.py /tmp/connect_full.ncu-rep 0 $O/render.o 30 > gpurun_out/connect_mem_by_source_r02.txt 2>&1
python tools/ncu_mem_by_source.py /tmp/extend_full.ncu-rep 0 $O/render.o 30 > gpurun_out/extend_mem_by_source_r02.txt 2>&1
python tools/ncu_summary.py /tmp/shade_full.ncu-rep gpurun_out/shade_full_r02b_summary.csv
python tools/ncu_summary.py /tmp/connect_full.ncu-rep gpurun_out/connect_full_r02b_summary.csv
tail -35 gpurun_out/shade_matte_mem_by_source_r02.txt
