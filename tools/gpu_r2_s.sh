#!/bin/bash
# round 2, GPU pass S: per-source-line memory tables (requests / sectors) of the two shade kernels and of k_extend — which gathers are the uncoalesced ones
set -x
R="python tools/profile_render.py plastic_dragon 0 1024 8 8 1"
O=rustraytracer_b200/csrc
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_shade --launch-skip 34 -c 2 -o /tmp/shade_full -f $R > gpurun_out/ncu_s.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_extend --launch-skip 17 -c 1 -o /tmp/extend_full -f $R > gpurun_out/ncu_e.log 2>&1
python tools/ncu_mem_by_source.py /tmp/shade_full.ncu-rep 0 $O/shade_f32.o 25 > gpurun_out/shade_matte_mem_by_source_r02.txt 2>&1
python tools/ncu_mem_by_source.py /tmp/shade_full.ncu-rep 1 $O/shade_f32.o 25 > gpurun_out/shade_plastic_mem_by_source_r02.txt 2>&1
python tools/ncu_mem_by_source.py /tmp/extend_full.ncu-rep 0 $O/render.o 25 > gpurun_out/extend_mem_by_source_r02.txt 2>&1
python tools/ncu_by_source.py /tmp/shade_full.ncu-rep 0 $O/shade_f32.o 40 > gpurun_out/shade_matte_by_source_r02.txt 2>&1
python tools/ncu_by_source.py /tmp/shade_full.ncu-rep 1 $O/shade_f32.o 40 > gpurun_out/shade_plastic_by_source_r02.txt 2>&1
head -50 gpurun_out/shade_matte_mem_by_source_r02.txt
