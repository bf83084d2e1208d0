#!/bin/bash
# round 2, GPU pass J: full-set ncu captures (second render, iteration 1), reduced ON THE BOX to the csv summaries / per-source-line tables (the three
# reports together exceed gpurun's 64 MiB return limit); only the shade report travels back
set -x
R="python tools/profile_render.py plastic_dragon 0 1024 8 8 1"
O=rustraytracer_b200/csrc
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_extend --launch-skip 17 -c 1 -o /tmp/extend_full_r02 -f $R > gpurun_out/ncu_e.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_connect --launch-skip 16 -c 1 -o /tmp/connect_full_r02 -f $R > gpurun_out/ncu_c.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_shade --launch-skip 34 -c 2 -o gpurun_out/shade_full_r02 -f $R > gpurun_out/ncu_s.log 2>&1
for k in extend connect; do
  python tools/ncu_summary.py /tmp/${k}_full_r02.ncu-rep gpurun_out/${k}_full_r02_summary.csv
  python tools/ncu_by_source.py /tmp/${k}_full_r02.ncu-rep 0 $O/render.o 60 > gpurun_out/${k}_by_source_r02.txt 2>&1
  python tools/ncu_local_mem.py /tmp/${k}_full_r02.ncu-rep 0 $O/render.o 30 > gpurun_out/${k}_local_mem_r02.txt 2>&1
done
python tools/ncu_summary.py gpurun_out/shade_full_r02.ncu-rep gpurun_out/shade_full_r02_summary.csv
python tools/ncu_by_source.py gpurun_out/shade_full_r02.ncu-rep 1 $O/shade_f32.o 60 > gpurun_out/shade_by_source_r02.txt 2>&1
python tools/ncu_local_mem.py gpurun_out/shade_full_r02.ncu-rep 1 $O/shade_f32.o 30 > gpurun_out/shade_local_mem_r02.txt 2>&1
du -sh gpurun_out; head -30 gpurun_out/connect_local_mem_r02.txt
