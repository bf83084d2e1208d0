#!/bin/bash
# round 2, GPU pass AM: what runs under shade.cuh:685 (19 % of the f64 matte kernel's instructions)?
set -x
R="python tools/profile_render.py plastic_dragon 0 1024 8 8 0"
O=rustraytracer_b200/csrc
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_shade --launch-skip 33 -c 3 -o /tmp/shade64 -f $R > gpurun_out/ncu_s64.log 2>&1
python tools/ncu_line_ops.py /tmp/shade64.ncu-rep "(int)0, (int)0" $O/render.o shade.cuh 685
python tools/ncu_line_ops.py /tmp/shade64.ncu-rep "(int)0, (int)0" $O/render.o geom.cuh 30
python tools/ncu_line_ops.py /tmp/shade64.ncu-rep "(int)0, (int)0" $O/render.o wavefront.cuh 289
