#!/bin/bash
for w in 2097152 4194304 8388608 16777216 33554432; do
  echo "W=$w"; RT_WAVEFRONT=$w timeout 300 python tools/profile_render.py plastic_dragon 0 1024 128 8 1 2>&1 | tail -n 1
done > gpurun_out/wave.log 2>&1
cat gpurun_out/wave.log
