#!/bin/bash
# round 2, GPU pass Y: static lobe slots in the f64 kernels of Matte .. Mirror: timing (previous build, pass W: dragon f64 16 spp shade 16.2 ms, 1,806 Mrays/s), parity suite, bench
set -x
for sc in "plastic_dragon 0 1024 16 8 0" "cornell_box_statue 0 1024 16 8 0" "two_dragons 0 1024 16 16 0" "plastic_dragon 0 1024 64 8 1"; do
  for rep in 1 2; do echo "== $sc"; timeout 300 python tools/profile_render.py $sc 2>&1 | tail -1; done
done
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/bench_y.json 2> gpurun_out/bench_y.err; python -c "
import json; d=json.load(open('gpurun_out/bench_y.json')); print(d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d.get('exact_f64_shading',{}).get('value'))"
