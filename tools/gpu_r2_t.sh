#!/bin/bash
# round 2, GPU pass T: A/B of (kernel parameters as __grid_constant__ + the RNG kept in registers) against the previous build, then parity + bench
set -x
for lib in librtcore_base.so librtcore.so; do
  for rep in 1 2; do
    echo "== $lib f32 dragon 64 spp"; RT_LIBRTCORE=$PWD/rustraytracer_b200/$lib timeout 300 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 2>&1 | tail -1
  done
  echo "== $lib f64 dragon 16 spp"; RT_LIBRTCORE=$PWD/rustraytracer_b200/$lib timeout 300 python tools/profile_render.py plastic_dragon 0 1024 16 8 0 2>&1 | tail -1
  echo "== $lib f32 cornell statue 64 spp"; RT_LIBRTCORE=$PWD/rustraytracer_b200/$lib timeout 300 python tools/profile_render.py cornell_box_statue 0 1024 64 8 1 2>&1 | tail -1
done
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/bench_t.json 2> gpurun_out/bench_t.err; python -c "
import json; d=json.load(open('gpurun_out/bench_t.json')); print(d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d.get('exact_f64_shading',{}).get('value'))"
