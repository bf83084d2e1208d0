#!/bin/bash
# round 2, GPU pass AK/AN: A/B against the previous build (f64 and f32), f64 films must be unchanged, suite, bench
set -x
for lib in librtcore_base.so librtcore.so; do
  for sc in "plastic_dragon 0 1024 64 8 1" "plastic_dragon 0 1024 16 8 0" "cornell_box_statue 0 1024 16 8 0"; do
    for rep in 1 2; do echo "== $lib $sc"; RT_LIBRTCORE=$PWD/rustraytracer_b200/$lib timeout 300 python tools/profile_render.py $sc 2>&1 | tail -1; done
  done
done
for sc in "plastic_dragon 0 128 8 8 0" "sphere_roughness 0 128 8 8 0" "plastic_dragon 0 256 16 8 1"; do
  RT_LIBRTCORE=$PWD/rustraytracer_b200/librtcore_base.so timeout 300 python tools/film_dump.py $sc /tmp/a.npy
  timeout 300 python tools/film_dump.py $sc /tmp/b.npy
  python -c "
import numpy as np
a, b = np.load('/tmp/a.npy'), np.load('/tmp/b.npy')
d = np.abs(a - b)[..., :3] / np.maximum(np.abs(a[..., :3]), 1e-6)
print('FILM $sc: max rel diff', float(d.max()))"
done
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/bench_ak.json 2> gpurun_out/bench_ak.err; python -c "
import json; d=json.load(open('gpurun_out/bench_ak.json')); print(d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d.get('exact_f64_shading',{}).get('value'))"
