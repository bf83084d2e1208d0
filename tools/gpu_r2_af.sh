#!/bin/bash
# round 2, GPU pass AF: branch-free BVH4 node step (child select by masks, pushes to a spare slot, FFMA pads) against the previous build
set -x
for lib in librtcore_base.so librtcore.so; do
  for sc in "plastic_dragon 0 1024 64 8 1" "cornell_box_statue 0 1024 64 8 1" "two_dragons 0 1024 32 16 1"; do
    for rep in 1 2; do echo "== $lib $sc"; RT_LIBRTCORE=$PWD/rustraytracer_b200/$lib timeout 300 python tools/profile_render.py $sc 2>&1 | tail -1; done
  done
  echo "== $lib soup"; RT_LIBRTCORE=$PWD/rustraytracer_b200/$lib timeout 600 python tools/soup_sweep.py --sizes 1,10 2>&1 | python -c "
import sys, json
for line in sys.stdin:
    try: d = json.loads(line); print(d['tris'], round(d['coherent_Mrays_s']), round(d['incoherent_Mrays_s']), d['id_checksum_coherent'], d['id_checksum_incoherent'])
    except Exception: pass"
done
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/bench_af.json 2> gpurun_out/bench_af.err; python -c "
import json; d=json.load(open('gpurun_out/bench_af.json')); print(d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d.get('exact_f64_shading',{}).get('value'))"
