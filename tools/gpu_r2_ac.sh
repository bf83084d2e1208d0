#!/bin/bash
# round 2, GPU pass AC: resident shade blocks per SM re-measured after the local-memory fixes (3 / 4 / 5 = 168 / 128 / 96 registers); new partial-lobe parity test; whole suite
set -x
for lib in librtcore.so librtcore_b3.so librtcore_b5.so; do
  for sc in "plastic_dragon 0 1024 64 8 1" "cornell_box_statue 0 1024 64 8 1"; do
    for rep in 1 2; do echo "== $lib $sc"; RT_LIBRTCORE=$PWD/rustraytracer_b200/$lib timeout 300 python tools/profile_render.py $sc 2>&1 | tail -1; done
  done
done
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
