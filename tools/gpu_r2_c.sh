#!/bin/bash
# round 2, GPU pass C: direct shapes on/off A/B, full parity suite, bench, soup parity at 50 M triangles
set -x
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/pytest_gpu.log 2>&1; cat gpurun_out/pytest_gpu.log
for d in 1 0; do
  RT_DIRECT=$d timeout 200 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 > gpurun_out/ab_direct$d.log 2>&1; tail -n 1 gpurun_out/ab_direct$d.log
  RT_DIRECT=$d timeout 200 python tools/profile_render.py cornell_box_statue 0 1024 16 25 1 > gpurun_out/ab_statue_direct$d.log 2>&1; tail -n 1 gpurun_out/ab_statue_direct$d.log
  RT_DIRECT=$d timeout 200 python tools/profile_render.py two_dragons 1 1024 16 16 1 > gpurun_out/ab_two_direct$d.log 2>&1; tail -n 1 gpurun_out/ab_two_direct$d.log
done
RT_LIBRTCORE=$PWD/rustraytracer_b200/librtcore_r01.so timeout 200 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 | tail -n 1
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_d.json 2> gpurun_out/bench_d.err; tail -c 1500 gpurun_out/bench_d.json
grep -q "passed" gpurun_out/pytest_gpu.log && ! grep -q "failed" gpurun_out/pytest_gpu.log && (timeout 1500 python tools/soup_parity.py --sizes 50 --out gpurun_out/soup_parity_50.json > gpurun_out/soup_parity_50.log 2>&1; tail -3 gpurun_out/soup_parity_50.log)
