#!/bin/bash
# round 2, GPU pass A: parity suite, bench, microbenchmark peaks, prefilter audit + A/B variants, soup parity at 1 M / 10 M triangles
set -x
nvidia-smi -L; free -g | head -2; nproc
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -25) > gpurun_out/pytest_gpu.log 2>&1
cat gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 3000 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/peaks tools/peaks.cu && timeout 300 /tmp/peaks > gpurun_out/peaks_r02.json; cat gpurun_out/peaks_r02.json
for so in rustraytracer_b200/librtcore.so rustraytracer_b200/librtcore_nopre.so rustraytracer_b200/librtcore_mb6.so rustraytracer_b200/librtcore_mb8.so; do
  tag=$(basename $so .so)
  RT_LIBRTCORE=$PWD/$so timeout 200 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 > gpurun_out/ab_$tag.log 2>&1
  tail -n 1 gpurun_out/ab_$tag.log
done
# prefilter audit: every triangle slot test of these runs is re-decided by the exact test; violations must be 0
RT_LIBRTCORE=$PWD/rustraytracer_b200/librtcore_check.so timeout 600 python tools/soup_parity.py --sizes 1 --res 2048 --out gpurun_out/soup_parity_check.json > gpurun_out/soup_parity_check.log 2>&1; tail -3 gpurun_out/soup_parity_check.log
RT_LIBRTCORE=$PWD/rustraytracer_b200/librtcore_check.so timeout 300 python tools/prefilter_audit.py > gpurun_out/prefilter_audit.json 2> gpurun_out/prefilter_audit.err; cat gpurun_out/prefilter_audit.json; tail -3 gpurun_out/prefilter_audit.err
timeout 1200 python tools/soup_parity.py --sizes 1,10 --out gpurun_out/soup_parity_1_10.json > gpurun_out/soup_parity_1_10.log 2>&1; tail -4 gpurun_out/soup_parity_1_10.log
