#!/bin/bash
# quick A/B on the GPU box: parity tests, then the 64-spp dragon render for every librtcore*.so variant present
set -x
(timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/ab_pytest.log 2>&1
for so in rustraytracer_b200/librtcore*.so; do
  tag=$(basename $so .so)
  RT_LIBRTCORE=$PWD/$so timeout 200 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 > gpurun_out/ab_$tag.log 2>&1
done
timeout 500 python bench.py --no-cpu-baseline > gpurun_out/ab_bench.json 2> gpurun_out/ab_bench.err
cat gpurun_out/ab_pytest.log; tail -n 1 gpurun_out/ab_librtcore*.log
python -c "
import json; d=json.load(open('gpurun_out/ab_bench.json')); print(d['value'], d['kernel_ms_per_step'], d['e2e']['value'])"
