#!/bin/bash
# quick A/B on the GPU box: parity tests, then the 8-spp and 64-spp dragon render with and without the L2 window
set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5) > gpurun_out/ab_pytest.log 2>&1
for p in 1 0; do
  RT_L2_PERSIST=$p timeout 200 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 > gpurun_out/ab_persist$p.log 2>&1
done
timeout 500 python bench.py --no-cpu-baseline > gpurun_out/ab_bench.json 2> gpurun_out/ab_bench.err
cat gpurun_out/ab_pytest.log gpurun_out/ab_persist1.log gpurun_out/ab_persist0.log
python -c "
import json; d=json.load(open('gpurun_out/ab_bench.json')); print(d['value'], d['kernel_ms_per_step'], d['e2e']['value'])"
