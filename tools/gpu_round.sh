#!/bin/bash
# One GPU-box pass: parity tests, bench (both arms) and the per-kernel timing of one 8-spp render.  ncu captures: tools/gpu_ncu_all.sh
nvidia-smi -L
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/pytest_gpu.log 2>&1
timeout 500 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>&1
timeout 200 python tools/profile_render.py plastic_dragon 0 1024 8 8 1 > gpurun_out/prof_plain.log 2>&1
cat gpurun_out/pytest_gpu.log gpurun_out/prof_plain.log
tail -c 2500 gpurun_out/bench.json
