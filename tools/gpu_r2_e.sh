#!/bin/bash
# round 2, GPU pass E: Disney parity tests, full suite, bench
set -x
(timeout 900 python -m pytest tests -m gpu -x -q -k "disney" 2>&1 | tail -30) > gpurun_out/pytest_disney.log 2>&1; cat gpurun_out/pytest_disney.log
(timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -15) > gpurun_out/pytest_gpu.log 2>&1; cat gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_e.json 2> gpurun_out/bench_e.err; tail -c 1500 gpurun_out/bench_e.json
