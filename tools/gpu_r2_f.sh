#!/bin/bash
# round 2, GPU pass F: the allocation-free resolve entry point, alone in a fresh process and after the whole suite
set -x
(timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -15) > gpurun_out/pytest_zz_alone.log 2>&1; cat gpurun_out/pytest_zz_alone.log
(timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/pytest_gpu.log 2>&1; cat gpurun_out/pytest_gpu.log
