#!/bin/bash
# round 2, GPU pass P (2 GPUs): the fused reduce + resolve over NVLink, the multi-GPU tests, timing against all-reduce + resolve
set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fused or render_multi or progressive" 2>&1 | tail -5
timeout 600 python tools/reduce_resolve_timing.py --out gpurun_out/reduce_resolve_r02.json 2>&1 | tail -6
