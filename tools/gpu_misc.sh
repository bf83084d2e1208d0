#!/bin/bash
set -x
(timeout 600 python -m pytest tests -m gpu -x -q -k "device_resident" 2>&1 | tail -12) > gpurun_out/misc_pytest.log 2>&1
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/misc_bench.json 2> gpurun_out/misc_bench.err
cat gpurun_out/misc_pytest.log | cut -c1-600
python -c "
import json; d=json.load(open('gpurun_out/misc_bench.json')); print(d['value'], d['kernel_ms_per_step'], d['e2e']['value'], d['config']['wavefront_paths'])"
tail -n 3 gpurun_out/misc_bench.err
