#!/bin/bash
# round 2, GPU pass AO (final build): whole GPU suite, default bench, smoke
set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/bench_r02_n1.json 2> gpurun_out/bench_r02_n1.err; python -c "
import json; d=json.load(open('gpurun_out/bench_r02_n1.json')); print(d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d.get('exact_f64_shading',{}).get('value'), d['clocks'])"
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
