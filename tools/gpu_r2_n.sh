#!/bin/bash
# round 2, GPU pass N: full GPU suite, ingest timings, default bench, soup parity at 1 / 10 / 50 M with the current device-tree defaults, smoke
set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
timeout 900 python tools/ingest_timing.py --sizes 10,50 --out gpurun_out/ingest_r02.json 2>&1 | tail -8
timeout 900 python bench.py > gpurun_out/bench_r02_n1.json 2> gpurun_out/bench_r02_n1.err; tail -3 gpurun_out/bench_r02_n1.err
python -c "
import json; d=json.load(open('gpurun_out/bench_r02_n1.json')); print(d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d.get('exact_f64_shading',{}).get('value')); print(json.dumps(d['roofline'])[:600]); print(d['clocks'])"
timeout 1500 python tools/soup_parity.py --sizes 1,10,50 --out gpurun_out/soup_parity_r02.json 2>&1 | tail -5
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
