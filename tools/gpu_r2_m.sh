#!/bin/bash
# round 2, GPU pass M: host SAH leaf cap A/B, bench of the current build
set -x
for l in 4 2 1; do
  echo "== SAH maxleaf $l"
  RT_SAH_MAXLEAF=$l RT_TRACE_BUILD=1 timeout 300 python tools/profile_render.py plastic_dragon 0 1024 64 8 1 2>&1 | grep -E "^bvh:|^f32" | tail -2
  RT_SAH_MAXLEAF=$l timeout 600 python tools/soup_sweep.py --sizes 1 2>&1 | python -c "
import sys, json
for line in sys.stdin:
    try: d = json.loads(line); print(d['tris'], round(d['coherent_Mrays_s']), round(d['incoherent_Mrays_s']), d['bvh4_nodes'], d['max_stack'])
    except Exception: pass"
done
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_m.json 2> gpurun_out/bench_m.err; python -c "
import json; d=json.load(open('gpurun_out/bench_m.json')); print(d['value'], d['e2e']['value'], d['kernel_ms_per_step']); print(d['exact_f64_shading']['value']); print(json.dumps(d['roofline'])[:900])"
