#!/bin/bash
# round 2, GPU pass Z (final build): per-ray work constants of k_extend / k_connect, launch list of the bench command, whole GPU suite, default bench, smoke
set -x
M=$(python tools/extend_work.py --metrics 2>/dev/null || python -c "import sys; sys.path.insert(0,'tools'); import extend_work as e; print(e.METRICS)")
R="python tools/profile_render.py plastic_dragon 0 1024 8 8 1"
RT_TRACE_ITER=1 timeout 200 $R > gpurun_out/trace_iter.log 2>&1; tail -3 gpurun_out/trace_iter.log | cut -c1-200
timeout 900 ncu --metrics $M --clock-control none -k regex:k_extend --launch-skip 16 -c 16 --csv --page raw --log-file gpurun_out/extend_metrics.csv $R > gpurun_out/ncu_em.log 2>&1
timeout 900 ncu --metrics $M --clock-control none -k regex:k_connect --launch-skip 16 -c 16 --csv --page raw --log-file gpurun_out/connect_metrics.csv $R > gpurun_out/ncu_cm.log 2>&1
python tools/extend_work.py gpurun_out/extend_metrics.csv 17629982 k_extend gpurun_out/extend_work_r02.json
python tools/extend_work.py gpurun_out/connect_metrics.csv 8360284 k_connect gpurun_out/connect_work_r02.json
timeout 900 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_short.json 2>/dev/null
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_r02_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/bench_r02_n1.json 2> gpurun_out/bench_r02_n1.err; python -c "
import json; d=json.load(open('gpurun_out/bench_r02_n1.json')); print(d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d.get('exact_f64_shading',{}).get('value')); print(json.dumps(d['roofline'])[:700]); print(d['clocks'], d.get('gpu_launches'))"
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
