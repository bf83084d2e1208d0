#!/bin/bash
# round 2, GPU pass H (gpurun --gpus 8): multi-GPU behind the ABI — rt_render_multi test, configs[3] on 8 ranks, strong scaling 1/2/4/8, weak N=2 through the ABI reduce
set -x
nvidia-smi -L | head -8; nproc
(timeout 600 python -m pytest tests -m gpu -x -q -k "render_multi or no_alloc" 2>&1 | tail -8) > gpurun_out/pytest_multi.log 2>&1; cat gpurun_out/pytest_multi.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 900 $TR --nproc-per-node 8 --master-port 29511 tools/config4_multi.py --out gpurun_out/config4_8gpu_r02.json > gpurun_out/config4.log 2>&1; tail -3 gpurun_out/config4.log | cut -c1-1500
timeout 400 python bench.py --scaling strong --no-cpu-baseline --no-f64-leg > gpurun_out/strong_n1.json 2> gpurun_out/strong_n1.err
for n in 2 4 8; do
  timeout 400 $TR --nproc-per-node $n --master-port $((29520 + n)) bench.py --gpus $n --scaling strong --no-cpu-baseline --no-f64-leg > gpurun_out/strong_n$n.json 2> gpurun_out/strong_n$n.err
done
timeout 400 $TR --nproc-per-node 2 --master-port 29531 bench.py --gpus 2 --no-cpu-baseline --no-f64-leg > gpurun_out/weak_n2.json 2> gpurun_out/weak_n2.err
timeout 400 $TR --nproc-per-node 8 --master-port 29532 bench.py --gpus 8 --no-cpu-baseline --no-f64-leg > gpurun_out/weak_n8.json 2> gpurun_out/weak_n8.err
for f in strong_n1 strong_n2 strong_n4 strong_n8 weak_n2 weak_n8; do python -c "
import json,sys
try:
    d=json.loads(open('gpurun_out/$f.json').read().strip().splitlines()[-1]); print('$f', d['n_gpus'], round(d['value'],1), round(d['e2e']['value'],1), round(d['ms_per_step'],2), d['kernel_ms_per_step'])
except Exception as e: print('$f FAILED', e); print(open('gpurun_out/$f.err').read()[-1500:])
"; done
