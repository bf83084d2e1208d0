#!/bin/bash
# round 2, GPU pass AA (2 GPUs, final build): the two-GPU tests, N=2 bench weak and strong
set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fused or render_multi or progressive" 2>&1 | tail -2
for mode in weak strong; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 2 --steps 3 --warmup 3 --scaling $mode --no-cpu-baseline > gpurun_out/bench_r02_n2_$mode.json 2> gpurun_out/bench_n2.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_r02_n2_$mode.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['scaling'], d['value'], d['e2e']['value'], d['ms_per_step'])"
done
