#!/bin/bash
# round 2, GPU pass AB (8 GPUs, final build): bench at N = 4 and 8, weak and strong
set -x
for n in 8; do for mode in weak strong; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2955$n bench.py --gpus $n --steps 3 --warmup 3 --scaling $mode --no-cpu-baseline > gpurun_out/bench_r02_n${n}_$mode.json 2> gpurun_out/bench_n$n.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_r02_n${n}_$mode.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['scaling'], d['value'], d['e2e']['value'], d['ms_per_step'])"
done; done
