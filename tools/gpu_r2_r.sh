#!/bin/bash
# round 2, GPU pass R (2 GPUs): the whole GPU suite where the two-GPU tests run, then the driver's N=2 launches of both bench arms
set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_r02_n2.json 2> gpurun_out/bench_r02_n2.err; tail -2 gpurun_out/bench_r02_n2.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_r02_n2.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['value'], d['e2e'], d['scaling'], d.get('gpu_launches'))"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>/dev/null | tail -1 | cut -c1-600
