#!/bin/bash
# BASELINE.json configs[4] across the GPUs of one box: every rank traces a contiguous 1/N slice of each ray batch against its own replica
# of the scene (no collective on the data path; max-over-ranks time).  tools/gpu_soup_scale.sh N
N=${1:-8}
set -x
nproc; free -g | head -2
for n in 8 4 2; do
  [ $n -le $N ] || continue
  RT_BUILD=lbvh timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2963$n tools/soup_sweep.py --sizes 1,50 --out gpurun_out/soup_scale_lbvh_n$n.json > gpurun_out/soup_scale_lbvh_n$n.log 2>&1
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2964$n tools/soup_sweep.py --sizes 1,5 --out gpurun_out/soup_scale_sah_n$n.json > gpurun_out/soup_scale_sah_n$n.log 2>&1
done
grep -h '^{"tris"' gpurun_out/soup_scale_*_n*.log | cut -c1-330
