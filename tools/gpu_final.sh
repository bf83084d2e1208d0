#!/bin/bash
# final pass of a round on a 2-GPU box: 1-GPU and 2-GPU bench lines of the same build, then the ncu launch list of the bench command itself
set -x
timeout 600 python bench.py > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/scale_n2.json 2> gpurun_out/scale_n2.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
tail -n 3 gpurun_out/scale_n2.err
python - <<'PY'
import json
for f in ("gpurun_out/scale_n1.json", "gpurun_out/scale_n2.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["n_gpus"], d["value"], d["e2e"]["value"], d["kernel_ms_per_step"], d["ms_per_step"], d["roofline"]["traffic"], d["gpu_launches"])
    except Exception as e: print(f, "FAILED", e)
PY
