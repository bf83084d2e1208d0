#!/bin/bash
# 2-GPU check of the sample-range partition + NCCL all-reduce, next to the 1-GPU line of the same build
set -x
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/scale_n2.json 2> gpurun_out/scale_n2.err
tail -n 3 gpurun_out/scale_n2.err
python - <<'PY'
import json
for f in ("gpurun_out/scale_n1.json", "gpurun_out/scale_n2.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["n_gpus"], d["value"], d["e2e"]["value"], d["kernel_ms_per_step"], d["ms_per_step"])
    except Exception as e: print(f, "FAILED", e)
PY
