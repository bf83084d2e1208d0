#!/bin/bash
# scaling run on an N-GPU box: bench.py at 1, 2, 4, ... N ranks of the same build (tools/gpu_scale.sh N)
N=${1:-4}
timeout 400 python bench.py --no-cpu-baseline > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
for n in 2 4 8; do
  [ $n -le $N ] || continue
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n bench.py --gpus $n --steps 3 --warmup 3 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/scale_n*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["n_gpus"], round(d["value"], 1), round(d["e2e"]["value"], 1), d["ms_per_step"], d["clocks"])
    except Exception as e: print(f, "FAILED", e)
PY
