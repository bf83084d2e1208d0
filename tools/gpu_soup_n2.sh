#!/bin/bash
# the coherent-batch checksum summed over 2 ranks must equal the 1-GPU checksum (same rays, same hits, sliced)
set -x
timeout 300 python tools/soup_sweep.py --sizes 1 > gpurun_out/soup_chk_n1.log 2>&1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655 tools/soup_sweep.py --sizes 1 > gpurun_out/soup_chk_n2.log 2>&1
python - <<'PY'
import json
r = [json.loads(l) for f in ("gpurun_out/soup_chk_n1.log", "gpurun_out/soup_chk_n2.log") for l in open(f) if l.startswith('{"tris"')]
for x in r: print(x["n_gpus"], x["id_checksum_coherent"], x["hit_rate_coherent"], round(x["coherent_Mrays_s"]), round(x["incoherent_Mrays_s"]))
print("coherent checksums agree:", r[0]["id_checksum_coherent"] == r[1]["id_checksum_coherent"])
PY
