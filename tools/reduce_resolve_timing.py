"""rt_film_reduce_resolve_rgb8 (one kernel per GPU over its slice of the pixels: peer loads over NVLink + sum + resolve + peer store) against rt_film_allreduce (NCCL) + rt_film_resolve_rgb8.

    python tools/reduce_resolve_timing.py [--sizes 1024,2048] [--out profiles/reduce_resolve_r02.json]

One process, every visible GPU.  Each GPU holds a film of random running sums (an external film over a torch tensor); both ways end with the RGB8
image in host memory, both are blocking calls, timed by the host clock around the call after two warm-up rounds, best of 7.  The fused way is also
checked byte for byte against the one-GPU resolve of the numpy sum taken in the same order.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from rustraytracer_b200 import Comm, Film  # noqa: E402
from rustraytracer_b200._native import check, lib  # noqa: E402


def best(fn, reps=7, warm=2):
    for _ in range(warm): fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    return min(ts) * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1024,2048")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    n = lib().rt_device_count()
    rows = []
    for side in [int(x) for x in a.sizes.split(",")]:
        for world in sorted({k for k in (1, 2, 4, 8) if k <= n}):
            g = torch.Generator(); g.manual_seed(side * 31 + world)
            host = [torch.rand(side, side, 4, generator=g, dtype=torch.float64) * 40 for _ in range(world)]
            for hst in host: hst[..., 3] = 32.0
            tens, films = [], []
            for d in range(world):
                check(lib().rt_set_device(d))
                t = host[d].to(f"cuda:{d}"); torch.cuda.synchronize(d)
                tens.append(t); films.append(Film(side, side, device_ptr=t.data_ptr()))
            check(lib().rt_set_device(0))
            fused = Film.reduce_resolve_rgb8(films)
            total = host[0].numpy().copy()
            for hst in host[1:]: total = total + hst.numpy()
            tt = torch.from_numpy(total).to("cuda:0"); torch.cuda.synchronize(0)
            ext = Film(side, side, device_ptr=tt.data_ptr())
            exact = bool(np.array_equal(fused, ext.resolve_rgb8()))
            ext.close()
            ms_fused = best(lambda: Film.reduce_resolve_rgb8(films))
            row = {"film": f"{side}x{side}", "gpus": world, "bytes_per_film": side * side * 32, "fused_ms": round(ms_fused, 3), "fused_equals_ordered_sum_resolve": exact}
            if world > 1:
                comm = Comm.local(list(range(world)))
                def two_step():
                    comm.allreduce(films); films[0].resolve_rgb8()
                row["allreduce_plus_resolve_ms"] = round(best(two_step), 3)
                row["allreduce_ms"] = round(best(lambda: comm.allreduce(films)), 3)
                comm.close()
            row["resolve_one_film_ms"] = round(best(lambda: films[0].resolve_rgb8()), 3)
            for f in films: f.close()
            del tens
            print(json.dumps(row), flush=True)
            rows.append(row)
    res = {"what": "N films -> one RGB8 image in host memory: fused peer-load kernel (every GPU takes a slice of the pixels) vs ncclAllReduce(f64 sum, in place on all N) + resolve on GPU 0; host clock, best of 7",
           "gpus_visible": n, "rows": rows, "ok": all(r["fused_equals_ordered_sum_resolve"] for r in rows)}
    if a.out: json.dump(res, open(a.out, "w"), indent=1)
    sys.exit(0 if res["ok"] else 1)


if __name__ == "__main__":
    main()
