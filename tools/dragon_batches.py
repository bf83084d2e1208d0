"""The dragon-scene ray batches of the parity contract as a RATE through rt_closest_device (the render's Mrays/s is
flattered by an open scene — most bounce rays leave it): batch A = 1024^2 x k primary rays through pixel centres, batch B = one cosine-distributed
bounce ray per primary HIT (origin = hit point; primary misses are dropped, so every B ray starts on a surface).  Device-resident rays, CUDA events,
best of 5.  Prints one JSON line; --out writes it."""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import soup_parity  # noqa: E402
from rustraytracer_b200 import _native as N, scenes  # noqa: E402
from rustraytracer_b200._native import check, lib  # noqa: E402


def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--res", type=int, default=4096)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    torch.cuda.set_device(0); check(lib().rt_set_device(0))
    dev, stream = "cuda:0", C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rows = []
    for name, kw in (("plastic_dragon", {}), ("two_dragons", {"variant": 1}), ("cornell_box_statue", {})):
        hs = scenes.HostScene(name, args.res, args.res, **kw)
        for build in ("sah", "lbvh"):
            gpu = hs.upload(build=build)
            rays_h = hs.camera().primary_rays(args.res, args.res)
            n = len(rays_h)
            a = torch.from_numpy(rays_h.view(np.float64).reshape(n, 6)).to(dev)
            b_all = soup_parity.incoherent_batch(gpu, a, dev, stream)
            prim = torch.zeros(n, dtype=torch.int32, device=dev); t = torch.zeros(n, dtype=torch.float64, device=dev)
            call = lambda r, m: check(lib().rt_closest_device(gpu.h, m, C.c_void_p(r.data_ptr()), 0.001, 1e308, C.c_void_p(prim.data_ptr()), C.c_void_p(t.data_ptr()), stream))
            ms_a = timed(lambda: call(a, n))
            hit = prim != -1
            b = b_all[hit].contiguous(); nb = b.shape[0]
            ms_b = timed(lambda: call(b, nb))
            hit_b = float((prim[:nb] != -1).float().mean())
            rows.append({"scene": name, **kw, "tree": build, "batch_A_rays": n, "batch_A_Mrays_s": n / ms_a / 1e3, "batch_A_hit_rate": float(hit.float().mean()),
                         "batch_B_rays": nb, "batch_B_Mrays_s": nb / ms_b / 1e3, "batch_B_hit_rate": hit_b})
            print(json.dumps(rows[-1]), flush=True)
            del gpu, a, b_all, b, prim, t
            torch.cuda.empty_cache()
    res = {"workload": f"SURVEY 8d dragon-scene batches through rt_closest_device, {args.res}^2 pixel-centre primaries (A) and one cosine bounce per primary hit (B)", "rows": rows}
    if args.out:
        json.dump(res, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
