"""BASELINE.json configs[4]: synthetic triangle-soup ray-cast sweep, coherent primary vs incoherent diffuse rays.

    python tools/soup_sweep.py [--sizes 1,2,5,10] [--res 4096] [--out profiles/soup_sweep_r01.json]

For every triangle count N (in millions): builds the soup of SURVEY.md §8d (Philox-generated, f32-exact vertices), commits
it (host binned-SAH build + upload, or the device LBVH with RT_BUILD=lbvh; timed), then times `rt_closest_device` (closest-hit ids + t, 16 B out per ray) on
  * the coherent batch: res x res pixel-centre rays of a pinhole camera at (0,0,-4), vfov 40;
  * the incoherent batch: one cosine-distributed bounce ray per coherent hit (origin = hit point), uniform-sphere rays
    from U[-1,1]^3 origins where the primary missed.
Rays and results stay device-resident (torch tensors); times are CUDA events around the launch on torch's current
stream, best of 3 after a warm-up.  Multi-GPU (torchrun): every rank traces a contiguous 1/G slice of each batch against
its own replica of the scene; the printed rate is total rays / max-over-ranks time.
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from rustraytracer_b200 import _native as N, scenes  # noqa: E402
from rustraytracer_b200._native import check, lib  # noqa: E402


def timed(fn, reps=3):
    fn(); torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1,2,5,10")
    ap.add_argument("--res", type=int, default=4096)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    check(lib().rt_set_device(local))
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = f"cuda:{local}"
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    res, rows = args.res, []
    for m in [float(x) for x in args.sizes.split(",")]:
        n_tris = int(m * 1e6)
        t0 = time.time()
        hs = scenes.HostScene("soup", res, res, n_tris=n_tris)
        gen_s = time.time() - t0
        t0 = time.time()
        gpu = hs.upload()
        build_s = time.time() - t0
        cam = hs.camera()
        rays_h = cam.primary_rays(res, res)
        n_all = len(rays_h)
        lo, hi = rank * n_all // world, (rank + 1) * n_all // world
        n = hi - lo
        rays = torch.from_numpy(rays_h[lo:hi].view(np.float64).reshape(n, 6)).to(dev)
        hits = torch.zeros(n * N.hit_dtype.itemsize, dtype=torch.uint8, device=dev)
        prim = torch.zeros(n, dtype=torch.int32, device=dev)
        t = torch.zeros(n, dtype=torch.float64, device=dev)
        check(lib().rt_intersect_device(gpu.h, n, C.c_void_p(rays.data_ptr()), 0.001, 1e308, C.c_void_p(hits.data_ptr()), stream))
        torch.cuda.synchronize()
        h = hits.view(torch.float64).reshape(n, 13)           # rt_hit: t, p[3], n[3], ns[3], uv[2], (prim, front)
        hit = h[:, 0] > 0
        g = torch.Generator(device=dev); g.manual_seed(0xD1FF + rank)
        nrm = h[:, 4:7]
        u1, u2 = torch.rand(n, generator=g, device=dev, dtype=torch.float64), torch.rand(n, generator=g, device=dev, dtype=torch.float64)
        r, phi = torch.sqrt(u1), 2 * np.pi * u2
        a = torch.where(nrm[:, 0:1].abs() > 0.9, torch.tensor([0., 1., 0.], device=dev, dtype=torch.float64), torch.tensor([1., 0., 0.], device=dev, dtype=torch.float64))
        tx = torch.linalg.cross(nrm, a.expand_as(nrm)); tx = tx / tx.norm(dim=1, keepdim=True).clamp_min(1e-30)
        ty = torch.linalg.cross(nrm, tx)
        d_hit = tx * (r * torch.cos(phi))[:, None] + ty * (r * torch.sin(phi))[:, None] + nrm * torch.sqrt((1 - u1).clamp_min(0))[:, None]
        d_miss = torch.randn(n, 3, generator=g, device=dev, dtype=torch.float64); d_miss = d_miss / d_miss.norm(dim=1, keepdim=True)
        o_miss = torch.rand(n, 3, generator=g, device=dev, dtype=torch.float64) * 2 - 1
        inc = torch.cat([torch.where(hit[:, None], h[:, 1:4], o_miss), torch.where(hit[:, None], d_hit, d_miss)], dim=1).contiguous()
        del hits, h
        ms_c = timed(lambda: check(lib().rt_closest_device(gpu.h, n, C.c_void_p(rays.data_ptr()), 0.001, 1e308, C.c_void_p(prim.data_ptr()), C.c_void_p(t.data_ptr()), stream)))
        hit_c = float((prim != -1).float().mean())
        # order-sensitive checksum of the closest-hit ids: equal between the host SAH tree and the device LBVH (and between 1 and N GPUs'
        # slices summed) iff both trees return the same primitive for every ray — the size-independent parity property at full size
        wgt = torch.arange(lo, hi, device=dev, dtype=torch.int64) % 65521 + 1
        chk_c = int(((prim.to(torch.int64) + 1) * wgt).sum())
        ms_i = timed(lambda: check(lib().rt_closest_device(gpu.h, n, C.c_void_p(inc.data_ptr()), 0.001, 1e308, C.c_void_p(prim.data_ptr()), C.c_void_p(t.data_ptr()), stream)))
        hit_i = float((prim != -1).float().mean())
        chk_i = int(((prim.to(torch.int64) + 1) * wgt).sum())
        if dist:
            tt = torch.tensor([ms_c, ms_i], device=dev, dtype=torch.float64); dist.all_reduce(tt, op=dist.ReduceOp.MAX); ms_c, ms_i = tt.tolist()
            # whole-batch checksum / hit rates: the ranks' slices summed (the coherent checksum then equals the 1-GPU one; the incoherent batch
            # is drawn per rank, so its checksum is only comparable at equal N)
            ck = torch.tensor([chk_c, chk_i], device=dev, dtype=torch.int64); dist.all_reduce(ck, op=dist.ReduceOp.SUM); chk_c, chk_i = [int(x) for x in ck.tolist()]
            hr = torch.tensor([hit_c * n, hit_i * n], device=dev, dtype=torch.float64); dist.all_reduce(hr, op=dist.ReduceOp.SUM); hit_c, hit_i = [x / n_all for x in hr.tolist()]
        b = gpu.device_bytes()
        bs = gpu.build_stats()
        row = {"tris": n_tris, "rays": n_all, "n_gpus": world, "coherent_Mrays_s": n_all / ms_c / 1e3, "incoherent_Mrays_s": n_all / ms_i / 1e3, "coherent_ms": ms_c, "incoherent_ms": ms_i,
               "hit_rate_coherent": hit_c, "hit_rate_incoherent": hit_i, "id_checksum_coherent": chk_c, "id_checksum_incoherent": chk_i, "bvh_nodes_MB": b["nodes"] / 1e6, "leaf_slots_MB": b["leaves"] / 1e6, "tables_MB": b["tables"] / 1e6,
               "scene_gen_s": round(gen_s, 2), "build_upload_s": round(build_s, 2),
               "builder": "device LBVH" if bs["mode"] == 1 else "host binned SAH", "tree_build_ms": round(bs["build_ms"], 2), "commit_ms": round(bs["commit_ms"], 1),
               "bvh4_nodes": bs["nodes"], "max_stack": bs["max_stack"]}
        if rank == 0:
            print(json.dumps(row), flush=True)
        rows.append(row)
        del gpu, hs, rays, inc, prim, t, wgt
        torch.cuda.empty_cache()
    if rank == 0 and args.out:
        json.dump({"workload": "SURVEY 8d config 5: triangle soup, rt_closest_device, device-resident rays", "rows": rows}, open(args.out, "w"), indent=1)
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
