"""BASELINE.json configs[3] AS SPECIFIED (VERDICT r01 missing #4): two dragons, glass + rough metal (scenes.rs:549-624 with the commented glass block
enabled), 2048 x 2048, max_depth 16, 1024 spp split over the ranks of one box (sample-range partition, scene replicated), films summed with
rt_film_allreduce — and the REDUCED film checked per pixel against the oracle on a subset of rows at the same (pixel, sample) RNG keys.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/config4_multi.py --out profiles/config4_8gpu_r02.json

Two renders: RT_SHADE_F32 (production mode; timed), then f64 radiometry (timed; this is the film that is compared).  Tolerance of the comparison: sample
counts exact (n = 1024 for every pixel); per pixel |G - O| / (|O| + 1e-6) <= 1e-7 except where one of the pixel's 1024 samples took the other side of a
knife-edge branch on a last-ulp libm difference (the per-sample flip rate of the small-image tests, ~5e-4, makes that about a third of the pixels here): those
must still be within 2e-2 of the pixel's value, i.e. a handful of its 1024 samples; the mean over the rows within 1e-4."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from rustraytracer_b200 import Comm, Film, IntType, _native as N, render_samples, scenes  # noqa: E402


def check(json_path):
    """CPU side: the oracle renders the kept rows at the same (pixel, sample) keys; the verdict is written into the JSON next to the GPU numbers."""
    import orc
    res = json.load(open(json_path))
    z = np.load(os.path.splitext(json_path)[0] + "_rows.npz")
    rows, g, f, W, spp, depth = [int(r) for r in z["rows"]], z["g64"], z["g32"], int(z["size"]), int(z["spp"]), int(z["depth"])
    hs = scenes.HostScene("two_dragons", W, W, variant=1)
    oracle = orc.OracleScene(hs.tables())
    cam = hs.camera()
    O = np.zeros((W, W, 4))
    t0 = time.time()
    for y in rows:
        oracle.render(cam, W, W, 0, spp, seed=4242, max_depth=depth, window=(0, y, W, y + 1), film=O, threads=os.cpu_count() or 8)
    o = O[rows]
    err = np.abs(g[..., :3] - o[..., :3]).max(-1) / (np.abs(o[..., :3]).max(-1) + 1e-6)
    chk = {"rows": rows, "pixels": int(err.size), "oracle_seconds": round(time.time() - t0, 1), "host_threads": os.cpu_count(),
           "sample_counts_exact": bool(np.array_equal(g[..., 3], o[..., 3]) and np.all(g[..., 3] == spp)),
           "frac_pixels_within_1e-7": float((err <= 1e-7).mean()), "frac_pixels_within_1e-3": float((err <= 1e-3).mean()), "max_rel_err": float(err.max()),
           "mean_rel_diff": float(abs(g[..., :3].mean() - o[..., :3].mean()) / o[..., :3].mean()),
           "f32_vs_oracle_mean_rel_diff": float(abs(f[..., :3].mean() - o[..., :3].mean()) / o[..., :3].mean())}
    chk["ok"] = bool(chk["sample_counts_exact"] and res.get("all_finite") and chk["max_rel_err"] <= 2e-2 and chk["mean_rel_diff"] <= 1e-4 and chk["frac_pixels_within_1e-7"] >= 0.3
                     and chk["f32_vs_oracle_mean_rel_diff"] <= 1e-2)
    res["parity_f64_reduced_film_vs_oracle"] = chk
    json.dump(res, open(json_path, "w"), indent=1)
    print(json.dumps(chk))
    sys.exit(0 if chk["ok"] else 1)


def main():
    if len(sys.argv) == 3 and sys.argv[1] == "--check":
        return check(sys.argv[2])
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=2048)
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--depth", type=int, default=16)
    ap.add_argument("--rows", type=int, default=4)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local); N.check(N.lib().rt_set_device(local))
    comm = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        ident = [Comm.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        comm = Comm.rank(ident[0], world, rank)
    W = H = a.size
    hs = scenes.HostScene("two_dragons", W, H, variant=1)
    gpu = hs.upload()
    gpu.set_wavefront_size(1 << 25)
    cam = hs.camera()
    lo, hi = a.spp * rank // world, a.spp * (rank + 1) // world
    res = {"workload": f"two_dragons glass + metal {W}x{H}, max_depth {a.depth}, {a.spp} spp over {world} GPU(s) (samples [{lo},{hi}) on rank {rank}), seed 4242", "n_gpus": world}
    films = {}
    for mode, integ in (("f32", IntType.Path(a.depth, shade_f32=True)), ("f64", IntType.Path(a.depth))):
        film = Film(W, H)
        render_samples(gpu, cam, integ, film, W, H, lo, min(hi, lo + 2), seed=4242)          # warm-up: module load, wavefront allocation
        film.clear()
        torch.cuda.synchronize()
        if comm: dist.barrier()
        t0 = time.perf_counter()
        render_samples(gpu, cam, integ, film, W, H, lo, hi, seed=4242)
        c = gpu.counters()
        if comm: comm.allreduce([film])
        torch.cuda.synchronize()
        secs = time.perf_counter() - t0
        st = torch.tensor([c["extend_rays"] + c["shadow_rays"] + c["probe_rays"], c["paths"], secs], dtype=torch.float64, device=f"cuda:{local}")
        if comm:
            s = st.clone(); dist.all_reduce(s); m = st.clone(); dist.all_reduce(m, op=dist.ReduceOp.MAX); st = torch.stack([s[0], s[1], m[2]])
        rays, paths, secs = (float(x) for x in st.tolist())
        res[mode] = {"Mrays_s": rays / secs / 1e6, "samples_per_s": paths / secs, "seconds_max_over_ranks": secs, "rays": rays, "paths": paths, "allreduce": "rt_film_allreduce (ncclAllReduce f64 sum)" if comm else None}
        films[mode] = film
    if rank == 0:
        G = films["f64"].read()
        F = films["f32"].read()
        rows = [int(H * (k + 0.5) / a.rows) for k in range(a.rows)]
        res["all_finite"] = bool(np.isfinite(G).all() and np.isfinite(F).all())
        res["sample_counts_all_pixels_exact"] = bool(np.all(G[..., 3] == a.spp) and np.all(F[..., 3] == a.spp))
        res["rows_kept_for_the_oracle_check"] = rows
        print(json.dumps(res), flush=True)
        if a.out:
            json.dump(res, open(a.out, "w"), indent=1)
            # the oracle side costs minutes of CPU: it runs afterwards on a CPU-only machine (--check) on these rows of the reduced films
            np.savez_compressed(os.path.splitext(a.out)[0] + "_rows.npz", rows=np.array(rows), g64=G[rows], g32=F[rows], size=W, spp=a.spp, depth=a.depth)
    if comm:
        dist.barrier()
        comm.close()
        dist.destroy_process_group()
    sys.exit(0)


if __name__ == "__main__":
    main()
