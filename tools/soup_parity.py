"""BASELINE.json configs[4] parity at its OWN sizes (VERDICT r01 "missing" #3).

    python tools/soup_parity.py [--sizes 1,10,50] [--res 4096] [--subset 1048576] [--out profiles/soup_parity_r02.json]

For every triangle count N (millions) the soup of SURVEY §8d is committed TWICE — host binned-SAH tree and device LBVH — and a 2^20-ray
subset of the coherent batch (pinhole primary rays) and of the incoherent batch (one cosine bounce per primary hit, uniform-sphere rays
elsewhere) is traced with `rt_closest_device` on each tree and compared with the ORACLE's hit set on the same rays:
  * closest-hit primitive id bit-exact, or a member of the oracle's exact-t tie set (the reference resolves ties by the order of its
    randomly built tree);  hit/miss flag exact;
  * t within rel 1e-12 (both sides run the reference's f64 operation order; the slack is for nothing but is stated).
The oracle side is `orc_intersect` mode 3: the reference's own predicate hit(prim) = leafbox_f64 AND primtest_f64, walked over a
median-split tree with the unchanged interval (every box that passes is visited), min-t record + tie set — identical to brute force
over all primitives (tests/test_oracle_kat.py::test_tree_modes_agree_on_random_scene), just not O(N) per ray.
The subset is every (rays / subset)-th ray of each batch rather than the first 2^20: the first 256 rows of the 4096^2 image look over the
top of the soup's cube and would all be misses.
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from rustraytracer_b200 import _native as N, scenes  # noqa: E402
from rustraytracer_b200._native import check, lib  # noqa: E402

NONE = N.RT_NONE


def mem_available_gb():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable"):
                return int(line.split()[1]) / 1e6
    except Exception:
        pass
    return 0.0


def incoherent_batch(gpu, rays_dev, dev, stream, seed=0xD1FF):
    """tools/soup_sweep.py's incoherent batch: one cosine-distributed direction about the geometric normal per primary hit, origin = hit
    point; primary misses become uniform-sphere directions from U[-1,1]^3 origins."""
    n = rays_dev.shape[0]
    hits = torch.zeros(n * N.hit_dtype.itemsize, dtype=torch.uint8, device=dev)
    check(lib().rt_intersect_device(gpu.h, n, C.c_void_p(rays_dev.data_ptr()), 0.001, 1e308, C.c_void_p(hits.data_ptr()), stream))
    torch.cuda.synchronize()
    h = hits.view(torch.float64).reshape(n, 13)
    hit = h[:, 0] > 0
    g = torch.Generator(device=dev); g.manual_seed(seed)
    nrm = h[:, 4:7]
    u1, u2 = torch.rand(n, generator=g, device=dev, dtype=torch.float64), torch.rand(n, generator=g, device=dev, dtype=torch.float64)
    r, phi = torch.sqrt(u1), 2 * np.pi * u2
    a = torch.where(nrm[:, 0:1].abs() > 0.9, torch.tensor([0., 1., 0.], device=dev, dtype=torch.float64), torch.tensor([1., 0., 0.], device=dev, dtype=torch.float64))
    tx = torch.linalg.cross(nrm, a.expand_as(nrm)); tx = tx / tx.norm(dim=1, keepdim=True).clamp_min(1e-30)
    ty = torch.linalg.cross(nrm, tx)
    d_hit = tx * (r * torch.cos(phi))[:, None] + ty * (r * torch.sin(phi))[:, None] + nrm * torch.sqrt((1 - u1).clamp_min(0))[:, None]
    d_miss = torch.randn(n, 3, generator=g, device=dev, dtype=torch.float64); d_miss = d_miss / d_miss.norm(dim=1, keepdim=True)
    o_miss = torch.rand(n, 3, generator=g, device=dev, dtype=torch.float64) * 2 - 1
    return torch.cat([torch.where(hit[:, None], h[:, 1:4], o_miss), torch.where(hit[:, None], d_hit, d_miss)], dim=1).contiguous()


def gpu_closest(gpu, rays_dev, dev, stream):
    n = rays_dev.shape[0]
    prim = torch.zeros(n, dtype=torch.int32, device=dev)
    t = torch.zeros(n, dtype=torch.float64, device=dev)
    check(lib().rt_closest_device(gpu.h, n, C.c_void_p(rays_dev.data_ptr()), 0.001, 1e308, C.c_void_p(prim.data_ptr()), C.c_void_p(t.data_ptr()), stream))
    torch.cuda.synchronize()
    return prim.cpu().numpy().view(np.uint32), t.cpu().numpy()


def compare(prim, t, want, ties):
    """-> dict of counts; `ok` iff ids are exact-or-tied and t within rel 1e-12"""
    same = prim == want["prim_index"]
    in_tie = (prim != NONE) & (ties == prim[:, None]).any(axis=1)
    bad = ~(same | in_tie)
    hit = same & (want["prim_index"] != NONE)
    rel = np.abs(t[hit] - want["t"][hit]) / np.abs(want["t"][hit])
    return {"rays": int(len(prim)), "hits": int((want["prim_index"] != NONE).sum()), "id_mismatch": int(bad.sum()), "resolved_by_tie_set": int((in_tie & ~same).sum()),
            "rays_with_ties": int((ties[:, 1] != NONE).sum()), "max_rel_t_err": float(rel.max()) if len(rel) else 0.0,
            "ok": bool(not bad.any() and (len(rel) == 0 or rel.max() <= 1e-12))}


def run_size(n_tris, res, subset, dev, stream, threads, builders=("sah", "lbvh")):
    t0 = time.time()
    hs = scenes.HostScene("soup", res, res, n_tris=n_tris)
    row = {"tris": n_tris, "scene_gen_s": round(time.time() - t0, 2)}
    import orc
    t0 = time.time()
    oracle = orc.OracleScene()
    oracle.load(hs.tables())
    lib_o = orc.lib()
    # boxes / areas are prepared by commit(); the faithful stable_sort tree it also builds is replaced by the quick one right away on big soups
    if n_tris <= 300000:
        oracle.commit(1)
    else:
        lib_o.orc_scene_commit_tables.argtypes = [C.c_void_p]
        lib_o.orc_scene_commit_tables(oracle.h)
    oracle.build_fast_tree(threads)
    row["oracle_build_s"] = round(time.time() - t0, 2)
    cam = hs.camera()
    rays_h = cam.primary_rays(res, res)
    n_all = len(rays_h)
    stride = max(1, n_all // subset)
    sel = np.arange(0, n_all, stride)[:subset]
    want = {}
    for b in builders:
        t0 = time.time()
        gpu = hs.upload(build=b)
        bs = gpu.build_stats()
        rays_dev = torch.from_numpy(rays_h.view(np.float64).reshape(n_all, 6)).to(dev)
        inc_dev = incoherent_batch(gpu, rays_dev, dev, stream)
        out = {"builder": "device LBVH" if bs["mode"] == 1 else "host binned SAH", "commit_s": round(time.time() - t0, 2), "bvh4_nodes": bs["nodes"], "max_stack": bs["max_stack"]}
        for name, batch in (("coherent", rays_dev), ("incoherent", inc_dev)):
            sub = batch[torch.from_numpy(sel).to(dev)].contiguous()
            prim, t = gpu_closest(gpu, sub, dev, stream)
            sub_h = np.ascontiguousarray(sub.cpu().numpy()).view(N.ray_dtype).reshape(-1)
            key = (name, sub_h.tobytes()[:4096])
            if name not in want or want[name][0] != key[1]:
                t1 = time.time()
                w, ties = oracle.intersect(sub_h, mode=3, threads=threads, want_ties=True)
                want[name] = (key[1], w, ties, round(time.time() - t1, 2))
            _, w, ties, secs = want[name]
            c = compare(prim, t, w, ties)
            c["oracle_s"] = secs
            out[name] = c
        row[b] = out
        del gpu, rays_dev, inc_dev
        torch.cuda.empty_cache()
    row["ok"] = all(row[b][k]["ok"] for b in builders for k in ("coherent", "incoherent"))
    return row


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1,10,50")
    ap.add_argument("--res", type=int, default=4096)
    ap.add_argument("--subset", type=int, default=1 << 20)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    torch.cuda.set_device(0)
    check(lib().rt_set_device(0))
    dev = "cuda:0"
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    threads = os.cpu_count() or 8
    rows = []
    for m in [float(x) for x in args.sizes.split(",")]:
        n_tris = int(m * 1e6)
        need_gb = n_tris * 900e-9 + 8                       # host tables twice + oracle boxes / nodes + the SAH builder's scratch
        if mem_available_gb() < need_gb:
            rows.append({"tris": n_tris, "skipped": f"needs ~{need_gb:.0f} GB of host memory, {mem_available_gb():.0f} GB available"})
            print(json.dumps(rows[-1]), flush=True)
            continue
        row = run_size(n_tris, args.res, args.subset, dev, stream, threads)
        print(json.dumps(row), flush=True)
        rows.append(row)
    stats = (C.c_uint64 * 4)()
    check(lib().rt_debug_prefilter_stats(stats))
    res = {"workload": "SURVEY 8d config 5 parity subset: rt_closest_device vs the oracle's hit set (mode 3) on both trees", "subset_rays_per_batch": args.subset,
           "tolerance": {"prim_id": "exact or in the oracle's tie set", "t_rel": 1e-12}, "host_threads": threads, "prefilter_audit": list(stats), "rows": rows,
           "ok": all(r.get("ok", True) for r in rows)}
    if args.out:
        json.dump(res, open(args.out, "w"), indent=1)
    print(json.dumps({"ok": res["ok"], "prefilter_audit": list(stats)}))
    sys.exit(0 if res["ok"] else 1)


if __name__ == "__main__":
    main()
