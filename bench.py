#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 path-tracing core (contract: see the task brief / DESIGN.md §6).

One "step" = one pass of the hot path (rt_render: generate → extend → shade → connect wavefront) over one batch
of synthetic input: BASELINE.json configs[2], the dragon scene (scenes.rs:310-375 `plastic_dragon`, rough-plastic
microfacet; stand-in mesh of 871,414 triangles because data/dragon/dragon.obj is a missing blob), 1024x1024,
max_depth 8, 256 samples per pixel per GPU.  Metric: Mrays/s = (extend + shadow + MIS-probe rays traced) / time.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--spp S] [--size PX]
Under torchrun (N > 1) every rank renders its own sample range of the same image (scene replicated, weak
scaling) and the (r,g,b,n) films are summed with an NCCL all-reduce inside the timed region.
`--impl reference` times the CPU restatement of the reference's own algorithm (the oracle port; the Rust
reference cannot be built in this image) on the host cores, on a bounded sample of the same workload.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SCENE, VARIANT, DEPTH = "plastic_dragon", 0, 8
METRIC, UNIT = "Mrays/s (dragon, 1024^2, 8 bounces)", "Mrays/s"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def load_json(name):
    p = os.path.join(ROOT, "profiles", name)
    return json.load(open(p)) if os.path.exists(p) else None


def extend_rooflines(tot, launches, clk, sms=148):
    """Rooflines of the dominant kernel, k_extend (closest-hit traversal).  Per-ray constants of the kernel (bytes its loads request, DRAM
    bytes, warp / thread instructions) come from ONE committed ncu capture of this build (profiles/extend_work_r02.json, made by
    tools/extend_work.py from `ncu --set full`); rays and kernel time are measured live (device counters + CUDA events on the render stream).
    Denominators: measured L2 bandwidth and issue rate from tools/peaks.cu (profiles/peaks_r02.json), measured HBM copy bandwidth from
    MEASURED_PEAKS.json.  The SURVEY §8d figure (bytes on the REFERENCE's median-split tree) is kept as `work_vs_reference_tree`: it says how
    much less the kernel moves than the reference's traversal would, it is not a fraction of any peak."""
    hbm_peak, peak_kind = load_peaks()
    work = load_json("extend_work_r02.json") or {}
    peaks = load_json("peaks_r02.json") or {}
    rays, ms = tot["extend_rays"], tot["extend_ms"]
    rays_per_s = rays / (ms * 1e-3) if ms > 0 else 0.0
    per_launch = rays / max(launches, 1)
    out = {"kernel": "k_extend (closest-hit BVH4 traversal, persistent)", "rays_per_launch": per_launch, "ms_per_launch": ms / max(launches, 1), "extend_Mrays_per_s": rays_per_s / 1e6,
           "per_ray_constants": "profiles/extend_work_r02.json" if work else None}
    l2_peak = (peaks.get("gather128_gbs") or {}).get("64MB")
    l2_bytes = work.get("l2_to_l1_bytes_per_ray")
    # headline roofline object: the bytes the kernel actually pulls out of L2 (L2 -> L1 sectors x 32 B, ncu) against the measured L2 bandwidth for its access
    # pattern — independent 128-byte node / slot records per lane (tools/peaks.cu gather128), not the 17.9-25 TB/s a coalesced stream reaches
    out.update({"bound": "l2", "unit": "GB/s", "achieved": l2_bytes * rays_per_s / 1e9 if l2_bytes else None, "peak": l2_peak,
                "peak_kind": "measured: tools/peaks.cu gather128 (independent random 128-byte records per lane, 64 MB footprint), profiles/peaks_r02.json",
                "traffic": work.get("dram_bytes_per_ray") * per_launch if work.get("dram_bytes_per_ray") else None,
                "l2_to_l1_bytes_per_ray": l2_bytes})
    out["frac"] = out["achieved"] / l2_peak if out["achieved"] and l2_peak else None
    ld_bytes = work.get("load_bytes_per_ray")
    stream_peak = (peaks.get("ld128_gbs") or {}).get("100MB")
    if ld_bytes and stream_peak:
        a = ld_bytes * rays_per_s / 1e9      # what the kernel's loads REQUEST at L1 (l1tex global-load sectors x 32 B): part of it is served by L1 hits
        out["l1_requests"] = {"achieved": a, "unit": "GB/s", "load_bytes_per_ray": ld_bytes, "vs_l2_stream_peak": a / stream_peak, "l2_stream_peak": stream_peak,
                              "note": "requested bytes incl. L1 hits; shown against the coalesced L2 -> SM stream bandwidth (ld128, 100 MB footprint) as an upper reference"}
    if work.get("dram_bytes_per_ray"):
        a = work["dram_bytes_per_ray"] * rays_per_s / 1e9
        out["hbm"] = {"achieved": a, "peak": hbm_peak, "unit": "GB/s", "frac": a / hbm_peak, "peak_kind": f"{peak_kind} HBM copy bandwidth (MEASURED_PEAKS.json)",
                      "dram_bytes_per_ray": work["dram_bytes_per_ray"]}
    if work.get("warp_inst_per_ray"):
        mhz = (clk or {}).get("sm_mhz") or (clk or {}).get("sm_max_mhz") or 1965
        peak_issue = sms * 4 * mhz * 1e6
        a = work["warp_inst_per_ray"] * rays_per_s
        lanes = work.get("lanes_per_inst")
        out["issue"] = {"achieved_warp_inst_per_s": a, "peak_warp_inst_per_s": peak_issue, "frac": a / peak_issue, "lanes_per_inst": lanes,
                        "useful_lane_issue_frac": a / peak_issue * lanes / 32 if lanes else None,
                        "peak_kind": f"{sms} SMs x 4 schedulers x {mhz} MHz (SM clock sampled during the timed region); measured FFMA-only rate: profiles/peaks_r02.json fp32_warp_inst_per_s"}
    alg = load_json("algorithmic_work_r01.json")
    if alg and f"{SCENE}_v{VARIANT}" in alg:
        w = alg[f"{SCENE}_v{VARIANT}"]
        bpr = 32 + 16 + w["n_box"] * 32 + w["n_tri"] * 48
        out["work_vs_reference_tree"] = {"algorithmic_bytes_per_ray": bpr, "n_box": w["n_box"], "n_tri": w["n_tri"], "GB_per_s_equivalent": bpr * rays_per_s / 1e9,
                                         "note": "SURVEY 8d bytes counted by the oracle's ordered traversal of the REFERENCE's median-split 1-prim-per-leaf tree; not a fraction of a peak"}
    return out


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(int(r[0]) for r in self.rows if len(r) >= 6 and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) >= 6 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def run_reference(args, rank, world):
    """The reference's own CPU implementation of the path (oracle port, all host threads), bounded sample per step."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import orc
    from rustraytracer_b200 import scenes
    cores = os.cpu_count() or 1
    hs = scenes.HostScene(SCENE, args.size, args.size, variant=VARIANT)
    oracle = orc.OracleScene(hs.tables())
    cam = hs.camera()
    # bounded sample: every 3rd pixel of the full image x 1 spp per pass (same scene, camera, depth, ray mix); a step is
    # as many passes (fresh sample index each) as fit ~TARGET_S seconds of all-core CPU work
    stride, TARGET_S = 3, 6.0
    rays = secs = paths = 0
    smp = 0
    for step in range(args.warmup + args.steps):
        s_rays = s_secs = s_paths = 0
        while s_secs < (TARGET_S if step >= args.warmup else 1.0):
            film, counts, s = oracle.render_ex(cam, args.size, args.size, smp, smp + 1, seed=1234, max_depth=DEPTH, threads=cores, pixel_stride=stride)
            smp += 1
            s_rays += int(counts[:3].sum()); s_secs += s; s_paths += int(film[..., 3].sum())
        if step >= args.warmup:
            rays += s_rays; secs += s_secs; paths += s_paths
    v = rays / secs / 1e6
    sample = (f"every {stride}rd pixel of the {args.size}^2 image, {smp / (args.warmup + args.steps):.0f} one-sample passes per step on average "
              f"({paths // max(args.steps, 1)} paths/step, ~{TARGET_S:.0f} s of {cores}-thread CPU work), max_depth {DEPTH}")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{SCENE} (stand-in dragon mesh, 871,414 tris) {args.size}x{args.size}, max_depth {DEPTH}", "sample": sample},
        "samples_per_s": paths / secs,
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "note": "C++ f64 restatement of the reference's algorithm (faithful tree, both-children traversal, closest-hit shadow rays); the Rust reference cannot be built here"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--spp", type=int, default=256)
    ap.add_argument("--size", type=int, default=1024)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--wavefront", type=int, default=0, help="paths in flight per GPU; 0 = 128 Mi when the GPU has 100 GB free (64 GB of queue state), else 32 Mi")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"], help="weak: --spp samples per GPU per step; strong: --spp samples per step in total, split over the GPUs")
    ap.add_argument("--no-f64-leg", action="store_true")
    ap.add_argument("--shade", default="f32", choices=["f32", "f64"], help="radiometric precision of the shade kernel (geometry is f64-exact in both)")
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import numpy as np
    import torch
    from rustraytracer_b200 import Film, IntType, _native as N, render_samples, scenes
    assert N.lib().rt_device_count() > 0, "no CUDA device: librtcore has no CPU path"
    torch.cuda.set_device(local)
    N.check(N.lib().rt_set_device(local))
    dist = comm = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        # the film reduce goes through the library's own entry point (rt_film_allreduce = ncclAllReduce f64 sum over NVLink, include/rtcore.h);
        # torch.distributed is the launcher's plumbing: it carries the communicator id to the ranks, the barriers and the timing statistics
        from rustraytracer_b200 import Comm
        ident = [Comm.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        comm = Comm.rank(ident[0], world, rank)
    W = H = args.size
    spp = int(N.lib().rt_round_spp(args.spp))            # sampler.rs:637-640
    hs = scenes.HostScene(SCENE, W, H, variant=VARIANT)
    t0 = time.time()
    objects = hs.upload()                                 # BVH build + flatten + H2D (outside the timed region: scene build is not the hot path)
    build_s = time.time() - t0
    if args.wavefront <= 0:
        # measured r01 (profiles/bench_r01_wavefront.json): 32 Mi 2,642, 64 Mi 2,683, 128 Mi 2,710 Mrays/s — fewer, larger launches;
        # the queue state is laid out for the 180 GB of a B200
        args.wavefront = 1 << 27 if torch.cuda.mem_get_info(local)[0] > 100e9 else 1 << 25
    objects.set_wavefront_size(args.wavefront)          # paths in flight; ~480 B of queue state each (64 GB at 128 Mi)
    cam = hs.camera()
    integ = IntType.Path(DEPTH, shade_f32=(args.shade == "f32"))
    film_t = torch.zeros(H * W * 4, dtype=torch.float64, device=f"cuda:{local}")      # torch owns the film so NCCL can reduce it in place
    film = Film(W, H, device_ptr=film_t.data_ptr())
    host_film = torch.empty(H * W * 4, dtype=torch.float64).pin_memory()

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
            torch.cuda.synchronize()

    strong = args.scaling == "strong"
    spp_rank = spp // world if strong else spp            # strong: `spp` samples in TOTAL per step, split over the ranks; weak: `spp` per GPU
    if strong and spp % world:
        raise SystemExit(f"--scaling strong needs spp ({spp}) divisible by the number of GPUs ({world})")

    def step(i, e2e, integ):
        """One pass of the hot path on this rank's sample range of the step's global sample space."""
        film_t.zero_()
        torch.cuda.synchronize()
        base = (i * world + rank) * spp_rank
        render_samples(objects, cam, integ, film, W, H, base, base + spp_rank, seed=1234)    # blocking; H2D = camera + descriptors
        c = objects.counters()
        if comm:
            comm.allreduce([film])                        # rt_film_allreduce: NCCL sum of the (r,g,b,n) accumulation buffers over NVLink, in place, blocking
        if e2e and rank == 0:
            host_film.copy_(film_t, non_blocking=False)   # D2H read of the step's result
        torch.cuda.synchronize()
        return c

    KEYS = ("extend_rays", "shadow_rays", "probe_rays", "paths", "iterations", "render_ms", "extend_ms", "shade_ms", "connect_ms")

    def leg(integ, first_step, e2e, sample_clocks=False):
        """W warm-up steps, then K timed steps bracketed by barrier + synchronize; returns (totals, wall seconds, clocks)."""
        for i in range(args.warmup):
            step(first_step + i, e2e, integ)
        clocks = ClockSampler(local) if (sample_clocks and rank == 0) else None
        barrier()
        t_start = time.perf_counter()
        tot = {k: 0.0 for k in KEYS}
        for i in range(args.steps):
            c = step(first_step + args.warmup + i, e2e, integ)
            for k in tot:
                tot[k] += c[k]
        barrier()
        wall = time.perf_counter() - t_start
        return tot, wall, (clocks.stop() if clocks else None)

    def reduce_leg(tot, wall):
        """whole-job rays / paths (sum over ranks) and the MAX over ranks of the wall and device times"""
        rays = tot["extend_rays"] + tot["shadow_rays"] + tot["probe_rays"]
        st = torch.tensor([rays, tot["paths"], wall, tot["render_ms"]], dtype=torch.float64, device=f"cuda:{local}")
        if dist:
            s_sum = st.clone(); dist.all_reduce(s_sum)
            s_max = st.clone(); dist.all_reduce(s_max, op=dist.ReduceOp.MAX)
            return s_sum[0].item(), s_sum[1].item(), s_max[2].item(), s_max[3].item()
        return rays, tot["paths"], wall, tot["render_ms"]

    n_leg = args.warmup + args.steps
    # ---- leg 1: device-resident (value) ; leg 2: the same through the public call with the host read-back every step (e2e)
    tot, wall, clk = leg(integ, 0, False, sample_clocks=True)
    _, wall_e2e, _ = leg(integ, n_leg, True)
    rays_all, paths_all, wall, dev_ms = reduce_leg(tot, wall)
    _, _, wall_e2e, _ = reduce_leg(tot, wall_e2e)
    # ---- leg 3 (f32 headline only): the same workload with the shade kernels in the reference's own f64 arithmetic (per-pixel parity mode),
    # warmed and timed exactly like leg 1 — r01's single un-warmed f64 render measured CUDA's lazy loading of the k_shade<double> kernels
    f64 = None
    if args.shade == "f32" and not args.no_f64_leg:
        integ64 = IntType.Path(DEPTH, shade_f32=False)
        tot64, wall64, _ = leg(integ64, 2 * n_leg, False)
        r64, p64, wall64, dev64 = reduce_leg(tot64, wall64)
        f64 = {"value": r64 / wall64 / 1e6, "unit": UNIT, "samples_per_s": p64 / wall64, "ms_per_step": 1e3 * wall64 / args.steps, "device_ms_per_step": dev64 / args.steps,
               "steps": args.steps, "warmup": args.warmup, "dtype": "f64 (geometry and radiometry: the reference's arithmetic; films match the oracle per pixel)",
               "kernel_ms_per_step": {"extend": tot64["extend_ms"] / args.steps, "shade+generate": tot64["shade_ms"] / args.steps, "connect": tot64["connect_ms"] / args.steps}}
    if rank != 0:
        if dist:
            comm.close()
            dist.destroy_process_group()
        return
    value = rays_all / wall / 1e6
    e2e = rays_all / wall_e2e / 1e6
    launches = int(tot["iterations"])                      # one k_extend launch per wavefront iteration
    tb = hs.tables()                                      # shading bins present = distinct (material type, triangle / parametric shape) pairs
    n_bins = len({(int(tb["materials"]["type"][m]), int(t == N.PRIM_TRIANGLE)) for m, t in set(zip(tb["prims"]["mat_index"].tolist(), tb["prims"]["type"].tolist()))})
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * wall / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f64 geometry (intersection, hit points, rays) + f32 radiometry" if args.shade == "f32" else "f64", "data": "synthetic",
        "config": {"workload": f"{SCENE} rough-plastic (stand-in dragon mesh, 871,414 tris; dragon.obj is a missing blob) {W}x{H}, max_depth {DEPTH}, "
                               + (f"{spp} spp per step in total = {spp_rank} per GPU" if strong else f"{spp} spp per GPU per step"),
                   "parallelism": f"scene replicated, sample-range partition x{world}, rt_film_allreduce (ncclAllReduce f64 sum, behind the C ABI) of the film" if world > 1 else "single GPU",
                   "l2": "no explicit flush: every wavefront iteration streams several GB of path/queue state (>> 126 MB L2); the BVH (33 MB nodes + 42 MB f32 prefilter records + 70 MB f64 leaf slots) is meant to stay L2-resident",
                   "wavefront_paths": args.wavefront, "bvh_build_s": round(build_s, 2),
                   "headline_precision": "RT_SHADE_F32: intersection decisions, hit points and rays are f64-exact (bit-exact ids vs the oracle); radiometry in f32, image parity within the relMSE / mean-bias bounds of tests/test_gpu_parity.py; the all-f64 figure of the same workload is `exact_f64_shading`"},
        "samples_per_s": paths_all / wall,
        "rays": {"extend": tot["extend_rays"], "shadow": tot["shadow_rays"], "probe": tot["probe_rays"]},
        "device_ms_per_step": dev_ms / args.steps,
        "kernel_ms_per_step": {"extend": tot["extend_ms"] / args.steps, "shade+generate": tot["shade_ms"] / args.steps, "connect": tot["connect_ms"] / args.steps},
        "gpu_launches": int(tot["iterations"] * (4 + n_bins) + 2 * args.steps),      # per iteration: prepare, generate, extend, one shade kernel per bin in the scene, connect; per render: closing prepare + film count
        "clocks": clk,
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": C.sizeof(N.Camera) + C.sizeof(N.IntegratorDesc) + C.sizeof(N.FilmDesc),
                "d2h_bytes_per_step": H * W * 4 * 8, "ms_per_step": 1e3 * wall_e2e / args.steps},
        "roofline": extend_rooflines(tot, launches, clk),
    }
    if f64:
        out["exact_f64_shading"] = f64
    if world == 1 and not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import orc                                         # the checker, used here ONLY as the timed CPU baseline
        cores = os.cpu_count() or 1
        oracle = orc.OracleScene(hs.tables())
        stride, c_rays, c_secs, smp = 3, 0, 0.0, 0
        while c_secs < 12.0:                              # ~12 s of all-core CPU work on the same image
            _, counts, secs = oracle.render_ex(cam, W, H, smp, smp + 1, seed=1234, max_depth=DEPTH, threads=cores, pixel_stride=stride)
            c_rays += int(counts[:3].sum()); c_secs += secs; smp += 1
        out["cpu_baseline"] = {"value": c_rays / c_secs / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
                               "sample": f"every {stride}rd pixel of the {W}x{H} image x {smp} spp ({c_secs:.1f} s of {cores}-thread CPU work)"}
    print(json.dumps(out))
    if dist:
        comm.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
