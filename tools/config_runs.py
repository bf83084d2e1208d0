"""Throughput of every BASELINE.json configuration on one GPU (device-resident; the headline config is bench.py's).

    python tools/config_runs.py [--out profiles/configs_r01.json] [--scale 1.0]

configs[0] sphere material-test scene 400x400, 16 spp          (max_depth 25, the reference default)
configs[1] Cornell box + statue (stand-in mesh) 1024x1024, 64 spp
configs[2] dragon rough-plastic AND rough-metal 1024x1024, 8 bounces, 256 spp
configs[3] two dragons, glass + metal, depth 16, 2048x2048 — 128 spp here (1/8 of the 8-GPU job's 1024 spp: one GPU's share)
Each line: Mrays/s = (extend + shadow + probe rays) / device time of rt_render, samples/s, per-kernel ms.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rustraytracer_b200 import Film, IntType, render_samples, scenes  # noqa: E402

CONFIGS = [
    ("configs[0] sphere_roughness 400x400 16spp depth25", "sphere_roughness", {}, 400, 16, 25),
    ("configs[1] cornell_box_statue (stand-in) 1024x1024 64spp depth25", "cornell_box_statue", {}, 1024, 64, 25),
    ("configs[2] plastic_dragon rough-plastic 1024x1024 256spp depth8", "plastic_dragon", {}, 1024, 256, 8),
    ("configs[2] plastic_dragon rough-metal 1024x1024 256spp depth8", "plastic_dragon", {"variant": 1}, 1024, 256, 8),
    ("configs[3] two_dragons glass+metal 2048x2048 128spp (one GPU's share of 1024) depth16", "two_dragons", {"variant": 1}, 2048, 128, 16),
    ("extra: material_hdr glass under the HDR environment 1024x1024 64spp depth12", "material_hdr", {"variant": 3}, 1024, 64, 12),
    ("extra: teapot_hdr plastic teapot (stand-in) on the checkered floor under the HDR environment 1024x1024 64spp depth8", "teapot_hdr", {}, 1024, 64, 8),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--scale", type=float, default=1.0, help="multiply every spp (quick runs)")
    args = ap.parse_args()
    rows = []
    for label, name, kw, size, spp, depth in CONFIGS:
        spp = max(1, int(spp * args.scale))
        hs = scenes.HostScene(name, size, size, **kw)
        gpu = hs.upload()
        cam = hs.camera()
        film = Film(size, size)
        for f32 in (True, False):
            if not f32 and spp > 64:
                continue                       # the f64-radiometry parity mode is reported on the smaller configs only
            for rep in range(2):               # first pass warms up
                film.clear()
                render_samples(gpu, cam, IntType.Path(depth, shade_f32=f32), film, size, size, 0, spp, seed=7)
            c = gpu.counters()
            rays = c["extend_rays"] + c["shadow_rays"] + c["probe_rays"]
            row = {"config": label, "shade": "f32" if f32 else "f64", "prims": gpu.num_primitives(), "standin": hs.uses_standin(), "Mrays_s": rays / c["render_ms"] / 1e3,
                   "samples_s": c["paths"] / c["render_ms"] * 1e3, "render_ms": c["render_ms"], "extend_ms": c["extend_ms"], "shade_ms": c["shade_ms"], "connect_ms": c["connect_ms"],
                   "rays": {"extend": c["extend_rays"], "shadow": c["shadow_rays"], "probe": c["probe_rays"]}, "iterations": c["iterations"]}
            print(json.dumps(row), flush=True)
            rows.append(row)
        film.close(); gpu.close(); hs.close()
    if args.out:
        json.dump(rows, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
