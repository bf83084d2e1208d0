"""Regenerates the frozen fixtures under tests/golden/ from the ORACLE (test infrastructure).

The reference (Rust) cannot be built or run in this image and ships no golden vectors of its own, so these
fixtures are NOT reference outputs: they freeze what the f64 oracle — pinned by the hand-derived KATs of
tests/test_oracle_kat.py — answers today on fixed, deterministic inputs.  They guard against drift of the oracle
(CPU test) and give the CUDA path a fixed target that does not depend on the oracle being rebuilt (GPU test).

    python tests/golden/make_golden.py          # rewrites tests/golden/*.npz
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import orc  # noqa: E402
from rustraytracer_b200 import _native as N, make_rays, scenes  # noqa: E402

NONE = N.RT_NONE
RAY_CASES = [("sphere_roughness", {}, 64), ("cornell_box", {}, 64), ("plastic_dragon", {}, 64), ("material_hdr", {}, 64), ("soup", {"n_tris": 20000}, 64)]
LOBE_SCENES = [("sphere_roughness", {}), ("plastic_dragon", {}), ("plastic_dragon", {"variant": 1}), ("two_dragons", {"variant": 1}), ("material_hdr", {"variant": 3})]
FILM_CASES = [("sphere_roughness", {}, 25, 24, 4), ("cornell_box", {}, 25, 24, 4), ("plastic_dragon", {}, 8, 24, 2), ("material_hdr", {"variant": 3}, 12, 24, 2)]


def ray_batches(hs, oracle, size):
    """batch A = pixel-centre primary rays; batch B = one seeded diffuse bounce ray per primary hit."""
    cam = hs.camera()
    a = cam.primary_rays(size, size)
    ha = oracle.intersect(a, mode=1)
    hit = ha["prim_index"] != NONE
    rng = np.random.default_rng(3)
    d = rng.normal(size=(int(hit.sum()), 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    d *= np.sign(np.sum(d * ha["n"][hit], axis=1, keepdims=True))
    b = make_rays(ha["p"][hit], d)
    return a, b


def main():
    for name, kw, size in RAY_CASES:
        hs = scenes.HostScene(name, size, size, **kw)
        oracle = orc.OracleScene(hs.tables())
        a, b = ray_batches(hs, oracle, size)
        rays = np.concatenate([a, b])
        hits, ties = oracle.intersect(rays, mode=1, want_ties=True)
        np.savez_compressed(os.path.join(HERE, f"rays_{name}.npz"), origin=rays["origin"], dir=rays["dir"], prim=hits["prim_index"], t=hits["t"],
                            n=hits["n"], front=hits["front"].astype(np.uint8), ties=ties)
        print(name, len(rays), "rays,", int((hits["prim_index"] != NONE).sum()), "hits")
    for name, kw, depth, size, spp in FILM_CASES:
        hs = scenes.HostScene(name, size, size, **kw)
        oracle = orc.OracleScene(hs.tables())
        film, counts, _ = oracle.render(hs.camera(), size, size, 0, spp, seed=2024, max_depth=depth)
        tag = name + ("_v%d" % kw["variant"] if "variant" in kw else "")
        np.savez_compressed(os.path.join(HERE, f"film_{tag}.npz"), film=film, counts=counts, depth=depth, spp=spp, seed=2024)
        print(tag, "film mean", film[..., :3].mean(), "rays", counts)
    # lobe-level vectors: f / pdf / sample_f of every material of the built-in scenes at fixed inputs
    rows = []
    rng = np.random.default_rng(8)
    for si, (name, kw) in enumerate(LOBE_SCENES):
        hs = scenes.HostScene(name, 32, 32, **kw)
        oracle = orc.OracleScene(hs.tables())
        for mat in range(hs.counts()["materials"]):
            for _ in range(8):
                wo = rng.normal(size=3) * rng.uniform(0.5, 10); wi = rng.normal(size=3); wi /= np.linalg.norm(wi)
                u = rng.uniform(size=2)
                f, pdf = oracle.bsdf_f_pdf(mat, wo, wi, 31 - 16)
                sf, swi, spdf, sty = oracle.bsdf_sample_f(mat, wo, u, 31, seed=5, pixel=len(rows), smp=1)
                rows.append(np.concatenate([[si, len(rows), mat], wo, wi, u, f, [pdf], sf, swi, [spdf, sty]]))
    np.savez_compressed(os.path.join(HERE, "lobes.npz"), rows=np.asarray(rows))
    print("lobes", len(rows))


if __name__ == "__main__":
    main()
