"""Frozen fixtures (tests/golden/*.npz, made by tests/golden/make_golden.py from the oracle — the reference has no golden
vectors and cannot run here): the oracle must still reproduce them (CPU), and the CUDA path must match them (GPU)."""
import glob
import os
import sys

import numpy as np
import pytest

import orc
from rustraytracer_b200 import _native as N, make_rays, scenes

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
sys.path.insert(0, GOLD)
import make_golden as G  # noqa: E402

NONE = N.RT_NONE


def load_rays(z):
    return make_rays(z["origin"], z["dir"])


@pytest.mark.parametrize("name,kw,size", G.RAY_CASES)
def test_oracle_reproduces_golden_rays(name, kw, size):
    z = np.load(os.path.join(GOLD, f"rays_{name}.npz"))
    hs = scenes.HostScene(name, size, size, **kw)
    oracle = orc.OracleScene(hs.tables())
    a, b = G.ray_batches(hs, oracle, size)
    rays = np.concatenate([a, b])
    assert np.array_equal(rays["origin"], z["origin"]) and np.array_equal(rays["dir"], z["dir"]), "ray batches are not deterministic"
    hits = oracle.intersect(rays, mode=1)
    assert np.array_equal(hits["prim_index"], z["prim"])
    assert np.array_equal(hits["t"], z["t"])
    # the reference-shaped tree (both children, ties to the right) answers the same up to exact ties
    h0 = oracle.intersect(rays[:2048], mode=0)
    assert np.all((h0["prim_index"] == z["prim"][:2048]) | (h0["t"] == z["t"][:2048]))


@pytest.mark.parametrize("name,kw,depth,size,spp", G.FILM_CASES)
def test_oracle_reproduces_golden_films(name, kw, depth, size, spp):
    tag = name + ("_v%d" % kw["variant"] if "variant" in kw else "")
    z = np.load(os.path.join(GOLD, f"film_{tag}.npz"))
    hs = scenes.HostScene(name, size, size, **kw)
    oracle = orc.OracleScene(hs.tables())
    film, counts, _ = oracle.render(hs.camera(), size, size, 0, spp, seed=int(z["seed"]), max_depth=depth)
    assert np.array_equal(counts, z["counts"])
    assert np.allclose(film, z["film"], rtol=1e-13, atol=0, equal_nan=True)     # same binary, same libm: identical up to nothing


def test_oracle_reproduces_golden_lobes():
    rows = np.load(os.path.join(GOLD, "lobes.npz"))["rows"]
    k = 0
    for si, (name, kw) in enumerate(G.LOBE_SCENES):
        hs = scenes.HostScene(name, 32, 32, **kw)
        oracle = orc.OracleScene(hs.tables())
        while k < len(rows) and int(rows[k][0]) == si:
            r = rows[k]; mat = int(r[2]); wo, wi, u = r[3:6], r[6:9], r[9:11]
            f, pdf = oracle.bsdf_f_pdf(mat, wo, wi, 31 - 16)
            sf, swi, spdf, sty = oracle.bsdf_sample_f(mat, wo, u, 31, seed=5, pixel=int(r[1]), smp=1)
            got = np.concatenate([f, [pdf], sf, swi, [spdf, sty]])
            assert np.allclose(got, r[11:], rtol=1e-13, atol=0, equal_nan=True), (name, kw, mat)
            k += 1
    assert k == len(rows)


@pytest.mark.gpu
@pytest.mark.parametrize("name,kw,size", G.RAY_CASES)
def test_gpu_matches_golden_rays(name, kw, size):
    from rustraytracer_b200 import get_intersection
    z = np.load(os.path.join(GOLD, f"rays_{name}.npz"))
    gpu = scenes.HostScene(name, size, size, **kw).upload()
    got = get_intersection(gpu, load_rays(z))
    same = got["prim_index"] == z["prim"]
    in_tie = np.array([g in t for g, t in zip(got["prim_index"], z["ties"])]) & (got["prim_index"] != NONE)
    assert (same | in_tie).all(), f"{(~(same | in_tie)).sum()} closest-hit ids differ from the golden batch"
    hit = same & (z["prim"] != NONE)
    assert np.allclose(got["t"][hit], z["t"][hit], rtol=1e-12, atol=0)
    assert np.allclose(got["n"][hit], z["n"][hit], rtol=0, atol=1e-10)
    assert np.array_equal(got["front"][hit], z["front"][hit])


@pytest.mark.gpu
@pytest.mark.parametrize("name,kw,depth,size,spp", G.FILM_CASES)
def test_gpu_matches_golden_films(name, kw, depth, size, spp):
    from rustraytracer_b200 import Film, IntType, render_samples
    tag = name + ("_v%d" % kw["variant"] if "variant" in kw else "")
    z = np.load(os.path.join(GOLD, f"film_{tag}.npz"))
    hs = scenes.HostScene(name, size, size, **kw)
    gpu = hs.upload()
    film = Film(size, size)
    render_samples(gpu, hs.camera(), IntType.Path(depth), film, size, size, 0, spp, seed=int(z["seed"]))
    g, o = film.read(), z["film"]
    assert np.array_equal(g[..., 3], o[..., 3])
    err = np.abs(g[..., :3] - o[..., :3]).max(-1) / (np.abs(o[..., :3]).max(-1) + 1e-6)
    assert (err > 1e-7).mean() <= 4e-3, f"{(err > 1e-7).sum()} pixels differ from the golden film"
    c = gpu.counters()
    for k, v in zip(("extend_rays", "shadow_rays", "probe_rays"), z["counts"]):
        assert abs(c[k] - int(v)) <= 0.003 * int(v) + 8
