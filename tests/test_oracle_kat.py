"""Known-answer tests that pin the oracle.  The reference has no tests or golden
vectors of its own ("parity unpinned"), so these are hand-derived from its source: analytic
ray/sphere, ray/rect, ray/triangle cases and the reference's documented quirks (§8.1)."""
import math

import numpy as np
import pytest

import orc
from rustraytracer_b200 import _native as N
from rustraytracer_b200.core import make_rays

NONE = N.RT_NONE


def one(scene, o, d, tmin=0.001, tmax=1e308, mode=1):
    return scene.intersect(make_rays([o], [d]), tmin, tmax, mode=mode, threads=1)[0]


def test_philox_known_answer():
    # Random123 kat_vectors: philox4x32-10, counter = 0, key = 0
    import ctypes as C
    c = (C.c_uint32 * 4)(0, 0, 0, 0)
    orc.lib().orc_philox(c, 0, 0)
    assert [hex(x) for x in c] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    c = (C.c_uint32 * 4)(0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF)
    orc.lib().orc_philox(c, 0xFFFFFFFF, 0xFFFFFFFF)
    assert [hex(x) for x in c] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    u = orc.lib().orc_rng_uniform(123, 4, 5, 6)
    assert 0.0 <= u < 1.0


def sphere_scene(c=(0, 0, 0), r=1.0):
    return orc.OracleScene(orc.simple_tables([orc.prim(N.PRIM_SPHERE, [*c, r])]))


def test_sphere_axis_ray_unnormalised_dir():
    s = sphere_scene()
    # |dir| = 2: t is in units of |dir| (intersects.rs:177-213): hit at distance 4 → t = 2
    h = one(s, (-5, 0, 0), (2, 0, 0))
    assert h["prim_index"] == 0 and h["t"] == pytest.approx(2.0, abs=1e-15)
    assert np.allclose(h["p"], (-1, 0, 0), atol=1e-12)
    # geometric normal faces the ray after set_front; shading normal of a sphere points INWARD (§8a row 7)
    assert np.allclose(h["n"], (-1, 0, 0), atol=1e-12)
    assert np.allclose(h["ns"], (1, 0, 0), atol=1e-9)
    assert h["front"] == 0
    # pole quirk: p.x == p.y == 0 → p.x = 1e-5 * r "so that atan2 exists" (intersects.rs:219-221)
    h = one(s, (0, 0, -5), (0, 0, 1))
    assert h["t"] == pytest.approx(4.0) and np.allclose(h["p"], (1e-5, 0, -1), atol=1e-15)
    # from inside: first root negative → second root
    h = one(s, (0, 0, 0), (1, 0, 0))
    assert h["t"] == pytest.approx(1.0) and np.allclose(h["n"], (-1, 0, 0), atol=1e-12) and h["front"] == 1
    # miss; the axis-parallel grazing ray has disc == 0 (a sphere hit) but dies in the leaf-box slab test:
    # (max - o) * (1/0) = 0 * inf = NaN, and Rust's f64::min/max drop the NaN → t1 = -inf (hittable.rs:494-508)
    assert one(s, (0, 2, -5), (0, 0, 1))["prim_index"] == NONE
    assert one(s, (0, 1, -5), (0, 0, 1))["prim_index"] == NONE
    # strict tmin/tmax
    assert one(s, (0, 0, -5), (0, 0, 1), tmin=4.0, tmax=1e308)["t"] == pytest.approx(6.0)
    assert one(s, (0, 0, -5), (0, 0, 1), tmin=0.001, tmax=4.0)["prim_index"] == NONE


def test_sphere_uv():
    s = sphere_scene()
    h = one(s, (5, 0, 0), (-1, 0, 0))        # hits (1,0,0): phi = 0, theta = acos(0) = pi/2
    assert h["uv"][0] == pytest.approx(0.0, abs=1e-12) and h["uv"][1] == pytest.approx(0.5 * math.pi / 3.14159265358979, rel=1e-12)
    h = one(s, (0, 5, 0.0), (0, -1, 0))      # hits (0,1,0): phi = pi/2 → u = (pi/2)/(2*PI_ref)
    assert h["uv"][0] == pytest.approx((math.pi / 2) / (2 * 3.14159265358979), rel=1e-12)


def rect_scene(ptype, v, flip=0, xf=None):
    transforms = [] if xf is None else [xf]
    return orc.OracleScene(orc.simple_tables([orc.prim(ptype, v, flip=flip, xf=0 if xf is not None else NONE)], transforms=transforms))


def test_rect_centre_edge_corner_inclusive():
    s = rect_scene(N.PRIM_XZ_RECT, [-1, -2, 1, 2, 0.5])      # x∈[-1,1], z∈[-2,2], y = 0.5
    h = one(s, (0, 3, 0), (0, -1, 0))
    assert h["prim_index"] == 0 and h["t"] == pytest.approx(2.5) and np.allclose(h["p"], (0, 0.5, 0))
    assert np.allclose(h["uv"], (0.5, 0.5))
    # n = normalize(dpdu x dpdv) = x̂ × ẑ = -ŷ, flipped toward the ray → +ŷ
    assert np.allclose(h["n"], (0, 1, 0)) and np.allclose(h["ns"], (0, -1, 0)) and h["front"] == 0
    assert one(s, (0, 3, 0), (1, -2.5, 2))["prim_index"] == 0        # exactly the corner (1, 0.5, 2): bounds are inclusive
    assert one(s, (1, 3, 2), (0, -1, 0))["prim_index"] == NONE       # same corner, axis-parallel: 0*inf = NaN kills the slab test
    assert one(s, (1.0000001, 3, 0), (0, -1, 0))["prim_index"] == NONE
    # parallel ray: t = ±inf/NaN → miss
    assert one(s, (0, 3, 0), (1, 0, 0))["prim_index"] == NONE
    # FlipFace toggles only `front`
    sf = rect_scene(N.PRIM_XZ_RECT, [-1, -2, 1, 2, 0.5], flip=1)
    hf = one(sf, (0, 3, 0), (0, -1, 0))
    assert hf["front"] == 1 and np.allclose(hf["n"], h["n"])


def test_rect_leaf_box_padding_and_tmin():
    s = rect_scene(N.PRIM_XY_RECT, [0, 0, 1, 1, 0.0])
    # box is z ∈ [-SMALL, SMALL] (primitive.rs:93-94); a hit at t just above tmin passes
    assert one(s, (0.5, 0.5, -0.01), (0, 0, 1))["prim_index"] == 0
    # rect test is inclusive in t (t < t0 rejects) but the leaf box needs tmax > tmin strictly
    assert one(s, (0.5, 0.5, -0.0005), (0, 0, 1), tmin=0.001)["prim_index"] == NONE       # t = 0.0005 < tmin
    b = s.prim_box(0)
    assert np.allclose(b, (0, 0, -0.001, 1, 1, 0.001))


def test_transformed_rect_cornell_cube_face():
    a = 15 * 3.14159265358979 / 180
    ca, sa = math.cos(a), math.sin(a)
    xf = [ca, 0, sa, 265, 0, 1, 0, 0, -sa, 0, ca, 295, 0, 0, 0, 1]       # translate · Ry(15°)  (scenes.rs:169-176)
    s = rect_scene(N.PRIM_XY_RECT, [0, 0, 165, 330, 165], xf=xf)           # z1 face of cube2
    # centre of the face in object space (82.5, 165, 165) → world
    pw = np.array([ca * 82.5 + sa * 165 + 265, 165, -sa * 82.5 + ca * 165 + 295])
    nw = np.array([sa, 0, ca])
    o = pw + 100 * nw
    h = one(s, o, -nw)
    assert h["prim_index"] == 0 and h["t"] == pytest.approx(100, rel=1e-12) and np.allclose(h["p"], pw, atol=1e-9)
    assert np.allclose(h["n"], nw, atol=1e-12) and np.allclose(h["uv"], (0.5, 0.5), atol=1e-12)


def tri_scene(p, n=None, uv=None):
    p = np.asarray(p, dtype=np.float64)
    mesh = (p, np.zeros((0, 3)) if n is None else np.asarray(n, float), np.zeros((0, 2)) if uv is None else np.asarray(uv, float), np.arange(len(p), dtype=np.uint32))
    prims = [orc.prim(N.PRIM_TRIANGLE, [], mesh=0, ind=3 * i) for i in range(len(p) // 3)]
    return orc.OracleScene(orc.simple_tables(prims, meshes=[mesh]))


TRI = [(0, 0, 0.1), (1, 0, 0.0), (0, 1, -0.1)]       # slightly tilted so its leaf box has thickness


def test_triangle_centre_edge_backface():
    s = tri_scene(TRI)
    h = one(s, (0.25, 0.25, -3), (0, 0, 1))
    zc = 0.1 * (1 - 0.25 - 0.25) + 0.0 * 0.25 - 0.1 * 0.25
    assert h["prim_index"] == 0 and h["t"] == pytest.approx(3 + zc, rel=1e-12) and np.allclose(h["p"], (0.25, 0.25, zc), atol=1e-12)
    # no uvs → fixed (0,0),(1,0),(1,1) (hittable.rs:455-460): uv = b1*(1,0) + b2*(1,1)
    assert np.allclose(h["uv"], (0.5, 0.25), atol=1e-12)
    n = np.cross(np.subtract(TRI[0], TRI[2]), np.subtract(TRI[1], TRI[2])); n /= np.linalg.norm(n)
    assert np.allclose(np.abs(h["n"]), np.abs(n), atol=1e-12) and np.dot(h["n"], (0, 0, 1)) < 0
    hb = one(s, (0.25, 0.25, 3), (0, 0, -1))          # back side: also a hit (no culling), normal flipped
    assert hb["prim_index"] == 0 and np.dot(hb["n"], (0, 0, -1)) < 0 and hb["front"] != h["front"]
    assert one(s, (0.8, 0.8, -3), (0, 0, 1))["prim_index"] == NONE     # outside
    # exactly on the edge y = 0 with an axis-parallel ray: the triangle test accepts (e == 0 is not "mixed sign")
    # but the leaf box's (min.y - o.y) * (1/0) = NaN makes the slab test fail → invisible
    assert one(s, (0.5, 0.0, -3), (0, 0, 1))["prim_index"] == NONE
    assert one(s, (0.5, 1e-9, -3), (0, 0, 1))["prim_index"] == 0


def test_triangle_rejects_small_t_and_small_det():
    s = tri_scene(TRI)
    # t < 1e-4 rejected whatever tmin is (hittable.rs:360); tmin is ignored by the triangle test
    zc = 0.025
    assert one(s, (0.25, 0.25, zc - 0.00005), (0, 0, 1), tmin=0.0)["prim_index"] == NONE
    h = one(s, (0.25, 0.25, zc - 0.0005), (0, 0, 1), tmin=0.001)   # t = 0.0005 ∈ [1e-4, tmin): still a hit — box straddles tmin
    assert h["prim_index"] == 0
    # |det| < 1e-7 is scale dependent: a tiny triangle is invisible (hittable.rs:339-343)
    tiny = [(0, 0, 1e-5), (1e-4, 0, 0), (0, 1e-4, -1e-5)]
    assert one(tri_scene(tiny), (2e-5, 2e-5, -1), (0, 0, 1))["prim_index"] == NONE


def test_axis_aligned_triangle_is_invisible():
    # zero-thickness leaf box: the slab test `tmax <= tmin` always fails (hittable.rs:503, §8.1 #3)
    flat = [(0, 0, 0), (1, 0, 0), (0, 1, 0)]
    s = tri_scene(flat)
    assert one(s, (0.25, 0.25, -3), (0, 0, 1), mode=1)["prim_index"] == NONE
    assert one(s, (0.25, 0.25, -3), (0, 0, 1), mode=0)["prim_index"] == NONE
    # but a slanted ray still misses: the box is empty in z for every ray
    assert one(s, (0.0, 0.0, -3), (0.1, 0.1, 1), mode=0)["prim_index"] == NONE


def test_triangle_shading_normal_interpolation_and_degenerate_uv():
    nrm = [(0, 0, -1), (0.6, 0, -0.8), (0, 0.6, -0.8)]
    s = tri_scene(TRI, n=nrm)
    h = one(s, (0.25, 0.25, -3), (0, 0, 1))
    # shading frame: ns = normalize(ss × ts) with ts = normalize(ns0 × ss), ss = normalize(ts × ns0) → equals ±ns0
    ns0 = 0.5 * np.array(nrm[0]) + 0.25 * np.array(nrm[1]) + 0.25 * np.array(nrm[2]); ns0 /= np.linalg.norm(ns0)
    assert np.allclose(h["ns"], ns0, atol=1e-12)
    # geometric normal is first faced toward the shading normal, then toward the ray
    assert np.dot(h["n"], (0, 0, 1)) < 0
    # degenerate uvs (all equal) + non-degenerate geometry: falls back to make_coordinate_system, still a hit
    s2 = tri_scene(TRI, uv=[(0.5, 0.5)] * 3)
    assert one(s2, (0.25, 0.25, -3), (0, 0, 1))["prim_index"] == 0
    # degenerate uvs + zero-area triangle → miss (hittable.rs:373-378); such a triangle fails `det` anyway
    s3 = tri_scene([(0, 0, 0.1), (1, 0, 0), (2, 0, -0.1)], uv=[(0.5, 0.5)] * 3)
    assert one(s3, (0.5, 0.0, -3), (0, 0, 1))["prim_index"] == NONE


def test_closest_of_two_and_ties():
    t2 = TRI + [(0, 0, 1.1), (1, 0, 1.0), (0, 1, 0.9)]
    s = tri_scene(t2)
    assert one(s, (0.25, 0.25, -3), (0, 0, 1))["prim_index"] == 0
    assert one(s, (0.25, 0.25, 3), (0, 0, -1))["prim_index"] == 1
    # two identical triangles: an exact tie; brute force reports both in the tie set
    s = tri_scene(TRI + TRI)
    hits, ties = s.intersect(make_rays([(0.25, 0.25, -3)], [(0, 0, 1)]), mode=1, threads=1, want_ties=True)
    assert set(ties[0][:2]) == {0, 1}
    h0 = s.intersect(make_rays([(0.25, 0.25, -3)], [(0, 0, 1)]), mode=0, threads=1)[0]
    assert h0["prim_index"] in (0, 1)


def test_unoccluded_semantics():
    # light rect above, blocker sphere between, receiver point on a floor
    prims = [orc.prim(N.PRIM_XZ_RECT, [-1, -1, 1, 1, 5.0]), orc.prim(N.PRIM_SPHERE, [0, 2.5, 0, 0.5]), orc.prim(N.PRIM_XZ_RECT, [-10, -10, 10, 10, 0.0])]
    s = orc.OracleScene(orc.simple_tables(prims))
    p0 = np.array([[0.0, 0, 0], [3.0, 0, 0], [3.0, 0, 0], [2.0, 0, 0]])
    p1 = np.array([[0.0, 5, 0], [0.5, 5, 0.5], [0.5, 5, 0.5], [-0.8, 7, 0.0]])
    prim = np.array([0, 0, 1, 0], dtype=np.uint32)
    out = s.occluded(p0, p1, prim, threads=1)
    # 0: blocked by the sphere; 1: clear, first hit is the light; 2: clear but p1.prim_index is wrong → "occluded";
    # 3: p1 above the light: the ray still hits the light rect first → visible (prim matches)
    assert list(out) == [False, True, False, True]
    # no hit at all ⇒ occluded for a finite light, visible for an infinite one (hittable.rs:31-33)
    out = s.occluded(np.array([[20.0, 1, 0]]), np.array([[30.0, 2, 0]]), np.array([0], dtype=np.uint32), threads=1)
    assert list(out) == [False]
    out = s.occluded(np.array([[20.0, 1, 0]]), np.array([[30.0, 2, 0]]), np.array([0], dtype=np.uint32), infinite=True, threads=1)
    assert list(out) == [True]


def test_tree_modes_agree_on_random_scene():
    rng = np.random.default_rng(5)
    n = 300
    c = rng.uniform(-1, 1, (n, 3)); e1 = rng.uniform(-0.3, 0.3, (n, 3)); e2 = rng.uniform(-0.3, 0.3, (n, 3))
    p = np.stack([c, c + e1, c + e2], 1).reshape(-1, 3)
    mesh = (p, np.zeros((0, 3)), np.zeros((0, 2)), np.arange(3 * n, dtype=np.uint32))
    prims = [orc.prim(N.PRIM_TRIANGLE, [], mesh=0, ind=3 * i) for i in range(n)]
    prims += [orc.prim(N.PRIM_SPHERE, [0.3, 0.2, 0.1, 0.4]), orc.prim(N.PRIM_XZ_RECT, [-2, -2, 2, 2, -1.0], flip=1)]
    s = orc.OracleScene(orc.simple_tables(prims, meshes=[mesh]), bvh_seed=3)
    o = rng.uniform(-3, 3, (4000, 3)); d = rng.normal(size=(4000, 3)) * rng.uniform(0.1, 10, (4000, 1))
    rays = make_rays(o, d)
    h0 = s.intersect(rays, mode=0)
    h1, ties = s.intersect(rays, mode=1, want_ties=True)
    h2 = s.intersect(rays, mode=2)
    assert (h0["prim_index"] != NONE).sum() > 500
    for h in (h0, h2):
        same = h["prim_index"] == h1["prim_index"]
        in_tie = np.array([h["prim_index"][i] in ties[i] for i in range(len(rays))])
        assert np.all(same | in_tie)
        assert np.array_equal(h["t"][same], h1["t"][same])      # bit-identical t: the same arithmetic ran
    # a different random tree gives the same hit set
    s2 = orc.OracleScene(orc.simple_tables(prims, meshes=[mesh]), bvh_seed=99)
    h3 = s2.intersect(rays, mode=0)
    assert np.array_equal(h3["prim_index"], h0["prim_index"]) or np.all((h3["prim_index"] == h0["prim_index"]) | (h3["t"] == h0["t"]))
