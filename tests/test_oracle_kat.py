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
    # the hit-set traversal over the tree (mode 3, what the 1 M - 50 M triangle soups are checked with) returns exactly what brute force
    # returns — same primitive, same t, same tie set — over the faithful tree and over the quick nth_element tree
    for fast in (False, True):
        if fast:
            s.build_fast_tree(threads=4)
        h4, ties4 = s.intersect(rays, mode=3, want_ties=True)
        assert np.array_equal(h4["prim_index"], h1["prim_index"]) and np.array_equal(h4["t"], h1["t"])
        assert np.array_equal(np.sort(ties4, axis=1), np.sort(ties, axis=1))
    # a different random tree gives the same hit set
    s2 = orc.OracleScene(orc.simple_tables(prims, meshes=[mesh]), bvh_seed=99)
    h3 = s2.intersect(rays, mode=0)
    assert np.array_equal(h3["prim_index"], h0["prim_index"]) or np.all((h3["prim_index"] == h0["prim_index"]) | (h3["t"] == h0["t"]))


# ---------------------------------------------------------------- Hdr / Image textures, Distribution, Light::Infinite
def env_tables(img, tex_type=N.TEX_HDR, prims=(), transform=None):
    """One environment texture (id 0) + a grey matte (texture 1, material 0) + one Infinite light."""
    tex = np.zeros(2, dtype=N.texture_dtype)
    tex[0]["type"] = tex_type; tex[0]["image"] = 0
    tex[1]["color"] = (0.5, 0.5, 0.5)
    mat = np.zeros(1, dtype=N.material_dtype); mat[0]["tex"] = N.RT_NONE; mat[0]["tex"][0] = 1
    lig = np.zeros(1, dtype=N.light_dtype)
    lig[0]["type"] = N.LIGHT_INFINITE; lig[0]["prim_index"] = N.RT_NONE; lig[0]["text_id"] = 0
    lig[0]["transform"] = N.RT_NONE if transform is None else 0
    return orc.simple_tables(list(prims), textures=tex, materials=mat, lights=lig, images=[img],
                             transforms=[] if transform is None else [transform])


def test_hdr_texture_requantisation_and_lookup():
    # 2x2 Hdr image; material.rs:570-587: x = round((1-u)*W) % W, y = round(v*H) % H, to_rgbe8, (c+0.5)*2^(e-128)/256
    img = np.zeros((2, 2, 3), dtype=np.float32)
    img[0, 0] = (1.0, 0.5, 0.25)        # mx = 1 -> exp = floor(log2 1)+1 = 1, mul = 2, c = (128, 64, 32), e = 129
    img[0, 1] = (0.75, 0.75, 0.75)      # mx = .75 -> exp = 0, mul = 1, c = 192, e = 128
    img[1, 0] = (0.0, 0.0, 0.0)         # mx <= 0 -> (0,0,0,0) -> (0.5) * 2^-128 / 256
    img[1, 1] = (3.0, 100.0, 7.0)       # mx = 100 -> exp = 7, mul = 128, c = (6, 200, 14), e = 135
    s = orc.OracleScene(env_tables(img))
    # u = 1 -> x = round(0) = 0 ; v = 0 -> y = 0
    assert np.array_equal(s.tex_value(0, 1.0, 0.0), [128.5 * 2 / 256, 64.5 * 2 / 256, 32.5 * 2 / 256])
    # u = 0.5 -> x = round(1.0) = 1
    assert np.array_equal(s.tex_value(0, 0.5, 0.0), [192.5 / 256] * 3)
    # u = 0 -> x = round(2) % 2 = 0 ; v = 0.5 -> y = 1
    assert np.array_equal(s.tex_value(0, 0.0, 0.5), [0.5 * 2.0 ** -128 / 256] * 3)
    # u = 0.26 -> (1-u)*2 = 1.48 -> 1 ; v = 0.74 -> 1.48 -> 1 ; v = 0.76 -> 1.52 -> 2 % 2 = 0
    assert np.array_equal(s.tex_value(0, 0.26, 0.74), [6.5 * 128 / 256, 200.5 * 128 / 256, 14.5 * 128 / 256])
    assert np.array_equal(s.tex_value(0, 0.26, 0.76), [192.5 / 256] * 3)
    # out-of-range coordinates: `as u32` saturates negatives to 0; large values wrap by the modulo
    assert np.array_equal(s.tex_value(0, 1.7, -0.3), s.tex_value(0, 1.0, 0.0))


def test_image_texture_lookup():
    # material.rs:601-611: x = round(u*W) % W, y = round((1-v)*H) % H, c / 255
    img = np.arange(2 * 3 * 3, dtype=np.uint8).reshape(2, 3, 3) * 10
    s = orc.OracleScene(env_tables(img, tex_type=N.TEX_IMAGE))
    assert np.array_equal(s.tex_value(0, 0.0, 1.0), img[0, 0] / 255.0)
    assert np.array_equal(s.tex_value(0, 0.4, 0.6), img[1, 1] / 255.0)      # round(1.2) = 1 ; round(0.8) = 1
    assert np.array_equal(s.tex_value(0, 0.9, 0.0), img[0, 0] / 255.0)      # round(2.7) = 3 % 3 = 0 ; round(2) % 2 = 0


def test_distribution_row_slicing_and_sampling():
    # A 1x1 SolidColor-like case is degenerate, so use a 1x2 Hdr image -> table 2x4 (nu = 2, nv = 4).
    img = np.zeros((2, 1, 3), dtype=np.float32); img[0, 0] = (1, 1, 1); img[1, 0] = (0.5, 0.5, 0.5)
    s = orc.OracleScene(env_tables(img))
    d = s.light_dist(0)
    assert (d["nu"], d["nv"]) == (2, 4)
    # f[u + v*2] = lum(texel(v)) * sin(PI*(v+.5)/4), texel row = round(vp*2) % 2 with vp = (v+.5)/4 -> rows 0,1,1,0
    PI = 3.14159265358979
    lum = lambda c: 0.2126 * c + 0.7152 * c + 0.0722 * c
    texel = [128.5 * 2 / 256, 128.5 / 256, 128.5 / 256, 128.5 * 2 / 256]       # (1,1,1) -> c=128,e=129 ; (.5,.5,.5) -> c=128,e=128
    f = np.array([lum(texel[v]) * math.sin(PI * (v + 0.5) / 4) for v in range(4) for _ in range(2)])
    # distribution.rs:120: row v uses the SLICE f[v .. v+nu), not f[v*nu ..]
    for v in range(4):
        row = f[v:v + 2]
        c1 = 0.0 + row[0] / 2; c2 = c1 + row[1] / 2
        assert d["row_int"][v] == c2
        assert np.array_equal(d["cdf"][v], [0.0, c1 / c2, c2 / c2])
    m = d["row_int"]
    acc = [0.0]
    for v in range(4):
        acc.append(acc[-1] + m[v] / 4)
    assert d["marg_int"] == acc[4]
    assert np.array_equal(d["marg_cdf"], [a / acc[4] if i else 0.0 for i, a in enumerate(acc)])


def test_infinite_light_sample_pdf_le_consistency():
    rng = np.random.default_rng(5)
    img = rng.uniform(0.05, 4.0, (8, 16, 3)).astype(np.float32)
    s = orc.OracleScene(env_tables(img))
    PI = 3.14159265358979
    # le: Y-up lookup (light.rs:499-513): direction +y -> theta = 0 -> v = 0 ; phi = atan2(z, x)
    assert np.array_equal(s.light_le(0, [0, 2, 0]), s.tex_value(0, 0.0, 0.0))
    assert np.array_equal(s.light_le(0, [0, 0, 3]), s.tex_value(0, 0.25, 0.5))       # phi = PI/2 (of the 15-digit PI), theta = PI/2
    assert np.array_equal(s.light_le(0, [-1, 0, 0]), s.tex_value(0, math.atan2(0.0, -1.0) / (2 * PI), math.acos(0.0) / PI))
    # sample_li builds the direction Z-up (light.rs:233): wi = (sin t cos p, sin t sin p, cos t), p1 = p + wi * 20000
    p = np.array([1.0, 2.0, 3.0])
    o = s.light_sample_li(0, p, [0.37, 0.81])
    assert o["pdf"] > 0 and abs(np.linalg.norm(o["wi"]) - 1) < 1e-15
    assert np.allclose(o["p1"], p + o["wi"] * 20000.0, rtol=1e-12)
    # the colour returned is the texture at the sampled (u, v), where v = acos(wi.z)/PI, u = atan2(wi.y, wi.x)/2PI
    theta = math.acos(o["wi"][2]); phi = math.atan2(o["wi"][1], o["wi"][0]) % (2 * PI)
    assert np.array_equal(o["color"], s.tex_value(0, phi / (2 * PI), theta / PI))
    # the solid-angle pdf of the sample is map_pdf / (2 PI^2 sin theta); dist.pdf at the same (u,v) equals map_pdf
    d = s.light_dist(0)
    # pdf_li reads the SAME table Y-up: feed it the direction whose Y-up angles are (theta, phi)
    w_yup = [math.sin(theta) * math.cos(phi), math.cos(theta), math.sin(theta) * math.sin(phi)]
    assert math.isclose(s.light_pdf_li(0, p, w_yup), o["pdf"], rel_tol=1e-9)
    # degenerate: straight up has sin(theta) = 0 -> pdf 0
    assert s.light_pdf_li(0, p, [0, 1, 0]) == 0.0
    # unoccluded(infinite = true): visible iff nothing is hit
    t = env_tables(img, prims=[orc.prim(N.PRIM_SPHERE, [0, 50, 0, 1.0]), orc.prim(N.PRIM_SPHERE, [0, 0, 5, 1.0])])
    s2 = orc.OracleScene(t)
    # §8.1 #14: |dir| = 2e4, so the shadow-ray origin is pushed 20 world units along wi — an occluder closer than that is skipped
    assert s2.occluded([(0, 0, 0)], [(0, 0, 20000)], [0], infinite=True)[0]
    # (OracleScene.occluded returns the reference's `unoccluded` boolean)
    assert not s2.occluded([(0, 0, 0)], [(0, 20000, 0)], [0], infinite=True)[0]      # the sphere is in the way
    assert s2.occluded([(0, 0, 0)], [(20000, 0, 0)], [0], infinite=True)[0]          # nothing hit -> visible
    assert not s2.occluded([(0, 0, 0)], [(20000, 0, 0)], [0], infinite=False)[0]     # area-light semantics: no hit -> occluded


def test_direct_and_whitted_integrators_oracle_properties():
    """IntType::Direct / Whitted (integrator.rs:131-165, 257-297, 464-528) in the oracle, against what their definitions imply:
    * on a scene without specular lobes, Direct{UniformOne} with max_depth 1 IS Path with max_depth 1 (emission + one
      uniform_sample_one_light at the first vertex, same random numbers);
    * Whitted and Direct{UniformAll} estimate the same direct light (no MIS vs MIS) — equal image means within noise;
    * recursion only adds non-negative radiance along specular lobes: deeper max_depth never darkens a pixel, and changes
      nothing on a scene that has no specular lobe."""
    from rustraytracer_b200 import scenes
    w = h = 40
    hs = scenes.HostScene("cornell_box", w, h)
    s = orc.OracleScene(hs.tables()); cam = hs.camera()
    path1, _, _ = s.render(cam, w, h, 0, 4, seed=7, max_depth=1, kind=0)
    dir1, c1, _ = s.render(cam, w, h, 0, 4, seed=7, max_depth=1, kind=1)
    assert np.allclose(dir1, path1, rtol=1e-12, atol=1e-14)
    assert c1[0] == w * h * 4                                          # one extend ray per camera sample at depth 1
    dir5, _, _ = s.render(cam, w, h, 0, 4, seed=7, max_depth=5, kind=1)
    assert np.array_equal(dir5, dir1)                                  # matte + light materials only: no specular lobe to follow
    whit, _, _ = s.render(cam, w, h, 0, 64, seed=3, max_depth=3, kind=3)
    dall, _, _ = s.render(cam, w, h, 0, 64, seed=3, max_depth=3, kind=2)
    mw, md = whit[..., :3].mean(), dall[..., :3].mean()
    assert mw > 0 and abs(mw - md) <= 0.03 * md
    # mirrors and glass: make_world (Lambertian + mirror + glass spheres)
    hs = scenes.HostScene("make_world", w, h)
    s = orc.OracleScene(hs.tables()); cam = hs.camera()
    for kind in (1, 3):
        d1, n1, _ = s.render(cam, w, h, 0, 2, seed=9, max_depth=1, kind=kind)
        d6, n6, _ = s.render(cam, w, h, 0, 2, seed=9, max_depth=6, kind=kind)
        assert np.all(d6[..., :3] >= d1[..., :3] - 1e-12)
        assert n6[0] > n1[0] and np.any(d6[..., :3] > d1[..., :3] + 1e-9)     # the mirror spheres do recurse


def test_basic_integrator_oracle_known_answers():
    """IntType::Basic → geometry::cast_ray (geometry.rs:242-277) in the oracle: a camera ray that hits the light returns its emission
    and nothing else; with max_depth 1 every other pixel is black (the walk ends before a light can be reached); a second vertex only
    ever adds light; the film divides by width - 1 (integrator.rs:197-200), so it differs from Path's first vertex only through that."""
    from rustraytracer_b200 import scenes
    w = h = 40
    hs = scenes.HostScene("cornell_box", w, h)
    s = orc.OracleScene(hs.tables()); cam = hs.camera()
    b1, n1, _ = s.render(cam, w, h, 0, 2, seed=5, max_depth=1, kind=4)
    assert n1[0] == w * h * 2 and n1[1] == 0 and n1[2] == 0
    rgb = b1[..., :3]
    lit = rgb.max(-1) > 0
    assert 0 < lit.sum() < 0.2 * w * h                                    # only the pixels that look straight at the ceiling light
    vals = np.unique(np.round(rgb[lit] / 2.0, 9), axis=0)               # 2 spp: one or two samples saw the light, each adds the same Le
    le = vals.max(0)
    assert np.all(le > 0) and all(np.allclose(v, le) or np.allclose(v * 2, le) for v in vals)
    b3, n3, _ = s.render(cam, w, h, 0, 2, seed=5, max_depth=3, kind=4)
    assert np.all(b3[..., :3] >= rgb - 1e-12) and (b3[..., :3].max(-1) > 0).sum() > lit.sum() and n3[0] > n1[0]


def _perlin_noise_py(scale, perm, vecs, p):
    """perlin.rs:36-85 restated with Python floats, for the oracle's Perlin to be checked against"""
    x, y, z = (c * scale for c in p)
    def cell(f): return 0 if not (math.floor(f) > 0) else int(math.floor(f))            # `as usize` saturates
    i0, j0, k0 = cell(x), cell(y), cell(z)
    def fract(f):
        r = f - math.trunc(f)
        return r + 1.0 if r < 0 else r
    u, v, w = fract(x), fract(y), fract(z)
    c = {}
    for di in range(2):
        for dj in range(2):
            for dk in range(2):
                c[(di, dj, dk)] = vecs[int(perm[0][(i0 + di) & 255]) ^ int(perm[1][(j0 + dj) & 255]) ^ int(perm[2][(k0 + dk) & 255])]
    u, v, w = (t * t * (3.0 - 2.0 * t) for t in (u, v, w))
    acc = 0.0
    for i in range(2):
        for j in range(2):
            for k in range(2):
                g = c[(i, j, k)]
                acc += (i * u + (1 - i) * (1 - u)) * (j * v + (1 - j) * (1 - v)) * (k * w + (1 - k) * (1 - w)) * (g[0] * (u - i) + g[1] * (v - j) + g[2] * (w - k))
    return acc


def test_perlin_texture_known_answers():
    """Texture::Perlin (material.rs:566-569) over perlin.rs: closed form for constant gradients, a Python restatement for random
    tables (negative coordinates included: the lattice index saturates at 0 while the fraction wraps), turbulence and the marble formula."""
    const = (2.0, np.tile(np.arange(256, dtype=np.uint32), (3, 1)), np.tile(np.array([0.0, 1.0, 0.0]), (256, 1)))
    rnd = orc.perlin_tables(3, 1.7)
    tex = np.zeros(3, dtype=N.texture_dtype)
    tex[0]["type"] = N.TEX_PERLIN; tex[0]["image"] = 0
    tex[1]["type"] = N.TEX_PERLIN; tex[1]["image"] = 1
    tex[2]["type"] = N.TEX_CHECKERED; tex[2]["odd"] = 0; tex[2]["even"] = 1; tex[2]["frequency"] = 3.0
    s = orc.OracleScene(orc.simple_tables([orc.prim(N.PRIM_SPHERE, [0, 0, 0, 1.0])], textures=tex, perlins=[const, rnd]))
    # every gradient = +y (v' = smoothed fraction of y): corners j = 0 carry weight (1 - v') and value v', corners j = 1 carry weight v'
    # and value v' - 1, so noise = (1 - v') v' + v' (v' - 1) = 0 everywhere — the weights use the SMOOTHED fraction (perlin.rs:71-79)
    for p in ([0.3, 0.4, 0.5], [1.25, -0.7, 3.3], [-2.1, 5.55, -0.2]):
        assert abs(s.perlin_noise(0, p)) < 1e-15
    # lattice points: every weight vector that matters is zero
    assert s.perlin_noise(1, [2.0 / 1.7, 3.0 / 1.7, 5.0 / 1.7]) == pytest.approx(0.0, abs=1e-14)
    rng = np.random.default_rng(9)
    pts = np.concatenate([rng.uniform(0, 40, (40, 3)), rng.uniform(-40, 40, (40, 3)), [[300.2, 0.1, 511.9]]])
    for p in pts:
        assert s.perlin_noise(1, p) == pytest.approx(_perlin_noise_py(rnd[0], rnd[1], rnd[2], p), abs=1e-14)
    assert max(abs(s.perlin_noise(1, p)) for p in pts) > 0.05
    for p in pts[:10]:
        turb = abs(sum(0.5 ** d * _perlin_noise_py(rnd[0], rnd[1], rnd[2], p * 2.0 ** d) for d in range(7)))
        want = 0.5 * (1.0 + math.sin(rnd[0] * p[2] + 10.0 * turb))
        got = s.tex_value_point(1, 0.2, 0.9, p)
        assert np.allclose(got, [want] * 3, atol=1e-12) and 0.0 <= got[0] <= 1.0
    # a checker over two Perlin textures picks by (u, v), each child sees the point
    for (u, v) in ((0.05, 0.05), (0.05, 0.25)):
        pick = 1 if math.sin(3.0 * u * 2 * orc_pi()) * math.sin(3.0 * v * 2 * orc_pi()) < 0 else 0
        assert np.array_equal(s.tex_value_point(2, u, v, pts[3]), s.tex_value_point(pick, u, v, pts[3]))


def orc_pi():
    return 3.14159265358979          # consts.rs:31


# ---------------------------------------------------------------- Light::Point / Spot / Distant (light.rs:142-175, 522-582)
def test_delta_light_sample_li_closed_form():
    t = orc.delta_light_tables()
    s = orc.OracleScene(t)
    p = np.array([0.5, 0.0, -2.0])
    # Point: position = to_world x origin = (-2, 3, 1); wi = normalised (pos - p), pdf 1, colour / |pos - p|^2, p1 = pos
    pos = np.array([-2.0, 3.0, 1.0])
    r = s.light_sample_li(0, p, (0.3, 0.7))
    d2 = float(((pos - p) ** 2).sum())
    assert r["pdf"] == 1.0 and np.array_equal(r["p1"], pos)
    assert np.allclose(r["wi"], (pos - p) / math.sqrt(d2), rtol=0, atol=1e-15)
    assert np.allclose(r["color"], np.array([30, 24, 18]) / d2, rtol=1e-15)
    # the sample does not depend on u
    assert all(np.array_equal(r[k], s.light_sample_li(0, p, (0.9, 0.1))[k]) for k in ("wi", "color", "p1"))
    # Spot at (1.5, 4, -1), axis (0, -cos .5, -sin .5); cosines are cos(x * 180 / PI) of the RADIAN arguments (light.rs:548-549)
    spos = np.array([1.5, 4.0, -1.0]); axis = np.array([0.0, -math.cos(0.5), -math.sin(0.5)])
    cw, cf = math.cos(0.009 * 180.0 * (1.0 / math.pi)), math.cos(0.004 * 180.0 * (1.0 / math.pi))
    assert abs(cw - 0.8699) < 1e-3 and abs(cf - 0.9738) < 1e-3
    seen = set()
    for q in ([1.5, 0.0, -3.18], [1.5, 0.0, -1.2], [1.5, 0.0, 2.0], [3.0, 0.0, -4.5], [0.2, 1.2, -1.6]):
        q = np.array(q)
        r = s.light_sample_li(1, q, (0.5, 0.5))
        w = (q - spos) / np.linalg.norm(q - spos)            # -wi, from the light to the point
        ct = float(w @ axis)
        fo = 0.0 if ct < cw else 1.0 if ct > cf else ((ct - cw) / (cf - cw)) ** 4
        seen.add(0 if fo == 0.0 else 2 if fo == 1.0 else 1)
        assert r["pdf"] == 1.0 and np.array_equal(r["p1"], spos)
        assert np.allclose(r["color"], np.array([60, 80, 100]) * fo / float(((q - spos) ** 2).sum()), rtol=1e-9, atol=1e-15)
    assert seen == {0, 1, 2}                                 # outside the cone, in the falloff band, inside the full-intensity cone
    # Distant: wi = normalised dir, colour as given, p1 = p + dir * 2e10
    dirv = np.array([0.3, 2.0, -0.4])
    r = s.light_sample_li(2, p, (0.5, 0.5))
    assert r["pdf"] == 1.0 and np.array_equal(r["color"], [0.8, 0.7, 0.5])
    assert np.allclose(r["wi"], dirv / np.linalg.norm(dirv), rtol=0, atol=1e-15)
    assert np.array_equal(r["p1"], p + dirv * 2e10)
    # Light::pdf_li of a delta light is 0 (light.rs:281-283)
    for li in (0, 1, 2):
        assert s.light_pdf_li(li, p, (0.0, 1.0, 0.0)) == 0.0


def test_point_and_distant_light_on_a_lambertian_floor():
    """Direct / Whitted / Path under ONE delta light: a floor pixel is (R / PI) * I / d^2 * cos(theta), emission 0, no MIS and no
    BSDF half for a delta light (integrator.rs:587-601).  Two behaviours of the reference that are restated as they are:
    (1) a Point light only arrives where the shadow ray's closest hit is primitive 0, here the enclosing sphere
    (Visibility::unoccluded compares prim_index with make_basic's 0, hittable.rs:37, 138) — with the floor and the sphere swapped in
    the primitive list the same scene renders black;  (2) a Distant light never arrives: its shadow ray is 2e10 x dir long and the
    origin is pushed SMALL x that, 2e7 units, far outside any scene, so nothing is hit and unoccluded() answers `infinite_light`."""
    from rustraytracer_b200 import Camera
    cam = Camera.new((0.0, 6.0, -0.001), (0.0, 0.0, 0.0), (0, 1, 0), 1.0, 2.0, 0.0, 10.0)
    want = 0.5 / math.pi * np.array([30.0, 24.0, 18.0]) / 9.0        # light 3 above the floor, seen (almost) straight from above
    for ltype in (N.LIGHT_POINT, N.LIGHT_DISTANT):
        for swap in (False, True):
            lig = np.zeros(1, dtype=N.light_dtype); lig["prim_index"] = N.RT_NONE
            lig[0]["type"] = ltype; lig[0]["color"] = (30, 24, 18); lig[0]["param"] = (0.3, 2.0, -0.4); lig[0]["transform"] = 0
            prims = [orc.prim(N.PRIM_SPHERE, [0, 0, 0, 30.0]), orc.prim(N.PRIM_XZ_RECT, [-8, -8, 8, 8, 0.0])]
            if swap:
                prims = prims[::-1]
            s = orc.OracleScene(orc.simple_tables(prims, lights=lig, transforms=[[1, 0, 0, 0, 0, 1, 0, 3.0, 0, 0, 1, 0, 0, 0, 0, 1]]))
            for kind in (0, 1, 2, 3):
                o, counts, _ = s.render(cam, 8, 8, 0, 2, seed=3, max_depth=1 if kind == 0 else 2, kind=kind)
                rgb = o[..., :3] / o[..., 3:4]
                if swap or ltype == N.LIGHT_DISTANT:
                    assert not rgb.any()
                else:
                    assert np.allclose(rgb, want[None, None, :], rtol=5e-3, atol=0), (kind, rgb[0, 0], want)
                assert int(counts[2]) == 0                       # no MIS probe rays for a delta light
                assert int(counts[1]) == 8 * 8 * 2               # one shadow ray per floor vertex


# ---------------------------------------------------------------- Material::Disney (material.rs:245-394) and its lobes
def disney_tables(thin=False, **over):
    """One Disney material (material 0) from solid textures; defaults = scenes.rs:867-904 `disney_mat`.  Scalar parameters go in the red
    channel (the reference reads `.x`), `color` and `scatter_distance` are full colours."""
    par = dict(color=(0.1608, 0.0014767, 0.4), metallic=0.0, eta=1.5, roughness=0.5, specular_tint=0.0, anisotropic=0.0, sheen=0.0, sheen_tint=0.5,
               clearcoat=0.0, clearcoat_gloss=1.0, spec_trans=0.0, scatter_distance=(0.0, 0.0, 0.0), flatness=0.0, diff_trans=1.0)
    par.update(over)
    names = N.DISNEY_FIELDS[:14]
    tex = np.zeros(14, dtype=N.texture_dtype)
    for i, k in enumerate(names):
        v = par[k]
        tex[i]["color"] = v if isinstance(v, tuple) else (v, 0.0, 0.0)
    d = np.zeros(1, dtype=N.disney_dtype)
    for i, k in enumerate(names):
        d[0][k] = i
    d[0]["thin"] = 1 if thin else 0
    mat = np.zeros(1, dtype=N.material_dtype); mat["tex"] = N.RT_NONE
    mat[0]["type"] = N.MAT_DISNEY; mat[0]["tex"][0] = 0
    return par, dict(textures=tex, materials=mat, disneys=d)


def py_disney_f_pdf(par, thin, wo_w, wi_w, flags=31):
    """Independent numpy restatement of Bsdf::f / Bsdf::pdf for the Disney BSDF, straight from bxdf.rs / material.rs / microfacet.rs
    (lobes: DisneyDiffuse, FakeSS, Retro, Sheen, MicrofacetReflection{DisneyFresnel, DisneyMicrofacet}, ClearCoat, MicrofacetTransmission,
    LambertianTransmission).  The frame is Bsdf::empty(): every axis is (1,1,1)/sqrt(3)."""
    PI = 3.14159265358979
    n = np.ones(3) / np.linalg.norm(np.ones(3))
    wo = np.full(3, np.dot(wo_w, n)); wi = np.full(3, np.dot(wi_w, n))
    reflect = np.dot(wi_w, n) * np.dot(wo_w, n) > 0
    sw = lambda u: np.clip(1 - u, 0, 1) ** 5
    lerp = lambda t, a, b: (1 - t) * a + t * b
    color = np.array(par["color"]); metallic = par["metallic"]; e = par["eta"]; strans = par["spec_trans"]
    dw = (1 - metallic) * (1 - strans); dt = par["diff_trans"] / 2; rough = par["roughness"]
    lum = 0.2126 * color[0] + 0.7152 * color[1] + 0.0722 * color[2]
    c_tint = color / lum if lum > 0 else np.ones(3)
    R, T, D, G = 1, 2, 4, 8
    lobes = []      # (type bits, f, pdf)
    ac = lambda w: abs(w[2])
    same = wo[2] * wi[2] > 0
    dpdf = ac(wi) / PI if same else 0.0
    wh = wi + wo
    whn = wh / np.linalg.norm(wh) if np.any(wh != 0) else None
    fo, fi = sw(ac(wo)), sw(ac(wi))

    def diffuse(c): return c / PI * (1 - fo / 2) * (1 - fi / 2)

    def fakess(c):
        if whn is None: return np.zeros(3)
        cd = np.dot(wi, whn); f90 = cd * cd * rough
        fss = lerp(fo, 1, f90) * lerp(fi, 1, f90)
        return c / PI * 1.25 * (fss * (1 / (ac(wo) + ac(wi)) - 0.5) + 0.5)

    def retro(c):
        if whn is None: return np.zeros(3)
        cd = np.dot(wi, whn); rr = 2 * rough * cd * cd
        return c / PI * rr * (fo + fi + fo * fi * (rr - 1))
    if dw > 0:
        if thin:
            flat = par["flatness"]
            lobes.append((D | R, diffuse(dw * (1 - flat) * (1 - dt) * color), dpdf))
            lobes.append((D | R, fakess(dw * flat * (1 - dt) * color), dpdf))
        elif tuple(par["scatter_distance"]) == (0.0, 0.0, 0.0):
            lobes.append((D | R, diffuse(dw * color), dpdf))
        else:
            lobes.append((T | 16, np.zeros(3), 0.0))            # SpecularTransmission: f = 0, pdf = 0
        lobes.append((D | R, retro(dw * color), dpdf))
        if par["sheen"] > 0:
            c_sheen = lerp(par["sheen_tint"], np.ones(3), c_tint)
            f = sw(np.dot(wi, whn)) * dw * par["sheen"] * c_sheen if whn is not None else np.zeros(3)
            lobes.append((D | R, f, dpdf))
    aspect = np.sqrt(1 - par["anisotropic"] * 0.9)
    ax, ay = max(0.001, rough * rough / aspect), max(0.001, rough * rough * aspect)

    def tr_d(h, ax, ay):
        c2 = h[2] * h[2]; s2 = max(0.0, 1 - c2); t2 = s2 / c2
        if np.isinf(t2): return 0.0
        s = np.sqrt(s2); cp = 1.0 if s == 0 else np.clip(h[0] / s, -1, 1); sp = 1.0 if s == 0 else np.clip(h[1] / s, -1, 1)
        ee = (cp * cp / (ax * ax) + sp * sp / (ay * ay)) * t2
        return 1 / (PI * ax * ay * c2 * c2 * (1 + ee) ** 2)

    def lam(w, ax, ay):
        c2 = w[2] * w[2]; s2 = max(0.0, 1 - c2); s = np.sqrt(s2); att = abs(s / w[2])
        if np.isinf(att): return 0.0
        cp = 1.0 if s == 0 else np.clip(w[0] / s, -1, 1); sp = 1.0 if s == 0 else np.clip(w[1] / s, -1, 1)
        al = np.sqrt(cp * cp * ax * ax + sp * sp * ay * ay)
        return (-1 + np.sqrt(1 + (al * att) ** 2)) / 2

    def frd(c, ei, et):
        c = np.clip(c, -1, 1)
        if c < 0: ei, et, c = et, ei, abs(c)
        si = np.sqrt(max(0.0, 1 - c * c)); st = ei / et * si
        ct = np.sqrt(max(0.0, 1 - st * st))
        if st >= 1: return 1.0
        rp = (et * c - ei * ct) / (et * c + ei * ct); rs = (ei * c - et * ct) / (ei * c + et * ct)
        return (rp * rp + rs * rs) / 2
    # MicrofacetReflection(white, DisneyFresnel, DisneyMicrofacet)
    r0 = lerp(metallic, ((e - 1) ** 2 / (e + 1) ** 2) * lerp(par["specular_tint"], np.ones(3), c_tint), color)
    f_mr, p_mr = np.zeros(3), 0.0
    if ac(wi) != 0 and ac(wo) != 0 and whn is not None:
        hh = whn if whn[2] >= 0 else -whn               # face_forward(wh, +z)
        cth = np.dot(wi, hh)
        fres = lerp(metallic, frd(cth, 1.0, e), r0 + sw(cth) * (1 - r0))
        g = (1 / (1 + lam(wo, ax, ay))) * (1 / (1 + lam(wi, ax, ay)))            # separable
        f_mr = tr_d(whn, ax, ay) * g * fres / (4 * ac(wi) * ac(wo))
    if same:
        h2 = (wo + wi) / np.linalg.norm(wo + wi)
        p_mr = tr_d(h2, ax, ay) * (1 / (1 + lam(wo, ax, ay))) * abs(np.dot(wo, h2)) / ac(wo) / (4 * np.dot(wo, h2))
    lobes.append((G | R, f_mr, p_mr))
    if par["clearcoat"] > 0:
        gloss = lerp(par["clearcoat_gloss"], 0.1, 0.001)
        f_cc, p_cc = np.zeros(3), 0.0
        if whn is not None:
            a2 = gloss * gloss
            ch = ac(whn)
            dr = (a2 - 1) / (PI * np.log(a2) * (1 + (a2 - 1) * ch * ch))
            fr = 0.04 + sw(np.dot(wo, whn)) * (1 - 0.04)
            sg = lambda c: 2 / (1 + np.sqrt(0.0625 + (1 - 0.0625) * c * c))
            f_cc = np.full(3, par["clearcoat"] / 4 * sg(ac(wo)) * sg(ac(wi)) * dr * fr)
            if same: p_cc = dr * ch / (4 * np.dot(wo, whn))
        lobes.append((G | R, f_cc, p_cc))
    if strans > 0:
        lobes.append((G | T, None, None))               # not evaluated by this helper (tested through the renders)
    if thin:
        lobes.append((D | T, dt * color / PI, ac(wi) / PI if not same else 0.0))
    f = np.zeros(3); pdf = 0.0
    for ty, lf, lp in lobes:
        m = (ty & flags) == ty
        if lf is None:
            raise NotImplementedError
        if (m and reflect and (ty & R)) or (not reflect and (ty & T)):      # Bsdf::f precedence quirk (bsdf.rs:90-92)
            f = f + lf
        if m:
            pdf += lp
    return f, pdf


@pytest.mark.parametrize("thin,over", [(False, {}), (False, dict(sheen=0.7, sheen_tint=0.3, clearcoat=0.8, clearcoat_gloss=0.4, metallic=0.3, specular_tint=0.6, anisotropic=0.5)),
                                       (True, dict(flatness=0.4, diff_trans=0.6, sheen=0.2)), (False, dict(scatter_distance=(0.1, 0.0, 0.0), clearcoat=0.3))])
def test_disney_bsdf_f_and_pdf_against_a_python_restatement(thin, over):
    par, t = disney_tables(thin, **over)
    s = orc.OracleScene(orc.simple_tables([orc.prim(N.PRIM_SPHERE, [0, 0, 0, 1.0])], **t))
    rng = np.random.default_rng(12)
    checked = 0
    for _ in range(200):
        wo = rng.normal(size=3); wo /= np.linalg.norm(wo)
        wi = rng.normal(size=3); wi /= np.linalg.norm(wi)
        for flags in (31, 31 - 16):
            f, pdf = s.bsdf_f_pdf(0, wo, wi, flags=flags)
            pf, ppdf = py_disney_f_pdf(par, thin, wo, wi, flags)
            assert np.allclose(f, pf, rtol=1e-9, atol=1e-14), (wo, wi, f, pf)
            assert pdf == pytest.approx(ppdf, rel=1e-9, abs=1e-14)
            checked += int(np.any(f != 0))
    assert checked > 100


def test_disney_sample_f_returns_well_formed_samples():
    """Bsdf::sample_f on the 4-lobe default disney_mat, on a thin clear-coated one and on a transmissive one.  The Disney arm never calls
    Bsdf::new, so the BSDF keeps Bsdf::empty()'s degenerate frame (every axis (1,1,1)/sqrt(3), bsdf.rs:37-47): local -> world -> local is not
    the identity there and sample_f cannot be cross-checked against f() at the returned world direction; what can be pinned is that the sample
    is well-formed, that its type is one of the lobes the material built, and that every wi is a multiple of the frame axis (local_to_world
    of that frame, bsdf.rs:71-77)."""
    n = np.ones(3) / np.sqrt(3.0)
    for thin, over, types in ((False, {}, {4 | 1, 8 | 1}), (True, dict(flatness=0.5, clearcoat=1.0, clearcoat_gloss=0.2, sheen=0.5, diff_trans=0.8), {4 | 1, 8 | 1, 4 | 2}),
                              (False, dict(spec_trans=0.6, roughness=0.3), {4 | 1, 8 | 1, 8 | 2})):
        par, t = disney_tables(thin, **over)
        s = orc.OracleScene(orc.simple_tables([orc.prim(N.PRIM_SPHERE, [0, 0, 0, 1.0])], **t))
        rng = np.random.default_rng(3)
        seen = set()
        for k in range(400):
            wo = rng.normal(size=3); wo /= np.linalg.norm(wo)
            f, wi, pdf, st = s.bsdf_sample_f(0, wo, rng.uniform(size=2), seed=5, pixel=k, smp=0)
            if pdf == 0:
                continue
            seen.add(st)
            assert st in types and pdf > 0 and np.isfinite(pdf) and np.isfinite(f).all() and (f >= 0).all()
            assert np.allclose(np.cross(wi, n), 0, atol=1e-12)
        assert len(seen) >= 2
