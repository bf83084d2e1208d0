"""ctypes binding of the test-side oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

from rustraytracer_b200 import _native as N

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liboracle.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])
        L = C.CDLL(LIB)
        vp, u32, u64, f64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_double, C.c_int
        L.orc_scene_create.restype = vp
        L.orc_scene_destroy.argtypes = [vp]
        L.orc_scene_add_mesh.argtypes = [vp, u64, vp, u64, vp, u64, vp, u64, vp]
        L.orc_scene_add_mesh.restype = u32
        L.orc_scene_add_transform.argtypes = [vp, vp]
        L.orc_scene_add_transform.restype = u32
        L.orc_scene_add_primitives.argtypes = [vp, u64, vp]
        L.orc_scene_set_textures.argtypes = [vp, u64, vp]
        L.orc_scene_set_materials.argtypes = [vp, u64, vp]
        L.orc_scene_set_lights.argtypes = [vp, u64, vp]
        L.orc_scene_add_disney.argtypes = [vp, vp]
        L.orc_scene_add_disney.restype = u32
        L.orc_scene_commit.argtypes = [vp, u64]
        L.orc_scene_build_fast_tree.argtypes = [vp, i32]
        L.orc_scene_num_primitives.argtypes = [vp]
        L.orc_scene_num_primitives.restype = u64
        L.orc_scene_prim_box.argtypes = [vp, u32, vp]
        L.orc_intersect.argtypes = [vp, u64, vp, f64, f64, i32, i32, vp, vp, vp]
        L.orc_occluded.argtypes = [vp, u64, vp, i32, vp]
        L.orc_camera_look_at.argtypes = [vp, vp, vp, f64, f64, f64, f64, f64, f64, vp]
        L.orc_render.argtypes = [vp, vp, u32, u32, vp, i32, vp, vp]
        L.orc_render.restype = f64
        L.orc_render_kind.argtypes = [vp, vp, u32, u32, u32, vp, i32, vp, vp]
        L.orc_render_kind.restype = f64
        L.orc_render_ex.argtypes = [vp, vp, u32, u32, vp, i32, vp, vp, u32, i32]
        L.orc_render_ex.restype = f64
        L.orc_single_sample.argtypes = [vp, vp, u32, u32, u32, u32, u32, u64, u32, u32, vp]
        L.orc_resolve_rgb8.argtypes = [vp, u64, vp]
        L.orc_philox.argtypes = [vp, u32, u32]
        L.orc_rng_uniform.argtypes = [u64, u32, u32, u32]
        L.orc_rng_uniform.restype = f64
        L.orc_bsdf_f_pdf.argtypes = [vp, u32, f64, f64, vp, vp, u32, vp, vp]
        L.orc_bsdf_sample_f.argtypes = [vp, u32, f64, f64, vp, vp, u32, u64, u32, u32, vp]
        L.orc_tex_value.argtypes = [vp, u32, f64, f64, vp]
        L.orc_fr_dielectric.argtypes = [f64, f64, f64]
        L.orc_fr_dielectric.restype = f64
        L.orc_fr_conductor.argtypes = [f64, vp, vp, vp]
        L.orc_scene_add_image_hdr.argtypes = [vp, u32, u32, vp]
        L.orc_scene_add_image_hdr.restype = u32
        L.orc_scene_add_image_rgb8.argtypes = [vp, u32, u32, vp]
        L.orc_scene_add_image_rgb8.restype = u32
        L.orc_scene_add_perlin.argtypes = [vp, f64, vp, vp]
        L.orc_scene_add_perlin.restype = u32
        L.orc_tex_value_point.argtypes = [vp, u32, f64, f64, vp, vp]
        L.orc_perlin_noise.argtypes = [vp, u32, vp]
        L.orc_perlin_noise.restype = f64
        L.orc_light_sample_li.argtypes = [vp, u32, vp, vp, vp]
        L.orc_light_pdf_li.argtypes = [vp, u32, vp, vp]
        L.orc_light_pdf_li.restype = f64
        L.orc_light_le.argtypes = [vp, u32, vp, vp]
        L.orc_light_dist.argtypes = [vp, u32, vp, vp, vp, vp, vp]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleScene:
    """Oracle-side scene fed from the SAME plain tables the CUDA library gets."""

    def __init__(self, tables=None, bvh_seed=1):
        self.h = C.c_void_p(lib().orc_scene_create())
        if tables is not None:
            self.load(tables)
            self.commit(bvh_seed)

    def __del__(self):
        try:
            if self.h:
                lib().orc_scene_destroy(self.h)
        except Exception:
            pass

    def load(self, t):
        L = lib()
        for (p, n, uv, ind) in t["meshes"]:
            p = np.ascontiguousarray(p, dtype=np.float64); n = np.ascontiguousarray(n, dtype=np.float64)
            uv = np.ascontiguousarray(uv, dtype=np.float64); ind = np.ascontiguousarray(ind, dtype=np.uint32)
            L.orc_scene_add_mesh(self.h, len(p), _p(p), len(n), _p(n) if len(n) else None, len(uv), _p(uv) if len(uv) else None, len(ind), _p(ind))
        for m in np.asarray(t["transforms"], dtype=np.float64).reshape(-1, 16):
            m = np.ascontiguousarray(m)
            L.orc_scene_add_transform(self.h, _p(m))
        prims = np.ascontiguousarray(t["prims"], dtype=N.prim_dtype)
        if len(prims):
            L.orc_scene_add_primitives(self.h, len(prims), _p(prims))
        for im in t.get("images", []):
            if im.dtype == np.uint8:
                im = np.ascontiguousarray(im); L.orc_scene_add_image_rgb8(self.h, im.shape[1], im.shape[0], _p(im))
            else:
                im = np.ascontiguousarray(im, dtype=np.float32); L.orc_scene_add_image_hdr(self.h, im.shape[1], im.shape[0], _p(im))
        for (scale, perm, vecs) in t.get("perlins", []):
            pm = np.ascontiguousarray(perm, dtype=np.uint32).reshape(768); vc = np.ascontiguousarray(vecs, dtype=np.float64).reshape(768)
            L.orc_scene_add_perlin(self.h, float(scale), _p(pm), _p(vc))
        tex = np.ascontiguousarray(t["textures"], dtype=N.texture_dtype)
        L.orc_scene_set_textures(self.h, len(tex), _p(tex) if len(tex) else None)
        for d in np.ascontiguousarray(t.get("disneys", np.zeros(0, dtype=N.disney_dtype)), dtype=N.disney_dtype):
            rec = np.ascontiguousarray(d).reshape(1)
            L.orc_scene_add_disney(self.h, _p(rec))
        mat = np.ascontiguousarray(t["materials"], dtype=N.material_dtype)
        L.orc_scene_set_materials(self.h, len(mat), _p(mat) if len(mat) else None)
        lig = np.ascontiguousarray(t["lights"], dtype=N.light_dtype)
        L.orc_scene_set_lights(self.h, len(lig), _p(lig) if len(lig) else None)

    def commit(self, bvh_seed=1):
        lib().orc_scene_commit(self.h, bvh_seed)

    def build_fast_tree(self, threads=8):
        """median-split tree built with nth_element in parallel: for the hit-set traversal (mode 3) on million-triangle soups"""
        lib().orc_scene_build_fast_tree(self.h, threads)

    def num_primitives(self):
        return lib().orc_scene_num_primitives(self.h)

    # ---- Infinite-light probes (light.rs:204-245, 285-294, 499-513; distribution.rs)
    def light_sample_li(self, light, p, u):
        p = np.ascontiguousarray(p, dtype=np.float64); u = np.ascontiguousarray(u, dtype=np.float64); out = np.zeros(10)
        lib().orc_light_sample_li(self.h, light, _p(p), _p(u), _p(out))
        return dict(wi=out[0:3], pdf=out[3], color=out[4:7], p1=out[7:10])

    def light_pdf_li(self, light, p, wi):
        p = np.ascontiguousarray(p, dtype=np.float64); wi = np.ascontiguousarray(wi, dtype=np.float64)
        return lib().orc_light_pdf_li(self.h, light, _p(p), _p(wi))

    def light_le(self, light, d):
        d = np.ascontiguousarray(d, dtype=np.float64); out = np.zeros(3)
        lib().orc_light_le(self.h, light, _p(d), _p(out))
        return out

    def light_dist(self, light):
        n = np.zeros(2, dtype=np.uint64)
        lib().orc_light_dist(self.h, light, _p(n), None, None, None, None)
        nu, nv = int(n[0]), int(n[1])
        cdf = np.zeros((nv, nu + 1)); row_int = np.zeros(nv); mcdf = np.zeros(nv + 1); mint = np.zeros(1)
        lib().orc_light_dist(self.h, light, _p(n), _p(cdf), _p(row_int), _p(mcdf), _p(mint))
        return dict(nu=nu, nv=nv, cdf=cdf, row_int=row_int, marg_cdf=mcdf, marg_int=float(mint[0]))

    def prim_box(self, i):
        out = np.zeros(6)
        lib().orc_scene_prim_box(self.h, i, _p(out))
        return out

    def intersect(self, rays, tmin=0.001, tmax=1e308, mode=0, threads=8, want_ties=False, want_stats=False):
        """mode 0 = reference-faithful tree, 1 = brute force (+tie sets), 2 = ordered (work counting), 3 = hit set + tie set over the tree."""
        rays = np.ascontiguousarray(rays, dtype=N.ray_dtype)
        hits = np.zeros(len(rays), dtype=N.hit_dtype)
        ties = np.full((len(rays), 4), N.RT_NONE, dtype=np.uint32) if want_ties else None
        stats = np.zeros(2, dtype=np.uint64)
        lib().orc_intersect(self.h, len(rays), _p(rays), tmin, tmax, mode, threads, _p(hits), _p(ties), _p(stats))
        out = [hits]
        if want_ties:
            out.append(ties)
        if want_stats:
            out.append(stats)
        return out[0] if len(out) == 1 else tuple(out)

    def occluded(self, p0, p1, prim, infinite=False, threads=8):
        q = np.zeros(len(p0), dtype=N.shadow_dtype)
        q["p0"] = p0; q["p1"] = p1; q["p1_prim_index"] = prim; q["infinite_light"] = 1 if infinite else 0
        out = np.zeros(len(q), dtype=np.uint8)
        lib().orc_occluded(self.h, len(q), _p(q), threads, _p(out))
        return out.astype(bool)

    def render_ex(self, camera, width, height, sample_begin, sample_end, seed=0, max_depth=25, threads=8, pixel_stride=1, count_ordered=False):
        """Bounded-sample render: counts = extend, shadow, probe rays, ordered box tests, ordered prim tests."""
        fd = N.FilmDesc(width, height, sample_begin, sample_end, 0, 0, width, height, seed)
        film = np.zeros((height, width, 4), dtype=np.float64)
        counts = np.zeros(5, dtype=np.uint64)
        secs = lib().orc_render_ex(self.h, C.byref(camera.c), max_depth, 0, C.byref(fd), threads, _p(film), _p(counts), pixel_stride, 1 if count_ordered else 0)
        return film, counts, secs

    def render(self, camera, width, height, sample_begin, sample_end, seed=0, max_depth=25, invisible_light=False, threads=8, window=None, film=None, kind=0):
        """kind: 0 Path, 1 Direct{UniformOne}, 2 Direct{UniformAll}, 3 Whitted (IntType, integrator.rs:25-41)"""
        x0, y0, x1, y1 = window if window is not None else (0, 0, width, height)
        fd = N.FilmDesc(width, height, sample_begin, sample_end, x0, y0, x1, y1, seed)
        if film is None:
            film = np.zeros((height, width, 4), dtype=np.float64)
        counts = np.zeros(3, dtype=np.uint64)
        secs = lib().orc_render_kind(self.h, C.byref(camera.c), kind, max_depth, 1 if invisible_light else 0, C.byref(fd), threads, _p(film), _p(counts))
        return film, counts, secs

    def bsdf_f_pdf(self, mat, wo, wi, flags=31, u=0.3, v=0.6):
        wo = np.ascontiguousarray(wo, dtype=np.float64); wi = np.ascontiguousarray(wi, dtype=np.float64)
        f = np.zeros(3); pdf = C.c_double()
        lib().orc_bsdf_f_pdf(self.h, mat, u, v, _p(wo), _p(wi), flags, _p(f), C.byref(pdf))
        return f, pdf.value

    def bsdf_sample_f(self, mat, wo, sample, flags=31, seed=0, pixel=0, smp=0, u=0.3, v=0.6):
        wo = np.ascontiguousarray(wo, dtype=np.float64); s = np.ascontiguousarray(sample, dtype=np.float64)
        out = np.zeros(8)
        lib().orc_bsdf_sample_f(self.h, mat, u, v, _p(wo), _p(s), flags, seed, pixel, smp, _p(out))
        return out[0:3], out[3:6], out[6], int(out[7])

    def tex_value_point(self, tid, u, v, p):
        out = np.zeros(3); p = np.ascontiguousarray(p, dtype=np.float64)
        lib().orc_tex_value_point(self.h, tid, u, v, _p(p), _p(out))
        return out

    def perlin_noise(self, perlin, p):
        p = np.ascontiguousarray(p, dtype=np.float64)
        return lib().orc_perlin_noise(self.h, perlin, _p(p))

    def tex_value(self, tid, u, v):
        out = np.zeros(3)
        lib().orc_tex_value(self.h, tid, u, v, _p(out))
        return out


def resolve_rgb8(film):
    film = np.ascontiguousarray(film, dtype=np.float64)
    out = np.zeros(film.shape[:2] + (3,), dtype=np.uint8)
    lib().orc_resolve_rgb8(_p(film), film.shape[0] * film.shape[1], _p(out))
    return out


def simple_tables(prims, textures=None, materials=None, lights=None, meshes=(), transforms=(), images=(), perlins=(), disneys=None):
    """Plain tables for hand-built KAT scenes."""
    if textures is None:
        textures = np.zeros(1, dtype=N.texture_dtype); textures[0]["color"] = (0.5, 0.5, 0.5)
    if materials is None:
        materials = np.zeros(1, dtype=N.material_dtype); materials[0]["tex"] = N.RT_NONE; materials[0]["tex"][0] = 0
    if lights is None:
        lights = np.zeros(0, dtype=N.light_dtype)
    return dict(meshes=list(meshes), prims=np.asarray(prims, dtype=N.prim_dtype), transforms=np.asarray(transforms, dtype=np.float64).reshape(-1, 16),
                textures=textures, materials=materials, lights=lights, images=list(images), perlins=list(perlins),
                disneys=np.zeros(0, dtype=N.disney_dtype) if disneys is None else np.asarray(disneys, dtype=N.disney_dtype))


def delta_light_tables(area_light=True):
    """Hand-built scene for Light::Point / Spot / Distant (light.rs:30-58, 142-175, 522-582).  Primitive 0 is the enclosing sphere: the
    delta lights' visibility record is HitRecord::make_basic (prim_index 0), so Visibility::unoccluded (hittable.rs:25-39) lets their
    light through only where the unbounded shadow ray's closest hit is primitive 0 — here wherever nothing stands between the shaded
    point, the light and the wall behind it."""
    tex = np.zeros(4, dtype=N.texture_dtype)
    tex[0]["color"] = (0.7, 0.7, 0.7); tex[1]["color"] = (0.9, 0.5, 0.2); tex[2]["color"] = (1, 1, 1); tex[3]["color"] = (0.3, 0.6, 0.9)
    mat = np.zeros(5, dtype=N.material_dtype); mat["tex"] = N.RT_NONE
    mat[0]["type"] = N.MAT_MATTE; mat[0]["tex"][0] = 0                                    # Lambertian walls / floor
    mat[1]["type"] = N.MAT_MATTE; mat[1]["tex"][0] = 1; mat[1]["f"][0] = 25.0             # Oren-Nayar ball
    mat[2]["type"] = N.MAT_MIRROR; mat[2]["tex"][0] = 2
    mat[3]["type"] = N.MAT_GLASS; mat[3]["tex"][0] = 2; mat[3]["tex"][1] = 2; mat[3]["f"] = (0.0, 0.0, 1.5)
    mat[4]["type"] = N.MAT_PLASTIC; mat[4]["tex"][0] = 3; mat[4]["tex"][1] = 2; mat[4]["f"][0] = 0.2; mat[4]["remap_roughness"] = 1
    a = 0.5                                                                               # spot: tilted about x, then lifted to (1.5, 4, -1)
    ca, sa = np.cos(a), np.sin(a)
    # falloff reads the object-space z of the direction (light.rs:17): the object z axis maps onto (0, -cos a, -sin a), down and towards -z
    spot_xf = [1, 0, 0, 1.5, 0, -sa, -ca, 4.0, 0, ca, -sa, -1.0, 0, 0, 0, 1]
    point_xf = [1, 0, 0, -2.0, 0, 1, 0, 3.0, 0, 0, 1, 1.0, 0, 0, 0, 1]
    lig = np.zeros(4 if area_light else 3, dtype=N.light_dtype)
    lig["prim_index"] = N.RT_NONE; lig["transform"] = N.RT_NONE
    lig[0]["type"] = N.LIGHT_POINT; lig[0]["color"] = (30, 24, 18); lig[0]["transform"] = 1
    # total_width / falloff go through cos(x * 180 / PI) (light.rs:548-549): 0.009 -> cos(0.5157) = 0.870, 0.004 -> cos(0.2292) = 0.974
    lig[1]["type"] = N.LIGHT_SPOT; lig[1]["color"] = (60, 80, 100); lig[1]["transform"] = 0; lig[1]["param"] = (0.009, 0.004, 0.0)
    lig[2]["type"] = N.LIGHT_DISTANT; lig[2]["color"] = (0.8, 0.7, 0.5); lig[2]["param"] = (0.3, 2.0, -0.4)    # dir is not normalised
    prims = [prim(N.PRIM_SPHERE, [0, 0, 0, 30.0], mat=0), prim(N.PRIM_XZ_RECT, [-8, -8, 8, 8, 0.0], mat=0),
             prim(N.PRIM_SPHERE, [-1.3, 1.0, 0.0, 1.0], mat=1), prim(N.PRIM_SPHERE, [1.2, 0.8, 0.6, 0.8], mat=2),
             prim(N.PRIM_SPHERE, [0.2, 0.6, -1.6, 0.6], mat=3), prim(N.PRIM_SPHERE, [2.8, 0.7, -1.2, 0.7], mat=4)]
    if area_light:
        lig[3]["type"] = N.LIGHT_DIFFUSE; lig[3]["prim_index"] = 6; lig[3]["color"] = (5, 5, 5)
        prims.append(prim(N.PRIM_XZ_RECT, [-1, -1, 1, 1, 6.0], mat=0, flip=1, light=3))
    return simple_tables(prims, textures=tex, materials=mat, lights=lig, transforms=[spot_xf, point_xf])


def perlin_tables(seed, scale):
    """What Perlin::new(256, scale) draws (perlin.rs:17-34): three shuffled permutations and 256 normalised vectors from [-1, 1]^3."""
    rng = np.random.default_rng(seed)
    perm = np.stack([rng.permutation(256) for _ in range(3)]).astype(np.uint32)
    vecs = rng.uniform(-1, 1, (256, 3)); vecs /= np.linalg.norm(vecs, axis=1, keepdims=True)
    return (scale, perm, vecs)


def prim(ptype, v, mat=0, flip=0, light=N.RT_NONE, mesh=0, ind=0, xf=N.RT_NONE):
    p = np.zeros(1, dtype=N.prim_dtype)[0]
    p["type"] = ptype; p["flip_face"] = flip; p["mat_index"] = mat; p["light_index"] = light; p["mesh_index"] = mesh; p["ind"] = ind; p["transform"] = xf
    vv = np.zeros(5); vv[:len(v)] = v; p["v"] = vv
    return p


def upload_tables(t):
    """The same tables → the CUDA library through the public ABI (needs a GPU to commit)."""
    from rustraytracer_b200 import Objects
    o = Objects()
    for (p, n, uv, ind) in t["meshes"]:
        o.add_mesh(p, n, uv, ind)
    for m in t["transforms"]:
        o.add_transform(m)
    for im in t.get("images", []):
        o.add_image_rgb8(im) if im.dtype == np.uint8 else o.add_image_hdr(im)
    for (scale, perm, vecs) in t.get("perlins", []):
        o.add_perlin(scale, perm, vecs)
    for d in t.get("disneys", []):
        o.add_disney(d)
    o.set_textures(t["textures"]); o.set_materials(t["materials"]); o.set_lights(t["lights"])
    if len(t["prims"]):
        o.add_primitives(t["prims"])
    return o
