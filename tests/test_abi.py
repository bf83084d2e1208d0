"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every declared symbol,
fails loudly without a GPU (no CPU fallback), and the host scene layer restates scenes.rs."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import orc
from rustraytracer_b200 import Objects, _native as N, scenes
from rustraytracer_b200._native import RtError

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_every_declared_symbol_is_exported():
    L = N.lib()
    for header, listed in (("rtcore.h", N.RTCORE_SYMBOLS), ("rthost.h", N.RTHOST_SYMBOLS)):
        text = open(os.path.join(ROOT, "include", header)).read()
        declared = set(re.findall(r"\b(rt[h]?_[a-z0-9_]+)\s*\(", text)) - {"rt_status"}
        assert declared == set(listed), declared ^ set(listed)
        for name in declared:
            assert getattr(L, name) is not None


def test_struct_layouts_match_header():
    assert C.sizeof(N.Camera) == 24 * 8 and C.sizeof(N.FilmDesc) == 40 and C.sizeof(N.IntegratorDesc) == 16
    assert C.sizeof(N.Counters) == 5 * 8 + 4 * 8


def test_round_spp_and_version():
    L = N.lib()
    assert [L.rt_round_spp(x) for x in (1, 2, 3, 16, 50, 64, 65, 256, 1000)] == [1, 2, 4, 16, 64, 64, 128, 256, 1024]   # sampler.rs:637-640
    assert b"sm_100a" in L.rt_version_string()


def test_camera_matches_oracle():
    from rustraytracer_b200 import Camera
    cam = Camera.new_motion_blur((278., 278., -800.), (278., 278., 0.), (0., 1., 0.), 16 / 9, 40., 0.2, 10., 0., 1.)
    ref = N.Camera()
    f, t, u = (np.array(x, dtype=np.float64) for x in ((278., 278., -800.), (278., 278., 0.), (0., 1., 0.)))
    orc.lib().orc_camera_look_at(orc._p(f), orc._p(t), orc._p(u), 16 / 9, 40., 0.2, 10., 0., 1., C.byref(ref))
    assert bytes(cam.c) == bytes(ref)
    # upper-left origin, +u to the right, v down (geometry.rs:157-159,181-182)
    assert cam.c.upper_left_corner[1] > 278 and cam.c.lens_radius == 0.1


@pytest.mark.skipif(N.lib().rt_device_count() > 0, reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    hs = scenes.HostScene("sphere_roughness", 32, 32)
    o = Objects()
    N.check(N.lib().rth_scene_upload(hs.h, o.h))
    with pytest.raises(RtError) as e:
        o.commit()
    assert e.value.status == N.RT_ERR_CUDA
    from rustraytracer_b200 import get_intersection, make_rays
    with pytest.raises(RtError) as e:
        get_intersection(o, make_rays([(0, 0, 0)], [(0, 0, 1)]))
    assert e.value.status == N.RT_ERR_STATE


def test_argument_errors():
    o = Objects()
    with pytest.raises(RtError):
        o.add_mesh(np.zeros((3, 3)), indices=[0, 1, 5])          # index out of range
    with pytest.raises(RtError):
        o.add_mesh(np.zeros((3, 3)), indices=[0, 1])             # not a multiple of 3
    with pytest.raises(RtError):
        o.add_primitives([orc.prim(N.PRIM_TRIANGLE, [], mesh=3)])
    with pytest.raises(RtError):
        o.add_transform(np.zeros(16))                            # singular
    t = np.zeros(1, dtype=N.texture_dtype); t["type"] = 7
    with pytest.raises(RtError):
        o.set_textures(t)
    assert N.lib().rt_scene_commit(None) == N.RT_ERR_INVALID


def test_scene_tables_sphere_roughness():
    # scenes.rs:474-546
    hs = scenes.HostScene("sphere_roughness", 400, 400)
    t = hs.tables()
    assert hs.counts() == dict(meshes=0, prims=10, transforms=0, textures=28, materials=9, lights=1)
    assert [tuple(p["v"][:4]) for p in t["prims"][:8]] == [(0., 1., -2.8 * 4.5 + 2.8 * i, 1.) for i in range(1, 9)]
    assert list(t["prims"]["mat_index"]) == [1, 2, 3, 4, 5, 6, 7, 8, 0, 8]
    assert t["prims"][9]["flip_face"] == 1 and t["prims"][9]["light_index"] == 0 and t["lights"][0]["prim_index"] == 9
    assert tuple(t["prims"][8]["v"]) == (-10000., -10000., 10000., 10000., -0.01)
    m = t["materials"][3]      # third metal sphere: roughness texture (i-1)/90 + SMALL
    assert m["type"] == N.MAT_METAL and m["remap_roughness"] == 1 and m["tex"][3] == N.RT_NONE and m["tex"][4] == N.RT_NONE
    assert t["textures"][m["tex"][2]]["color"][0] == 2 / 90. + 0.001
    assert tuple(t["textures"][m["tex"][0]]["color"]) == (0., 0., 0.) and tuple(t["textures"][m["tex"][1]]["color"]) == (1., 1., 1.)
    chk = t["textures"][2]
    assert chk["type"] == N.TEX_CHECKERED and chk["even"] == 0 and chk["odd"] == 1 and chk["frequency"] == 0.1
    assert hs.png_name() == "metal_spheres.png"


def test_scene_tables_cornell_and_dragon():
    hs = scenes.HostScene("cornell_box", 64, 64)
    t = hs.tables()
    assert hs.counts()["prims"] == 18 and hs.counts()["transforms"] == 2      # 6 walls + 2 cubes x 6 (scenes.rs:114-190)
    assert list(t["prims"]["flip_face"][:6]) == [1, 0, 1, 0, 1, 1]
    assert t["materials"][3]["type"] == N.MAT_LIGHT and t["prims"][2]["mat_index"] == 3
    # cube sides order z0,z1,y0,y1,x0,x1 (hittable.rs:846)
    assert list(t["prims"]["type"][6:12]) == [N.PRIM_XY_RECT, N.PRIM_XY_RECT, N.PRIM_XZ_RECT, N.PRIM_XZ_RECT, N.PRIM_YZ_RECT, N.PRIM_YZ_RECT]
    xf = t["transforms"][1].reshape(4, 4)          # first_transform = translate(265,0,295) * Ry(15°)
    a = 15 * 3.14159265358979 / 180
    assert np.allclose(xf, [[np.cos(a), 0, np.sin(a), 265], [0, 1, 0, 0], [-np.sin(a), 0, np.cos(a), 295], [0, 0, 0, 1]])
    hd = scenes.HostScene("plastic_dragon", 64, 64)
    c = hd.counts()
    assert c["prims"] == 871414 + 2 and hd.uses_standin()                   # floor + dragon + light (scenes.rs:348-358)
    td = hd.tables()
    assert td["prims"][0]["type"] == N.PRIM_XZ_RECT and td["prims"][-1]["flip_face"] == 1 and td["lights"][0]["prim_index"] == 871415
    assert td["materials"][1]["type"] == N.MAT_PLASTIC and td["materials"][1]["f"][0] == 0.001
    p = td["meshes"][0][0]
    assert p.min(0)[1] == pytest.approx(-2.83, abs=1e-5) and np.all(np.abs(p).max(0) < 4.0)
    assert len(td["meshes"][0][1]) == len(p) and len(td["meshes"][0][2]) == 0        # normals, no uvs


def test_obj_parser_tobj_semantics(tmp_path):
    # parser.rs:8-87: triangulated, single-index, models[0] only, transform on points and normals
    obj = tmp_path / "quad.obj"
    obj.write_text("""# two objects; only the first is used
o first
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
vn 0 0 1
vt 0 0
vt 1 0
vt 1 1
vt 0 1
f 1/1/1 2/2/1 3/3/1 4/4/1
f -4/1/1 -3/2/1 -2/3/1
o second
v 5 5 5
f 1 2 5
""")
    h = C.c_void_p()
    N.check(N.lib().rth_scene_create_empty(C.byref(h)))
    xf = np.array([2, 0, 0, 10, 0, 2, 0, 0, 0, 0, 2, 0, 0, 0, 0, 1], dtype=np.float64)
    out = C.c_uint32()
    N.check(N.lib().rth_scene_add_obj(h, str(obj).encode(), N.ptr(xf), C.byref(out)))
    sz = (C.c_uint64 * 4)()
    N.check(N.lib().rth_scene_mesh_sizes(h, 0, sz))
    assert list(sz) == [4, 4, 4, 9]                 # 4 unique (v,vt,vn) triples; quad → 2 tris, + 1 tri
    p = np.zeros((4, 3)); n = np.zeros((4, 3)); uv = np.zeros((4, 2)); ind = np.zeros(9, dtype=np.uint32)
    N.check(N.lib().rth_scene_mesh_copy(h, 0, N.ptr(p), N.ptr(n), N.ptr(uv), N.ptr(ind)))
    assert np.array_equal(p, [[10, 0, 0], [12, 0, 0], [12, 2, 0], [10, 2, 0]])
    assert np.array_equal(n, [[0, 0, 2]] * 4)       # transform_vector: scaled, not renormalised (parser.rs:45)
    assert list(ind) == [0, 1, 2, 0, 2, 3, 0, 1, 2]
    assert N.lib().rth_scene_add_obj(h, b"/nonexistent.obj", N.ptr(xf), None) == N.RT_ERR_IO
    N.lib().rth_scene_destroy(h)


REF_DATA = "/root/reference/data"


@pytest.mark.skipif(not os.path.exists(REF_DATA + "/material/models/Mesh000.obj"), reason="reference assets only exist in the build container")
def test_obj_parser_on_reference_asset():
    h = C.c_void_p()
    N.check(N.lib().rth_scene_create_empty(C.byref(h)))
    xf = np.eye(4).reshape(16)
    N.check(N.lib().rth_scene_add_obj(h, (REF_DATA + "/material/models/Mesh000.obj").encode(), N.ptr(xf), None))
    sz = (C.c_uint64 * 4)()
    N.check(N.lib().rth_scene_mesh_sizes(h, 0, sz))
    assert sz[3] == 3 * 13312 and sz[0] == sz[1] == sz[2]       # 13,312 triangles, v/vt/vn
    N.lib().rth_scene_destroy(h)


def _write_hdr(path, img_rgbe, rle):
    """Radiance .hdr writer for the reader test: img_rgbe (h, w, 4) uint8; new-style RLE scanlines or flat quadruples."""
    h, w, _ = img_rgbe.shape
    with open(path, "wb") as f:
        f.write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n")
        f.write(f"-Y {h} +X {w}\n".encode())
        for y in range(h):
            row = img_rgbe[y]
            if not rle:
                f.write(row.tobytes()); continue
            f.write(bytes([2, 2, w >> 8, w & 255]))
            for c in range(4):
                x = 0
                comp = row[:, c]
                while x < w:
                    run = 1
                    while x + run < w and run < 127 and comp[x + run] == comp[x]:
                        run += 1
                    if run >= 3:
                        f.write(bytes([128 + run, int(comp[x])])); x += run
                    else:
                        lit = 1
                        while x + lit < w and lit < 128 and not (x + lit + 2 < w and comp[x + lit] == comp[x + lit + 1] == comp[x + lit + 2]):
                            lit += 1
                        f.write(bytes([lit])); f.write(comp[x:x + lit].tobytes()); x += lit


@pytest.mark.parametrize("rle", [False, True])
def test_hdr_reader_matches_rgbe_definition(tmp_path, rle):
    """Texture::new_hdr's decoder (image 0.23 HdrDecoder): pixel = c * 2^(e-136) as f32, 0 when e == 0."""
    rng = np.random.default_rng(3)
    w, h = 40, 7
    img = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    img[:, 5:20, :] = img[:, 5:6, :]            # runs for the RLE path
    img[2, 3, 3] = 0                            # e == 0 -> black
    img[..., 3] = np.clip(img[..., 3].astype(int) % 40 + 110, 0, 255).astype(np.uint8); img[2, 3, 3] = 0
    p = str(tmp_path / "t.hdr")
    _write_hdr(p, img, rle)
    ww, hh = C.c_uint32(), C.c_uint32()
    N.check(N.lib().rth_load_hdr(p.encode(), C.byref(ww), C.byref(hh), None, 0))
    assert (ww.value, hh.value) == (w, h)
    out = np.zeros((h, w, 3), dtype=np.float32)
    N.check(N.lib().rth_load_hdr(p.encode(), C.byref(ww), C.byref(hh), N.ptr(out), out.size))
    e = img[..., 3].astype(np.float32)
    want = img[..., :3].astype(np.float32) * np.exp2(e - 136.0)[..., None]
    want[img[..., 3] == 0] = 0
    assert np.array_equal(out, want.astype(np.float32))
    assert N.lib().rth_load_hdr(str(tmp_path / "missing.hdr").encode(), C.byref(ww), C.byref(hh), None, 0) == N.RT_ERR_IO


@pytest.mark.skipif(not os.path.exists(REF_DATA + "/material/textures/envmap.hdr"), reason="reference assets only exist in the build container")
def test_hdr_reader_on_reference_asset():
    ww, hh = C.c_uint32(), C.c_uint32()
    N.check(N.lib().rth_load_hdr((REF_DATA + "/material/textures/envmap.hdr").encode(), C.byref(ww), C.byref(hh), None, 0))
    assert (ww.value, hh.value) == (1024, 512)          # `-Y 512 +X 1024`
    out = np.zeros((512, 1024, 3), dtype=np.float32)
    N.check(N.lib().rth_load_hdr((REF_DATA + "/material/textures/envmap.hdr").encode(), C.byref(ww), C.byref(hh), N.ptr(out), out.size))
    assert np.isfinite(out).all() and out.min() >= 0 and out.max() > 10
    hs = scenes.HostScene("material_hdr", 32, 32, data_dir=REF_DATA)
    assert hs.tables()["images"][0].shape == (512, 1024, 3)


def test_image_and_infinite_light_validation():
    """Errors are status codes, never a crash; these are all caught on the host before any device work."""
    from rustraytracer_b200 import Objects
    o = Objects()
    tex = np.zeros(1, dtype=N.texture_dtype); tex[0]["type"] = N.TEX_HDR; tex[0]["image"] = 0
    with pytest.raises(N.RtError):                       # image index out of range: images must be added first
        o.set_textures(tex)
    assert o.add_image_hdr(np.ones((2, 4, 3), dtype=np.float32)) == 0
    assert o.add_image_rgb8(np.zeros((3, 3, 3), dtype=np.uint8)) == 1
    o.set_textures(tex)
    with pytest.raises(N.RtError):
        N.check(N.lib().rt_scene_add_image_hdr(o.h, 0, 4, None, None))
    lig = np.zeros(1, dtype=N.light_dtype); lig[0]["type"] = 7
    with pytest.raises(N.RtError):
        o.set_lights(lig)
    # rt_light is 72 bytes (type .. _pad, param[3]); Point / Spot / Distant are accepted, the first type past Distant is not
    assert N.light_dtype.itemsize == 72 and N.light_dtype.fields["param"][1] == 48
    for ltype in (N.LIGHT_POINT, N.LIGHT_SPOT, N.LIGHT_DISTANT):
        lig[0]["type"] = ltype; lig[0]["transform"] = N.RT_NONE
        o.set_lights(lig)
    lig[0]["type"] = N.LIGHT_DISTANT + 1
    with pytest.raises(N.RtError):
        o.set_lights(lig)
    # an Infinite light over a Checkered texture has a (0, 0) table in the reference (Texture::get_size) — refused at commit
    tex = np.zeros(3, dtype=N.texture_dtype); tex[0]["type"] = N.TEX_CHECKERED; tex[0]["odd"] = 1; tex[0]["even"] = 2
    o.set_textures(tex)
    mat = np.zeros(1, dtype=N.material_dtype); mat[0]["tex"] = N.RT_NONE; mat[0]["tex"][0] = 1
    o.set_materials(mat)
    lig = np.zeros(1, dtype=N.light_dtype); lig[0]["type"] = N.LIGHT_INFINITE; lig[0]["text_id"] = 0; lig[0]["transform"] = N.RT_NONE; lig[0]["prim_index"] = N.RT_NONE
    o.set_lights(lig)
    with pytest.raises(N.RtError) as ei:
        o.commit()
    assert ei.value.status == N.RT_ERR_INVALID


def test_multi_gpu_entry_points_argument_errors():
    """rt_comm_* / rt_render_multi / rt_scene_replicate (SURVEY §8e behind the ABI): argument checking happens before any CUDA / NCCL work."""
    L = N.lib()
    h = C.c_void_p()
    assert L.rt_comm_create_local(None, 0, C.byref(h)) == N.RT_ERR_INVALID
    assert L.rt_comm_unique_id(None) == N.RT_ERR_INVALID
    ident = np.zeros(128, dtype=np.uint8)
    assert L.rt_comm_create_rank(N.ptr(ident), 2, 2, C.byref(h)) == N.RT_ERR_INVALID      # rank >= world
    assert L.rt_comm_create_rank(None, 2, 0, C.byref(h)) == N.RT_ERR_INVALID
    assert L.rt_film_allreduce(None, None, 0) == N.RT_ERR_INVALID
    assert L.rt_film_reduce_resolve_rgb8(None, 0, None) == N.RT_ERR_INVALID
    assert L.rt_render_multi(None, None, None, 0, None, None, None) == N.RT_ERR_INVALID
    assert L.rt_scene_replicate(None, 0, C.byref(h)) == N.RT_ERR_INVALID
    assert L.rt_comm_size(None, None, None) == N.RT_ERR_INVALID
    assert L.rt_comm_destroy(None) == N.RT_OK
    if L.rt_device_count() == 0:
        dev = np.zeros(1, dtype=np.int32)
        assert L.rt_comm_create_local(N.ptr(dev), 1, C.byref(h)) == N.RT_ERR_CUDA        # no CPU path here either
        o = Objects()
        assert L.rt_scene_replicate(o.h, 0, C.byref(h)) == N.RT_ERR_CUDA


def test_rust_ffi_crate_declares_exactly_the_header():
    """ffi/src/lib.rs (the thin C-ABI FFI crate north_star names; no rustc in this image, so it is kept honest here): its extern block lists
    exactly the functions include/rtcore.h declares, and its #[repr(C)] structs have the header's fields in the header's order."""
    hdr = open(os.path.join(ROOT, "include", "rtcore.h")).read()
    rs = open(os.path.join(ROOT, "ffi", "src", "lib.rs")).read()
    declared = set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", hdr)) - {"rt_status"}
    bound = set(re.findall(r"pub fn (rt_[a-z0-9_]+)\s*\(", rs))
    assert declared == bound, declared ^ bound
    for name in ("rt_primitive", "rt_texture", "rt_material", "rt_disney", "rt_light", "rt_camera", "rt_ray", "rt_hit", "rt_shadow_query", "rt_integrator_desc", "rt_film_desc",
                 "rt_counters", "rt_build_stats"):
        c_body = re.search(r"typedef struct " + name + r" \{(.*?)\} " + name + r";", hdr, re.S).group(1)
        c_body = re.sub(r"/\*.*?\*/", "", c_body, flags=re.S)
        c_fields = []
        for decl in c_body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            for part in decl.split(",")[0:1] + [p for p in decl.split(",")[1:]]:
                c_fields.append(re.sub(r"\[.*\]", "", part.strip().split()[-1]).lstrip("*"))
        r_body = re.search(r"pub struct " + name + r" \{(.*?)\n\}", rs, re.S).group(1)
        r_fields = [f.rstrip("_") if f in ("type_",) else f for f in re.findall(r"pub (\w+):", r_body)]
        assert c_fields == r_fields, (name, c_fields, r_fields)


def _py_obj(path, xf):
    """Plain-Python restatement of tobj 2.0.2's load_obj(triangulate = true) as parser.rs:8-87 uses it: models[0] only, (v, vt, vn) triples
    re-indexed to a single index in order of first appearance, f32 values, fan triangulation, `transform` on points (with w) and normals."""
    v, vt, vn, pos, tex, nor, ind, remap = [], [], [], [], [], [], [], {}
    have_faces = False
    for raw in open(path, "rb").read().split(b"\n"):
        line = raw.decode("latin1").lstrip(" \t")
        flt = lambda toks, k: [np.float32(float(t)) if i < len(toks) else np.float32(0) for i, t in list(enumerate(toks + ["0"] * k))[:k]]
        if line[:2] in ("v ", "v\t"):
            v.append(flt(line[2:].split(), 3))
        elif line[:2] == "vn":
            vn.append(flt(line[2:].split(), 3))
        elif line[:2] == "vt":
            vt.append(flt(line[2:].split(), 2))
        elif line[:2] in ("f ", "f\t"):
            poly = []
            for tok in line[1:].split("#")[0].split():
                parts = tok.split("/")
                iv = int(parts[0]); it = int(parts[1]) if len(parts) > 1 and parts[1] else None; inn = int(parts[2]) if len(parts) > 2 and parts[2] else None
                iv = len(v) + iv if iv < 0 else iv - 1
                it = -1 if it is None else (len(vt) + it if it < 0 else it - 1)
                inn = -1 if inn is None else (len(vn) + inn if inn < 0 else inn - 1)
                key = (iv, it, inn)
                if key not in remap:
                    remap[key] = len(pos); pos.append(v[iv])
                    if 0 <= it < len(vt): tex.append(vt[it])
                    if 0 <= inn < len(vn): nor.append(vn[inn])
                poly.append(remap[key])
            for k in range(1, len(poly) - 1):
                ind += [poly[0], poly[k], poly[k + 1]]
            have_faces = True
        elif line[:1] in ("o", "g") and (len(line.rstrip("\r")) == 1 or line[1] in " \t\r") and have_faces:
            break
    m = np.asarray(xf, dtype=np.float64).reshape(4, 4)
    P = np.asarray(pos, dtype=np.float64).reshape(-1, 3)
    Ph = np.stack([m[r, 0] * P[:, 0] + m[r, 1] * P[:, 1] + m[r, 2] * P[:, 2] + m[r, 3] for r in range(4)], 1)
    out_p = Ph[:, :3] / Ph[:, 3:4]
    Nn = np.asarray(nor, dtype=np.float64).reshape(-1, 3)
    out_n = np.stack([m[r, 0] * Nn[:, 0] + m[r, 1] * Nn[:, 1] + m[r, 2] * Nn[:, 2] for r in range(3)], 1) if len(nor) == len(pos) else np.zeros((0, 3))
    out_uv = np.asarray(tex, dtype=np.float64).reshape(-1, 2) if len(tex) == len(pos) else np.zeros((0, 2))
    return out_p, out_n, out_uv, np.asarray(ind, dtype=np.uint32)


def _load_obj_native(path, xf):
    h = C.c_void_p()
    N.check(N.lib().rth_scene_create_empty(C.byref(h)))
    N.check(N.lib().rth_scene_add_obj(h, str(path).encode(), N.ptr(np.asarray(xf, dtype=np.float64)), None))
    sz = (C.c_uint64 * 4)()
    N.check(N.lib().rth_scene_mesh_sizes(h, 0, sz))
    p = np.zeros((sz[0], 3)); n = np.zeros((sz[1], 3)); uv = np.zeros((sz[2], 2)); ind = np.zeros(sz[3], dtype=np.uint32)
    N.check(N.lib().rth_scene_mesh_copy(h, 0, N.ptr(p), N.ptr(n), N.ptr(uv), N.ptr(ind)))
    N.lib().rth_scene_destroy(h)
    return p, n, uv, ind


def test_parallel_obj_parser_matches_a_sequential_restatement(tmp_path):
    """The OBJ reader parses the file in one chunk per host thread (f-2: parallel ingest).  A synthetic file large enough for several chunks, with the
    features that make chunking delicate — relative (negative) indices that count back across chunk boundaries, polygons, `v//vn` and `v/vt` corners,
    tabs, CRLF lines, comments, a second object that must be ignored — must give exactly what a line-by-line restatement of tobj gives."""
    rng = np.random.default_rng(4)
    lines = ["# synthetic", "o first"]
    nv = nvt = nvn = 0
    for block in range(400):
        k = int(rng.integers(3, 9))
        for _ in range(k):
            x, y, z = rng.normal(size=3)
            lines.append(("v\t%.7g %.7g %.7g" if block % 7 == 0 else "v %.7g %.7g %.7g") % (x, y, z)); nv += 1
            lines.append("vn %.6f %.6f %.6f" % tuple(rng.normal(size=3))); nvn += 1
            lines.append("vt %.5f %.5f" % tuple(rng.uniform(size=2))); nvt += 1
        for _ in range(20):
            m = int(rng.integers(3, 6))
            style = int(rng.integers(0, 4))
            toks = []
            for _ in range(m):
                if style == 0:      # absolute, all three, possibly far back
                    i = int(rng.integers(1, nv + 1)); toks.append(f"{i}/{int(rng.integers(1, nvt + 1))}/{int(rng.integers(1, nvn + 1))}")
                elif style == 1:    # relative
                    i = -int(rng.integers(1, min(nv, 50) + 1)); toks.append(f"{i}/{i}/{i}")
                elif style == 2:
                    i = int(rng.integers(1, nv + 1)); toks.append(f"{i}//{i}")
                else:
                    i = int(rng.integers(1, nv + 1)); toks.append(f"{i}/{i}/{i}")
            lines.append("f " + " ".join(toks) + ("\r" if block % 5 == 0 else "") + (" # c" if block % 11 == 0 and not block % 5 == 0 else ""))
    lines += ["o second", "v 9 9 9", "f 1 2 -1"]
    path = tmp_path / "big.obj"
    path.write_bytes(("\n".join(lines) + "\n").encode())
    assert path.stat().st_size > 3 * 256 * 1024 // 2          # several 256 KB chunks
    xf = [2, 0, 0, 1, 0, 3, 0, -2, 0.5, 0, 1, 0, 0, 0, 0, 2]
    got, want = _load_obj_native(path, xf), _py_obj(path, xf)
    for g, w, what in zip(got, want, ("points", "normals", "uvs", "indices")):
        assert g.shape == w.shape and np.array_equal(g, w), what


@pytest.mark.skipif(not os.path.exists(REF_DATA + "/stormtrooper/shape_0.obj"), reason="reference assets only exist in the build container")
@pytest.mark.parametrize("asset", ["material/models/Mesh001.obj", "stormtrooper/shape_0.obj", "stormtrooper/shape_mesh.obj"])
def test_parallel_obj_parser_on_reference_assets(asset):
    if not os.path.exists(REF_DATA + "/" + asset):
        pytest.skip("asset absent")
    xf = np.eye(4).reshape(16)
    got, want = _load_obj_native(REF_DATA + "/" + asset, xf), _py_obj(REF_DATA + "/" + asset, xf)
    for g, w, what in zip(got, want, ("points", "normals", "uvs", "indices")):
        assert g.shape == w.shape and np.array_equal(g, w), (asset, what)


@pytest.mark.parametrize("text,ok", [
    (b"", False), (b"# only comments\n", False), (b"v 0 0 0\nv 1 0 0\n", False),                      # no faces: parser.rs unwraps models[0]
    (b"v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 4\n", False), (b"v 0 0 0\nv 1 0 0\nv 0 1 0\nf -1 -2 -4\n", False),   # index past the vertices read so far
    (b"v 0 0 0\nv 1 0 0\nv 0 1 0\nf 0 1 2\n", False),                                                     # OBJ indices start at 1
    (b"v 0 0 0\nv 1 0 0\nf 1 2 3\nv 0 1 0\n", False),                                                     # a face may only look back
    (b"v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3", True),                                                        # no newline at the end of the file
    (b"v\t0 0 0\r\nv 1 0 0\r\nv 0 1 0\r\nf\t1 2 3\r\n", True),
    (b"v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2\nf 1 2 3\n", True),                                             # a two-corner face yields no triangle
])
def test_obj_reader_edge_cases(tmp_path, text, ok):
    path = tmp_path / "case.obj"
    path.write_bytes(text)
    h = C.c_void_p()
    N.check(N.lib().rth_scene_create_empty(C.byref(h)))
    rc = N.lib().rth_scene_add_obj(h, str(path).encode(), N.ptr(np.eye(4).reshape(16)), None)
    if ok:
        assert rc == 0
        sz = (C.c_uint64 * 4)()
        N.check(N.lib().rth_scene_mesh_sizes(h, 0, sz))
        assert (sz[0], sz[3]) == (3, 3)
    else:
        assert rc != 0
        N.lib().rth_last_error_string.restype = C.c_char_p
        assert b"Failed to parse obj" in N.lib().rth_last_error_string()
    N.lib().rth_scene_destroy(h)
