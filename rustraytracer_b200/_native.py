"""ctypes binding of librtcore.so (include/rtcore.h, include/rthost.h).

This is harness plumbing only: every computation happens inside the CUDA library.  There is no
Python or CPU fallback — if the library (or a CUDA device) is missing, calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_LIBRTCORE", os.path.join(_HERE, "librtcore.so"))   # override only for A/B builds of the same library

RT_NONE = 0xFFFFFFFF
RT_OK, RT_ERR_INVALID, RT_ERR_CUDA, RT_ERR_STATE, RT_ERR_CANCELLED, RT_ERR_IO = range(6)
PRIM_SPHERE, PRIM_TRIANGLE, PRIM_XY_RECT, PRIM_XZ_RECT, PRIM_YZ_RECT = range(5)
TEX_SOLID, TEX_CHECKERED, TEX_HDR, TEX_IMAGE, TEX_PERLIN = 0, 1, 2, 3, 4
LIGHT_DIFFUSE, LIGHT_INFINITE, LIGHT_POINT, LIGHT_SPOT, LIGHT_DISTANT = range(5)
SHADE_F32 = 1
MAT_MATTE, MAT_LIGHT, MAT_PLASTIC, MAT_GLASS, MAT_METAL, MAT_MIRROR, MAT_DISNEY = range(7)

# numpy mirrors of the POD structs (align=True reproduces the C layout)
prim_dtype = np.dtype([("type", "<u4"), ("flip_face", "<u4"), ("mat_index", "<u4"), ("light_index", "<u4"),
                       ("mesh_index", "<u4"), ("ind", "<u4"), ("transform", "<u4"), ("_pad", "<u4"), ("v", "<f8", (5,))], align=True)
texture_dtype = np.dtype([("type", "<u4"), ("odd", "<u4"), ("even", "<u4"), ("image", "<u4"), ("color", "<f8", (3,)), ("frequency", "<f8")], align=True)
material_dtype = np.dtype([("type", "<u4"), ("remap_roughness", "<u4"), ("tex", "<u4", (5,)), ("_pad", "<u4"), ("f", "<f8", (3,))], align=True)
DISNEY_FIELDS = ("color", "metallic", "eta", "roughness", "specular_tint", "anisotropic", "sheen", "sheen_tint", "clearcoat", "clearcoat_gloss",
                 "spec_trans", "scatter_distance", "flatness", "diff_trans", "thin", "_pad")
disney_dtype = np.dtype([(f, "<u4") for f in DISNEY_FIELDS], align=True)
light_dtype = np.dtype([("type", "<u4"), ("prim_index", "<u4"), ("two_sided", "<u4"), ("text_id", "<u4"), ("color", "<f8", (3,)),
                        ("transform", "<u4"), ("_pad", "<u4"), ("param", "<f8", (3,))], align=True)
ray_dtype = np.dtype([("origin", "<f8", (3,)), ("dir", "<f8", (3,))], align=True)
hit_dtype = np.dtype([("t", "<f8"), ("p", "<f8", (3,)), ("n", "<f8", (3,)), ("ns", "<f8", (3,)), ("uv", "<f8", (2,)),
                      ("prim_index", "<u4"), ("front", "<u4")], align=True)
shadow_dtype = np.dtype([("p0", "<f8", (3,)), ("p1", "<f8", (3,)), ("p1_prim_index", "<u4"), ("infinite_light", "<u4")], align=True)
assert prim_dtype.itemsize == 72 and texture_dtype.itemsize == 48 and material_dtype.itemsize == 56
assert disney_dtype.itemsize == 64 and light_dtype.itemsize == 72 and ray_dtype.itemsize == 48 and hit_dtype.itemsize == 104 and shadow_dtype.itemsize == 56


class Camera(C.Structure):
    _fields_ = [("origin", C.c_double * 3), ("upper_left_corner", C.c_double * 3), ("horizontal_offset", C.c_double * 3),
                ("vertical_offset", C.c_double * 3), ("u", C.c_double * 3), ("v", C.c_double * 3), ("w", C.c_double * 3),
                ("lens_radius", C.c_double), ("t0", C.c_double), ("t1", C.c_double)]


class IntegratorDesc(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("max_depth", C.c_uint32), ("invisible_light", C.c_uint32), ("flags", C.c_uint32)]


class FilmDesc(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("sample_begin", C.c_uint32), ("sample_end", C.c_uint32),
                ("x0", C.c_uint32), ("y0", C.c_uint32), ("x1", C.c_uint32), ("y1", C.c_uint32), ("seed", C.c_uint64)]


class Counters(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("extend_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("probe_rays", C.c_uint64),
                ("iterations", C.c_uint64), ("render_ms", C.c_double), ("extend_ms", C.c_double), ("shade_ms", C.c_double),
                ("connect_ms", C.c_double)]


class HostParams(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("variant", C.c_uint32), ("_pad", C.c_uint32),
                ("n_tris", C.c_uint64), ("seed", C.c_uint64), ("data_dir", C.c_char_p)]


BUILD_HOST_SAH, BUILD_DEVICE_LBVH = 0, 1


class BuildStats(C.Structure):
    _fields_ = [("mode", C.c_uint32), ("max_stack", C.c_uint32), ("nodes", C.c_uint64), ("build_ms", C.c_double), ("commit_ms", C.c_double)]


# every symbol include/rtcore.h and include/rthost.h declare (tests check that they all resolve)
RTCORE_SYMBOLS = [
    "rt_last_error_string", "rt_version_string", "rt_device_count", "rt_set_device", "rt_round_spp", "rt_camera_look_at",
    "rt_scene_create", "rt_scene_destroy", "rt_scene_add_mesh", "rt_scene_add_transform", "rt_scene_add_primitives",
    "rt_scene_add_mesh_triangles", "rt_scene_set_light_index", "rt_scene_add_image_hdr", "rt_scene_add_image_rgb8", "rt_scene_add_perlin", "rt_scene_add_disney", "rt_scene_set_textures", "rt_scene_set_materials",
    "rt_scene_set_lights", "rt_scene_commit", "rt_scene_num_primitives", "rt_scene_device_bytes", "rt_intersect",
    "rt_intersect_range", "rt_intersect_device", "rt_closest_device", "rt_occluded", "rt_occluded_device", "rt_film_create",
    "rt_film_create_external", "rt_film_destroy", "rt_film_clear", "rt_film_read", "rt_film_resolve_rgb8", "rt_render",
    "rt_render_cancel", "rt_get_counters", "rt_set_wavefront_size", "rt_scene_set_build_mode", "rt_scene_build_stats",
    "rt_debug_prefilter_stats", "rt_film_resolve_rgb8_static",
    "rt_comm_create_local", "rt_comm_unique_id", "rt_comm_create_rank", "rt_comm_destroy", "rt_comm_size", "rt_film_allreduce", "rt_scene_replicate", "rt_render_multi",
    "rt_film_reduce_resolve_rgb8",
]
RTHOST_SYMBOLS = [
    "rth_scene_build", "rth_scene_destroy", "rth_scene_camera", "rth_scene_png_name", "rth_scene_uses_standin", "rth_scene_counts",
    "rth_scene_mesh_sizes", "rth_scene_mesh_copy", "rth_scene_copy_primitives", "rth_scene_copy_transforms", "rth_scene_copy_textures",
    "rth_scene_copy_materials", "rth_scene_copy_lights", "rth_scene_upload", "rth_scene_add_obj", "rth_scene_create_empty", "rth_primary_rays",
    "rth_scene_num_images", "rth_scene_image_info", "rth_scene_image_copy", "rth_load_hdr", "rth_last_error_string",
    "rth_scene_num_disney", "rth_scene_copy_disney",
]

HOST_LIB_PATH = os.environ.get("RT_LIBRTHOST", os.path.join(_HERE, "librthost.so"))


class RtError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"rtcore status {status}: {message}")
        self.status = status


def _missing(path):
    return ImportError(f"{path} is missing: build it with __graft_entry__.build(); there is no fallback path")


def _load_core():
    """librtcore.so: the CUDA library (include/rtcore.h)."""
    if not os.path.exists(LIB_PATH):
        raise _missing(LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, u32, u64, f64 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_double
    L.rt_last_error_string.restype = C.c_char_p
    L.rt_version_string.restype = C.c_char_p
    L.rt_round_spp.restype = u32
    L.rt_round_spp.argtypes = [u32]
    sigs = {
        "rt_set_device": [C.c_int32],
        "rt_camera_look_at": [vp, vp, vp, f64, f64, f64, f64, f64, f64, vp],
        "rt_scene_create": [vp], "rt_scene_destroy": [vp],
        "rt_scene_add_mesh": [vp, u64, vp, u64, vp, u64, vp, u64, vp, vp],
        "rt_scene_add_transform": [vp, vp, vp], "rt_scene_add_primitives": [vp, u64, vp, vp],
        "rt_scene_add_mesh_triangles": [vp, u32, u32, vp], "rt_scene_set_light_index": [vp, u32, u32],
        "rt_scene_add_image_hdr": [vp, u32, u32, vp, vp], "rt_scene_add_image_rgb8": [vp, u32, u32, vp, vp], "rt_scene_add_perlin": [vp, f64, vp, vp, vp], "rt_scene_add_disney": [vp, vp, vp],
        "rt_scene_set_textures": [vp, u64, vp], "rt_scene_set_materials": [vp, u64, vp], "rt_scene_set_lights": [vp, u64, vp],
        "rt_scene_commit": [vp], "rt_scene_num_primitives": [vp, vp], "rt_scene_device_bytes": [vp, vp, vp, vp],
        "rt_intersect": [vp, u64, vp, vp], "rt_intersect_range": [vp, u64, vp, f64, f64, vp],
        "rt_intersect_device": [vp, u64, vp, f64, f64, vp, vp], "rt_closest_device": [vp, u64, vp, f64, f64, vp, vp, vp],
        "rt_occluded": [vp, u64, vp, vp], "rt_occluded_device": [vp, u64, vp, vp, vp],
        "rt_film_create": [u32, u32, vp], "rt_film_create_external": [u32, u32, vp, vp], "rt_film_destroy": [vp],
        "rt_film_clear": [vp], "rt_film_read": [vp, vp], "rt_film_resolve_rgb8": [vp, vp], "rt_film_resolve_rgb8_static": [vp, vp],
        "rt_render": [vp, vp, vp, vp, vp], "rt_render_cancel": [vp], "rt_get_counters": [vp, vp], "rt_set_wavefront_size": [vp, u32],
        "rt_scene_set_build_mode": [vp, u32], "rt_scene_build_stats": [vp, vp], "rt_debug_prefilter_stats": [vp],
        "rt_comm_create_local": [vp, u32, vp], "rt_comm_unique_id": [vp], "rt_comm_create_rank": [vp, u32, u32, vp], "rt_comm_destroy": [vp],
        "rt_comm_size": [vp, vp, vp], "rt_film_allreduce": [vp, vp, u32], "rt_scene_replicate": [vp, C.c_int32, vp], "rt_render_multi": [vp, vp, vp, u32, vp, vp, vp],
        "rt_film_reduce_resolve_rgb8": [vp, u32, vp],
    }
    for name, args in sigs.items():
        if not hasattr(L, name):
            continue          # an older A/B build of the library (RT_LIBRTCORE); tests/test_abi.py checks the real one exports everything
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = C.c_int32
    L.rt_device_count.restype = C.c_int32
    return L


def _load_host():
    """librthost.so: the host scene layer (include/rthost.h).  Pure host code: loading it does not map the CUDA library."""
    if not os.path.exists(HOST_LIB_PATH):
        raise _missing(HOST_LIB_PATH)
    L = C.CDLL(HOST_LIB_PATH)
    vp, u32, u64 = C.c_void_p, C.c_uint32, C.c_uint64
    L.rth_scene_png_name.restype = C.c_char_p
    L.rth_scene_png_name.argtypes = [vp]
    L.rth_last_error_string.restype = C.c_char_p
    sigs = {
        "rth_scene_build": [C.c_char_p, vp, vp], "rth_scene_destroy": [vp], "rth_scene_camera": [vp, vp],
        "rth_scene_uses_standin": [vp], "rth_scene_counts": [vp, vp], "rth_scene_mesh_sizes": [vp, u32, vp],
        "rth_scene_mesh_copy": [vp, u32, vp, vp, vp, vp], "rth_scene_copy_primitives": [vp, vp], "rth_scene_copy_transforms": [vp, vp],
        "rth_scene_copy_textures": [vp, vp], "rth_scene_copy_materials": [vp, vp], "rth_scene_copy_lights": [vp, vp],
        "rth_scene_upload": [vp, vp], "rth_scene_add_obj": [vp, C.c_char_p, vp, vp], "rth_scene_create_empty": [vp],
        "rth_primary_rays": [vp, u32, u32, vp],
        "rth_scene_num_images": [vp, vp], "rth_scene_image_info": [vp, u32, vp], "rth_scene_image_copy": [vp, u32, vp],
        "rth_load_hdr": [C.c_char_p, vp, vp, vp, u64],
        "rth_scene_num_disney": [vp, vp], "rth_scene_copy_disney": [vp, vp],
    }
    for name, args in sigs.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = None if name in ("rth_scene_destroy",) else C.c_int32
    return L


class _Libs:
    """`lib().rt_*` resolves in librtcore.so, `lib().rth_*` in librthost.so; each library is loaded on first use, so a process that only
    builds scene tables (bench.py --impl reference: host layer + oracle) never maps the CUDA library."""
    core = None
    host = None

    def __getattr__(self, name):
        if name.startswith("rth_"):
            if _Libs.host is None:
                _Libs.host = _load_host()
            return getattr(_Libs.host, name)
        if _Libs.core is None:
            _Libs.core = _load_core()
        return getattr(_Libs.core, name)


_libs = _Libs()


def lib():
    return _libs


def check(status):
    if status != RT_OK:
        msgs = []
        if _Libs.core is not None:
            msgs.append(_Libs.core.rt_last_error_string().decode())
        if _Libs.host is not None and _Libs.host.rth_last_error_string():
            msgs.append("host layer: " + _Libs.host.rth_last_error_string().decode())
        raise RtError(status, " | ".join(m for m in msgs if m))


def ptr(a):
    """void* of a C-contiguous numpy array (or None)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)
