"""The reference's scene functions (scenes.rs) behind the C++ host scene layer (include/rthost.h).

Each function returns what the reference's does — `(png path, Camera, Samplers)` — plus the `Objects`
it filled (the reference fills a process global).  `HostScene` also exports the plain tables so that
tests can hand the *same* inputs to the oracle.
"""
import ctypes as C
import os

import numpy as np

from . import _native as N
from ._native import check, lib, ptr
from .core import Camera, Objects, Samplers


class HostScene:
    def __init__(self, name, width=720, height=720, variant=0, n_tris=0, seed=0, data_dir=None):
        self.name = name
        self.h = C.c_void_p()
        if data_dir is None:
            data_dir = os.environ.get("RT_DATA_DIR")
        p = N.HostParams(width, height, variant, 0, n_tris, seed, data_dir.encode() if data_dir else None)
        check(lib().rth_scene_build(name.encode(), C.byref(p), C.byref(self.h)))
        self.width, self.height = width, height

    def close(self):
        if self.h:
            lib().rth_scene_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def camera(self):
        cam = Camera()
        check(lib().rth_scene_camera(self.h, C.byref(cam.c)))
        return cam

    def png_name(self):
        return lib().rth_scene_png_name(self.h).decode()

    def uses_standin(self):
        return bool(lib().rth_scene_uses_standin(self.h))

    def counts(self):
        c = (C.c_uint64 * 6)()
        check(lib().rth_scene_counts(self.h, c))
        return dict(zip(("meshes", "prims", "transforms", "textures", "materials", "lights"), [int(x) for x in c]))

    def tables(self):
        """Plain tables: dict(meshes=[(p,n,uv,ind)], prims, transforms, textures, materials, lights)."""
        cnt = self.counts()
        meshes = []
        for m in range(cnt["meshes"]):
            sz = (C.c_uint64 * 4)()
            check(lib().rth_scene_mesh_sizes(self.h, m, sz))
            p = np.zeros((sz[0], 3)); n = np.zeros((sz[1], 3)); uv = np.zeros((sz[2], 2)); ind = np.zeros(sz[3], dtype=np.uint32)
            check(lib().rth_scene_mesh_copy(self.h, m, ptr(p), ptr(n), ptr(uv), ptr(ind)))
            meshes.append((p, n, uv, ind))
        prims = np.zeros(cnt["prims"], dtype=N.prim_dtype)
        if len(prims):
            check(lib().rth_scene_copy_primitives(self.h, ptr(prims)))
        xf = np.zeros((cnt["transforms"], 16))
        if len(xf):
            check(lib().rth_scene_copy_transforms(self.h, ptr(xf)))
        tex = np.zeros(cnt["textures"], dtype=N.texture_dtype)
        check(lib().rth_scene_copy_textures(self.h, ptr(tex)))
        mat = np.zeros(cnt["materials"], dtype=N.material_dtype)
        check(lib().rth_scene_copy_materials(self.h, ptr(mat)))
        lig = np.zeros(cnt["lights"], dtype=N.light_dtype)
        if len(lig):
            check(lib().rth_scene_copy_lights(self.h, ptr(lig)))
        images = []
        n_img = C.c_uint64()
        check(lib().rth_scene_num_images(self.h, C.byref(n_img)))
        for i in range(n_img.value):
            info = (C.c_uint32 * 3)()
            check(lib().rth_scene_image_info(self.h, i, info))
            px = np.zeros((info[1], info[0], 3), dtype=np.float32 if info[2] else np.uint8)
            check(lib().rth_scene_image_copy(self.h, i, ptr(px)))
            images.append(px)
        n_dis = C.c_uint64()
        check(lib().rth_scene_num_disney(self.h, C.byref(n_dis)))
        dis = np.zeros(n_dis.value, dtype=N.disney_dtype)
        if len(dis):
            check(lib().rth_scene_copy_disney(self.h, ptr(dis)))
        return dict(meshes=meshes, prims=prims, transforms=xf, textures=tex, materials=mat, lights=lig, images=images, disneys=dis)

    def upload(self, commit=True, build="sah"):
        """Push the tables through the public rt_scene_* ABI and commit (BVH build + device upload).
        build = "sah" (host binned SAH) or "lbvh" (device LBVH)."""
        obj = Objects()
        check(lib().rth_scene_upload(self.h, obj.h))
        obj.set_build_mode(build)
        if commit:
            obj.commit()
        return obj


def _scene_kw(name, width, height, spp, seed, kw):
    hs = HostScene(name, width, height, **kw)
    objects = hs.upload()
    return hs.png_name(), hs.camera(), Samplers.new_zero_two_sequence_sampler(spp, 0, seed), objects, hs


def sphere_roughness(width=720, height=720, spp=50, seed=0):          # scenes.rs:474
    return _scene_kw("sphere_roughness", width, height, spp, seed, {})


def make_world(width=720, height=720, spp=50, seed=0, scene_seed=1):  # scenes.rs:15-86 (commented), seeded
    return _scene_kw("make_world", width, height, spp, seed, dict(seed=scene_seed))


def cornell_box(width=720, height=720, spp=50, seed=0):               # scenes.rs:89
    return _scene_kw("cornell_box", width, height, spp, seed, {})


def cornell_box_statue(width=720, height=720, spp=50, seed=0, variant=0, n_tris=0, data_dir=None):   # scenes.rs:200
    return _scene_kw("cornell_box_statue", width, height, spp, seed, dict(variant=variant, n_tris=n_tris, data_dir=data_dir))


def plastic_dragon(width=720, height=720, spp=50, seed=0, variant=0, data_dir=None):                # scenes.rs:310
    return _scene_kw("plastic_dragon", width, height, spp, seed, dict(variant=variant, data_dir=data_dir))


def two_dragons(width=720, height=720, spp=50, seed=0, variant=1, data_dir=None):                   # scenes.rs:549
    return _scene_kw("two_dragons", width, height, spp, seed, dict(variant=variant, data_dir=data_dir))


def material_hdr(mat_num=0, width=720, height=720, spp=50, seed=0, data_dir=None):                   # scenes.rs:627
    return _scene_kw("material_hdr", width, height, spp, seed, dict(variant=mat_num, data_dir=data_dir))


def teapot_hdr(width=1280, height=720, spp=50, seed=0, data_dir=None):                                     # scenes.rs:744
    return _scene_kw("teapot_hdr", width, height, spp, seed, dict(data_dir=data_dir))


def soup(n_tris, width=4096, height=4096, seed=0):
    return _scene_kw("soup", width, height, 1, 0, dict(n_tris=n_tris, seed=seed))
