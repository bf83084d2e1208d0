"""Host-side mirror of the reference's interface for the hot path, over the C ABI.

Names follow the reference (calvin-godfrey/RustRaytracer, `src/`):
  Objects            ~ `struct Objects` / `get_objects_mut()`        geometry.rs:13-55
  Camera             ~ `Camera::new / new_motion_blur`               geometry.rs:110-175
  IntType            ~ `enum IntType`                                integrator.rs:26-41
  Samplers           ~ `Samplers::new_zero_two_sequence_sampler`     sampler.rs:633-642
  get_intersection   ~ `geometry::get_intersection`                  geometry.rs:238-240
  unoccluded         ~ `Visibility::unoccluded`                      hittable.rs:25-39
  tile_multithread   ~ `render::tile_multithread`                    render.rs:13-159
Everything computes in librtcore.so on the GPU; this file only marshals buffers.
"""
import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _native as N
from ._native import Camera as _CCamera, check, lib, ptr

MAX_DEPTH = 25        # consts.rs:7
TILE_SIZE = 16        # consts.rs:10
SMALL = 0.001         # consts.rs:32
INFINITY = 1e308      # consts.rs:30


class Camera:
    """`struct Camera` (geometry.rs:95-108)."""

    def __init__(self, c=None):
        self.c = c if c is not None else _CCamera()

    @staticmethod
    def new(frm, to, up, aspect_ratio, vfov, aperture, focus_dist):
        return Camera.new_motion_blur(frm, to, up, aspect_ratio, vfov, aperture, focus_dist, 0.0, 0.0)

    @staticmethod
    def new_motion_blur(frm, to, up, aspect_ratio, vfov, aperture, focus_dist, t0, t1):
        cam = Camera()
        f, t, u = (np.ascontiguousarray(x, dtype=np.float64) for x in (frm, to, up))
        check(lib().rt_camera_look_at(ptr(f), ptr(t), ptr(u), aspect_ratio, vfov, aperture, focus_dist, t0, t1, C.byref(cam.c)))
        return cam

    def primary_rays(self, width, height):
        """Pixel-centre rays (SURVEY.md §8d batch A), `Camera::get_ray` with lens_radius = 0."""
        rays = np.zeros(width * height, dtype=N.ray_dtype)
        check(lib().rth_primary_rays(C.byref(self.c), width, height, ptr(rays)))
        return rays


@dataclass
class IntType:
    """`IntType` (integrator.rs:25-41).  `Path { max_depth, invisible_light }` is the only integrator the reference ever
    selects (main.rs:262-265); `Direct { max_depth, strategy }` and `Whitted { max_depth }` run on the same wavefront
    (f64 radiometry only), as does `Basic { max_depth }` (geometry::cast_ray, geometry.rs:242-277)."""
    max_depth: int = MAX_DEPTH
    invisible_light: bool = False
    # radiometric precision of the shade kernel: False = f64 in the reference's operation order (per-pixel parity
    # with the f64 oracle), True = f32 (production speed).  Intersection decisions are f64-exact in both.
    shade_f32: bool = False
    kind: int = 0                # rtcore.h RT_INTEGRATOR_*: 0 Path, 1 Direct{UniformOne}, 2 Direct{UniformAll}, 3 Whitted, 4 Basic

    @staticmethod
    def Path(max_depth=MAX_DEPTH, invisible_light=False, shade_f32=False):
        return IntType(max_depth, invisible_light, shade_f32, 0)

    @staticmethod
    def Direct(max_depth=MAX_DEPTH, strategy="UniformOne"):
        """`LightStrategy::UniformOne` / `UniformAll` (integrator.rs:19-23)"""
        if strategy not in ("UniformOne", "UniformAll"):
            raise ValueError("strategy must be 'UniformOne' or 'UniformAll'")
        return IntType(max_depth, False, False, 1 if strategy == "UniformOne" else 2)

    @staticmethod
    def Whitted(max_depth=MAX_DEPTH):
        return IntType(max_depth, False, False, 3)

    @staticmethod
    def Basic(max_depth=MAX_DEPTH):
        return IntType(max_depth, False, False, 4)


@dataclass
class Samplers:
    """Effective behaviour of `Samplers::new_zero_two_sequence_sampler(spp, 0)`: next_power_of_two(spp)
    iid uniforms per pixel (sampler.rs:633-642, 306-339); here from a counter-based RNG keyed by `seed`."""
    samples_per_pixel: int
    seed: int = 0

    @staticmethod
    def new_zero_two_sequence_sampler(samples_per_pixel, n_sampled_dimensions=0, seed=0):
        return Samplers(int(lib().rt_round_spp(int(samples_per_pixel))), seed)


class Objects:
    """The scene tables + committed BVH (`struct Objects`, geometry.rs:13-21) held by the library."""

    def __init__(self):
        self.h = C.c_void_p()
        check(lib().rt_scene_create(C.byref(self.h)))

    def close(self):
        if self.h:
            lib().rt_scene_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- pushes (scenes.rs style)
    def add_mesh(self, points, normals=None, uvs=None, indices=None):
        p = np.ascontiguousarray(points, dtype=np.float64).reshape(-1, 3)
        n = None if normals is None or len(normals) == 0 else np.ascontiguousarray(normals, dtype=np.float64).reshape(-1, 3)
        uv = None if uvs is None or len(uvs) == 0 else np.ascontiguousarray(uvs, dtype=np.float64).reshape(-1, 2)
        ind = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1)
        out = C.c_uint32()
        check(lib().rt_scene_add_mesh(self.h, len(p), ptr(p), 0 if n is None else len(n), ptr(n), 0 if uv is None else len(uv), ptr(uv),
                                      len(ind), ptr(ind), C.byref(out)))
        return out.value

    def add_transform(self, m4):
        m = np.ascontiguousarray(m4, dtype=np.float64).reshape(16)
        out = C.c_uint32()
        check(lib().rt_scene_add_transform(self.h, ptr(m), C.byref(out)))
        return out.value

    def add_primitives(self, prims):
        prims = np.ascontiguousarray(prims, dtype=N.prim_dtype)
        out = C.c_uint32()
        check(lib().rt_scene_add_primitives(self.h, len(prims), ptr(prims), C.byref(out)))
        return out.value

    def add_mesh_triangles(self, mesh_index, mat_index):
        out = C.c_uint32()
        check(lib().rt_scene_add_mesh_triangles(self.h, mesh_index, mat_index, C.byref(out)))
        return out.value

    def set_light_index(self, prim, light):
        check(lib().rt_scene_set_light_index(self.h, prim, light))

    def add_image_hdr(self, rgb_f32):
        """`Texture::new_hdr`'s decoded pixels (material.rs:626-642): (h, w, 3) float32."""
        a = np.ascontiguousarray(rgb_f32, dtype=np.float32)
        out = C.c_uint32()
        check(lib().rt_scene_add_image_hdr(self.h, a.shape[1], a.shape[0], ptr(a), C.byref(out)))
        return out.value

    def add_image_rgb8(self, rgb_u8):
        """`Texture::new_texture`'s RgbImage (material.rs:613-616): (h, w, 3) uint8."""
        a = np.ascontiguousarray(rgb_u8, dtype=np.uint8)
        out = C.c_uint32()
        check(lib().rt_scene_add_image_rgb8(self.h, a.shape[1], a.shape[0], ptr(a), C.byref(out)))
        return out.value

    def add_perlin(self, scale, perm, vecs):
        """`Texture::new_perlin(scale)` → `Perlin::new(256, scale)` (material.rs:628-631, perlin.rs:17-34): perm = (3, 256) permutation
        tables x / y / z, vecs = (256, 3) unit gradient vectors (the reference draws both from its global RNG)."""
        pm = np.ascontiguousarray(perm, dtype=np.uint32).reshape(768)
        vc = np.ascontiguousarray(vecs, dtype=np.float64).reshape(768)
        out = C.c_uint32()
        check(lib().rt_scene_add_perlin(self.h, float(scale), ptr(pm), ptr(vc), C.byref(out)))
        return out.value

    def set_textures(self, t):
        t = np.ascontiguousarray(t, dtype=N.texture_dtype)
        check(lib().rt_scene_set_textures(self.h, len(t), ptr(t)))

    def set_materials(self, m):
        m = np.ascontiguousarray(m, dtype=N.material_dtype)
        check(lib().rt_scene_set_materials(self.h, len(m), ptr(m)))

    def set_lights(self, l):
        l = np.ascontiguousarray(l, dtype=N.light_dtype)
        check(lib().rt_scene_set_lights(self.h, len(l), ptr(l)))

    def add_disney(self, params):
        """`Material::make_disney(..)` (material.rs:480-516): one rt_disney record (a numpy record of _native.disney_dtype); returns its index,
        which an RT_MAT_DISNEY material carries in tex[0]."""
        d = np.ascontiguousarray(params, dtype=N.disney_dtype).reshape(1)
        out = C.c_uint32()
        check(lib().rt_scene_add_disney(self.h, ptr(d), C.byref(out)))
        return out.value

    def commit(self):
        """`objects.node = BvhNode::new(..)` (e.g. scenes.rs:368-371): build + flatten + upload."""
        check(lib().rt_scene_commit(self.h))

    def num_primitives(self):
        out = C.c_uint64()
        check(lib().rt_scene_num_primitives(self.h, C.byref(out)))
        return out.value

    def set_build_mode(self, mode):
        """"sah" (host binned SAH, default) or "lbvh" (Morton LBVH built on the GPU): which builder commit() uses."""
        m = {"sah": N.BUILD_HOST_SAH, "lbvh": N.BUILD_DEVICE_LBVH}.get(mode, mode)
        check(lib().rt_scene_set_build_mode(self.h, int(m)))

    def build_stats(self):
        b = N.BuildStats()
        check(lib().rt_scene_build_stats(self.h, C.byref(b)))
        return {k: getattr(b, k) for k, _ in N.BuildStats._fields_}

    def device_bytes(self):
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        check(lib().rt_scene_device_bytes(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return {"nodes": a.value, "leaves": b.value, "tables": c.value}

    def counters(self):
        c = N.Counters()
        check(lib().rt_get_counters(self.h, C.byref(c)))
        return {k: getattr(c, k) for k, _ in N.Counters._fields_}

    def set_wavefront_size(self, paths):
        check(lib().rt_set_wavefront_size(self.h, int(paths)))


def make_rays(origins, dirs):
    o = np.asarray(origins, dtype=np.float64).reshape(-1, 3)
    d = np.asarray(dirs, dtype=np.float64).reshape(-1, 3)
    rays = np.zeros(len(o), dtype=N.ray_dtype)
    rays["origin"] = o
    rays["dir"] = d
    return rays


def get_intersection(objects, rays, tmin=SMALL, tmax=INFINITY):
    """`geometry::get_intersection(&Ray) -> Option<HitRecord>` (geometry.rs:238-240) for a batch of rays.
    Returns a structured array (rt_hit); prim_index == RT_NONE is `None`."""
    rays = np.ascontiguousarray(rays, dtype=N.ray_dtype)
    hits = np.zeros(len(rays), dtype=N.hit_dtype)
    check(lib().rt_intersect_range(objects.h, len(rays), ptr(rays), float(tmin), float(tmax), ptr(hits)))
    return hits


def unoccluded(objects, p0, p1, p1_prim_index, infinite_light=False):
    """`Visibility{p0,p1}.unoccluded(infinite_light)` (hittable.rs:25-39) for a batch of point pairs."""
    p0 = np.asarray(p0, dtype=np.float64).reshape(-1, 3)
    q = np.zeros(len(p0), dtype=N.shadow_dtype)
    q["p0"] = p0
    q["p1"] = np.asarray(p1, dtype=np.float64).reshape(-1, 3)
    q["p1_prim_index"] = np.asarray(p1_prim_index, dtype=np.uint32)
    q["infinite_light"] = 1 if infinite_light else 0
    out = np.zeros(len(q), dtype=np.uint8)
    check(lib().rt_occluded(objects.h, len(q), ptr(q), ptr(out)))
    return out.astype(bool)


class Film:
    """The reference's film `Vec<Vec<(f64,f64,f64,u32)>>` (render.rs:18-22) as (h, w, 4) f64 sums on the device."""

    def __init__(self, width, height, device_ptr=None):
        self.width, self.height = width, height
        self.h = C.c_void_p()
        if device_ptr is None:
            check(lib().rt_film_create(width, height, C.byref(self.h)))
        else:
            check(lib().rt_film_create_external(width, height, C.c_void_p(device_ptr), C.byref(self.h)))

    def clear(self):
        check(lib().rt_film_clear(self.h))

    def read(self):
        out = np.zeros((self.height, self.width, 4), dtype=np.float64)
        check(lib().rt_film_read(self.h, ptr(out)))
        return out

    def resolve_rgb8(self):
        """`util::point_to_color` (util.rs:441-471) on the device."""
        out = np.zeros((self.height, self.width, 3), dtype=np.uint8)
        check(lib().rt_film_resolve_rgb8(self.h, ptr(out)))
        return out

    def resolve_rgb8_static(self):
        """rt_film_resolve_rgb8_static: the same image through the entry point that allocates nothing (static device scratch, films up to 2048 x 2048)."""
        out = np.zeros((self.height, self.width, 3), dtype=np.uint8)
        check(lib().rt_film_resolve_rgb8_static(self.h, ptr(out)))
        return out

    @staticmethod
    def reduce_resolve_rgb8(films):
        """rt_film_reduce_resolve_rgb8: the films of one process's GPUs summed (in list order) and resolved to RGB8 by one kernel on films[0]'s GPU,
        which reads the others over NVLink; no NCCL, the films stay as they are."""
        out = np.zeros((films[0].height, films[0].width, 3), dtype=np.uint8)
        arr = (C.c_void_p * len(films))(*[f.h for f in films])
        check(lib().rt_film_reduce_resolve_rgb8(arr, len(films), ptr(out)))
        return out

    def close(self):
        if self.h:
            lib().rt_film_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def render_samples(objects, camera, int_type, film, width, height, sample_begin, sample_end, seed=0, window=None):
    """`Integrator::render` over a pixel window and a sample range (integrator.rs:336-348): adds
    samples [sample_begin, sample_end) of every pixel in `window` = (x0, y0, x1, y1) to `film`."""
    x0, y0, x1, y1 = window if window is not None else (0, 0, width, height)
    integ = N.IntegratorDesc(int(int_type.kind), int(int_type.max_depth), 1 if int_type.invisible_light else 0, N.SHADE_F32 if int_type.shade_f32 else 0)
    fd = N.FilmDesc(width, height, int(sample_begin), int(sample_end), x0, y0, x1, y1, int(seed))
    check(lib().rt_render(objects.h, C.byref(camera.c), C.byref(integ), C.byref(fd), film.h))


class Comm:
    """NCCL communicator behind the C ABI (include/rtcore.h).  `Comm.local(devices)`: one process driving several GPUs;
    `Comm.rank(id, world, rank)`: one process per GPU (torchrun), `id` = Comm.unique_id() of rank 0 handed round out of band."""

    def __init__(self, h):
        self.h = h

    @staticmethod
    def local(devices):
        d = np.ascontiguousarray(devices, dtype=np.int32)
        h = C.c_void_p()
        check(lib().rt_comm_create_local(ptr(d), len(d), C.byref(h)))
        return Comm(h)

    @staticmethod
    def unique_id():
        out = np.zeros(128, dtype=np.uint8)
        check(lib().rt_comm_unique_id(ptr(out)))
        return out.tobytes()

    @staticmethod
    def rank(unique_id, world, rank):
        buf = np.frombuffer(unique_id, dtype=np.uint8).copy()
        h = C.c_void_p()
        check(lib().rt_comm_create_rank(ptr(buf), int(world), int(rank), C.byref(h)))
        return Comm(h)

    def size(self):
        w, n = C.c_uint32(), C.c_uint32()
        check(lib().rt_comm_size(self.h, C.byref(w), C.byref(n)))
        return w.value, n.value

    def allreduce(self, films):
        """in-place sum of the films over all ranks (one film per local device of the communicator)"""
        arr = (C.c_void_p * len(films))(*[f.h for f in films])
        check(lib().rt_film_allreduce(self.h, arr, len(films)))

    def close(self):
        if self.h:
            lib().rt_comm_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def replicate(objects, device):
    """the same scene committed on another GPU (rt_scene_replicate)"""
    o = Objects.__new__(Objects)
    o.h = C.c_void_p()
    check(lib().rt_scene_replicate(objects.h, int(device), C.byref(o.h)))
    return o


def render_multi(comm, objects_per_gpu, films_per_gpu, camera, int_type, width, height, sample_begin, sample_end, seed=0, window=None):
    """`render::tile_multithread` on the GPUs of a local communicator: the sample range is split across them and the films are summed
    (rt_render_multi).  Films are added to: clear them first."""
    x0, y0, x1, y1 = window if window is not None else (0, 0, width, height)
    integ = N.IntegratorDesc(int(int_type.kind), int(int_type.max_depth), 1 if int_type.invisible_light else 0, N.SHADE_F32 if int_type.shade_f32 else 0)
    fd = N.FilmDesc(width, height, int(sample_begin), int(sample_end), x0, y0, x1, y1, int(seed))
    n = len(objects_per_gpu)
    sc = (C.c_void_p * n)(*[o.h for o in objects_per_gpu])
    fl = (C.c_void_p * n)(*[f.h for f in films_per_gpu])
    check(lib().rt_render_multi(comm.h, sc, fl, n, C.byref(camera.c), C.byref(integ), C.byref(fd)))


def tile_multithread(path, camera, sampler, int_type, objects, width, height, film=None, write_png=True):
    """`render::tile_multithread(path, camera, sampler, int_type)` (render.rs:13-159): renders
    `sampler.samples_per_pixel` samples for every pixel and (optionally) writes the tonemapped PNG to `path`.
    The 6 worker threads / 16x16 tile queue of the reference become one wavefront on the GPU.
    Returns the film as an (h, w, 4) array of (r, g, b, n) sums."""
    own = film is None
    if own:
        film = Film(width, height)
    render_samples(objects, camera, int_type, film, width, height, 0, sampler.samples_per_pixel, sampler.seed)
    if write_png and path:
        from PIL import Image
        Image.fromarray(film.resolve_rgb8(), "RGB").save(path)
    out = film.read()
    if own:
        film.close()
    return out
