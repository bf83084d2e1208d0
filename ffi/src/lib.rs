//! rtcore-sys — Rust declarations of `include/rtcore.h`, the C ABI of the B200 path-tracing core (`librtcore.so`).
//!
//! Every item mirrors the header one to one (same names, same field order, `#[repr(C)]`); the header cites, per entry point, the item of
//! calvin-godfrey/RustRaytracer it stands behind (`geometry::get_intersection`, `Visibility::unoccluded`, `render::tile_multithread`, the
//! `get_objects_mut()` pushes of `scenes.rs`, ...).  INTEGRATION.md shows the safe wrappers a maintainer adds on top.
//! Nothing unwinds across this boundary: every call returns an `RtStatus`, the message of the last failure on the calling thread is
//! `rt_last_error_string()`.
#![allow(non_camel_case_types)]

use std::os::raw::{c_char, c_void};

pub const RT_NONE: u32 = 0xFFFF_FFFF;

pub type RtStatus = i32;
pub const RT_OK: RtStatus = 0;
pub const RT_ERR_INVALID: RtStatus = 1;
pub const RT_ERR_CUDA: RtStatus = 2;
pub const RT_ERR_STATE: RtStatus = 3;
pub const RT_ERR_CANCELLED: RtStatus = 4;
pub const RT_ERR_IO: RtStatus = 5;

// enum Primitive (primitive.rs:10-61)
pub const RT_PRIM_SPHERE: u32 = 0;
pub const RT_PRIM_TRIANGLE: u32 = 1;
pub const RT_PRIM_XY_RECT: u32 = 2;
pub const RT_PRIM_XZ_RECT: u32 = 3;
pub const RT_PRIM_YZ_RECT: u32 = 4;
// enum Texture (material.rs:520-539)
pub const RT_TEX_SOLID: u32 = 0;
pub const RT_TEX_CHECKERED: u32 = 1;
pub const RT_TEX_HDR: u32 = 2;
pub const RT_TEX_IMAGE: u32 = 3;
pub const RT_TEX_PERLIN: u32 = 4;
// enum Material (material.rs:17-73)
pub const RT_MAT_MATTE: u32 = 0;
pub const RT_MAT_LIGHT: u32 = 1;
pub const RT_MAT_PLASTIC: u32 = 2;
pub const RT_MAT_GLASS: u32 = 3;
pub const RT_MAT_METAL: u32 = 4;
pub const RT_MAT_MIRROR: u32 = 5;
pub const RT_MAT_DISNEY: u32 = 6;
// enum Light (light.rs:30-80)
pub const RT_LIGHT_DIFFUSE: u32 = 0;
pub const RT_LIGHT_INFINITE: u32 = 1;
pub const RT_LIGHT_POINT: u32 = 2;
pub const RT_LIGHT_SPOT: u32 = 3;
pub const RT_LIGHT_DISTANT: u32 = 4;
// enum IntType (integrator.rs:25-41)
pub const RT_INTEGRATOR_PATH: u32 = 0;
pub const RT_INTEGRATOR_DIRECT: u32 = 1;
pub const RT_INTEGRATOR_DIRECT_ALL: u32 = 2;
pub const RT_INTEGRATOR_WHITTED: u32 = 3;
pub const RT_INTEGRATOR_BASIC: u32 = 4;
pub const RT_SHADE_F32: u32 = 1;
pub const RT_BUILD_HOST_SAH: u32 = 0;
pub const RT_BUILD_DEVICE_LBVH: u32 = 1;
pub const RT_RESOLVE_STATIC_MAX_PIXELS: u32 = 2048 * 2048;

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_primitive {
    pub type_: u32,
    pub flip_face: u32,
    pub mat_index: u32,
    pub light_index: u32,
    pub mesh_index: u32,
    pub ind: u32,
    pub transform: u32,
    pub _pad: u32,
    pub v: [f64; 5],
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_texture {
    pub type_: u32,
    pub odd: u32,
    pub even: u32,
    pub image: u32,
    pub color: [f64; 3],
    pub frequency: f64,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_material {
    pub type_: u32,
    pub remap_roughness: u32,
    pub tex: [u32; 5],
    pub _pad: u32,
    pub f: [f64; 3],
}

/// Material::Disney (material.rs:55-73): the 14 texture ids + `thin`, referenced from an `rt_material` of type RT_MAT_DISNEY through `tex[0]`.
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_disney {
    pub color: u32,
    pub metallic: u32,
    pub eta: u32,
    pub roughness: u32,
    pub specular_tint: u32,
    pub anisotropic: u32,
    pub sheen: u32,
    pub sheen_tint: u32,
    pub clearcoat: u32,
    pub clearcoat_gloss: u32,
    pub spec_trans: u32,
    pub scatter_distance: u32,
    pub flatness: u32,
    pub diff_trans: u32,
    pub thin: u32,
    pub _pad: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_light {
    pub type_: u32,
    pub prim_index: u32,
    pub two_sided: u32,
    pub text_id: u32,
    pub color: [f64; 3],
    pub transform: u32,
    pub _pad: u32,
    pub param: [f64; 3],
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_camera {
    pub origin: [f64; 3],
    pub upper_left_corner: [f64; 3],
    pub horizontal_offset: [f64; 3],
    pub vertical_offset: [f64; 3],
    pub u: [f64; 3],
    pub v: [f64; 3],
    pub w: [f64; 3],
    pub lens_radius: f64,
    pub t0: f64,
    pub t1: f64,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_ray {
    pub origin: [f64; 3],
    pub dir: [f64; 3],
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_hit {
    pub t: f64,
    pub p: [f64; 3],
    pub n: [f64; 3],
    pub ns: [f64; 3],
    pub uv: [f64; 2],
    pub prim_index: u32,
    pub front: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_shadow_query {
    pub p0: [f64; 3],
    pub p1: [f64; 3],
    pub p1_prim_index: u32,
    pub infinite_light: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_integrator_desc {
    pub kind: u32,
    pub max_depth: u32,
    pub invisible_light: u32,
    pub flags: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_film_desc {
    pub width: u32,
    pub height: u32,
    pub sample_begin: u32,
    pub sample_end: u32,
    pub x0: u32,
    pub y0: u32,
    pub x1: u32,
    pub y1: u32,
    pub seed: u64,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_counters {
    pub paths: u64,
    pub extend_rays: u64,
    pub shadow_rays: u64,
    pub probe_rays: u64,
    pub iterations: u64,
    pub render_ms: f64,
    pub extend_ms: f64,
    pub shade_ms: f64,
    pub connect_ms: f64,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt_build_stats {
    pub mode: u32,
    pub max_stack: u32,
    pub nodes: u64,
    pub build_ms: f64,
    pub commit_ms: f64,
}

#[repr(C)]
pub struct rt_scene {
    _private: [u8; 0],
}
#[repr(C)]
pub struct rt_film {
    _private: [u8; 0],
}
#[repr(C)]
pub struct rt_comm {
    _private: [u8; 0],
}

extern "C" {
    // ---- library
    pub fn rt_last_error_string() -> *const c_char;
    pub fn rt_version_string() -> *const c_char;
    pub fn rt_device_count() -> i32;
    pub fn rt_set_device(device: i32) -> RtStatus;
    pub fn rt_round_spp(spp: u32) -> u32;
    pub fn rt_camera_look_at(from: *const f64, to: *const f64, up: *const f64, aspect_ratio: f64, vfov: f64, aperture: f64, focus_dist: f64,
                             t0: f64, t1: f64, out: *mut rt_camera) -> RtStatus;
    // ---- scene build  ⇔  get_objects_mut() pushes + `objects.node = BvhNode::new(..)`
    pub fn rt_scene_create(out: *mut *mut rt_scene) -> RtStatus;
    pub fn rt_scene_destroy(scene: *mut rt_scene) -> RtStatus;
    pub fn rt_scene_add_mesh(scene: *mut rt_scene, n_points: u64, points_xyz: *const f64, n_normals: u64, normals_xyz: *const f64, n_uvs: u64,
                             uvs: *const f64, n_indices: u64, indices: *const u32, mesh_index_out: *mut u32) -> RtStatus;
    pub fn rt_scene_add_transform(scene: *mut rt_scene, m_row_major: *const f64, index_out: *mut u32) -> RtStatus;
    pub fn rt_scene_add_primitives(scene: *mut rt_scene, n: u64, prims: *const rt_primitive, first_index_out: *mut u32) -> RtStatus;
    pub fn rt_scene_add_mesh_triangles(scene: *mut rt_scene, mesh_index: u32, mat_index: u32, first_index_out: *mut u32) -> RtStatus;
    pub fn rt_scene_set_light_index(scene: *mut rt_scene, prim_index: u32, light_index: u32) -> RtStatus;
    pub fn rt_scene_add_image_hdr(scene: *mut rt_scene, width: u32, height: u32, rgb: *const f32, image_index_out: *mut u32) -> RtStatus;
    pub fn rt_scene_add_image_rgb8(scene: *mut rt_scene, width: u32, height: u32, rgb: *const u8, image_index_out: *mut u32) -> RtStatus;
    pub fn rt_scene_add_perlin(scene: *mut rt_scene, scale: f64, perm: *const u32, vecs: *const f64, perlin_index_out: *mut u32) -> RtStatus;
    pub fn rt_scene_add_disney(scene: *mut rt_scene, params: *const rt_disney, disney_index_out: *mut u32) -> RtStatus;
    pub fn rt_scene_set_textures(scene: *mut rt_scene, n: u64, textures: *const rt_texture) -> RtStatus;
    pub fn rt_scene_set_materials(scene: *mut rt_scene, n: u64, materials: *const rt_material) -> RtStatus;
    pub fn rt_scene_set_lights(scene: *mut rt_scene, n: u64, lights: *const rt_light) -> RtStatus;
    pub fn rt_scene_commit(scene: *mut rt_scene) -> RtStatus;
    pub fn rt_scene_set_build_mode(scene: *mut rt_scene, mode: u32) -> RtStatus;
    pub fn rt_scene_build_stats(scene: *const rt_scene, out: *mut rt_build_stats) -> RtStatus;
    pub fn rt_scene_num_primitives(scene: *const rt_scene, out: *mut u64) -> RtStatus;
    pub fn rt_scene_device_bytes(scene: *const rt_scene, nodes: *mut u64, leaves: *mut u64, tables: *mut u64) -> RtStatus;
    // ---- intersection  ⇔  geometry::get_intersection / BvhNode::intersects / Visibility::unoccluded
    pub fn rt_intersect(scene: *mut rt_scene, n: u64, rays: *const rt_ray, hits: *mut rt_hit) -> RtStatus;
    pub fn rt_intersect_range(scene: *mut rt_scene, n: u64, rays: *const rt_ray, tmin: f64, tmax: f64, hits: *mut rt_hit) -> RtStatus;
    pub fn rt_intersect_device(scene: *mut rt_scene, n: u64, d_rays: *const rt_ray, tmin: f64, tmax: f64, d_hits: *mut rt_hit, cuda_stream: *mut c_void) -> RtStatus;
    pub fn rt_closest_device(scene: *mut rt_scene, n: u64, d_rays: *const rt_ray, tmin: f64, tmax: f64, d_prim: *mut u32, d_t: *mut f64, cuda_stream: *mut c_void) -> RtStatus;
    pub fn rt_occluded(scene: *mut rt_scene, n: u64, queries: *const rt_shadow_query, unoccluded_out: *mut u8) -> RtStatus;
    pub fn rt_occluded_device(scene: *mut rt_scene, n: u64, d_queries: *const rt_shadow_query, d_unoccluded_out: *mut u8, cuda_stream: *mut c_void) -> RtStatus;
    pub fn rt_debug_prefilter_stats(out: *mut u64) -> RtStatus;
    // ---- film + render  ⇔  the `Vec<Vec<(f64,f64,f64,u32)>>` film, render::tile_multithread, GLOBAL_STATE.stop_render
    pub fn rt_film_create(width: u32, height: u32, out: *mut *mut rt_film) -> RtStatus;
    pub fn rt_film_create_external(width: u32, height: u32, device_rgba_f64: *mut c_void, out: *mut *mut rt_film) -> RtStatus;
    pub fn rt_film_destroy(film: *mut rt_film) -> RtStatus;
    pub fn rt_film_clear(film: *mut rt_film) -> RtStatus;
    pub fn rt_film_read(film: *mut rt_film, host_rgba: *mut f64) -> RtStatus;
    pub fn rt_film_resolve_rgb8(film: *mut rt_film, host_rgb: *mut u8) -> RtStatus;
    /// no counterpart in the reference: allocation-free variant, films up to RT_RESOLVE_STATIC_MAX_PIXELS
    pub fn rt_film_resolve_rgb8_static(film: *mut rt_film, host_rgb: *mut u8) -> RtStatus;
    pub fn rt_render(scene: *mut rt_scene, camera: *const rt_camera, integrator: *const rt_integrator_desc, desc: *const rt_film_desc, film: *mut rt_film) -> RtStatus;
    pub fn rt_render_cancel(scene: *mut rt_scene) -> RtStatus;
    pub fn rt_get_counters(scene: *const rt_scene, out: *mut rt_counters) -> RtStatus;
    pub fn rt_set_wavefront_size(scene: *mut rt_scene, paths: u32) -> RtStatus;
    // ---- several GPUs of one box
    pub fn rt_comm_create_local(devices: *const i32, n: u32, out: *mut *mut rt_comm) -> RtStatus;
    pub fn rt_comm_unique_id(id_out: *mut u8) -> RtStatus;
    pub fn rt_comm_create_rank(id: *const u8, world: u32, rank: u32, out: *mut *mut rt_comm) -> RtStatus;
    pub fn rt_comm_destroy(comm: *mut rt_comm) -> RtStatus;
    pub fn rt_comm_size(comm: *const rt_comm, world_out: *mut u32, n_local_out: *mut u32) -> RtStatus;
    pub fn rt_film_allreduce(comm: *mut rt_comm, films: *const *mut rt_film, n_films: u32) -> RtStatus;
    /// sum + resolve in one kernel over NVLink peer loads (one process, no NCCL); the films are not modified
    pub fn rt_film_reduce_resolve_rgb8(films: *const *mut rt_film, n_films: u32, host_rgb: *mut u8) -> RtStatus;
    pub fn rt_scene_replicate(src: *const rt_scene, device: i32, out: *mut *mut rt_scene) -> RtStatus;
    pub fn rt_render_multi(comm: *mut rt_comm, scenes: *const *mut rt_scene, films: *const *mut rt_film, n: u32, camera: *const rt_camera,
                           integrator: *const rt_integrator_desc, desc: *const rt_film_desc) -> RtStatus;
}

/// `rt_last_error_string()` as an owned `String`.
pub fn last_error() -> String {
    unsafe {
        let p = rt_last_error_string();
        if p.is_null() { String::new() } else { std::ffi::CStr::from_ptr(p).to_string_lossy().into_owned() }
    }
}

/// `Result` adapter used by the wrappers in INTEGRATION.md.
pub fn check(status: RtStatus) -> Result<(), (RtStatus, String)> {
    if status == RT_OK { Ok(()) } else { Err((status, last_error())) }
}

#[cfg(test)]
mod tests {
    use super::*;
    use std::mem::size_of;
    #[test]
    fn layouts_match_the_header() {
        // the sizes tests/test_abi.py checks against the ctypes / numpy mirrors
        assert_eq!(size_of::<rt_primitive>(), 72);
        assert_eq!(size_of::<rt_texture>(), 48);
        assert_eq!(size_of::<rt_material>(), 56);
        assert_eq!(size_of::<rt_light>(), 72);
        assert_eq!(size_of::<rt_ray>(), 48);
        assert_eq!(size_of::<rt_hit>(), 104);
        assert_eq!(size_of::<rt_shadow_query>(), 56);
        assert_eq!(size_of::<rt_camera>(), 192);
        assert_eq!(size_of::<rt_film_desc>(), 40);
        assert_eq!(size_of::<rt_integrator_desc>(), 16);
        assert_eq!(size_of::<rt_counters>(), 72);
        assert_eq!(size_of::<rt_disney>(), 64);
    }
}
