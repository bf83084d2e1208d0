/*
 * rthost.h — host scene layer, librthost.so (C ABI).  Pure host code with no link-time dependency on the CUDA library librtcore.so:
 * only rth_scene_upload talks to it, through the public rt_scene_* calls it looks up at run time.
 *
 * The reference keeps scene construction on the host (scenes.rs, parser.rs; north_star: "the Rust
 * host keeps scene parsing, material setup ..."); no Rust toolchain exists in this image, so the
 * scene builders are restated in C++ behind this small C interface.  They produce plain tables
 * (the POD structs of rtcore.h) that are (a) pushed into an rt_scene through the public rt_scene_*
 * calls — exactly what a Rust host would do over FFI — and (b) exported to the test-side oracle so
 * that both sides see identical inputs.  Nothing here is on the hot path.
 */
#ifndef RTHOST_H
#define RTHOST_H
#include "rtcore.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct rth_scene rth_scene;

typedef struct rth_params {
    uint32_t    width, height;   /* GLOBAL_STATE aspect ratio = width/height (main.rs:84-99)          */
    uint32_t    variant;         /* plastic_dragon: 0 = rough plastic (scenes.rs:310), 1 = rough metal
                                    (commented metal_dragon, scenes.rs:424-471);
                                    cornell_box_statue: 0 = matte statue, 1 = metal (scenes.rs:243-246);
                                    two_dragons: 0 = as committed (metal dragon only), 1 = glass + metal
                                    (commented block scenes.rs:608-611 enabled);
                                    material_hdr: mat_num 0 smooth plastic, 1 rose-gold metal, 2 mirror,
                                    3 glass (scenes.rs:627,712-718), 4 disney_mat (scenes.rs:867-904)                                     */
    uint32_t    _pad;
    uint64_t    n_tris;          /* soup: triangle count; dragon/statue stand-ins: 0 = default          */
    uint64_t    seed;            /* generators                                                          */
    const char* data_dir;        /* directory holding the reference's data/ (may be NULL): real
                                    dragon/statue OBJ files are used when found there                  */
} rth_params;

/* name ∈ { "sphere_roughness" (scenes.rs:474), "make_world" (commented, scenes.rs:15-86, seeded),
 *          "cornell_box" (:89), "cornell_box_statue" (:200), "plastic_dragon" (:310),
 *          "two_dragons" (:549), "material_hdr" (:627, HDR environment = Light::Infinite),
 *          "teapot_hdr" (:744, plastic teapot on a transformed checkered floor rect under the HDR
 *          environment), "soup"  }                                                          */
/* message of the last failed rth_* call on the calling thread */
const char* rth_last_error_string(void);
rt_status   rth_scene_build(const char* name, const rth_params* params, rth_scene** out);
void        rth_scene_destroy(rth_scene* s);
/* the (String path, Camera, Samplers) a scene function returns */
rt_status   rth_scene_camera(const rth_scene* s, rt_camera* out);
const char* rth_scene_png_name(const rth_scene* s);
/* 1 when a procedural stand-in replaced a missing OBJ asset */
int32_t     rth_scene_uses_standin(const rth_scene* s);

/* table export (for the oracle / inspection): counts = meshes, prims, transforms, textures, materials, lights */
rt_status   rth_scene_counts(const rth_scene* s, uint64_t counts[6]);
rt_status   rth_scene_mesh_sizes(const rth_scene* s, uint32_t mesh, uint64_t sizes[4]); /* points, normals, uvs, indices */
rt_status   rth_scene_mesh_copy(const rth_scene* s, uint32_t mesh, double* points, double* normals, double* uvs, uint32_t* indices);
rt_status   rth_scene_copy_primitives(const rth_scene* s, rt_primitive* out);
rt_status   rth_scene_copy_transforms(const rth_scene* s, double* out16_each);
rt_status   rth_scene_copy_textures(const rth_scene* s, rt_texture* out);
rt_status   rth_scene_copy_materials(const rth_scene* s, rt_material* out);
rt_status   rth_scene_copy_lights(const rth_scene* s, rt_light* out);
/* Material::Disney records (rt_disney), referenced by tex[0] of the RT_MAT_DISNEY materials */
rt_status   rth_scene_num_disney(const rth_scene* s, uint64_t* n);
rt_status   rth_scene_copy_disney(const rth_scene* s, rt_disney* out);

/* Hdr / Image texel blocks (Texture::Hdr{data}, Texture::Image{img}); info = width, height, is_hdr;
 * copy writes width*height*3 floats (hdr) or bytes (rgb8)                                             */
rt_status   rth_scene_num_images(const rth_scene* s, uint64_t* n);
rt_status   rth_scene_image_info(const rth_scene* s, uint32_t image, uint32_t info[3]);
rt_status   rth_scene_image_copy(const rth_scene* s, uint32_t image, void* out);
/* Texture::new_hdr's decoder (material.rs:626-642): Radiance RGBE file -> Rgb<f32> pixels.  rgb_out may
 * be NULL to query the size.                                                                          */
rt_status   rth_load_hdr(const char* path, uint32_t* width, uint32_t* height, float* rgb_out, uint64_t capacity_floats);

/* push every table into `dst` through the public rt_scene_* calls (does not commit) */
rt_status   rth_scene_upload(const rth_scene* s, rt_scene* dst);

/* parser::parse_obj (parser.rs:8-87): tobj::load_obj(path, triangulate = true), models[0] only,
 * single-index, f32 values widened to f64, `transform` (row-major 4x4) applied to points and normals.
 * Appends the mesh to `s` and returns its index. */
rt_status   rth_scene_add_obj(rth_scene* s, const char* path, const double transform[16], uint32_t* mesh_index_out);
/* an empty scene to append to */
rt_status   rth_scene_create_empty(rth_scene** out);

/* ray batches of SURVEY.md §8d: primary rays through pixel centres of a width x height image */
rt_status   rth_primary_rays(const rt_camera* cam, uint32_t width, uint32_t height, rt_ray* out);

#ifdef __cplusplus
}
#endif
#endif
