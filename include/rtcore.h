/*
 * rtcore.h — C ABI of the B200-native path-tracing core (librtcore.so).
 *
 * Drop-in boundary for the one data-parallel hot path of calvin-godfrey/RustRaytracer
 * (SURVEY.md §8b).  The reference has no FFI of its own; these entry points sit directly
 * behind the Rust items its render driver and scene code call, and are what a Rust
 * `extern "C"` shim would bind.  Each entry point cites the reference
 * item it replaces as `file:line` relative to the reference's `src/`.
 *
 * Conventions
 *   - plain pointers and sizes only; all host arrays are borrowed for the duration of the call;
 *   - every function returns an rt_status (0 = OK); the message of the last failure on the
 *     calling thread is available from rt_last_error_string(); nothing unwinds across the ABI;
 *   - all arithmetic inputs are f64, as in the reference; indices are uint32_t with
 *     RT_NONE (0xFFFFFFFF) standing for the reference's `usize::MAX` sentinel;
 *   - there is NO CPU fallback: every compute entry point fails with RT_ERR_CUDA when no
 *     CUDA device is usable.
 */
#ifndef RTCORE_H
#define RTCORE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_NONE 0xFFFFFFFFu

typedef int32_t rt_status;
enum {
    RT_OK = 0,
    RT_ERR_INVALID = 1,     /* bad argument / inconsistent scene tables            */
    RT_ERR_CUDA = 2,        /* CUDA runtime error or no device                      */
    RT_ERR_STATE = 3,       /* call order (e.g. render before commit)               */
    RT_ERR_CANCELLED = 4,   /* rt_render stopped by rt_render_cancel                */
    RT_ERR_IO = 5           /* file could not be read / parsed (host scene layer)   */
};

/* ---------------------------------------------------------------- scene tables
 * Mirror of `struct Objects { meshes, objs, lights, materials, textures, node }`
 * (geometry.rs:13-21).  Cross indices keep the reference's meaning.               */

/* `enum Primitive` (primitive.rs:10-61).  FlipFace{obj} is the `flip_face` flag on the
 * wrapped primitive (it only toggles `front` and negates the sampled normal,
 * primitive.rs:300-310,533-537).                                                     */
enum { RT_PRIM_SPHERE = 0, RT_PRIM_TRIANGLE = 1, RT_PRIM_XY_RECT = 2, RT_PRIM_XZ_RECT = 3, RT_PRIM_YZ_RECT = 4 };

typedef struct rt_primitive {
    uint32_t type;         /* RT_PRIM_*                                                     */
    uint32_t flip_face;    /* 1 = wrapped in Primitive::FlipFace                            */
    uint32_t mat_index;    /* index into materials                                          */
    uint32_t light_index;  /* index into lights, RT_NONE = not emissive (usize::MAX)        */
    uint32_t mesh_index;   /* triangle: mesh                                                */
    uint32_t ind;          /* triangle: offset of its first index in mesh.ind (multiple of 3) */
    uint32_t transform;    /* rect: index into transforms, RT_NONE = None                   */
    uint32_t _pad;
    double   v[5];         /* sphere: cx,cy,cz,r,-   rect: a0,b0,a1,b1,k  (xy: x0,y0,x1,y1,k;
                              xz: x0,z0,x1,z1,k; yz: y0,z0,y1,z1,k)                        */
} rt_primitive;

/* `enum Texture` (material.rs:520-539).  Perlin{perlin} (perlin.rs): 0.5 * (1 + sin(scale * p.z + 10 * turb(p, 7))) in every channel,
 * a function of the hit POINT; its permutation tables and gradient vectors are added with rt_scene_add_perlin (the reference draws
 * them from its global RNG in Perlin::new; here the host passes them, so a render is reproducible).  A scene that holds a Perlin
 * texture is shaded in f64 even when RT_SHADE_F32 is asked for.
 * Hdr{meta,data} and Image{img} keep their texels in an image added with rt_scene_add_image_*:
 * Hdr looks up the nearest texel at (round((1-u)*W) % W, round(v*H) % H), re-quantises it to RGBE8
 * and returns (c+0.5)*2^(e-128)/256 (material.rs:570-587); Image looks up (round(u*W) % W,
 * round((1-v)*H) % H) and returns c/255 (material.rs:601-611).                          */
enum { RT_TEX_SOLID = 0, RT_TEX_CHECKERED = 1, RT_TEX_HDR = 2, RT_TEX_IMAGE = 3, RT_TEX_PERLIN = 4 };
typedef struct rt_texture {
    uint32_t type;
    uint32_t odd, even;    /* checkered: texture ids (material.rs:528-532)                  */
    uint32_t image;        /* hdr / image: index returned by rt_scene_add_image_* ; perlin: by rt_scene_add_perlin */
    double   color[3];     /* solid colour                                                  */
    double   frequency;    /* checkered                                                     */
} rt_texture;

/* `enum Material` (material.rs:17-73).  Disney (material.rs:55-73, 245-394) carries 14 texture ids + `thin`: they live in an rt_disney record
 * added with rt_scene_add_disney, and the material's tex[0] is that record's index.  A scene that holds a Disney material is shaded by the f64
 * kernels even when RT_SHADE_F32 is asked for (up to 8 lobes per BSDF).                                                                          */
enum { RT_MAT_MATTE = 0, RT_MAT_LIGHT = 1, RT_MAT_PLASTIC = 2, RT_MAT_GLASS = 3, RT_MAT_METAL = 4, RT_MAT_MIRROR = 5, RT_MAT_DISNEY = 6 };
typedef struct rt_material {
    uint32_t type;
    uint32_t remap_roughness;
    /* texture ids:  Matte {k_d}            Light {texture}      Plastic {k_d,k_s}
     *               Glass {k_r,k_t}        Mirror {color}
     *               Metal {eta,k,rough,urough,vrough}  (urough/vrough may be RT_NONE)
     *               Disney {index of its rt_disney record}                                  */
    uint32_t tex[5];
    uint32_t _pad;
    /* scalars:      Matte {sigma}   Plastic {roughness}   Glass {u_roughness,v_roughness,index} */
    double   f[3];
} rt_material;

/* Material::Disney's fields (material.rs:55-73) in declaration order, `bump_id` (never read) dropped: texture ids, and the `thin` flag */
typedef struct rt_disney {
    uint32_t color, metallic, eta, roughness, specular_tint, anisotropic, sheen, sheen_tint;
    uint32_t clearcoat, clearcoat_gloss, spec_trans, scatter_distance, flatness, diff_trans;
    uint32_t thin;
    uint32_t _pad;
} rt_disney;

/* `enum Light` (light.rs:30-80): Diffuse (area) lights, the Infinite (environment map) light and
 * the delta lights Point / Spot / Distant (light.rs:522-582; sample_li light.rs:142-175, pdf_li = 0,
 * estimate_direct takes them without MIS and without the BSDF half, integrator.rs:587-601).
 * Point / Spot: `transform` = to_world, the position is to_world x origin; Spot: param = {total_width,
 * falloff} exactly as handed to Light::make_spot_light (the reference takes cos(x * 180 / PI) of each,
 * light.rs:548-549); Distant: param = dir (not normalised by the constructor), world_radius = 1e10.
 * Infinite: Light::make_infinite_light(to_world, n_samples, text_id) (light.rs:608-638) — the
 * (2W x 2H) luminance*sin(theta) table and its Distribution2D (distribution.rs:98-148) are built
 * at commit from the texture; world_radius is the reference's fixed 10000 (light.rs:636).   */
enum { RT_LIGHT_DIFFUSE = 0, RT_LIGHT_INFINITE = 1, RT_LIGHT_POINT = 2, RT_LIGHT_SPOT = 3, RT_LIGHT_DISTANT = 4 };
typedef struct rt_light {
    uint32_t type;
    uint32_t prim_index;   /* diffuse: the emitting primitive (light.rs:62)                 */
    uint32_t two_sided;    /* diffuse                                                       */
    uint32_t text_id;      /* infinite: environment texture (light.rs:74)                   */
    double   color[3];     /* diffuse, point, spot, distant                                 */
    uint32_t transform;    /* infinite, point, spot: to_world as an index into transforms, RT_NONE = identity */
    uint32_t _pad;
    double   param[3];     /* spot: {total_width, falloff, -}; distant: dir                 */
    /* `area` (light.rs:63) is derived from the primitive at commit (primitive.rs:339-359)  */
} rt_light;

/* `struct Camera` (geometry.rs:95-108) */
typedef struct rt_camera {
    double origin[3];
    double upper_left_corner[3];
    double horizontal_offset[3];
    double vertical_offset[3];
    double u[3], v[3], w[3];
    double lens_radius, t0, t1;
} rt_camera;

/* `struct Ray` (geometry.rs:211-236): direction is NOT normalised, t is in units of |dir|. */
typedef struct rt_ray {
    double origin[3];
    double dir[3];
} rt_ray;

/* The fields of `HitRecord` (hittable.rs:51-72) the path integrator consumes.
 * prim_index == RT_NONE encodes `Option::None`.                                     */
typedef struct rt_hit {
    double   t;
    double   p[3];
    double   n[3];          /* geometric normal after set_front (hittable.rs:186-189)        */
    double   ns[3];         /* shading.n                                                     */
    double   uv[2];
    uint32_t prim_index;
    uint32_t front;         /* HitRecord::front (after FlipFace toggling)                    */
} rt_hit;

/* one `Visibility{p0,p1}` query (hittable.rs:15-40) */
typedef struct rt_shadow_query {
    double   p0[3];
    double   p1[3];
    uint32_t p1_prim_index; /* p1.prim_index                                                 */
    uint32_t infinite_light;
} rt_shadow_query;

/* `IntType::Path{max_depth, invisible_light}` (integrator.rs:37-40) */
/* IntType (integrator.rs:25-41).  PATH is the one the reference's main.rs selects.  DIRECT = Direct{strategy: UniformOne},
 * DIRECT_ALL = Direct{strategy: UniformAll}, WHITTED = Whitted: emission + direct light at every vertex, recursion only along
 * perfectly specular lobes (integrator.rs:131-165, 257-297, 464-528).  BASIC = Basic → geometry::cast_ray (geometry.rs:242-277):
 * no light sampling, a hit light ends the walk, one BSDF sample per vertex.  All but PATH: f64 radiometry only, max_depth >= 1. */
enum { RT_INTEGRATOR_PATH = 0, RT_INTEGRATOR_DIRECT = 1, RT_INTEGRATOR_DIRECT_ALL = 2, RT_INTEGRATOR_WHITTED = 3, RT_INTEGRATOR_BASIC = 4 };
/* flags: radiometric precision of the shade kernel.  0 = f64 in the reference's operation order (films match the
 * f64 oracle per pixel); RT_SHADE_F32 = same formulas in f32 (production speed; image parity within the stated
 * relative MSE).  Geometry — intersection decisions, hit points, ray origins — is f64-exact in both. */
enum { RT_SHADE_F32 = 1 };
typedef struct rt_integrator_desc {
    uint32_t kind;
    uint32_t max_depth;
    uint32_t invisible_light;
    uint32_t flags;
} rt_integrator_desc;

/* What `GLOBAL_STATE` width/height/spp (main.rs:74-88) and the tile/sample partition of
 * render.rs carry.  spp is used as given: the caller applies next_power_of_two
 * (sampler.rs:637-640) — rt_round_spp() does it.                                      */
typedef struct rt_film_desc {
    uint32_t width, height;
    uint32_t sample_begin, sample_end;   /* samples [begin,end) of every pixel in the window */
    uint32_t x0, y0, x1, y1;             /* pixel window [x0,x1) x [y0,y1) (a tile, or the image) */
    uint64_t seed;                       /* counter-based RNG key                            */
} rt_film_desc;

typedef struct rt_counters {
    uint64_t paths;          /* camera samples generated                                     */
    uint64_t extend_rays;    /* ray type A: get_intersection in li (integrator.rs:388)       */
    uint64_t shadow_rays;    /* ray type B: Visibility::unoccluded (integrator.rs:584)       */
    uint64_t probe_rays;     /* ray type C: MIS BSDF probe (integrator.rs:615)               */
    uint64_t iterations;     /* wavefront iterations                                         */
    double   render_ms;      /* device time of the last rt_render (CUDA events on its stream) */
    double   extend_ms;      /* summed device time of the extend (closest-hit) launches       */
    double   shade_ms;       /* summed device time of the generate + shade launches           */
    double   connect_ms;     /* summed device time of the shadow-connect + probe launches     */
} rt_counters;

typedef struct rt_scene rt_scene;
typedef struct rt_film rt_film;

/* ---------------------------------------------------------------- library */
const char* rt_last_error_string(void);
const char* rt_version_string(void);
/* number of usable CUDA devices (0 ⇒ every compute call fails with RT_ERR_CUDA) */
int32_t     rt_device_count(void);
rt_status   rt_set_device(int32_t device);
uint32_t    rt_round_spp(uint32_t spp);                      /* sampler.rs:637-640 */

/* Camera::new_motion_blur (geometry.rs:133-175) */
rt_status rt_camera_look_at(const double from[3], const double to[3], const double up[3],
                            double aspect_ratio, double vfov, double aperture, double focus_dist,
                            double t0, double t1, rt_camera* out);

/* ---------------------------------------------------------------- scene build
 * Replaces the `get_objects_mut()` pushes of scenes.rs and the final
 * `objects.node = BvhNode::new(..)` (e.g. scenes.rs:368-371).                        */
rt_status rt_scene_create(rt_scene** out);
rt_status rt_scene_destroy(rt_scene* scene);                 /* geometry.rs:42-55 clear_objects */

/* objects.meshes.push(Mesh{ind,p,n,uv}) (hittable.rs:242-250).  normals / uvs may be NULL
 * (n_normals / n_uvs = 0).  Positions are already transformed (parser.rs:29,45).       */
rt_status rt_scene_add_mesh(rt_scene* scene, uint64_t n_points, const double* points_xyz,
                            uint64_t n_normals, const double* normals_xyz,
                            uint64_t n_uvs, const double* uvs,
                            uint64_t n_indices, const uint32_t* indices, uint32_t* mesh_index_out);
/* rect transforms: a row-major 4x4 `Projective3` (primitive.rs:36,46,56) */
rt_status rt_scene_add_transform(rt_scene* scene, const double m_row_major[16], uint32_t* index_out);
/* objects.objs.push / extend */
rt_status rt_scene_add_primitives(rt_scene* scene, uint64_t n, const rt_primitive* prims, uint32_t* first_index_out);
/* Mesh::generate_triangles (hittable.rs:257-288) + objs.extend: one Triangle per face */
rt_status rt_scene_add_mesh_triangles(rt_scene* scene, uint32_t mesh_index, uint32_t mat_index, uint32_t* first_index_out);
/* Primitive::set_light_index (primitive.rs:318-337) */
rt_status rt_scene_set_light_index(rt_scene* scene, uint32_t prim_index, uint32_t light_index);
/* Texture::new_hdr (material.rs:626-642): the decoded `Vec<Rgb<f32>>`, row-major, width*height*3
 * floats.  The RGBE8 re-quantisation of material.rs:578 (image 0.23 `to_rgbe8`) is applied once here. */
rt_status rt_scene_add_image_hdr(rt_scene* scene, uint32_t width, uint32_t height, const float* rgb, uint32_t* image_index_out);
/* Texture::new_texture (material.rs:613-616): an `RgbImage`, row-major, width*height*3 bytes       */
rt_status rt_scene_add_image_rgb8(rt_scene* scene, uint32_t width, uint32_t height, const uint8_t* rgb, uint32_t* image_index_out);
/* Perlin::new(256, scale) (perlin.rs:17-34, material.rs:628-631): perm = perm_x | perm_y | perm_z, 3 x 256 values < 256;
 * vecs = 256 unit gradient vectors (xyz).  Texture.image of an RT_TEX_PERLIN texture is the index returned here. */
rt_status rt_scene_add_perlin(rt_scene* scene, double scale, const uint32_t perm[768], const double vecs[768], uint32_t* perlin_index_out);
/* Material::make_disney (material.rs:480-516): the record an RT_MAT_DISNEY material points at through tex[0] */
rt_status rt_scene_add_disney(rt_scene* scene, const rt_disney* params, uint32_t* disney_index_out);
rt_status rt_scene_set_textures(rt_scene* scene, uint64_t n, const rt_texture* textures);
rt_status rt_scene_set_materials(rt_scene* scene, uint64_t n, const rt_material* materials);
rt_status rt_scene_set_lights(rt_scene* scene, uint64_t n, const rt_light* lights);
/* Builds the acceleration structure, flattens every table to its SoA device layout and
 * uploads it to the current device.  ⇔ BvhNode::new (hittable.rs:637-723).            */
rt_status rt_scene_commit(rt_scene* scene);
/* Which builder rt_scene_commit uses for the acceleration structure that replaces BvhNode::new (hittable.rs:637-752).
 * The hit set does not depend on the tree, so both give bit-identical results; they trade build time for traversal speed.
 *   RT_BUILD_HOST_SAH    (default) binned-SAH build on the host threads, ~0.7 s per million triangles, fastest traversal
 *   RT_BUILD_DEVICE_LBVH Morton-code LBVH built on the GPU in milliseconds (time to first pixel; scenes that change often).
 *                        Scenes of fewer than 8 primitives, or an LBVH deeper than the traversal stack allows, fall back to
 *                        the host build (rt_build_stats.mode tells which one ran).
 * The environment variable RT_BUILD=lbvh|sah overrides the call (benchmark tools).                                        */
enum { RT_BUILD_HOST_SAH = 0, RT_BUILD_DEVICE_LBVH = 1 };
rt_status rt_scene_set_build_mode(rt_scene* scene, uint32_t mode);
typedef struct rt_build_stats {
    uint32_t mode;           /* builder that produced the committed tree                       */
    uint32_t max_stack;      /* worst-case traversal stack entries of the tree (limit 64)       */
    uint64_t nodes;          /* BVH4 nodes                                                      */
    double   build_ms;       /* tree build alone (host wall clock, or CUDA events for the LBVH) */
    double   commit_ms;      /* whole rt_scene_commit: tables, build, upload                    */
} rt_build_stats;
rt_status rt_scene_build_stats(const rt_scene* scene, rt_build_stats* out);
rt_status rt_scene_num_primitives(const rt_scene* scene, uint64_t* out);
/* bytes of device memory held by the committed scene: nodes, leaf slots, tables */
rt_status rt_scene_device_bytes(const rt_scene* scene, uint64_t* nodes, uint64_t* leaves, uint64_t* tables);

/* ---------------------------------------------------------------- intersection
 * geometry::get_intersection(&Ray) -> Option<HitRecord> (geometry.rs:238-240):
 * closest hit in (SMALL, INFINITY) over the whole scene.  Host buffers.               */
rt_status rt_intersect(rt_scene* scene, uint64_t n, const rt_ray* rays, rt_hit* hits);
/* same, caller-chosen [tmin,tmax] ⇔ BvhNode::intersects(ray,tmin,tmax,0) (hittable.rs:591) */
rt_status rt_intersect_range(rt_scene* scene, uint64_t n, const rt_ray* rays, double tmin, double tmax, rt_hit* hits);
/* device-resident variant: `rays`/`hits` are device pointers on the scene's device; the
 * launch is enqueued on `cuda_stream` (a cudaStream_t, NULL = default) and not synchronised. */
rt_status rt_intersect_device(rt_scene* scene, uint64_t n, const rt_ray* d_rays, double tmin, double tmax,
                              rt_hit* d_hits, void* cuda_stream);
/* closest-hit ids only (prim_index, t): 16 bytes per ray out — the ray-cast sweep entry */
rt_status rt_closest_device(rt_scene* scene, uint64_t n, const rt_ray* d_rays, double tmin, double tmax,
                            uint32_t* d_prim, double* d_t, void* cuda_stream);
/* Visibility::unoccluded(infinite_light) -> bool (hittable.rs:25-39).  Host buffers.   */
rt_status rt_occluded(rt_scene* scene, uint64_t n, const rt_shadow_query* queries, uint8_t* unoccluded_out);
rt_status rt_occluded_device(rt_scene* scene, uint64_t n, const rt_shadow_query* d_queries,
                             uint8_t* d_unoccluded_out, void* cuda_stream);

/* Audit of the conservative f32 triangle prefilter that runs in front of the exact f64 test (Mesh::intersects_triangle,
 * hittable.rs:292-362).  Only a library built with -DRT_PREFILTER_CHECK=1 counts: out = {triangle slot tests, skipped by the
 * prefilter, exact hits, VIOLATIONS (a skipped slot whose exact hit the traversal would have used; must stay 0)}; zeros otherwise. */
rt_status rt_debug_prefilter_stats(uint64_t out[4]);

/* ---------------------------------------------------------------- render
 * Film = the reference's `Vec<Vec<(f64,f64,f64,u32)>>` running sums (render.rs:18-22),
 * flattened row-major as 4 doubles (r,g,b,n) per pixel, pixel (0,0) = upper-left.      */
rt_status rt_film_create(uint32_t width, uint32_t height, rt_film** out);
/* film over caller-owned DEVICE memory of width*height*4 doubles (e.g. a torch tensor, so
 * that the multi-GPU sum can be an NCCL all-reduce on that tensor)                     */
rt_status rt_film_create_external(uint32_t width, uint32_t height, void* device_rgba_f64, rt_film** out);
rt_status rt_film_destroy(rt_film* film);
rt_status rt_film_clear(rt_film* film);
/* util::thread_safe_* read side (util.rs:300-371): copy sums to host                   */
rt_status rt_film_read(rt_film* film, double* host_rgba);
/* point_to_color (util.rs:441-471) on the device: (sum/n) -> ACES approx -> gamma -> u8 */
rt_status rt_film_resolve_rgb8(rt_film* film, uint8_t* host_rgb);
/* rt_film_resolve_rgb8 for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * NO COUNTERPART IN THE REFERENCE.  Same result as rt_film_resolve_rgb8, byte for byte, but the
 * 3-bytes-per-pixel device staging buffer is a static device array inside the library instead of a cudaMalloc per call: the call allocates
 * and frees nothing.  Limit: width * height <= RT_RESOLVE_STATIC_MAX_PIXELS (2048 x 2048; 12 MiB held by every GPU that loads the library until
 * exit); a larger film is refused with RT_ERR_INVALID.  All calls on a device share that array (it keeps the last call's bytes; every byte that
 * is read back is rewritten first): calls from host threads of one process are serialised, any other use that overlaps on the GPU races. */
#define RT_RESOLVE_STATIC_MAX_PIXELS (2048u * 2048u)
rt_status rt_film_resolve_rgb8_static(rt_film* film, uint8_t* host_rgb);

/* render::tile_multithread / Integrator::render / single_sample
 * (render.rs:13, integrator.rs:336-369): adds `sample_end-sample_begin` samples for
 * every pixel of the window to `film`.  Blocking.                                     */
rt_status rt_render(rt_scene* scene, const rt_camera* camera, const rt_integrator_desc* integrator,
                    const rt_film_desc* desc, rt_film* film);
/* GLOBAL_STATE.stop_render (main.rs:75,317-318; render.rs:93) */
rt_status rt_render_cancel(rt_scene* scene);
rt_status rt_get_counters(const rt_scene* scene, rt_counters* out);
/* size of the wavefront (paths in flight, ~480 bytes of device queue state each); default 1<<24 */
rt_status rt_set_wavefront_size(rt_scene* scene, uint32_t paths);

/* ---------------------------------------------------------------- several GPUs of one box
 * Replaces render::tile_multithread's NUM_THREADS = 6 workers + Mutex-guarded film merge (render.rs:13-159, consts.rs:8): the scene is replicated
 * on every GPU, the sample range is split across them (the RNG is keyed by the global sample index, so the partition does not change the image)
 * and the (r,g,b,n) films are summed in place with one ncclAllReduce(f64, sum) over NVLink.  NCCL is bound at run time (dlopen libnccl.so.2);
 * without it these calls fail with RT_ERR_STATE.                                                                                                */
typedef struct rt_comm rt_comm;
/* one process, one host thread per GPU: ncclCommInitAll over `devices` */
rt_status rt_comm_create_local(const int32_t* devices, uint32_t n, rt_comm** out);
/* one process per GPU (torchrun / MPI): rank 0 creates the id and hands it to every rank out of band; the communicator is formed on the
 * calling thread's current device (rt_set_device)                                                                                              */
rt_status rt_comm_unique_id(uint8_t id_out[128]);
rt_status rt_comm_create_rank(const uint8_t id[128], uint32_t world, uint32_t rank, rt_comm** out);
rt_status rt_comm_destroy(rt_comm* comm);
rt_status rt_comm_size(const rt_comm* comm, uint32_t* world_out, uint32_t* n_local_out);
/* in-place sum over all ranks; films[i] lives on the communicator's i-th local device (n_films = 1 for a rank communicator).  Blocking. */
rt_status rt_film_allreduce(rt_comm* comm, rt_film* const* films, uint32_t n_films);
/* Sum AND resolve in one kernel, for the GPUs of ONE process: films[0]'s GPU reads the other accumulators in place over NVLink (peer access,
 * enabled on first use), adds them in the order given — the same bits whatever a ring or a tree would have done — applies point_to_color
 * (util.rs:441-471) and returns width*height*3 bytes.  No NCCL, no communicator, no reduced film in memory; the films are not modified.  This is
 * what follows a multi-GPU render when only the picture is wanted (render.rs:117-153 merges the workers' tiles, then saves it); rt_film_allreduce
 * stays for callers that go on accumulating.  At most 16 films; RT_ERR_STATE when the GPUs have no peer access to each other.            */
rt_status rt_film_reduce_resolve_rgb8(rt_film* const* films, uint32_t n_films, uint8_t* host_rgb);
/* a new scene with src's tables, committed on `device` (build mode and wavefront size carry over) */
rt_status rt_scene_replicate(const rt_scene* src, int32_t device, rt_scene** out);
/* rt_render on the n GPUs of a LOCAL communicator: scenes[g] / films[g] on its g-th device; GPU g adds samples
 * [begin + g S / n, begin + (g+1) S / n) of desc's range to films[g], then the films are all-reduced (clear them first).  Blocking.             */
rt_status rt_render_multi(rt_comm* comm, rt_scene* const* scenes, rt_film* const* films, uint32_t n, const rt_camera* camera,
                          const rt_integrator_desc* integrator, const rt_film_desc* desc);

#ifdef __cplusplus
}
#endif
#endif /* RTCORE_H */
