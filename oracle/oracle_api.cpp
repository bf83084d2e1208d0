// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.h).  PARITY UNPINNED.
// C interface (ctypes) over the f64 restatement.  Scene tables use the POD structs of
// include/rtcore.h so that tests feed the oracle and the CUDA library the same inputs.
#include "orc_shade.h"
#include <atomic>
#include <chrono>
#include <thread>

using namespace orc;

template <class F> static void parallel_for(uint64_t n, int threads, F f) {
    if (threads <= 1 || n < 64) { for (uint64_t i = 0; i < n; i++) f(i, 0); return; }
    std::atomic<uint64_t> next(0);
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++) pool.emplace_back([&, t]() {
        for (;;) { uint64_t b = next.fetch_add(256); if (b >= n) break; uint64_t e = std::min(n, b + 256); for (uint64_t i = b; i < e; i++) f(i, t); }
    });
    for (auto& th : pool) th.join();
}

extern "C" {

void* orc_scene_create() { return new Scene(); }
void orc_scene_destroy(void* s) { delete (Scene*)s; }

uint32_t orc_scene_add_mesh(void* sp, uint64_t n_points, const double* p, uint64_t n_normals, const double* n,
                            uint64_t n_uvs, const double* uv, uint64_t n_ind, const uint32_t* ind) {
    Scene& s = *(Scene*)sp;
    Mesh m;
    m.p.resize(n_points); for (uint64_t i = 0; i < n_points; i++) m.p[i] = V3(p[3 * i], p[3 * i + 1], p[3 * i + 2]);
    m.n.resize(n_normals); for (uint64_t i = 0; i < n_normals; i++) m.n[i] = V3(n[3 * i], n[3 * i + 1], n[3 * i + 2]);
    m.uv.resize(n_uvs); for (uint64_t i = 0; i < n_uvs; i++) m.uv[i] = V2(uv[2 * i], uv[2 * i + 1]);
    m.ind.assign(ind, ind + n_ind);
    s.meshes.push_back(std::move(m));
    return (uint32_t)s.meshes.size() - 1;
}
uint32_t orc_scene_add_transform(void* sp, const double m[16]) {
    Scene& s = *(Scene*)sp; Xform x; memcpy(x.fwd.m, m, sizeof(double) * 16);
    if (!inverse(x.fwd, x.inv)) x.inv = M4::identity();
    s.xforms.push_back(x); return (uint32_t)s.xforms.size() - 1;
}
void orc_scene_add_primitives(void* sp, uint64_t n, const rt_primitive* prims) { Scene& s = *(Scene*)sp; s.prims.insert(s.prims.end(), prims, prims + n); }
void orc_scene_set_textures(void* sp, uint64_t n, const rt_texture* t) { ((Scene*)sp)->textures.assign(t, t + n); }
uint32_t orc_scene_add_image_hdr(void* sp, uint32_t w, uint32_t h, const float* rgb) {
    Scene& s = *(Scene*)sp; Image im; im.w = w; im.h = h; im.hdr = true; im.f.assign(rgb, rgb + (size_t)w * h * 3); s.images.push_back(std::move(im)); return (uint32_t)s.images.size() - 1;
}
uint32_t orc_scene_add_image_rgb8(void* sp, uint32_t w, uint32_t h, const uint8_t* rgb) {
    Scene& s = *(Scene*)sp; Image im; im.w = w; im.h = h; im.hdr = false; im.b.assign(rgb, rgb + (size_t)w * h * 3); s.images.push_back(std::move(im)); return (uint32_t)s.images.size() - 1;
}
void orc_scene_set_materials(void* sp, uint64_t n, const rt_material* m) { ((Scene*)sp)->materials.assign(m, m + n); }
uint32_t orc_scene_add_disney(void* sp, const rt_disney* d) { Scene& s = *(Scene*)sp; s.disneys.push_back(*d); return (uint32_t)s.disneys.size() - 1; }
void orc_scene_set_lights(void* sp, uint64_t n, const rt_light* l) {
    Scene& s = *(Scene*)sp; s.lights.clear();
    for (uint64_t i = 0; i < n; i++) { Light L; L.d = l[i]; L.area = 0; s.lights.push_back(L); }
}
// boxes and light tables, no tree (follow with orc_scene_build_fast_tree): the million-triangle soups, where the faithful builder takes minutes
void orc_scene_commit_tables(void* sp) {
    Scene& s = *(Scene*)sp;
    s.boxes.resize(s.prims.size());
    for (size_t i = 0; i < s.prims.size(); i++) s.boxes[i] = prim_box(s, s.prims[i]);
    for (auto& L : s.lights) {
        if (L.d.type == RT_LIGHT_INFINITE) make_infinite_light(s, L);           // light.rs:608-638
        else if (is_delta_light(L)) make_delta_light(s, L);                    // light.rs:522-582
        else L.area = prim_area(s, s.prims[L.d.prim_index]);                    // light.rs:603
    }
    // Light::eq (derived PartialEq, light.rs:28): same variant and every field equal.  The fields this ABI carries: type, prim_index, two_sided, text_id,
    // colour, transform, param (area / tables are functions of those).  A NaN field makes a light unequal to everything, itself included — kept.
    for (size_t i = 0; i < s.lights.size(); i++) {
        const rt_light& a = s.lights[i].d;
        auto same = [](const rt_light& x, const rt_light& y) {
            if (x.type != y.type || x.prim_index != y.prim_index || x.two_sided != y.two_sided || x.text_id != y.text_id || x.transform != y.transform) return false;
            for (int k = 0; k < 3; k++) if (!(x.color[k] == y.color[k]) || !(x.param[k] == y.param[k])) return false;
            return true;
        };
        s.lights[i].eq_class = same(a, a) ? (uint32_t)i : RT_NONE - 1 - (uint32_t)i;     // NaN: a class of its own that matches nothing (see below)
        for (size_t j = 0; j < i; j++) if (same(s.lights[j].d, a)) { s.lights[i].eq_class = s.lights[j].eq_class; break; }
    }
    s.nodes.clear(); s.root = -1;
}
// boxes + the faithful tree (hittable.rs:637-723) with a seeded axis choice
void orc_scene_commit(void* sp, uint64_t bvh_seed) {
    Scene& s = *(Scene*)sp;
    orc_scene_commit_tables(sp);
    if (!s.prims.empty()) {
        s.nodes.reserve(2 * s.prims.size());
        std::vector<uint32_t> idx(s.prims.size());
        for (size_t i = 0; i < idx.size(); i++) idx[i] = (uint32_t)i;
        BuildRng rng{bvh_seed * 2 + 1};
        s.root = bvh_build(s, idx, 0, idx.size(), rng);
    }
}
// The quick median-split tree for the hit-set traversal (orc_intersect mode 3) on large soups; replaces the faithful tree of orc_scene_commit
// (modes 0 and 2 then walk this tree: still a valid tree of the reference's shape, just not the seeded random-axis one).
void orc_scene_build_fast_tree(void* sp, int threads) { bvh_build_fast(*(Scene*)sp, threads < 1 ? 1 : threads); }
uint64_t orc_scene_num_primitives(void* sp) { return ((Scene*)sp)->prims.size(); }
void orc_scene_prim_box(void* sp, uint32_t i, double out[6]) {
    const Box& b = ((Scene*)sp)->boxes[i];
    out[0] = b.mn.x; out[1] = b.mn.y; out[2] = b.mn.z; out[3] = b.mx.x; out[4] = b.mx.y; out[5] = b.mx.z;
}

static void fill_hit(rt_hit& h, bool hit, const HitRecord& r) {
    memset(&h, 0, sizeof(h));
    if (!hit) { h.prim_index = RT_NONE; return; }
    h.t = r.t; h.p[0] = r.p.x; h.p[1] = r.p.y; h.p[2] = r.p.z;
    h.n[0] = r.n.x; h.n[1] = r.n.y; h.n[2] = r.n.z;
    h.ns[0] = r.shading.n.x; h.ns[1] = r.shading.n.y; h.ns[2] = r.shading.n.z;
    h.uv[0] = r.uv.x; h.uv[1] = r.uv.y; h.prim_index = r.prim_index; h.front = r.front ? 1 : 0;
}

// mode 0: faithful tree (both children, no shrinking); 1: brute force over all prims (+ tie set);
// 2: ordered traversal (SURVEY.md §8d work counting); 3: hit set + tie set over the tree (same answer as 1, for large scenes).  stats[0..1] += box tests, prim tests.
// ties_out (mode 1, may be NULL): per ray up to 4 prim ids sharing t_min (RT_NONE padded).
void orc_intersect(void* sp, uint64_t n, const rt_ray* rays, double tmin, double tmax, int mode, int threads,
                   rt_hit* hits, uint32_t* ties_out, uint64_t* stats) {
    const Scene& s = *(Scene*)sp;
    std::vector<TravStats> st(std::max(threads, 1));
    parallel_for(n, threads, [&](uint64_t i, int t) {
        Ray ray{V3(rays[i].origin[0], rays[i].origin[1], rays[i].origin[2]), V3(rays[i].dir[0], rays[i].dir[1], rays[i].dir[2]), 0.};
        HitRecord rec; bool hit;
        if (mode == 1) {
            std::vector<uint32_t> ties;
            hit = closest_brute(s, ray, tmin, tmax, rec, &ties);
            if (ties_out) for (int k = 0; k < 4; k++) ties_out[4 * i + k] = (hit && k < (int)ties.size()) ? ties[k] : RT_NONE;
        } else if (mode == 3) {
            std::vector<uint32_t> ties;
            hit = closest_hitset(s, ray, tmin, tmax, rec, &ties);
            if (ties_out) for (int k = 0; k < 4; k++) ties_out[4 * i + k] = (hit && k < (int)ties.size()) ? ties[k] : RT_NONE;
        } else if (mode == 2) {
            closest_ordered(s, ray, tmin, tmax, hit, rec, st[t]);
        } else {
            hit = intersect_range(s, ray, tmin, tmax, rec, &st[t]);
        }
        fill_hit(hits[i], hit, rec);
    });
    if (stats) for (auto& x : st) { stats[0] += x.box_tests; stats[1] += x.prim_tests; }
}

void orc_occluded(void* sp, uint64_t n, const rt_shadow_query* q, int threads, uint8_t* out) {
    const Scene& s = *(Scene*)sp;
    parallel_for(n, threads, [&](uint64_t i, int) {
        out[i] = unoccluded(s, V3(q[i].p0[0], q[i].p0[1], q[i].p0[2]), V3(q[i].p1[0], q[i].p1[1], q[i].p1[2]), q[i].p1_prim_index, 0., q[i].infinite_light != 0) ? 1 : 0;
    });
}

void orc_camera_look_at(const double from[3], const double to[3], const double up[3], double aspect, double vfov,
                        double aperture, double focus, double t0, double t1, rt_camera* out) {
    *out = camera_new(V3(from[0], from[1], from[2]), V3(to[0], to[1], to[2]), V3(up[0], up[1], up[2]), aspect, vfov, aperture, focus, t0, t1);
}

// render::tile_multithread (render.rs:13-159) semantics: worker threads pull 16x16 tiles, every pixel gets
// samples [sample_begin, sample_end).  film: width*height*4 doubles (r,g,b,n) running sums.
// counts[0..2] += extend, shadow, probe rays.  Returns wall seconds.
// pixel_stride > 1 renders only pixels whose index py*width+px is a multiple of it (bounded CPU-baseline samples).
// counts[3..4] += box / primitive tests of the extend rays under the ordered traversal when count_ordered != 0.
static double render_impl(void* sp, const rt_camera* cam, uint32_t kind, uint32_t max_depth, uint32_t invisible_light, const rt_film_desc* d,
                          int threads, double* film, uint64_t* counts, uint32_t pixel_stride, int count_ordered) {
    const Scene& s = *(Scene*)sp;
    const uint32_t TILE = 16;   // consts.rs:10
    uint32_t tx0 = d->x0 / TILE, ty0 = d->y0 / TILE, tx1 = (d->x1 + TILE - 1) / TILE, ty1 = (d->y1 + TILE - 1) / TILE;
    uint64_t ntiles = (uint64_t)(tx1 - tx0) * (ty1 - ty0);
    std::atomic<uint64_t> next(0);
    std::vector<RayCounts> rcs(std::max(threads, 1));
    for (auto& r : rcs) r.count_ordered = count_ordered != 0;
    if (pixel_stride == 0) pixel_stride = 1;
    auto t_start = std::chrono::steady_clock::now();
    auto worker = [&](int t) {
        for (;;) {
            uint64_t tile = next.fetch_add(1); if (tile >= ntiles) break;
            uint32_t tx = tx0 + (uint32_t)(tile % (tx1 - tx0)), ty = ty0 + (uint32_t)(tile / (tx1 - tx0));
            for (uint32_t py = std::max(ty * TILE, d->y0); py < std::min((ty + 1) * TILE, d->y1); py++)
                for (uint32_t px = std::max(tx * TILE, d->x0); px < std::min((tx + 1) * TILE, d->x1); px++) {
                    if (((uint64_t)py * d->width + px) % pixel_stride != 0) continue;
                    double* cell = film + 4 * ((uint64_t)py * d->width + px);
                    for (uint32_t smp = d->sample_begin; smp < d->sample_end; smp++) {
                        V3 c = single_sample(s, *cam, d->width, d->height, px, py, smp, d->seed, max_depth, invisible_light != 0, rcs[t], kind);
                        cell[0] += c.x; cell[1] += c.y; cell[2] += c.z; cell[3] += 1.;   // util::increment_color (util.rs:208-232)
                    }
                }
        }
    };
    if (threads <= 1) worker(0);
    else { std::vector<std::thread> pool; for (int t = 0; t < threads; t++) pool.emplace_back(worker, t); for (auto& th : pool) th.join(); }
    double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count();
    if (counts) for (auto& r : rcs) {
        counts[0] += r.extend; counts[1] += r.shadow; counts[2] += r.probe;
        if (count_ordered) { counts[3] += r.ordered.box_tests; counts[4] += r.ordered.prim_tests; }
    }
    return secs;
}
double orc_render_ex(void* sp, const rt_camera* cam, uint32_t max_depth, uint32_t invisible_light, const rt_film_desc* d,
                     int threads, double* film, uint64_t* counts, uint32_t pixel_stride, int count_ordered) {
    return render_impl(sp, cam, 0, max_depth, invisible_light, d, threads, film, counts, pixel_stride, count_ordered);
}
// kind = rtcore.h RT_INTEGRATOR_*: 0 Path, 1 Direct{UniformOne}, 2 Direct{UniformAll}, 3 Whitted, 4 Basic (max_depth >= 1 for 1..4)
double orc_render_kind(void* sp, const rt_camera* cam, uint32_t kind, uint32_t max_depth, uint32_t invisible_light, const rt_film_desc* d,
                       int threads, double* film, uint64_t* counts) {
    return render_impl(sp, cam, kind, max_depth, invisible_light, d, threads, film, counts, 1, 0);
}
double orc_render(void* sp, const rt_camera* cam, uint32_t max_depth, uint32_t invisible_light, const rt_film_desc* d,
                  int threads, double* film, uint64_t* counts) {
    return orc_render_ex(sp, cam, max_depth, invisible_light, d, threads, film, counts, 1, 0);
}

// one camera sample, for per-sample parity checks; out rgb
void orc_single_sample(void* sp, const rt_camera* cam, uint32_t width, uint32_t height, uint32_t px, uint32_t py, uint32_t sample,
                       uint64_t seed, uint32_t max_depth, uint32_t invisible_light, double out[3]) {
    RayCounts rc;
    V3 c = single_sample(*(Scene*)sp, *cam, width, height, px, py, sample, seed, max_depth, invisible_light != 0, rc);
    out[0] = c.x; out[1] = c.y; out[2] = c.z;
}

void orc_resolve_rgb8(const double* film, uint64_t n_pixels, uint8_t* rgb) { for (uint64_t i = 0; i < n_pixels; i++) point_to_color(film + 4 * i, rgb + 3 * i); }

void orc_philox(uint32_t c[4], uint32_t k0, uint32_t k1) { philox4x32_10(c, k0, k1); }
double orc_rng_uniform(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t dim) { Rng r(seed, pixel, sample); r.dim = dim; return r.next(); }

// ---- lobe-level probes for KATs and device-function parity.
// BSDF of material `mat` on a canonical frame: ns = ng = (0,0,1), ss = (1,0,0), ts = (0,1,0).
static HitRecord canonical_record(uint32_t mat, double u, double v) {
    HitRecord r = record_new(V3(0, 0, 0), V2(u, v), V3(0, 0, 1), V3(1, 0, 0), V3(0, 1, 0), 1., 0, mat);
    return r;
}
void orc_bsdf_f_pdf(void* sp, uint32_t mat, double u, double v, const double wo[3], const double wi[3], uint32_t flags, double f_out[3], double* pdf_out) {
    HitRecord r = canonical_record(mat, u, v);
    compute_scattering(*(Scene*)sp, r, RADIANCE, true);
    V3 f = bsdf_f(r.bsdf, V3(wo[0], wo[1], wo[2]), V3(wi[0], wi[1], wi[2]), (uint8_t)flags);
    f_out[0] = f.x; f_out[1] = f.y; f_out[2] = f.z;
    *pdf_out = bsdf_pdf(r.bsdf, V3(wo[0], wo[1], wo[2]), V3(wi[0], wi[1], wi[2]), (uint8_t)flags);
}
// out: f[3], wi[3], pdf, sampled_type; rng keyed (seed, pixel, sample) starting at dim 0
void orc_bsdf_sample_f(void* sp, uint32_t mat, double u, double v, const double wo[3], const double sample[2], uint32_t flags,
                       uint64_t seed, uint32_t pixel, uint32_t smp, double out[8]) {
    HitRecord r = canonical_record(mat, u, v);
    compute_scattering(*(Scene*)sp, r, RADIANCE, true);
    Rng rng(seed, pixel, smp);
    V3 f, wi; double pdf; uint8_t st;
    bsdf_sample_f(r.bsdf, V3(wo[0], wo[1], wo[2]), V2(sample[0], sample[1]), (uint8_t)flags, rng, f, wi, pdf, st);
    out[0] = f.x; out[1] = f.y; out[2] = f.z; out[3] = wi.x; out[4] = wi.y; out[5] = wi.z; out[6] = pdf; out[7] = (double)st;
}
// Infinite-light probes: sample_li → out[0..2] wi, [3] pdf, [4..6] colour, [7..9] p1 ; pdf_li ; le
void orc_light_sample_li(void* sp, uint32_t light, const double p[3], const double u[2], double out[10]) {
    const Scene& s = *(Scene*)sp; HitRecord rec = record_basic(V3(p[0], p[1], p[2]), 0.);
    LiSample o = sample_li(s, s.lights[light], rec, V2(u[0], u[1]));
    out[0] = o.wi.x; out[1] = o.wi.y; out[2] = o.wi.z; out[3] = o.pdf; out[4] = o.color.x; out[5] = o.color.y; out[6] = o.color.z; out[7] = o.p1.x; out[8] = o.p1.y; out[9] = o.p1.z;
}
double orc_light_pdf_li(void* sp, uint32_t light, const double p[3], const double wi[3]) {
    const Scene& s = *(Scene*)sp; HitRecord rec = record_basic(V3(p[0], p[1], p[2]), 0.);
    return pdf_li(s, s.lights[light], rec, V3(wi[0], wi[1], wi[2]));
}
void orc_light_le(void* sp, uint32_t light, const double d[3], double out[3]) {
    const Scene& s = *(Scene*)sp; Ray r{V3(0, 0, 0), V3(d[0], d[1], d[2]), 0.};
    V3 c = light_le(s, s.lights[light], r); out[0] = c.x; out[1] = c.y; out[2] = c.z;
}
// the Distribution2D of an Infinite light: sizes → n[0] = nu, n[1] = nv; tables (any may be NULL): func of row 0..nv-1 is not
// exported (it is the slice f[v..v+nu]); cdf rows nv*(nu+1), row integrals nv, marginal cdf nv+1, marginal integral
void orc_light_dist(void* sp, uint32_t light, uint64_t n[2], double* cdf_rows, double* row_int, double* marg_cdf, double* marg_int) {
    const Light& l = ((Scene*)sp)->lights[light];
    size_t nv = l.dist.p_cond_v.size(), nu = nv ? l.dist.p_cond_v[0].count() : 0;
    n[0] = nu; n[1] = nv;
    for (size_t v = 0; v < nv; v++) {
        if (cdf_rows) memcpy(cdf_rows + v * (nu + 1), l.dist.p_cond_v[v].cdf.data(), (nu + 1) * sizeof(double));
        if (row_int) row_int[v] = l.dist.p_cond_v[v].func_int;
    }
    if (marg_cdf) memcpy(marg_cdf, l.dist.p_marginal.cdf.data(), (nv + 1) * sizeof(double));
    if (marg_int) *marg_int = l.dist.p_marginal.func_int;
}
void orc_tex_value(void* sp, uint32_t id, double u, double v, double out[3]) { V3 c = tex_value(*(Scene*)sp, id, u, v); out[0] = c.x; out[1] = c.y; out[2] = c.z; }
void orc_tex_value_point(void* sp, uint32_t id, double u, double v, const double p[3], double out[3]) { V3 c = tex_value(*(Scene*)sp, id, u, v, V3(p[0], p[1], p[2])); out[0] = c.x; out[1] = c.y; out[2] = c.z; }
// Texture::new_perlin (material.rs:628-631): perm = perm_x | perm_y | perm_z (3 x 256), vecs = 256 unit vectors
uint32_t orc_scene_add_perlin(void* sp, double scale, const uint32_t* perm, const double* vecs) {
    Scene& s = *(Scene*)sp; size_t base = s.perlin.size(); s.perlin.resize(base + 1537); s.perlin[base] = scale;
    for (int i = 0; i < 768; i++) { s.perlin[base + 1 + i] = (double)perm[i]; s.perlin[base + 769 + i] = vecs[i]; }
    return (uint32_t)(base / 1537);
}
double orc_perlin_noise(void* sp, uint32_t perlin, const double p[3]) { Scene& s = *(Scene*)sp; return perlin_noise(s.perlin.data() + (size_t)perlin * 1537, V3(p[0], p[1], p[2])); }
double orc_fr_dielectric(double c, double ei, double et) { return fr_dielectric(c, ei, et); }
void orc_fr_conductor(double c, const double eta[3], const double k[3], double out[3]) { V3 r = fr_conductor(c, V3(eta[0], eta[1], eta[2]), V3(k[0], k[1], k[2])); out[0] = r.x; out[1] = r.y; out[2] = r.z; }

}  // extern "C"
