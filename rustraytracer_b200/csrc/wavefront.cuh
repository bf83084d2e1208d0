// Shared device declarations of the wavefront path tracer: path-state records, queues, the control block, and the shade
// kernel template.  Included by render.cu (everything except the f32 shade instantiations) and by shade_f32.cu, which
// instantiates k_shade<float, ...> in a translation unit of its own so that the production-mode radiometry can be compiled
// with fast f32 division / square root (see the Makefile) while every f64 expression keeps the reference's rounding.
#pragma once
#include "scene.h"
#include "shade.cuh"

namespace rt {

// Path state, one entry per path in flight, as two arrays of 48-byte records (three 16-byte words each).
// extend streams `ray` in order (fully coalesced); shade reaches both through the bin index, so each path costs
// 3 + 3 fully-used 32-byte sectors instead of the 13 quarter-used sectors of a field-per-array layout.
struct __align__(16) PRay { double ox, oy, oz, dx, dy, dz; };
struct __align__(16) PAux { double bx, by, bz; uint32_t pixel, sample, dim, flags; };   // flags: bits 0..15 bounces, bit 16 specular_bounce
static_assert(sizeof(PRay) == 48 && sizeof(PAux) == 48, "path records must be 48 bytes");
struct PathSet { PRay* ray; PAux* aux; };
__device__ __forceinline__ PRay load_ray(const PRay* p) {
    const double2* q = reinterpret_cast<const double2*>(p);
    double2 a = __ldcs(q), b = __ldcs(q + 1), c = __ldcs(q + 2);     // streaming: path state must not evict the BVH from L2
    PRay r; r.ox = a.x; r.oy = a.y; r.oz = b.x; r.dx = b.y; r.dy = c.x; r.dz = c.y; return r;
}
__device__ __forceinline__ void store_ray(PRay* p, D3 o, D3 d) {
    double2* q = reinterpret_cast<double2*>(p);
    __stcs(q, make_double2(o.x, o.y)); __stcs(q + 1, make_double2(o.z, d.x)); __stcs(q + 2, make_double2(d.y, d.z));
}
__device__ __forceinline__ PAux load_aux(const PAux* p) {      // words: (bx,by) (bz,[pixel,sample]) ([dim,flags],pad)
    const double2* q = reinterpret_cast<const double2*>(p);
    double2 a = __ldcs(q), b = __ldcs(q + 1), c = __ldcs(q + 2);
    PAux r; r.bx = a.x; r.by = a.y; r.bz = b.x;
    r.pixel = (uint32_t)__double2loint(b.y); r.sample = (uint32_t)__double2hiint(b.y);
    r.dim = (uint32_t)__double2loint(c.x); r.flags = (uint32_t)__double2hiint(c.x);
    return r;
}
__device__ __forceinline__ void store_aux(PAux* p, double bx, double by, double bz, uint32_t pixel, uint32_t sample, uint32_t dim, uint32_t flags) {
    double2* q = reinterpret_cast<double2*>(p);
    __stcs(q, make_double2(bx, by));
    __stcs(q + 1, make_double2(bz, __hiloint2double((int)sample, (int)pixel)));
    __stcs(q + 2, make_double2(__hiloint2double((int)flags, (int)dim), 0.));
}
struct ConnQ {             // shadow / probe queue: a ray (a→b or origin+dir), a pre-weighted contribution, a target
    double *ax, *ay, *az, *bx, *by, *bz, *cr, *cg, *cb;
    uint32_t *target, *pixel;                     // shadow: light primitive ; probe: light index
};
struct Ctrl {
    uint32_t n_cur, n_new, n_tot, n_next, n_shadow, n_probe, overflow, pad1;   // overflow: a queue append beyond W was dropped (branching integrators only)
    unsigned long long next_sample, sample_base, total_samples;
    unsigned long long c_paths, c_extend, c_shadow, c_probe, c_iter;
    uint32_t fetch[4];           // work counters of the persistent traversal kernels (extend, shadow, probe)
    uint32_t bin_count[16];      // paths per shading bin after extend (material type x primitive class, 15 = miss)
    uint32_t bin_base[16];       // RT_SORT_HITS: first entry of each bin in the sorted entry array
};
constexpr int N_BINS = 16;
// RT_SORT_HITS: the (path, primitive) entries extend hands to shade are bucket-sorted by (shading bin, Morton cell of the hit point)
// instead of being appended per bin in completion order: a warp of k_shade then shades neighbouring surface points (its triangle
// gathers share sectors), and the continuation and shadow rays it appends — the next extend / connect launches fetch them in queue
// order — start close together (render.cu: ExtendJob::finish, k_sort_scan, k_sort_scatter).
// MEASURED (dragon, 1024^2, 64 spp, one B200): extend 31.6 -> 30.3 ms, connect 20.2 -> 19.1 ms, but shade 30.1 -> 37.5 ms: the two sort
// kernels cost 0.03 + 0.1..0.34 ms per iteration, and the shade kernels themselves run 20-27 % slower although their triangle gathers
// are now coherent (ncu launch list: plastic/triangle kernel -6 % when every launch starts cold) — live they lose the near-sequential order
// of the 2 x 48 B path-state gathers that completion order gives for free.  82.1 -> 87.1 ms per render, 2,628 -> 2,429 Mrays/s on the
// bench: OFF.  What would make it pay: moving the path state along with the entry (another 192 B per path) — not with rays this short.
#ifndef RT_SORT_HITS
#define RT_SORT_HITS 0
#endif
#ifndef RT_SORT_BITS
#define RT_SORT_BITS 5          // Morton bits per axis of the hit cell
#endif
static_assert(RT_SORT_BITS >= 4 && RT_SORT_BITS <= 7, "RT_SORT_BITS out of range");
constexpr uint32_t SORT_CELLS = 1u << (3 * RT_SORT_BITS);
#ifndef SHADE_MIN_BLOCKS
#define SHADE_MIN_BLOCKS 4      // <= 128 registers: 16 warps/SM; measured against 3 and 5 in profiles/shade_r01_notes.md, and again after the register-BSDF change (r02 pass AC,
                                // dragon / Cornell 64 spp shade ms: 3 blocks 22.54 / 44.73, 4 blocks 22.53 / 44.79, 5 blocks 22.46 / 44.18): the kernel does not care between 12 and 20 warps
#endif
#ifndef SHADE_PREFETCH
#define SHADE_PREFETCH 0      // measured r01: no change (40.9 vs 40.8 ms shade at 64 spp) — the kernel does not wait on these gathers
#endif
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
struct RenderParams {
    rt_camera cam; uint32_t width, height, x0, y0, ww, wh, sample_begin; uint32_t max_depth, invisible_light; uint32_t k0, k1; uint32_t W;
    uint32_t kind;               // RT_INTEGRATOR_* ; the f32 kernels are compiled for RT_INTEGRATOR_PATH only
};
// first entry of shading bin `bin` in the entry array filled for k_shade
__device__ __forceinline__ size_t bin_first(const Ctrl* c, const RenderParams& rp, int bin) {
#if RT_SORT_HITS
    return c->bin_base[bin];
#else
    return (size_t)bin * rp.W;
#endif
}

__device__ __forceinline__ uint32_t warp_append(uint32_t* counter) {
    // warp-aggregated queue append: one atomic per warp (ballot + popc), everyone gets a unique slot
    unsigned mask = __activemask();
    int leader = __ffs(mask) - 1, lane = threadIdx.x & 31;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(counter, (uint32_t)__popc(mask));
    base = __shfl_sync(mask, base, leader);
    return base + __popc(mask & ((1u << lane) - 1u));
}
__device__ __forceinline__ void film_add(double* film, uint32_t pixel, D3 c) {
    // util::increment_color (util.rs:208-232): the NaN guards there compare with `== NAN` and never fire
    atomicAdd(film + 4 * (size_t)pixel + 0, c.x); atomicAdd(film + 4 * (size_t)pixel + 1, c.y); atomicAdd(film + 4 * (size_t)pixel + 2, c.z);
}

__device__ __forceinline__ D3 light_l(const DLight& l, D3 n, D3 w) {      // Light::l (light.rs:475-496)
    if (dot(n, w) > 0. || l.two_sided) return ld3(l.color);
    return mk3(0., 0., 0.);
}
template <class R> __device__ __forceinline__ R power_heuristic(R f_pdf, R g_pdf) { R f = R(1) * f_pdf, g = R(1) * g_pdf; return (f * f) / (f * f + g * g); }   // integrator.rs:655-659

// The body of PathIntegrator::li's loop (integrator.rs:386-443) for one path vertex.  R = double reproduces the
// reference's arithmetic; R = float is the production path (geometry stays f64 in both, see shade.cuh).
// One specialisation per shading bin (material type MAT x primitive class CLS): every warp runs one material's code.
// Paths in the "miss" bin need no kernel at all: without an infinite light a miss adds nothing (light.rs:499-513).
// ENVL = the scene has a Light::Infinite: only then are the environment-light arms compiled in (they cost the area-light-only
// scenes ~5 % of shade time when always present: 52.8 -> 55.6 ms on the dragon, r01).
template <class R, int MAT, int CLS, bool ENVL>
__global__ void __launch_bounds__(128, (sizeof(R) == 4 ? SHADE_MIN_BLOCKS : 2)) k_shade(Ctrl* c, const __grid_constant__ DScene sc, const __grid_constant__ RenderParams rp, PathSet in, PathSet out,
                                              const uint2* __restrict__ bin_idx, ConnQ shq, ConnQ prq, double* film) {
    constexpr uint32_t KM = mat_km<MAT>();
    constexpr int BIN = MAT * 2 + CLS;
    // Integrator: Path, or (f64 kernels only) Direct{UniformOne}, Direct{UniformAll}, Whitted — integrator.rs:131-165, 257-297.  These
    // shade EVERY vertex with emission + direct light and continue only along perfectly specular lobes, reflection and transmission
    // as two independent children (specular_reflect / specular_transmit, integrator.rs:464-504).
    const uint32_t kind = sizeof(R) == 8 ? rp.kind : (uint32_t)RT_INTEGRATOR_PATH;
    const bool path = kind == RT_INTEGRATOR_PATH;
    const bool basic = !path && kind == RT_INTEGRATOR_BASIC;       // geometry::cast_ray (geometry.rs:242-277)
    const uint32_t n = c->bin_count[BIN];
    const uint32_t stride = gridDim.x * blockDim.x, k0 = blockIdx.x * blockDim.x + threadIdx.x;
    bin_idx += bin_first(c, rp, BIN);
    uint2 entry = k0 < n ? bin_idx[k0] : make_uint2(0u, 0u);
    for (uint32_t k = k0; k < n; k += stride) {
        const uint32_t i = entry.x;
        const PRay pr = load_ray(in.ray + i);
        const PAux pa = load_aux(in.aux + i);
#if SHADE_PREFETCH
        // the next entry of this thread and its path record: fetched one iteration ahead so that the first two levels of the
        // dependent gather chain (bin entry -> 96 B of path state -> primitive) overlap this iteration's work
        const uint2 entry_next = k + stride < n ? bin_idx[k + stride] : make_uint2(0u, 0u);
#endif
        D3 o = mk3(pr.ox, pr.oy, pr.oz), d = mk3(pr.dx, pr.dy, pr.dz);
        V3<R> beta = v3<R>((R)pa.bx, (R)pa.by, (R)pa.bz);
        uint32_t pixel = pa.pixel, flags = pa.flags;
        Rng rng; rng.k0 = rp.k0; rng.k1 = rp.k1; rng.pixel = pixel; rng.sample = pa.sample; rng.dim = pa.dim; rng.w2 = 0xffffffffu; rng.w3 = 0;
        uint32_t bounces = flags & 0xffffu; bool specular_bounce = (flags >> 16) & 1u;
        uint32_t prim = entry.y;
        bool is_some = prim != RT_NONE;
#if SHADE_PREFETCH
        if (k + stride < n) {
            prefetch_l2(in.ray + entry_next.x); prefetch_l2(reinterpret_cast<const char*>(in.ray + entry_next.x) + 32);
            prefetch_l2(in.aux + entry_next.x); prefetch_l2(reinterpret_cast<const char*>(in.aux + entry_next.x) + 32);
            if (entry_next.y != RT_NONE) prefetch_l2(sc.prims + entry_next.y);
        }
#endif
        Rec rec;
        if (is_some) {
            if constexpr (sizeof(R) == 4 && CLS == 1) is_some = prim_record_tri_f32(sc, prim, o, d, SMALL, INF, rec);   // f32 shading frame (shade.cuh)
            // leaf_box = false: extend has already accepted this primitive with the reference's full predicate (f64 leaf box ∧
            // primitive test, traverse.cuh:slot_test); the record only needs t / barycentrics again, and `front` (the one field
            // FlipFace toggles) is not consumed by the path integrator
            else is_some = prim_record_t<CLS, sizeof(R) == 4>(sc, prim, o, d, SMALL, INF, false, rec);
        }
        bool alive = false;
        do {
            if (!path || bounces == 0 || specular_bounce) {
                if (is_some) {
                    uint32_t li = sc.prims[prim].light;
                    if (li != RT_NONE) {
                        if (path && rp.invisible_light) { o = rec.p; alive = true; break; }          // go straight through the light, same bounce
                        film_add(film, pixel, dbl3(cmul(cast3<R>(light_l(sc.lights[li], rec.n, -d)), beta)));   // HitRecord::le (hittable.rs:215-225)
                        if (basic) break;                                                 // cast_ray returns the emission (geometry.rs:251-253)
                    }
                }   // miss: only Infinite lights emit (light.rs:499-513) — none in scope
            }
            if (!is_some || (path && bounces >= rp.max_depth)) break;
            Bsdf<R> bsdf;
            compute_scattering<R, MAT>(sc, rec, bsdf);
            const V3<R> wo_w = cast3<R>(rec.wo), sn = cast3<R>(rec.sn);
            // uniform_sample_one_light (integrator.rs:530-557)
            uint32_t n_lights = sc.n_lights;
            // Path / Direct{UniformOne}: one light picked uniformly, estimate x n_lights.  Direct{UniformAll} (uniform_sample_all_lights,
            // integrator.rs:506-528): every light once, unscaled.  Whitted (integrator.rs:148-158): every light's sample_li, no MIS, f over ALL lobes.
            uint32_t l_begin = 0, l_end = basic ? 0u : n_lights;
            if (n_lights != 0 && !basic && (path || kind == RT_INTEGRATOR_DIRECT)) {
                l_begin = min(n_lights - 1, (uint32_t)(unsigned long long)(rng_next(rng) * (double)n_lights)); l_end = l_begin + 1;
            }
            const R l_scale = (path || kind == RT_INTEGRATOR_DIRECT) ? (R)n_lights : R(1);
            const bool whitted = !path && kind == RT_INTEGRATOR_WHITTED;
            // one body, called once (f32 kernels: Path only — the code the compiler sees is the plain `if (n_lights != 0) { ... }`) or per light
            auto light_estimate = [&](const uint32_t light_num) {
                D2 u_light; u_light.x = rng_next(rng); u_light.y = rng_next(rng);
                V2<R> u_scatter; u_scatter.x = R(0); u_scatter.y = R(0);
                if (!whitted) {
                    if constexpr (mat_ignores_sample2d<MAT>()) rng_skip(rng, 2);       // never read (shade.cuh); the values stay 0
                    else { u_scatter.x = rng_real<R>(rng); u_scatter.y = rng_real<R>(rng); }
                }
                const DLight& light = sc.lights[light_num];
                const uint32_t nsf = whitted ? (uint32_t)BSDF_ALL : (uint32_t)(BSDF_ALL - BSDF_SPECULAR);
                // estimate_direct, light half (integrator.rs:573-597); Light::sample_li Diffuse (light.rs:176-203), Primitive::sample
                // (primitive.rs:445-456) / Infinite (light.rs:204-245)
                const bool inf_light = ENVL && light.type == RT_LIGHT_INFINITE;
                // Point / Spot / Distant (light.rs:142-175): f64 kernels only (rt_render shades a scene that holds one in f64); no MIS weight
                // and no BSDF half (integrator.rs:587-601)
                bool delta = false;
                if constexpr (sizeof(R) == 8) delta = sc.has_delta != 0 && light.type >= RT_LIGHT_POINT;
                {
                    D3 p1; double light_pdf; D3 color, wid = mk3(0., 0., 0.);
                    if (inf_light) infinite_sample_li(sc, light, rec.p, u_light, wid, light_pdf, color, p1);
                    else if (delta) { if constexpr (sizeof(R) == 8) delta_sample_li(sc, light, rec.p, wid, light_pdf, color, p1); }
                    else {
                        D3 n1; sample_area(sc, light.prim, u_light, p1, n1);
                        if constexpr (sizeof(R) == 4) {
                            // f32 mode: the sampled POINT stays f64 (it is the shadow ray's end point); the direction handed to the
                            // BSDF and the area -> solid-angle factor are f32 (two f64 normalisations less)
                            const D3 wiu = p1 - rec.p; const double d2 = norm2(wiu);
                            if (d2 == 0.) { light_pdf = 0.; color = ld3(light.color); }
                            else {
                                const float inv = rsqrtf((float)d2);
                                const V3<float> w = v3<float>((float)wiu.x * inv, (float)wiu.y * inv, (float)wiu.z * inv);
                                const float cosl = fabsf((float)n1.x * w.x + (float)n1.y * w.y + (float)n1.z * w.z);
                                wid = mk3((double)w.x, (double)w.y, (double)w.z);
                                light_pdf = (double)((float)d2 / ((float)light.area * cosl));
                                color = light_l(light, n1, -wid);
                            }
                        } else {
                            double pdf = 1. / light.area;
                            D3 wiu = p1 - rec.p;
                            if (norm2(wiu) == 0.) pdf = 0.;
                            else { D3 wn = normalize(wiu); pdf = pdf * norm2(rec.p - p1) / fabs(dot(n1, -wn)); }
                            if (pdf == 0. || norm2(rec.p - p1) == 0.) { light_pdf = 0.; color = ld3(light.color); }
                            else { wid = normalize(p1 - rec.p); light_pdf = pdf; color = light_l(light, n1, -wid); }
                        }
                    }
                    if (light_pdf > 0. && !is_black(color)) {
                        V3<R> wi = cast3<R>(wid);
                        V3<R> f = bsdf_f<R, KM>(bsdf, wo_w, wi, nsf);
                        const bool f_some = !is_black(f);         // Whitted tests f itself, estimate_direct f x |cos| (integrator.rs:155, 578-583)
                        f = f * fabs(dot(wi, sn));
                        R scattering_pdf = whitted ? R(0) : bsdf_pdf<R, KM>(bsdf, wo_w, wi, nsf);
                        if (whitted ? f_some : !is_black(f)) {
                            R weight = (whitted || delta) ? R(1) : power_heuristic<R>((R)light_pdf, scattering_pdf);
                            V3<R> ld = cmul(f, cast3<R>(color)) * (weight / (R)light_pdf);
                            V3<R> contrib = cmul(ld * l_scale, beta);
                            // RT_SHADE_F32 only: a sample whose f32 radiometry overflowed (inf / inf at a grazing microfacet, ...) is dropped instead of
                            // poisoning its pixel for good; the f64 kernels propagate NaN as the reference does (util.rs:208-232: its guards never fire)
                            if constexpr (sizeof(R) == 4) { if (!finite3(contrib)) contrib = v3<R>(R(0), R(0), R(0)); }
                            uint32_t q = warp_append(&c->n_shadow);
                            if (!path && q >= rp.W) { c->overflow = 1u; return; }
                            __stcs(shq.ax + q, rec.p.x); __stcs(shq.ay + q, rec.p.y); __stcs(shq.az + q, rec.p.z); __stcs(shq.bx + q, p1.x); __stcs(shq.by + q, p1.y); __stcs(shq.bz + q, p1.z);
                            __stcs(shq.cr + q, (double)contrib.x); __stcs(shq.cg + q, (double)contrib.y); __stcs(shq.cb + q, (double)contrib.z); __stcs(shq.target + q, inf_light ? RT_NONE : light.prim); __stcs(shq.pixel + q, pixel);
                        }
                    }
                }
                // estimate_direct, BSDF half (integrator.rs:601-632)
                if (!whitted && !delta) {
                    V3<R> f, wi; R scattering_pdf; uint32_t st;
                    bsdf_sample_f<R, KM>(bsdf, wo_w, u_scatter, nsf, rng, f, wi, scattering_pdf, st);
                    f = f * fabs(dot(wi, sn));
                    if (!is_black(f) && scattering_pdf > R(0)) {
                        R weight = R(1); bool go = true;
                        D3 wid = dbl3(wi);
                        if ((st & BSDF_SPECULAR) == 0) {
                            // Light::pdf_li → Primitive::pdf (primitive.rs:462-473): re-intersect the ONE light primitive, tmin = 0, no leaf box
                            // Infinite: Light::pdf_li (light.rs:285-294)
                            D3 lp, ln; double light_pdf = 0.;
                            if (inf_light) light_pdf = infinite_pdf_li(sc, light, wid);
                            else if (prim_hit_pn<sizeof(R) == 4>(sc, light.prim, rec.p, wid, 0., INF, lp, ln)) {
                                if constexpr (sizeof(R) == 4) light_pdf = (double)((float)norm2(rec.p - lp) / ((float)light.area * fabsf((float)dot(ln, -wid))));   // an f32 MIS weight: f32 quotient
                                else light_pdf = norm2(rec.p - lp) / (light.area * fabs(dot(ln, -wid)));
                            }
                            if (light_pdf == 0.) go = false;
                            else weight = power_heuristic<R>(scattering_pdf, (R)light_pdf);
                        }
                        if (go) {
                            V3<R> contrib = cmul((f * (weight / scattering_pdf)) * l_scale, beta);   // × Le decided by the probe ray
                            if constexpr (sizeof(R) == 4) { if (!finite3(contrib)) go = false; }
                            if (go) {
                            uint32_t q = warp_append(&c->n_probe);
                            if (!path && q >= rp.W) { c->overflow = 1u; return; }
                            prq.ax[q] = rec.p.x; prq.ay[q] = rec.p.y; prq.az[q] = rec.p.z; prq.bx[q] = wid.x; prq.by[q] = wid.y; prq.bz[q] = wid.z;
                            prq.cr[q] = contrib.x; prq.cg[q] = contrib.y; prq.cb[q] = contrib.z; prq.target[q] = light_num; prq.pixel[q] = pixel;
                            }
                        }
                    }
                }
            };
            if constexpr (sizeof(R) == 4) { if (n_lights != 0) light_estimate(l_begin); }
            else { for (uint32_t ln = l_begin; ln < l_end; ln++) light_estimate(ln); }
            if (basic) {
                // cast_ray: one sample over all lobes; weight color / pdf x |incoming ray.dir . ns| (unnormalised, as the reference has it)
                if (bounces + 1 >= rp.max_depth) break;
                V2<R> u; u.x = rng_real<R>(rng); u.y = rng_real<R>(rng);
                V3<R> f, wi; R pdf; uint32_t sflags;
                bsdf_sample_f<R, KM>(bsdf, cast3<R>(-d), u, BSDF_ALL, rng, f, wi, pdf, sflags);
                if (pdf == R(0)) break;
                beta = cmul(beta, f * (R(1) / pdf)) * fabs(dot(cast3<R>(d), sn));
                o = rec.p; d = dbl3(wi); bounces++; alive = true;
                break;
            }
            if (!path) {
                // specular_reflect / specular_transmit (integrator.rs:464-504): `if depth - 1 > 0`, depth = max_depth - bounces.  Both samples are
                // drawn at the vertex; each surviving branch becomes a queue entry of its own that continues from this dimension counter.
                if (bounces + 1 >= rp.max_depth) break;
                V2<R> ub[2];
                ub[0].x = rng_real<R>(rng); ub[0].y = rng_real<R>(rng); ub[1].x = rng_real<R>(rng); ub[1].y = rng_real<R>(rng);
#pragma unroll
                for (int b = 0; b < 2; b++) {
                    V3<R> f, wi; R pdf; uint32_t sflags;
                    Rng scratch = rng;            // specular lobes draw nothing
                    bsdf_sample_f<R, KM>(bsdf, wo_w, ub[b], b == 0 ? (uint32_t)(BSDF_REFLECTION | BSDF_SPECULAR) : (uint32_t)(BSDF_TRANSMISSION | BSDF_SPECULAR), scratch, f, wi, pdf, sflags);
                    const R cs = fabs(dot(sn, wi));
                    if (pdf > R(0) && !is_black(f) && cs != R(0)) {
                        const V3<R> nb = cmul(beta, f) * (cs / pdf);
                        const uint32_t q = warp_append(&c->n_next);
                        if (q >= rp.W) { c->overflow = 1u; continue; }
                        store_ray(out.ray + q, rec.p, dbl3(wi));
                        store_aux(out.aux + q, (double)nb.x, (double)nb.y, (double)nb.z, pixel, rng.sample, rng.dim, ((bounces + 1u) & 0xffffu) | 0x10000u);
                    }
                }
                break;
            }
            // continuation (integrator.rs:421-441)
            V3<R> wo = cast3<R>(-d);
            V2<R> u; u.x = R(0); u.y = R(0);
            if constexpr (mat_ignores_sample2d<MAT>()) rng_skip(rng, 2);
            else { u.x = rng_real<R>(rng); u.y = rng_real<R>(rng); }
            V3<R> f, wi; R pdf; uint32_t sflags;
            bsdf_sample_f<R, KM>(bsdf, wo, u, BSDF_ALL, rng, f, wi, pdf, sflags);
            if (is_black(f) || pdf == R(0)) break;
            beta = cmul(beta, f) * (fabs(dot(wi, sn)) / pdf);
            if constexpr (sizeof(R) == 4) { if (!finite3(beta)) break; }          // f32 mode: end a path whose throughput overflowed (see above)
            specular_bounce = (sflags & BSDF_SPECULAR) != 0;
            o = rec.p; d = dbl3(wi);
            if (bounces > 3) {
                R q = fmax(R(0.05), R(1) - fmax(beta.x, fmax(beta.y, beta.z)));
                if (rng_real<R>(rng) < q) break;
                beta = beta * (R(1) / (R(1) - q));
            }
            bounces++;
            alive = true;
        } while (false);
        if (alive) {
            uint32_t q = warp_append(&c->n_next);
            store_ray(out.ray + q, o, d);
            store_aux(out.aux + q, (double)beta.x, (double)beta.y, (double)beta.z, pixel, rng.sample, rng.dim, (bounces & 0xffffu) | (specular_bounce ? 0x10000u : 0u));
        }
#if SHADE_PREFETCH
        entry = entry_next;
#else
        if (k + stride < n) entry = bin_idx[k + stride];
#endif
    }
}

// one launch per shading bin present in the scene (render.cu: R = double ; shade_f32.cu: R = float)
struct ShadeLaunch { Ctrl* ctrl; const DScene* sc; const RenderParams* rp; PathSet in, out; const uint2* bin_idx; ConnQ shadow, probe; double* film; uint32_t bin_mask; int grid; cudaStream_t stream; };
template <class R, bool ENVL> inline void launch_shade_bins(const ShadeLaunch& L) {
#define RT_SHADE(M) \
    if (L.bin_mask & (1u << (M * 2))) k_shade<R, M, 0, ENVL><<<L.grid, 128, 0, L.stream>>>(L.ctrl, *L.sc, *L.rp, L.in, L.out, L.bin_idx, L.shadow, L.probe, L.film); \
    if (L.bin_mask & (1u << (M * 2 + 1))) k_shade<R, M, 1, ENVL><<<L.grid, 128, 0, L.stream>>>(L.ctrl, *L.sc, *L.rp, L.in, L.out, L.bin_idx, L.shadow, L.probe, L.film);
    RT_SHADE(RT_MAT_MATTE) RT_SHADE(RT_MAT_LIGHT) RT_SHADE(RT_MAT_PLASTIC) RT_SHADE(RT_MAT_GLASS) RT_SHADE(RT_MAT_METAL) RT_SHADE(RT_MAT_MIRROR)
    if constexpr (sizeof(R) == 8) { RT_SHADE(RT_MAT_DISNEY) }        // Material::Disney: f64 kernels only (rt_render shades such a scene in f64)
#undef RT_SHADE
}
void launch_shade_f32(const ShadeLaunch& L, bool envl);      // shade_f32.cu

}  // namespace rt
