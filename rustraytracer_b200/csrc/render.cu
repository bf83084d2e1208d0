// Device half of the C ABI, part 2: the wavefront MIS path tracer.
//
// Replaces render::tile_multithread → PathIntegrator::render/single_sample/li (reference render.rs:13-159,
// integrator.rs:336-445) and uniform_sample_one_light / estimate_direct (integrator.rs:530-634).
// The reference runs one recursive f64 path per CPU thread; here a wavefront of W paths lives in HBM as
// SoA queues and every iteration runs
//     prepare → generate (refill with new camera samples) → extend (closest hit)
//             → shade (emission, BSDF build, light sample, MIS BSDF sample, continuation, roulette)
//             → connect (shadow any-hit rays, MIS probe closest-hit rays; adds to the film)
// with queue compaction by warp-aggregated atomics.  No host synchronisation inside an iteration; the
// host polls the control block every few iterations.  Randomness: counter-based Philox keyed by
// (seed; pixel, sample, dimension) so a path's numbers do not depend on scheduling or on the GPU count.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include "wavefront.cuh"
#include "film.cuh"
#include "traverse.cuh"

namespace rt {

#define CK(call)                                                                                     \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) { set_error(std::string(#call) + ": " + cudaGetErrorString(e_)); return RT_ERR_CUDA; } \
    } while (0)

// RT_SORT_HITS: where the hit-cell bucket sort keeps its state, and the map from a hit point to its cell.  Scenes mix a 20,000-unit floor
// with a 6-unit dragon, so the grid is not laid over the scene bounds: a point is first compressed with u / (1 + |u|), u = (p - centre) / scale
// (centre = per-axis median of the primitive centroids, scale = 2 x the median centroid distance), which keeps most cells where the geometry is.
struct SortParams {
    float cx, cy, cz, inv_s;
    uint32_t n_keys;                              // (bins present in the scene + miss) x SORT_CELLS
    uint64_t slots;                               // shading bin -> key segment, 4 bits per bin, 0xf = bin absent from the scene (at most 13 segments exist)
    __host__ __device__ uint32_t slot(uint32_t bin) const { return (uint32_t)(slots >> (4u * bin)) & 15u; }
    uint32_t *hist, *offs;                        // per key: count (zeroed again by the scan) and running output position
    uint2* ent;                                   // per path of the wavefront: (key, hit primitive), written by extend
};
__device__ __forceinline__ uint32_t spread3(uint32_t v) {       // 10 bits -> every third bit
    v = (v | (v << 16)) & 0x030000ffu; v = (v | (v << 8)) & 0x0300f00fu; v = (v | (v << 4)) & 0x030c30c3u; v = (v | (v << 2)) & 0x09249249u;
    return v;
}
__device__ __forceinline__ uint32_t hit_cell(const SortParams& sp, float x, float y, float z) {
    const float G = (float)(1u << RT_SORT_BITS);
    float ux = (x - sp.cx) * sp.inv_s, uy = (y - sp.cy) * sp.inv_s, uz = (z - sp.cz) * sp.inv_s;
    ux = __fdividef(ux, 1.f + fabsf(ux)); uy = __fdividef(uy, 1.f + fabsf(uy)); uz = __fdividef(uz, 1.f + fabsf(uz));
    const uint32_t m = (1u << RT_SORT_BITS) - 1u;
    const uint32_t qx = min((uint32_t)fmaxf((ux * 0.5f + 0.5f) * G, 0.f), m), qy = min((uint32_t)fmaxf((uy * 0.5f + 0.5f) * G, 0.f), m), qz = min((uint32_t)fmaxf((uz * 0.5f + 0.5f) * G, 0.f), m);
    return spread3(qx) | (spread3(qy) << 1) | (spread3(qz) << 2);
}

struct Wavefront {
    uint32_t W = 0;
    PathSet set[2]{}; ConnQ shadow{}, probe{};
    uint2* bin_idx = nullptr;                     // (path index, hit primitive) per shading bin: N_BINS x W filled by extend, or W sorted entries (RT_SORT_HITS)
    SortParams sort{};
    Ctrl* ctrl = nullptr; Ctrl* h_ctrl = nullptr;
    cudaStream_t stream = nullptr;
    std::vector<void*> allocs;
    std::vector<cudaEvent_t> events;
    cudaEvent_t dbg[2] = {nullptr, nullptr};      // RT_TRACE_ITER: after k_sort_scan / k_sort_scatter
};

// `cap` = most paths allowed in flight: W for Path (one continuation per path at most); W / max(2, n_lights) for the branching integrators,
// whose vertices append up to two continuations and one shadow + one probe entry PER LIGHT
// Queue appends beyond W are dropped by the shade kernels (branching integrators only; `overflow` is raised and the host aborts the
// render), but the counters keep counting: every consumer clamps them to W so that nothing is ever read or filed past an allocation.
__global__ void k_prepare(Ctrl* c, uint32_t cap, uint32_t W) {
    c->c_extend += c->n_tot; c->c_shadow += min(c->n_shadow, W); c->c_probe += min(c->n_probe, W); c->c_iter += 1;
    uint32_t n_cur = min(c->n_next, W);
    unsigned long long remaining = c->total_samples - c->next_sample;
    uint32_t room = cap > n_cur ? cap - n_cur : 0u;
    uint32_t n_new = (uint32_t)(remaining < (unsigned long long)room ? remaining : (unsigned long long)room);
    c->n_cur = n_cur; c->n_new = n_new; c->n_tot = n_cur + n_new; c->n_next = 0; c->n_shadow = 0; c->n_probe = 0;
    c->sample_base = c->next_sample; c->next_sample += n_new; c->c_paths += n_new;
    for (int b = 0; b < N_BINS; b++) c->bin_count[b] = 0;
    for (int b = 0; b < 4; b++) c->fetch[b] = 0;
}

// Order in which the pixels of the render window receive camera samples.  The reference hands out 16x16 tiles (render.rs:49-66);
// any order gives the same film because the RNG is keyed by (pixel, sample).  Scanline order makes a warp a 32x1 strip; here, when the
// window is a multiple of 32 in both directions, a warp is an 8x4 pixel tile and 32 consecutive warps fill a 32x32 super-tile, so the camera
// rays of a warp — and the bounce-1 rays and shadow rays derived from them in queue order — start out close together in 2-D.
// MEASURED (dragon, 1024^2, 64 spp): 81.97 vs 82.16 ms per render, extend 31.41 vs 31.54 ms — no gain, so scanline order stays the default.
#ifndef RT_TILED_GEN
#define RT_TILED_GEN 0
#endif
__device__ __forceinline__ void window_pixel(const RenderParams& rp, uint32_t pi, uint32_t& px, uint32_t& py) {
#if RT_TILED_GEN
    if (((rp.ww | rp.wh) & 31u) == 0u) {
        const uint32_t st = pi >> 10, in = pi & 1023u, tile = in >> 5, l = in & 31u, stw = rp.ww >> 5;
        px = rp.x0 + (st % stw) * 32u + (tile & 3u) * 8u + (l & 7u);
        py = rp.y0 + (st / stw) * 32u + (tile >> 2) * 4u + (l >> 3);
        return;
    }
#endif
    px = rp.x0 + pi % rp.ww; py = rp.y0 + pi / rp.ww;
}

// PathIntegrator::single_sample up to the camera ray (integrator.rs:357-365), Samplers::get_camera_sample
// (sampler.rs:606-613), Camera::get_ray (geometry.rs:177-190)
__global__ void __launch_bounds__(256) k_generate(const Ctrl* __restrict__ c, RenderParams rp, PathSet ps, double* film) {
    uint32_t n_new = c->n_new, n_cur = c->n_cur; unsigned long long base = c->sample_base;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_new; i += gridDim.x * blockDim.x) {
        unsigned long long g = base + i;
        unsigned long long npix = (unsigned long long)rp.ww * rp.wh;
        uint32_t s = (uint32_t)(g / npix), pi = (uint32_t)(g % npix);
        uint32_t px, py; window_pixel(rp, pi, px, py);
        uint32_t pixel = py * rp.width + px;
        Rng rng; rng.k0 = rp.k0; rng.k1 = rp.k1; rng.pixel = pixel; rng.sample = rp.sample_begin + s; rng.dim = 0; rng.w2 = 0xffffffffu; rng.w3 = 0;
        double offx = rng_next(rng), offy = rng_next(rng);
        double fx = (double)px + offx, fy = (double)py + offy;
        rng_skip(rng, 3);                    // time, lens sample (2): drawn by get_camera_sample, read by nobody (get_ray draws its own lens point) — the
                                             // dimensions are consumed, their Philox block is not computed
        // BasicIntegrator::single_sample (integrator.rs:196-204) divides by width - 1 / height - 1
        const bool basic = rp.kind == RT_INTEGRATOR_BASIC;
        double u = fx / (double)(basic ? rp.width - 1u : rp.width), v = fy / (double)(basic ? rp.height - 1u : rp.height);
        D3 in_disk;
        for (;;) { double x = rng_next(rng), y = rng_next(rng); if (x * x + y * y < 1.) { in_disk = mk3(x, y, 0.); break; } }   // util::rand_in_disk
        in_disk = in_disk * rp.cam.lens_radius;
        D3 offset = ld3(rp.cam.u) * in_disk.x + ld3(rp.cam.v) * in_disk.y;
        D3 to = ld3(rp.cam.upper_left_corner) + ld3(rp.cam.horizontal_offset) * u - ld3(rp.cam.vertical_offset) * v;
        D3 dir = to - ld3(rp.cam.origin);
        rng_skip(rng, 1);                    // util::rand_range(t0, t1): ray time, unused by geometry
        D3 o = ld3(rp.cam.origin) + offset, d = dir - offset;
        uint32_t k = n_cur + i;
        store_ray(ps.ray + k, o, d);
        store_aux(ps.aux + k, 1., 1., 1., pixel, rng.sample, rng.dim, 0u);
        // the sample count `n += 1` of util::increment_color (util.rs:231) is added once per pixel by k_film_count at the end of the
        // render: camera samples are generated in the fixed order g = sample * npix + pixel, so it is a function of the total alone
    }
}

// cell.n += (samples of this pixel among the first `generated` camera samples) — replaces one f64 atomic per path in k_generate
__global__ void k_film_count(RenderParams rp, unsigned long long generated, double* film) {
    const unsigned long long npix = (unsigned long long)rp.ww * rp.wh;
    const unsigned long long full = generated / npix, rest = generated % npix;
    for (unsigned long long pi = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; pi < npix; pi += (unsigned long long)gridDim.x * blockDim.x) {
        uint32_t px, py; window_pixel(rp, (uint32_t)pi, px, py);
        film[4 * ((size_t)py * rp.width + px) + 3] += (double)(full + (pi < rest ? 1ull : 0ull));
    }
}

// ray type A: geometry::get_intersection (geometry.rs:238-240).  Besides the hit, extend files every path into a
// shading bin keyed by (material type, triangle / parametric shape) — or "miss" — with one atomic per distinct key
// per warp (__match_any_sync), so that the shade kernel runs warps of like work instead of diverging on the
// reference's `match material` / `match primitive` arms.
struct ExtendJob {
    static constexpr bool shared_leaf_ray = false;      // traverse.cuh:trace_jobs
    Ctrl* c; const DScene* sc; PathSet ps; uint2* bin_idx; uint32_t W; const SortParams* sp;
    __device__ __forceinline__ double tmin(const TState&) const { return SMALL; }
    __device__ __forceinline__ double tmax(const TState&) const { return INF; }
    __device__ __forceinline__ void ray(uint32_t i, D3& o, D3& d) const {
        const double2* q = reinterpret_cast<const double2*>(ps.ray + i);
        const double2 a = __ldcs(q), b = __ldcs(q + 1), c = __ldcs(q + 2);     // streaming: path state must not evict the BVH from L2
        o = mk3(a.x, a.y, b.x); d = mk3(b.y, c.x, c.y);
    }
    __device__ __forceinline__ bool load(uint32_t i, TState& st, int32_t* stack) {
        D3 o, d; ray(i, o, d);
        tstate_init(*sc, st, stack, o, d, SMALL, INF, false, 0., 0u);
        return true;
    }
    __device__ __forceinline__ void finish(uint32_t i, const TState& st) {
        uint32_t prim = st.best.prim, key = N_BINS - 1;
        if (prim != RT_NONE) key = __ldg(sc->prim_bin + prim);
        unsigned active = __activemask();
#if RT_SORT_HITS
        // count the (bin, hit cell) key — one atomic per distinct key per warp — and leave (key, primitive) for k_sort_scatter
        uint32_t cell = 0;
        if (prim != RT_NONE) cell = hit_cell(*sp, 0.f, 0.f, 0.f /* RT_SORT_HITS is retired: the state no longer carries the f64 ray */);
        const uint32_t skey = sp->slot(key) * SORT_CELLS + cell;
        const unsigned same = __match_any_sync(active, skey);
        if ((int)(threadIdx.x & 31) == __ffs(same) - 1) atomicAdd(&sp->hist[skey], (uint32_t)__popc(same));
        sp->ent[i] = make_uint2(skey, prim);
#else
        unsigned peers = __match_any_sync(active, key);
        int leader = __ffs(peers) - 1, lane = threadIdx.x & 31;
        uint32_t base = 0;
        if (lane == leader) base = atomicAdd(&c->bin_count[key], (uint32_t)__popc(peers));
        base = __shfl_sync(peers, base, leader);
        bin_idx[(size_t)key * W + base + __popc(peers & ((1u << lane) - 1u))] = make_uint2(i, prim);   // path index + hit primitive
#endif
    }
};
#ifndef TRACE_MIN_BLOCKS
#define TRACE_MIN_BLOCKS 5      // resident 128-thread blocks per SM of the persistent traversal kernels (register cap 65536 / (128 * N))
#endif
#ifndef TRACE_GRID_BLOCKS
#define TRACE_GRID_BLOCKS 8     // persistent grid = SMs x this
#endif
__global__ void __launch_bounds__(128, TRACE_MIN_BLOCKS) k_extend(Ctrl* c, const __grid_constant__ DScene sc, PathSet ps, uint2* __restrict__ bin_idx, uint32_t W, SortParams sp) {
    ExtendJob job{c, &sc, ps, bin_idx, W, &sp};
    trace_jobs(sc, job, c->n_tot, &c->fetch[0]);
}

// RT_SORT_HITS, step 2 of 3: exclusive prefix sum of the key counts (one block; every warp owns a contiguous run of keys and reads it
// 128 keys at a time), the counts zeroed for the next iteration, and the bins' extents published for the shade kernels.
__global__ void __launch_bounds__(1024) k_sort_scan(Ctrl* c, SortParams sp) {
    __shared__ uint32_t wsum[32];
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t per_warp4 = sp.n_keys / 128;               // uint4 words per warp (n_keys is a multiple of 4096)
    uint4* h = reinterpret_cast<uint4*>(sp.hist) + (size_t)warp * per_warp4;
    uint4* o = reinterpret_cast<uint4*>(sp.offs) + (size_t)warp * per_warp4;
    uint32_t s = 0;
    for (uint32_t t = lane; t < per_warp4; t += 32) { const uint4 v = h[t]; s += v.x + v.y + v.z + v.w; }
    for (int d = 16; d; d >>= 1) s += __shfl_xor_sync(FULL, s, d);
    if (lane == 0) wsum[warp] = s;
    __syncthreads();
    if (warp == 0) {
        const uint32_t v = wsum[lane]; uint32_t inc = v;
        for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(FULL, inc, d); if (lane >= d) inc += y; }
        wsum[lane] = inc - v;
    }
    __syncthreads();
    uint32_t carry = wsum[warp];
    for (uint32_t t0 = 0; t0 < per_warp4; t0 += 32) {
        const uint32_t t = t0 + lane;
        const uint4 v = h[t];
        const uint32_t loc = v.x + v.y + v.z + v.w; uint32_t inc = loc;
        for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(FULL, inc, d); if (lane >= d) inc += y; }
        const uint32_t ex = carry + inc - loc;
        o[t] = make_uint4(ex, ex + v.x, ex + v.x + v.y, ex + v.x + v.y + v.z);
        h[t] = make_uint4(0u, 0u, 0u, 0u);
        carry += __shfl_sync(FULL, inc, 31);
    }
    __syncthreads();
    if (threadIdx.x < N_BINS) {
        const uint32_t b = threadIdx.x, slot = sp.slot(b);
        uint32_t base = 0, count = 0;
        if (slot != 15u) {
            base = sp.offs[(size_t)slot * SORT_CELLS];
            const uint32_t end = ((slot + 1u) * SORT_CELLS < sp.n_keys) ? sp.offs[(size_t)(slot + 1u) * SORT_CELLS] : c->n_tot;
            count = end - base;
        }
        c->bin_base[b] = base; c->bin_count[b] = count;
    }
}
// step 3 of 3: every path's entry goes to the next free position of its key
__global__ void __launch_bounds__(256) k_sort_scatter(const Ctrl* c, SortParams sp, uint2* __restrict__ sorted) {
    const uint32_t n = c->n_tot;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint2 e = sp.ent[i];
        const unsigned active = __activemask();
        const unsigned peers = __match_any_sync(active, e.x);
        const int leader = __ffs(peers) - 1, lane = threadIdx.x & 31;
        uint32_t base = 0;
        if (lane == leader) base = atomicAdd(&sp.offs[e.x], (uint32_t)__popc(peers));
        base = __shfl_sync(peers, base, leader);
        sorted[base + __popc(peers & ((1u << lane) - 1u))] = make_uint2(i, e.y);
    }
}

// Paths whose extend ray left the scene (bin 15).  Only launched when the scene has an Infinite light: at bounce 0 or after a
// specular bounce every light's `le(ray)` is added (integrator.rs:406-410; Light::le is black for area lights).
__global__ void __launch_bounds__(128) k_shade_miss(Ctrl* c, const __grid_constant__ DScene sc, const __grid_constant__ RenderParams rp, PathSet in, const uint2* __restrict__ bin_idx, double* film) {
    const uint32_t n = c->bin_count[N_BINS - 1];
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        const uint32_t i = bin_idx[bin_first(c, rp, N_BINS - 1) + k].x;
        const PAux pa = load_aux(in.aux + i);
        if (rp.kind == RT_INTEGRATOR_PATH && !((pa.flags & 0xffffu) == 0 || ((pa.flags >> 16) & 1u))) continue;   // Whitted / Direct: every level (integrator.rs:135-140)
        const PRay pr = load_ray(in.ray + i);
        D3 d = mk3(pr.dx, pr.dy, pr.dz), beta = mk3(pa.bx, pa.by, pa.bz), l = mk3(0., 0., 0.);
        for (uint32_t li = 0; li < sc.n_lights; li++)
            if (sc.lights[li].type == RT_LIGHT_INFINITE) {
                D3 le = infinite_le(sc, sc.lights[li], d);
                if (rp.kind == RT_INTEGRATOR_BASIC) le = le * (1. / (4. * PI));     // cast_ray's miss arm (geometry.rs:269-275)
                l = l + cmul(le, beta);
            }
        film_add(film, pa.pixel, l);
    }
}

// Ray types B and C are traced by ONE persistent launch over the concatenated queue [shadow | probe]: the probe queue is
// ~5 % of the shadow queue, and as a launch of its own it paid a whole persistent-kernel tail for it (ncu launch list r01:
// 11 % of the step for 2.4 % of the rays).
// B: Visibility::unoccluded (hittable.rs:25-39) as an any-hit query: exact hit of the light primitive first (a miss
//    means "occluded"), then search for anything closer, stopping at the first one.
// C: the MIS BSDF probe (integrator.rs:614-630): closest hit; counts only if it is the sampled light.
struct ConnectJob {
    static constexpr bool shared_leaf_ray = !RT_PREFILTER;      // traverse.cuh:trace_jobs
    const DScene* sc; ConnQ sh, pr; double* film; uint32_t n_shadow;
    __device__ __forceinline__ double tmin(const TState& st) const { return st.any ? 0. : SMALL; }     // shadow rays (any-hit): (0, INF); probes: (SMALL, INF)
    __device__ __forceinline__ double tmax(const TState&) const { return INF; }
    __device__ __forceinline__ void ray(uint32_t i, D3& o, D3& d) const {
        if (i >= n_shadow) {
            const uint32_t k = i - n_shadow;
            o = mk3(__ldcs(pr.ax + k), __ldcs(pr.ay + k), __ldcs(pr.az + k)); d = mk3(__ldcs(pr.bx + k), __ldcs(pr.by + k), __ldcs(pr.bz + k));
            return;
        }
        const D3 p0 = mk3(__ldcs(sh.ax + i), __ldcs(sh.ay + i), __ldcs(sh.az + i)), p1 = mk3(__ldcs(sh.bx + i), __ldcs(sh.by + i), __ldcs(sh.bz + i));
        d = p1 - p0;                    // Visibility::unoccluded (hittable.rs:25-30): dir = p1 - p0, origin pushed SMALL x dir along it
        o = p0 + d * SMALL;
    }
    __device__ __forceinline__ bool load(uint32_t i, TState& st, int32_t* stack) {
        D3 org, dir; ray(i, org, dir);
        // one tstate_init for the three kinds of entry (probe: closest hit over (SMALL, INF); shadow ray to an area light: any hit below the light
        // primitive's own t; shadow ray to the Infinite light: any hit at all, hittable.rs:31-33), so that its body is inlined once
        bool any = false; double t_limit = 0.; uint32_t skip = 0u;
        if (i < n_shadow) {
            const uint32_t target = __ldcs(sh.target + i);
            any = true; t_limit = INF; skip = RT_NONE;
            if (target != RT_NONE) {
                const DPrim prm = sc->prims[target];
                if ((prm.info & 7u) == RT_PRIM_TRIANGLE) {
                    Rec r;
                    if (!prim_record_t<1>(*sc, target, org, dir, 0., INF, true, r)) return false;
                    t_limit = r.t;
                } else if (!shape_test(*sc, prm, org, dir, 0., INF, true, t_limit)) return false;
                skip = target;
            }
        }
        tstate_init(*sc, st, stack, org, dir, any ? 0. : SMALL, INF, any, t_limit, skip);
        return true;
    }
    __device__ __forceinline__ void finish(uint32_t i, const TState& st) {
        if (i < n_shadow) {
            if (!st.occluded) film_add(film, __ldcs(sh.pixel + i), mk3(__ldcs(sh.cr + i), __ldcs(sh.cg + i), __ldcs(sh.cb + i)));
            return;
        }
        uint32_t k = i - n_shadow;
        uint32_t prim = st.best.prim;
        if (prim == RT_NONE) {         // the probe left the scene: `color = light.le(&new_ray)` — non-black only for an Infinite light
            if (!sc->has_infinite) return;
            const DLight& l = sc->lights[pr.target[k]];
            if (l.type != RT_LIGHT_INFINITE) return;
            D3 le = infinite_le(*sc, l, mk3(pr.bx[k], pr.by[k], pr.bz[k]));
            if (is_black(le)) return;
            film_add(film, pr.pixel[k], cmul(mk3(pr.cr[k], pr.cg[k], pr.cb[k]), le));
            return;
        }
        uint32_t li = sc->prims[prim].light;
        if (li == RT_NONE) return;
        {   // `lights[index].eq(light)` (integrator.rs:621): field equality (classes computed at commit), not identity; NaN colours never match
            const DLight& a = sc->lights[li]; const DLight& b = sc->lights[pr.target[k]];
            if (a.eq_class != b.eq_class || !(a.color[0] == a.color[0] && a.color[1] == a.color[1] && a.color[2] == a.color[2])) return;
        }
        D3 o = mk3(pr.ax[k], pr.ay[k], pr.az[k]), d = mk3(pr.bx[k], pr.by[k], pr.bz[k]);
        Rec r;
        if (!prim_record(*sc, prim, o, d, SMALL, INF, true, r)) return;
        D3 le = light_l(sc->lights[li], r.n, -d);                  // new_record.le(&-wi)
        if (is_black(le)) return;
        film_add(film, pr.pixel[k], cmul(mk3(pr.cr[k], pr.cg[k], pr.cb[k]), le));
    }
};
#ifndef CONNECT_MIN_BLOCKS
#define CONNECT_MIN_BLOCKS TRACE_MIN_BLOCKS     // resident blocks per SM of k_connect alone (its load() / finish() carry more state than k_extend's)
#endif
__global__ void __launch_bounds__(128, CONNECT_MIN_BLOCKS) k_connect(Ctrl* c, const __grid_constant__ DScene sc, ConnQ shq, ConnQ prq, double* film, uint32_t W) {
    const uint32_t n_shadow = min(c->n_shadow, W), n_probe = min(c->n_probe, W);      // see k_prepare: counters may run past W on overflow
    ConnectJob job{&sc, shq, prq, film, n_shadow};
    trace_jobs(sc, job, n_shadow + n_probe, &c->fetch[1]);
}

// util::point_to_color (util.rs:441-471): film.cuh
__global__ void k_resolve(const double* __restrict__ film, uint64_t n, uint8_t* __restrict__ rgb) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        resolve_pixel(film[4 * i], film[4 * i + 1], film[4 * i + 2], film[4 * i + 3], rgb + 3 * i);
}

}  // namespace rt

namespace rt {

template <class T> static bool dalloc(Wavefront& w, T*& p, size_t n) {
    void* q = nullptr;
    if (cudaMalloc(&q, n * sizeof(T)) != cudaSuccess) return false;
    w.allocs.push_back(q); p = (T*)q; return true;
}
static bool alloc_set(Wavefront& w, PathSet& s, size_t n) {
    return dalloc(w, s.ray, n) && dalloc(w, s.aux, n);
}
static bool alloc_q(Wavefront& w, ConnQ& q, size_t n) {
    return dalloc(w, q.ax, n) && dalloc(w, q.ay, n) && dalloc(w, q.az, n) && dalloc(w, q.bx, n) && dalloc(w, q.by, n) && dalloc(w, q.bz, n) &&
           dalloc(w, q.cr, n) && dalloc(w, q.cg, n) && dalloc(w, q.cb, n) && dalloc(w, q.target, n) && dalloc(w, q.pixel, n);
}
static void destroy_wavefront(Wavefront* w) {
    if (!w) return;
    for (void* p : w->allocs) cudaFree(p);
    for (auto e : w->events) cudaEventDestroy(e);
    for (auto e : w->dbg) if (e) cudaEventDestroy(e);
    if (w->h_ctrl) cudaFreeHost(w->h_ctrl);
    if (w->stream) cudaStreamDestroy(w->stream);
    delete w;
}
void render_prefilter_stats(unsigned long long out[4]) { prefilter_stats_add(out); }
void free_wavefront(rt_scene& s) { destroy_wavefront((Wavefront*)s.wavefront); s.wavefront = nullptr; }

static Wavefront* get_wavefront(rt_scene& s) {
    if (s.wavefront) return (Wavefront*)s.wavefront;
    Wavefront* w = new Wavefront();
    w->W = s.wavefront_size;
    size_t n = w->W;
#if RT_SORT_HITS
    // key segments: one per shading bin that occurs in the scene, and the miss bin
    SortParams& sp = w->sort;
    uint32_t n_slots = 0;
    sp.slots = 0;
    for (int b = 0; b < N_BINS; b++) sp.slots |= (uint64_t)(((s.bin_mask >> b) & 1u) || b == N_BINS - 1 ? n_slots++ : 15u) << (4 * b);
    sp.n_keys = n_slots * SORT_CELLS;
    {   // centre and scale of the cell map: medians over the primitive centroids
        const size_t np = s.prims.size();
        std::vector<double> c[3], dev(np);
        for (int a = 0; a < 3; a++) c[a].resize(np);
        for (size_t i = 0; i < np; i++) { const Box64 b = prim_box(s, s.prims[i]); for (int a = 0; a < 3; a++) c[a][i] = 0.5 * (b.mn[a] + b.mx[a]); }
        double med[3] = {0., 0., 0.}, scale = 1.;
        if (np) {
            for (int a = 0; a < 3; a++) { std::vector<double> t = c[a]; std::nth_element(t.begin(), t.begin() + np / 2, t.end()); med[a] = t[np / 2]; }
            for (size_t i = 0; i < np; i++) dev[i] = std::max(std::fabs(c[0][i] - med[0]), std::max(std::fabs(c[1][i] - med[1]), std::fabs(c[2][i] - med[2])));
            std::nth_element(dev.begin(), dev.begin() + np / 2, dev.end());
            scale = 2. * dev[np / 2];
            if (!(scale > 0.) || !std::isfinite(scale)) scale = 1.;
        }
        sp.cx = (float)med[0]; sp.cy = (float)med[1]; sp.cz = (float)med[2]; sp.inv_s = (float)(1. / scale);
    }
    const size_t n_entries = n;
    bool ok = dalloc(*w, sp.hist, sp.n_keys) && dalloc(*w, sp.offs, sp.n_keys) && dalloc(*w, sp.ent, n) && cudaMemset(sp.hist, 0, sp.n_keys * sizeof(uint32_t)) == cudaSuccess;
#else
    const size_t n_entries = n * N_BINS;
    bool ok = true;
#endif
    ok = ok && alloc_set(*w, w->set[0], n) && alloc_set(*w, w->set[1], n) && alloc_q(*w, w->shadow, n) && alloc_q(*w, w->probe, n) &&
              dalloc(*w, w->bin_idx, n_entries) && dalloc(*w, w->ctrl, 1) &&
              cudaMallocHost((void**)&w->h_ctrl, sizeof(Ctrl)) == cudaSuccess && cudaStreamCreateWithFlags(&w->stream, cudaStreamNonBlocking) == cudaSuccess;
    if (!ok) { set_error(std::string("wavefront allocation failed: ") + cudaGetErrorString(cudaGetLastError())); destroy_wavefront(w); return nullptr; }
    // Keep the inner nodes of the BVH resident in L2 while ~1.4 GB of path state streams through every iteration:
    // a persisting access-policy window over `nodes` on the render stream.  Measured on the dragon (r01, 64 spp): no gain
    // for extend/connect (37.4 vs 37.8 ms, 26.6 vs 26.7 ms) and a 4 % loss for shade (55.1 vs 52.8 ms: less L2 left for the
    // path-state gathers), so it is OFF unless RT_L2_PERSIST=1.
    const char* pe = getenv("RT_L2_PERSIST");
    if (pe && pe[0] == '1' && s.d.nodes && s.bytes_nodes) {
        int max_persist = 0, max_window = 0;
        cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, s.device);
        cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, s.device);
        if (max_persist > 0 && max_window > 0) {
            size_t want = std::min<size_t>(s.bytes_nodes, (size_t)max_persist);
            cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want);
            cudaStreamAttrValue av{};
            av.accessPolicyWindow.base_ptr = const_cast<DNodeT*>(s.d.nodes);
            av.accessPolicyWindow.num_bytes = std::min<size_t>(s.bytes_nodes, (size_t)max_window);
            av.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)want / (double)av.accessPolicyWindow.num_bytes);
            av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
            av.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
            cudaStreamSetAttribute(w->stream, cudaStreamAttributeAccessPolicyWindow, &av);
            cudaGetLastError();
        }
    }
    s.wavefront = w;
    return w;
}

static cudaEvent_t ev(Wavefront& w, size_t i) {
    while (w.events.size() <= i) { cudaEvent_t e; cudaEventCreate(&e); w.events.push_back(e); }
    return w.events[i];
}

}  // namespace rt

using namespace rt;

extern "C" {

rt_status rt_film_create(uint32_t w, uint32_t h, rt_film** out) {
    if (!out || !w || !h) { set_error("rt_film_create: bad argument"); return RT_ERR_INVALID; }
    rt_film* f = new rt_film(); f->w = w; f->h = h; f->owned = true;
    if (cudaGetDevice(&f->device) != cudaSuccess || cudaMalloc((void**)&f->d, (size_t)w * h * 4 * sizeof(double)) != cudaSuccess) {
        set_error(std::string("rt_film_create: ") + cudaGetErrorString(cudaGetLastError())); delete f; return RT_ERR_CUDA;
    }
    // rt_render accumulates on its own non-blocking stream, which has no implicit ordering with the legacy stream the memset runs on:
    // the film must be zero before the call returns
    if (cudaMemset(f->d, 0, (size_t)w * h * 4 * sizeof(double)) != cudaSuccess || cudaStreamSynchronize(cudaStreamLegacy) != cudaSuccess) {
        set_error(std::string("rt_film_create: ") + cudaGetErrorString(cudaGetLastError())); cudaFree(f->d); delete f; return RT_ERR_CUDA;
    }
    *out = f; return RT_OK;
}
rt_status rt_film_create_external(uint32_t w, uint32_t h, void* dev, rt_film** out) {
    if (!out || !w || !h || !dev) { set_error("rt_film_create_external: bad argument"); return RT_ERR_INVALID; }
    rt_film* f = new rt_film(); f->w = w; f->h = h; f->owned = false; f->d = (double*)dev;
    if (cudaGetDevice(&f->device) != cudaSuccess) { set_error("rt_film_create_external: no CUDA device"); delete f; return RT_ERR_CUDA; }
    cudaPointerAttributes pa{};                    // the device that owns the caller's allocation, not whichever happens to be current
    if (cudaPointerGetAttributes(&pa, dev) == cudaSuccess && (pa.type == cudaMemoryTypeDevice || pa.type == cudaMemoryTypeManaged)) f->device = pa.device;
    else cudaGetLastError();
    *out = f; return RT_OK;
}
rt_status rt_film_destroy(rt_film* f) { if (!f) return RT_OK; if (f->owned && f->d) cudaFree(f->d); delete f; return RT_OK; }
rt_status rt_film_clear(rt_film* f) {
    if (!f) { set_error("rt_film_clear: null film"); return RT_ERR_INVALID; }
    CK(cudaSetDevice(f->device));
    CK(cudaMemset(f->d, 0, (size_t)f->w * f->h * 4 * sizeof(double)));
    CK(cudaStreamSynchronize(cudaStreamLegacy));          // ordered before any later rt_render (see rt_film_create)
    return RT_OK;
}
rt_status rt_film_read(rt_film* f, double* host) {
    if (!f || !host) { set_error("rt_film_read: null argument"); return RT_ERR_INVALID; }
    CK(cudaMemcpy(host, f->d, (size_t)f->w * f->h * 4 * sizeof(double), cudaMemcpyDeviceToHost)); return RT_OK;
}
rt_status rt_film_resolve_rgb8(rt_film* f, uint8_t* host_rgb) {
    if (!f || !host_rgb) { set_error("rt_film_resolve_rgb8: null argument"); return RT_ERR_INVALID; }
    uint64_t n = (uint64_t)f->w * f->h; uint8_t* d = nullptr;
    CK(cudaMalloc((void**)&d, n * 3));
    k_resolve<<<(unsigned)std::min<uint64_t>((n + 255) / 256, 148 * 16), 256>>>(f->d, n, d);
    cudaError_t e = cudaMemcpy(host_rgb, d, n * 3, cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (e != cudaSuccess) { set_error(std::string("rt_film_resolve_rgb8: ") + cudaGetErrorString(e)); return RT_ERR_CUDA; }
    return RT_OK;
}

// ---- rt_film_resolve_rgb8_static: rt_film_resolve_rgb8 for callers that may not allocate (code under CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.  The 3-bytes-per-pixel staging
// buffer that rt_film_resolve_rgb8 cudaMallocs and frees on every call lives here in a file-scope device array of RT_RESOLVE_STATIC_MAX_PIXELS * 3
// bytes (12 MiB) that every GPU which loads the library holds until exit.  In host code `g_resolve_scratch` is only a stand-in for the symbol, not a
// device address: its address is looked up once per device, with that device current, through cudaGetSymbolAddress, and that pointer is what the
// kernel and the copy get.  The entry point calls no cudaMalloc / cudaFree / allocating library routine.  All calls on a device share the array and
// it keeps the last call's bytes; nothing has to start at zero because k_resolve writes every byte of [0, 3 w h) before the copy reads it.  Host
// threads of this process are serialised by a mutex; anything else that overlaps on the GPU with a call (another stream using the same array) races.
static __device__ __align__(256) uint8_t g_resolve_scratch[(size_t)RT_RESOLVE_STATIC_MAX_PIXELS * 3];
namespace {
constexpr int RESOLVE_MAX_DEVICES = 64;
uint8_t* g_resolve_ptr[RESOLVE_MAX_DEVICES] = {};      // device address of g_resolve_scratch, per device
std::mutex g_resolve_mutex;
}
rt_status rt_film_resolve_rgb8_static(rt_film* f, uint8_t* host_rgb) {
    if (!f || !host_rgb) { set_error("rt_film_resolve_rgb8_static: null argument"); return RT_ERR_INVALID; }
    const uint64_t n = (uint64_t)f->w * f->h;
    if (n > (uint64_t)RT_RESOLVE_STATIC_MAX_PIXELS) {
        set_error("rt_film_resolve_rgb8_static: film of " + std::to_string(f->w) + " x " + std::to_string(f->h) + " pixels exceeds the static scratch of " +
                  std::to_string((unsigned long long)RT_RESOLVE_STATIC_MAX_PIXELS) + " pixels (RT_RESOLVE_STATIC_MAX_PIXELS); use rt_film_resolve_rgb8");
        return RT_ERR_INVALID;
    }
    if (f->device < 0 || f->device >= RESOLVE_MAX_DEVICES) { set_error("rt_film_resolve_rgb8_static: device index out of range"); return RT_ERR_INVALID; }
    std::lock_guard<std::mutex> lock(g_resolve_mutex);
    int prev = 0;
    CK(cudaGetDevice(&prev));
    if (prev != f->device) CK(cudaSetDevice(f->device));
    rt_status st = RT_OK;
    if (!g_resolve_ptr[f->device]) {                    // set-up, once per device, with that device current
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_resolve_scratch);
        if (e != cudaSuccess || !p) { set_error(std::string("rt_film_resolve_rgb8_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e)); st = RT_ERR_CUDA; }
        else g_resolve_ptr[f->device] = (uint8_t*)p;
    }
    if (st == RT_OK) {
        uint8_t* d = g_resolve_ptr[f->device];
        k_resolve<<<(unsigned)std::min<uint64_t>((n + 255) / 256, 148 * 16), 256>>>(f->d, n, d);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaMemcpy(host_rgb, d, n * 3, cudaMemcpyDeviceToHost);      // same (legacy default) stream as the kernel: ordered after it, and synchronous
        if (e != cudaSuccess) { set_error(std::string("rt_film_resolve_rgb8_static: ") + cudaGetErrorString(e)); st = RT_ERR_CUDA; }
    }
    if (prev != f->device) cudaSetDevice(prev);
    return st;
}

rt_status rt_render(rt_scene* s, const rt_camera* cam, const rt_integrator_desc* integ, const rt_film_desc* fd, rt_film* film) {
    if (!s || !cam || !integ || !fd || !film) { set_error("rt_render: null argument"); return RT_ERR_INVALID; }
    if (!s->committed) { set_error("rt_render: scene not committed"); return RT_ERR_STATE; }
    if (integ->kind > RT_INTEGRATOR_BASIC) { set_error("rt_render: unknown integrator kind"); return RT_ERR_INVALID; }
    if (integ->kind != RT_INTEGRATOR_PATH && (integ->flags & RT_SHADE_F32)) { set_error("rt_render: RT_SHADE_F32 is implemented for IntType::Path only"); return RT_ERR_INVALID; }
    if (integ->kind != RT_INTEGRATOR_PATH && integ->max_depth == 0) { set_error("rt_render: Whitted / Direct / Basic need max_depth >= 1"); return RT_ERR_INVALID; }
    if (fd->width != film->w || fd->height != film->h) { set_error("rt_render: film size mismatch"); return RT_ERR_INVALID; }
    if (fd->x1 > fd->width || fd->y1 > fd->height || fd->x0 > fd->x1 || fd->y0 > fd->y1 || fd->sample_end < fd->sample_begin) { set_error("rt_render: bad window / sample range"); return RT_ERR_INVALID; }
    if (integ->max_depth > 0xfff0u) { set_error("rt_render: max_depth too large"); return RT_ERR_INVALID; }
    if (film->device != s->device) { set_error("rt_render: the film lives on another device than the scene"); return RT_ERR_INVALID; }
    CK(cudaSetDevice(s->device));
    Wavefront* wp = get_wavefront(*s);
    if (!wp) return RT_ERR_CUDA;
    Wavefront& w = *wp;
    s->cancel.store(0);
    RenderParams rp{};
    rp.cam = *cam; rp.width = fd->width; rp.height = fd->height; rp.x0 = fd->x0; rp.y0 = fd->y0; rp.ww = fd->x1 - fd->x0; rp.wh = fd->y1 - fd->y0;
    rp.sample_begin = fd->sample_begin; rp.max_depth = integ->max_depth; rp.invisible_light = integ->invisible_light;
    rp.k0 = (uint32_t)fd->seed; rp.k1 = (uint32_t)(fd->seed >> 32); rp.W = w.W; rp.kind = integ->kind;
    uint32_t cap = (integ->kind == RT_INTEGRATOR_PATH || integ->kind == RT_INTEGRATOR_BASIC) ? w.W : w.W / std::max<uint32_t>(2u, (uint32_t)s->lights.size());
    // test hook (tests/test_gpu_parity.py::test_wavefront_queue_overflow_is_contained): admit more camera paths than the branching integrators'
    // queues can take, so that the overflow path — clamped counters, clean RT_ERR_STATE — is exercised; never set in production
    if (const char* e = getenv("RT_TEST_FORCE_CAP")) { const long v = atol(e); if (v > 0) cap = (uint32_t)std::min<long>(v, (long)w.W); }
    Ctrl init{}; init.total_samples = (unsigned long long)rp.ww * rp.wh * (fd->sample_end - fd->sample_begin);
    s->counters = rt_counters{};
    if (init.total_samples == 0) return RT_OK;
    cudaStream_t st = w.stream;
    CK(cudaMemcpyAsync(w.ctrl, &init, sizeof(Ctrl), cudaMemcpyHostToDevice, st));
#if RT_SORT_HITS
    CK(cudaMemsetAsync(w.sort.hist, 0, w.sort.n_keys * sizeof(uint32_t), st));      // a cancelled render may have left counts behind
#endif
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device);
    int g128 = (int)std::min<uint64_t>((w.W + 127) / 128, (uint64_t)sms * 32), gtrace = sms * TRACE_GRID_BLOCKS, g256 = (int)std::min<uint64_t>((w.W + 255) / 256, (uint64_t)sms * 8);
    cudaEvent_t e_begin = ev(w, 0), e_end = ev(w, 1);
    size_t n_ev = 2;
    CK(cudaEventRecord(e_begin, st));
    int cur = 0; uint64_t iter = 0; bool done = false; rt_status result = RT_OK;
    const bool f32 = (integ->flags & RT_SHADE_F32) != 0 && !s->d.has_perlin && !s->d.has_delta && !s->d.has_disney;   // Texture::Perlin and the delta lights are evaluated by the f64 kernels only (shade.cuh:tex_value_at, delta_sample_li)
    const bool trace = getenv("RT_TRACE_ITER") != nullptr;     // debugging aid: per-iteration sizes and kernel times on stderr
    if (trace && !w.dbg[0]) { cudaEventCreate(&w.dbg[0]); cudaEventCreate(&w.dbg[1]); }
    while (!done) {
        for (int k = 0; k < (trace ? 1 : 4); k++, iter++) {
            k_prepare<<<1, 1, 0, st>>>(w.ctrl, cap, w.W);
            if (k == 0) cudaMemcpyAsync(w.h_ctrl, w.ctrl, sizeof(Ctrl), cudaMemcpyDeviceToHost, st);
            cudaEventRecord(ev(w, n_ev + 0), st);
            k_generate<<<g256, 256, 0, st>>>(w.ctrl, rp, w.set[cur], film->d);
            cudaEventRecord(ev(w, n_ev + 1), st);
            k_extend<<<gtrace, 128, 0, st>>>(w.ctrl, s->d, w.set[cur], w.bin_idx, w.W, w.sort);
            cudaEventRecord(ev(w, n_ev + 2), st);
#if RT_SORT_HITS
            k_sort_scan<<<1, 1024, 0, st>>>(w.ctrl, w.sort);             // timed with the shade kernels, their consumer
            if (trace) cudaEventRecord(w.dbg[0], st);
            k_sort_scatter<<<g256, 256, 0, st>>>(w.ctrl, w.sort, w.bin_idx);
            if (trace) cudaEventRecord(w.dbg[1], st);
#endif
{
                ShadeLaunch L{w.ctrl, &s->d, &rp, w.set[cur], w.set[cur ^ 1], w.bin_idx, w.shadow, w.probe, film->d, s->bin_mask, g128, st};
                if (f32) launch_shade_f32(L, s->d.has_infinite != 0);
                else if (s->d.has_infinite) launch_shade_bins<double, true>(L);
                else launch_shade_bins<double, false>(L);
            }
            if (s->d.has_infinite) k_shade_miss<<<g128, 128, 0, st>>>(w.ctrl, s->d, rp, w.set[cur], w.bin_idx, film->d);
            cudaEventRecord(ev(w, n_ev + 3), st);
            k_connect<<<gtrace, 128, 0, st>>>(w.ctrl, s->d, w.shadow, w.probe, film->d, w.W);
            cudaEventRecord(ev(w, n_ev + 4), st);
            n_ev += 5;
            cur ^= 1;
            { cudaError_t le = cudaPeekAtLastError(); if (le != cudaSuccess) { cudaGetLastError(); set_error(std::string("rt_render: kernel launch failed: ") + cudaGetErrorString(le)); return RT_ERR_CUDA; } }
        }
        CK(cudaStreamSynchronize(st));
        if (trace) {
            float a = 0, b = 0, cc = 0, d = 0; size_t i = n_ev - 5;
            cudaEventElapsedTime(&a, w.events[i], w.events[i + 1]); cudaEventElapsedTime(&b, w.events[i + 1], w.events[i + 2]);
            cudaEventElapsedTime(&cc, w.events[i + 2], w.events[i + 3]); cudaEventElapsedTime(&d, w.events[i + 3], w.events[i + 4]);
            Ctrl after; cudaMemcpy(&after, w.ctrl, sizeof(Ctrl), cudaMemcpyDeviceToHost);
            fprintf(stderr, "iter %3llu n_tot %8u (new %8u) next %8u shadow %8u probe %7u | gen %.3f extend %.3f shade %.3f connect %.3f ms | bins", (unsigned long long)iter - 1,
                    w.h_ctrl->n_tot, w.h_ctrl->n_new, after.n_next, after.n_shadow, after.n_probe, a, b, cc, d);
            for (int q = 0; q < N_BINS; q++) if (after.bin_count[q]) fprintf(stderr, " %d:%u", q, after.bin_count[q]);
#if RT_SORT_HITS
            { float s1 = 0, s2 = 0; cudaEventElapsedTime(&s1, w.events[i + 2], w.dbg[0]); cudaEventElapsedTime(&s2, w.dbg[0], w.dbg[1]); fprintf(stderr, " | scan %.3f scatter %.3f", s1, s2); }
#endif
            fprintf(stderr, "\n");
        }
        if (w.h_ctrl->n_tot == 0) done = true;       // snapshot taken right after a prepare: nothing in flight, nothing left to generate
        if (w.h_ctrl->overflow) done = true;          // a queue append was dropped: the film is already incomplete, stop here (RT_ERR_STATE below)
        if (s->cancel.load()) { result = RT_ERR_CANCELLED; set_error("rt_render: cancelled"); done = true; }
        if (n_ev > 200000) n_ev = 2;                  // very long renders: stop growing the event pool (timing breakdown becomes partial)
    }
    k_prepare<<<1, 1, 0, st>>>(w.ctrl, cap, w.W);         // account the last iteration
    CK(cudaMemcpyAsync(w.h_ctrl, w.ctrl, sizeof(Ctrl), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    // camera samples actually generated: all of them unless cancelled (the closing k_prepare claims n_new more that never run)
    if (w.h_ctrl->overflow) { set_error("rt_render: wavefront queue overflow (both specular branches taken at many vertices); raise rt_set_wavefront_size"); result = RT_ERR_STATE; }
    const unsigned long long generated = w.h_ctrl->c_paths - w.h_ctrl->n_new;
    k_film_count<<<g256, 256, 0, st>>>(rp, generated, film->d);
    CK(cudaEventRecord(e_end, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    rt_counters& c = s->counters;
    c.paths = generated; c.extend_rays = w.h_ctrl->c_extend; c.shadow_rays = w.h_ctrl->c_shadow; c.probe_rays = w.h_ctrl->c_probe; c.iterations = iter;
    float ms = 0; cudaEventElapsedTime(&ms, e_begin, e_end); c.render_ms = ms;
    for (size_t i = 2; i + 4 < n_ev; i += 5) {
        float a = 0, b = 0, cc = 0, d = 0;
        cudaEventElapsedTime(&a, w.events[i], w.events[i + 1]); cudaEventElapsedTime(&b, w.events[i + 1], w.events[i + 2]);
        cudaEventElapsedTime(&cc, w.events[i + 2], w.events[i + 3]); cudaEventElapsedTime(&d, w.events[i + 3], w.events[i + 4]);
        c.shade_ms += a + cc; c.extend_ms += b; c.connect_ms += d;
    }
    return result;
}

}  // extern "C"
