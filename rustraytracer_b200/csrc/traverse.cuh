// Software BVH traversal for B200 (no RT cores): one thread per ray, while-while loop, short
// per-thread stack, node and leaf-slot fetches as 16-byte vector loads through the read-only path.
//
// Replaces BvhNode::intersects (reference hittable.rs:591-634), which visits both children with an
// unchanged [tmin,tmax] and builds a full HitRecord at every primitive hit.  Here:
//   * inner nodes are culled with a conservative f32 slab test (FMA form, per-ray padding that
//     covers the f32 rounding of the ray origin) — it may only ever over-accept;
//   * every surviving candidate primitive is decided by the reference's own f64 arithmetic
//     (geom.cuh) including the f64 leaf-box slab test, so hit(prim) ≡ leafbox_f64 ∧ primtest_f64
//     exactly as in the reference; the interval given to those tests is
//     the ray's ORIGINAL [tmin,tmax] — only the f32 culling interval shrinks;
//   * the closest hit is chosen on the exact f64 t.  On an exact tie the reference keeps the right
//     subtree's record (`if l.t < r.t {left} else {right}`), which depends on its randomly built
//     tree; here the smaller primitive index wins, and the oracle reports the tie set.
#pragma once
#include "geom.cuh"

namespace rt {

struct RayF {                       // f32 culling constants
    float idx, idy, idz;            // 1/d
    float nx, ny, nz;               // o/d + pad  (subtracted from near-plane products)
    float fx, fy, fz;               // o/d - pad  (subtracted from far-plane products)
    int sx, sy, sz;                 // direction sign (1 = negative)
};
__device__ __forceinline__ float safe_inv(float d) {
    const float tiny = 1e-30f;
    if (fabsf(d) < tiny) d = copysignf(tiny, d);
    return 1.0f / d;
}
__device__ __forceinline__ RayF make_rayf(D3 o, D3 d) {
    RayF r;
    float ox = (float)o.x, oy = (float)o.y, oz = (float)o.z;
    r.idx = safe_inv((float)d.x); r.idy = safe_inv((float)d.y); r.idz = safe_inv((float)d.z);
    // absolute pad: f32 rounding of the origin (2^-24 relative) and of o*inv_d, with margin
    float pad = (fmaxf(fabsf(ox), fmaxf(fabsf(oy), fabsf(oz))) + 1e-30f) * 4.8e-7f;
    float ex = pad * fabsf(r.idx), ey = pad * fabsf(r.idy), ez = pad * fabsf(r.idz);
    float qx = ox * r.idx, qy = oy * r.idy, qz = oz * r.idz;
    // relative slack on the products themselves
    ex += fabsf(qx) * 2.4e-7f; ey += fabsf(qy) * 2.4e-7f; ez += fabsf(qz) * 2.4e-7f;
    r.nx = qx + ex; r.ny = qy + ey; r.nz = qz + ez;
    r.fx = qx - ex; r.fy = qy - ey; r.fz = qz - ez;
    r.sx = d.x < 0.; r.sy = d.y < 0.; r.sz = d.z < 0.;
    return r;
}

// conservative slab test of one child box; returns entry distance in tn
__device__ __forceinline__ bool box_test(const RayF& r, float minx, float maxx, float miny, float maxy, float minz, float maxz,
                                         float tmin, float tmax, float& tn) {
    float nx = r.sx ? maxx : minx, fx = r.sx ? minx : maxx;
    float ny = r.sy ? maxy : miny, fy = r.sy ? miny : maxy;
    float nz = r.sz ? maxz : minz, fz = r.sz ? minz : maxz;
    float t0x = __fmaf_rn(nx, r.idx, -r.nx), t1x = __fmaf_rn(fx, r.idx, -r.fx);
    float t0y = __fmaf_rn(ny, r.idy, -r.ny), t1y = __fmaf_rn(fy, r.idy, -r.fy);
    float t0z = __fmaf_rn(nz, r.idz, -r.nz), t1z = __fmaf_rn(fz, r.idz, -r.fz);
    float t0 = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, tmin));
    float t1 = fminf(fminf(t1x, t1y), fminf(t1z, tmax));
    // relative slack for the rounding of inv_d (2 ulp) on both ends
    t0 = t0 - fabsf(t0) * 4.8e-7f; t1 = t1 + fabsf(t1) * 4.8e-7f;
    tn = t0;
    return t0 <= t1;
}

__device__ __forceinline__ float f32_up(double x) { return __double2float_ru(x); }
__device__ __forceinline__ float f32_down(double x) { return __double2float_rd(x); }


// ------------------------------------------------------------------------------------------------------------
// Conservative f32 prefilter in front of the exact f64 triangle test (SURVEY.md §7 "hard parts").
//
// Every candidate slot used to go straight to the reference's f64 watertight test (geom.cuh:tri_test): five 16-byte loads and ~60 f64
// operations, of which nine in ten end in "mixed edge-function signs -> miss".  The prefilter evaluates the SAME three edge functions
// in f32 on a 48-byte f32 copy of the triangle (DScene::tri32, leaf order) together with a rigorous bound B on |e_f32 - e_exact|, and
// lets a slot skip the f64 test only when the f64 test is certain to reject it — or certain to return a hit that cannot matter:
//   (1) min(e0,e1,e2) < -B  and  max(e0,e1,e2) > +B   =>  the f64 edge functions have mixed signs (hittable.rs:336)          -> miss
//   (2) every sheared z_k * (1/d[kz]) lies beyond the culling interval's far end, or all lie behind the origin: an accepted t is a
//       convex combination of those three values (barycentric weights e_i/det >= 0), so t > best-so-far / t_limit, or t < 0.
// Error budget, u = 2^-24 (f32 unit roundoff), coordinates |p| <= pmax (stored per triangle, rounded up), |o| <= omax:
//   a   = fl(fl(p) - fl(o))                         |a - a*|   <= u(|p| + |o| + |a|)            <= D0 := 2u(pmax + omax)
//   x   = fma(Sx, az, ax), |Sx*| <= 1 (kz = argmax|d|), Sx = fl(Sx*)
//                                                   |x - x*|   <= 2 D0 + u|az| + u|x| (+ O(u D0))  <= E := 2 D0 + u(zm + 2 xm)
//   e0  = fma(x1, y2, -fl(y1 x2))                   |e0 - e0*| <= 4 xm E + 2 E^2 + 3u xm^2     =: B0      (xm = max|x_k|,|y_k|, zm = max|az_k|)
//   z_k = fl(az_k Sz), Sz = fl(1/d[kz])             |z - z*|   <= |Sz| D0 + 2u|z|              =: Ez0
// and the tests use B = 2 B0, Ez = 2 Ez0.  The f64 test's own rounding (2^-53 of the same magnitudes) disappears in that factor.  Any
// NaN / inf makes a comparison false, i.e. "go on to the exact test".  RT_PREFILTER_CHECK builds count — on the device, over every
// slot test of a run — the cases where the prefilter would have skipped a hit the traversal then used (rt_debug_prefilter_violations()).
#ifndef RT_PREFILTER
#define RT_PREFILTER 0
#endif
#ifndef RT_PREFILTER_CHECK
#define RT_PREFILTER_CHECK 0
#endif
#ifndef RT_LEAF_PREFETCH
#define RT_LEAF_PREFETCH 0
#endif
struct RayP { float ox, oy, oz, shx, shy, shz, eo; int kz; };     // f32 origin, shear (sx, sy, 1/d[kz]) and 2u * max|o|
constexpr float F32_U = 5.9604645e-8f;                            // 2^-24
__device__ __forceinline__ RayP make_rayp(const RayX& rx) {
    RayP r;
    r.ox = (float)rx.o.x; r.oy = (float)rx.o.y; r.oz = (float)rx.o.z;
    r.shx = (float)rx.sx; r.shy = (float)rx.sy; r.shz = (float)comp(rx.inv, rx.kz);
    r.eo = 2.f * F32_U * 1.0000002f * fmaxf(fabsf(r.ox), fmaxf(fabsf(r.oy), fabsf(r.oz)));
    r.kz = rx.kz;
    return r;
}
// true = run the exact test; false = the exact test cannot produce a hit that matters for a culling interval ending at `ftmax`
__device__ __forceinline__ bool tri_prefilter(const RayP& r, const float4 w0, const float4 w1, const float4 w2, float ftmax) {
    const float d0 = __fmaf_rn(2.f * F32_U, w2.w, r.eo);                       // D0 = 2u(pmax + omax)
    float ax0 = w0.x - r.ox, ay0 = w0.y - r.oy, az0 = w0.z - r.oz;
    float ax1 = w0.w - r.ox, ay1 = w1.x - r.oy, az1 = w1.y - r.oz;
    float ax2 = w1.z - r.ox, ay2 = w1.w - r.oy, az2 = w2.x - r.oz;
    // permute to (kx, ky, kz) = ((kz+1)%3, (kz+2)%3, kz) as geom.cuh:permute does
    if (r.kz == 0) { float t; t = ax0; ax0 = ay0; ay0 = az0; az0 = t; t = ax1; ax1 = ay1; ay1 = az1; az1 = t; t = ax2; ax2 = ay2; ay2 = az2; az2 = t; }
    else if (r.kz == 1) { float t; t = ay0; ay0 = ax0; ax0 = az0; az0 = t; t = ay1; ay1 = ax1; ax1 = az1; az1 = t; t = ay2; ay2 = ax2; ax2 = az2; az2 = t; }
    const float x0 = __fmaf_rn(r.shx, az0, ax0), y0 = __fmaf_rn(r.shy, az0, ay0);
    const float x1 = __fmaf_rn(r.shx, az1, ax1), y1 = __fmaf_rn(r.shy, az1, ay1);
    const float x2 = __fmaf_rn(r.shx, az2, ax2), y2 = __fmaf_rn(r.shy, az2, ay2);
    const float e0 = __fmaf_rn(x1, y2, -(y1 * x2)), e1 = __fmaf_rn(x2, y0, -(y2 * x0)), e2 = __fmaf_rn(x0, y1, -(y0 * x1));
    const float zm = fmaxf(fabsf(az0), fmaxf(fabsf(az1), fabsf(az2)));
    const float xm = fmaxf(fmaxf(fmaxf(fabsf(x0), fabsf(y0)), fmaxf(fabsf(x1), fabsf(y1))), fmaxf(fabsf(x2), fabsf(y2)));
    const float E = __fmaf_rn(F32_U, zm + 2.f * xm, 2.f * d0);
    const float B = 2.f * (4.f * xm * E + 2.f * E * E + 3.f * F32_U * xm * xm);
    const float emin = fminf(e0, fminf(e1, e2)), emax = fmaxf(e0, fmaxf(e1, e2));
    if (emin < -B && emax > B) return false;
    const float z0 = az0 * r.shz, z1 = az1 * r.shz, z2 = az2 * r.shz;
    const float zlo = fminf(z0, fminf(z1, z2)), zhi = fmaxf(z0, fmaxf(z1, z2));
    const float Ez = 2.f * (fabsf(r.shz) * d0 + 2.f * F32_U * fmaxf(fabsf(zlo), fabsf(zhi)));
    if (zlo - Ez > ftmax || zhi + Ez < 0.f) return false;
    return true;
}
#if RT_PREFILTER_CHECK
static __device__ unsigned long long g_prefilter_stats[4];      // [0] triangle slot tests, [1] skipped by the prefilter, [2] exact hits, [3] VIOLATIONS: skipped, yet the exact hit would have been used
#endif
// host side: adds this translation unit's counters (every TU that includes this header has its own copy) to out[4]
static inline void prefilter_stats_add(unsigned long long out[4]) {
#if RT_PREFILTER_CHECK
    unsigned long long h[4] = {0, 0, 0, 0};
    if (cudaMemcpyFromSymbol(h, g_prefilter_stats, sizeof(h)) == cudaSuccess) for (int k = 0; k < 4; k++) out[k] += h[k];
#else
    (void)out;
#endif
}

struct Hit { double t; uint32_t prim; };

constexpr int STACK = TRAVERSAL_STACK;
constexpr int32_t SENTINEL = 0x7fffffff;

#if RT_BVH_WIDTH == 4
// One BVH4 node: the conservative slab test of box_test() on all four children.  Returns the hit mask; key[k] packs child
// k's entry distance with its slot number in the two low mantissa bits (non-negative floats order like ints), +inf for a
// miss — so four min/max compare-exchanges sort the children front to back without moving (distance, index) pairs around.
__device__ __forceinline__ unsigned node4_test(const float4* __restrict__ np, const RayF& r, float tmin, float tmax, int32_t key[4], int4& ch) {
    // near / far planes picked by the load address (rows are min-x, max-x, min-y, ...; s* = 1 for a negative direction) instead of 24 selects
    const float4 nx = __ldg(np + r.sx), fx = __ldg(np + (r.sx ^ 1)), ny = __ldg(np + 2 + r.sy), fy = __ldg(np + (3 - r.sy)), nz = __ldg(np + 4 + r.sz), fz = __ldg(np + (5 - r.sz));
    ch = __ldg(reinterpret_cast<const int4*>(np + 6));
    unsigned mask = 0;
#define RT_CHILD(c, k)                                                                                                              \
    {                                                                                                                               \
        float t0 = fmaxf(fmaxf(__fmaf_rn(nx.c, r.idx, -r.nx), __fmaf_rn(ny.c, r.idy, -r.ny)), fmaxf(__fmaf_rn(nz.c, r.idz, -r.nz), tmin)); \
        float t1 = fminf(fminf(__fmaf_rn(fx.c, r.idx, -r.fx), __fmaf_rn(fy.c, r.idy, -r.fy)), fminf(__fmaf_rn(fz.c, r.idz, -r.fz), tmax)); \
        t0 = __fmaf_rn(-fabsf(t0), 4.8e-7f, t0); t1 = __fmaf_rn(fabsf(t1), 4.8e-7f, t1);   /* one FFMA each: the library is built with -fmad=false */ \
        const bool h = t0 <= t1;                                                                                                    \
        key[k] = h ? ((__float_as_int(t0) & ~3) | k) : (0x7f800000 | k);                                                            \
        mask |= h ? (1u << k) : 0u;                                                                                                 \
    }
    RT_CHILD(x, 0) RT_CHILD(y, 1) RT_CHILD(z, 2) RT_CHILD(w, 3)
#undef RT_CHILD
    return mask;
}
// child word of the slot number packed in a key's two low bits — as bit masks, not as `k == 0 ? ... : ...`: nvcc turned the nested conditional into
// three divergent branches with a reconvergence barrier (13 instructions, the lanes of a warp hold different slot numbers and walked all four arms one
// after the other), four times per node step (r02 SASS: 52 of the step's ~190 instructions, 14 of them branches).  Seven instructions, no branch.
__device__ __forceinline__ int32_t child_of(const int4& ch, int32_t key) {
    const int32_t m0 = (key << 31) >> 31, m1 = (key << 30) >> 31;        // all ones where bit 0 / bit 1 of the slot number is set
    const int32_t lo = (ch.x & ~m0) | (ch.y & m0), hi = (ch.z & ~m0) | (ch.w & m0);
    return (lo & ~m1) | (hi & m1);
}
__device__ __forceinline__ void sort4(int32_t k[4]) {
#define RT_CSWAP(a, b) { int32_t lo = min(k[a], k[b]), hi = max(k[a], k[b]); k[a] = lo; k[b] = hi; }
    RT_CSWAP(0, 1) RT_CSWAP(2, 3) RT_CSWAP(0, 2) RT_CSWAP(1, 3) RT_CSWAP(1, 2)
#undef RT_CSWAP
}
#ifndef RT_PREFETCH_PUSHED
#define RT_PREFETCH_PUSHED 0      // measured r02 (pass AG): extend 29.5 -> 30.2 ms, soup casts -3...4 %: the extra L1 fills cost more than the early fetch saves
#endif
// closest-hit descent: nearest hit child next, the others pushed far to near
__device__ __forceinline__ int32_t node4_closest(const DScene& sc, int32_t cur, const RayF& rf, float ftmin, float ftmax, int32_t* stack, int& sp) {
    int32_t key[4]; int4 ch;
    const unsigned mask = node4_test(reinterpret_cast<const float4*>(sc.nodes) + 8 * (size_t)cur, rf, ftmin, ftmax, key, ch);
    const int nh = __popc(mask);
    if (nh == 0) return stack[--sp];
    sort4(key);
    // three stores without a branch: a push that does not happen (fewer hit children; a full stack, which the commit-time bound rules out) goes to
    // the spare entry behind the stack instead of around a divergent `if` (r02 SASS: each conditional push was a branch + reconvergence barrier)
    const bool v3 = nh > 3 && sp < STACK; stack[v3 ? sp : STACK] = child_of(ch, key[3]); sp += v3;
    const bool v2 = nh > 2 && sp < STACK; stack[v2 ? sp : STACK] = child_of(ch, key[2]); sp += v2;
    const int32_t c1 = child_of(ch, key[1]);
    const bool v1 = nh > 1 && sp < STACK; stack[v1 ? sp : STACK] = c1; sp += v1;
#if RT_PREFETCH_PUSHED
    // the entry now on top of the stack is the next one popped unless the near child's subtree ends the ray: start its fetch now
    if (v1) {
        const char* a = c1 >= 0 ? reinterpret_cast<const char*>(sc.nodes) + 128 * (size_t)c1 : reinterpret_cast<const char*>(sc.slots) + sizeof(DSlot) * (size_t)((uint32_t)~c1 >> 2);
        asm volatile("prefetch.global.L1 [%0];" ::"l"(a));
    }
#endif
    return child_of(ch, key[0]);
}
// any-hit descent: order does not matter
__device__ __forceinline__ int32_t node4_any(const DScene& sc, int32_t cur, const RayF& rf, float ftmin, float ftmax, int32_t* stack, int& sp) {
    int32_t key[4]; int4 ch;
    const unsigned mask = node4_test(reinterpret_cast<const float4*>(sc.nodes) + 8 * (size_t)cur, rf, ftmin, ftmax, key, ch);
    if (mask == 0u) return stack[--sp];
    const int32_t c[4] = {ch.x, ch.y, ch.z, ch.w};
    int32_t next = SENTINEL; bool have = false;
#pragma unroll
    for (int k = 0; k < 4; k++) if ((mask >> k) & 1u) { if (have) { if (sp < STACK) stack[sp++] = c[k]; } else { next = c[k]; have = true; } }
    return next;
}
#endif


// ---- per-ray data of the leaf tests
// RT_PREFILTER = 0 (default): the exact test's f64 constants (RayX: origin, direction, shear, 1/d) live in the traversal state, every candidate
//                 slot goes straight to the reference's f64 test — round 1's scheme, and still the fastest on B200 (see below).
// RT_PREFILTER = 1: the f32 prefilter's constants (RayP) live in the state; the f64 ray is kept too (RT_STATE_RAY = 1) or read back from the
//                 queue the ray came from (Job::ray) for the slots that reach the exact test.
// MEASURED (r02, dragon 1024^2 64 spp, extend / connect ms; profiles/prefilter_ab_r02.json): no prefilter 31.6 / 20.2; prefilter with the ray read
// back + exact test as a call 36.9 / 24.3; ray in the state 36.0 / 24.0; ... + exact test inlined 34.9 / 23.0; + L1 prefetch of the leaf's records: no
// change.  Why it loses here: B200's FP64 pipe is half-rate, not vestigial, so the f64 test's early-out path (~40 instructions up to the mixed-sign
// reject) costs about what the f32 test with its error bound does (~60), the survivors pay both plus a second dependent fetch, and this scene runs
// only ~0.5 triangle tests per ray (most rays meet the floor or leave).  The prefilter itself is sound: 0 violations in 214 M audited slot tests
// (profiles/prefilter_audit_r02.json) with 80-88 % of the triangle slots skipped.  It stays in the tree as a build option for FP64-poor parts.
#ifndef RT_STATE_RAY
#define RT_STATE_RAY 1
#endif
#if RT_PREFILTER
struct RayL {
    RayP rp;
#if RT_STATE_RAY
    D3 o, d;
#endif
};
__device__ __forceinline__ RayL make_rayl(const RayX& rx) {
    RayL r; r.rp = make_rayp(rx);
#if RT_STATE_RAY
    r.o = rx.o; r.d = rx.d;
#endif
    return r;
}
#else
struct RayL { RayX rx; };
__device__ __forceinline__ RayL make_rayl(const RayX& rx) { RayL r; r.rx = rx; return r; }
#endif

// The exact test as a CALL (RT_EXACT_CALL, prefilter builds only): as a function of its own it keeps its ~60 f64 temporaries out of the register
// allocation of the traversal loop.
#ifndef RT_EXACT_CALL
#define RT_EXACT_CALL 0
#endif
#if RT_EXACT_CALL
#define RT_EXACT_FN static __device__ __noinline__
#else
#define RT_EXACT_FN __device__ __forceinline__
#endif
RT_EXACT_FN bool tri_exact(const DSlot* __restrict__ slot, D3 o, D3 d, double tmin, double tmax, double& t) {
    const double2* q = reinterpret_cast<const double2*>(slot);
    const double2 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3), q4 = __ldg(q + 4);
    const D3 p0 = mk3(q0.x, q0.y, q1.x), p1 = mk3(q1.y, q2.x, q2.y), p2 = mk3(q3.x, q3.y, q4.x);
    const RayX rx = make_rayx(o, d);
    double b0, b1, b2;
    if (!tri_test(rx, p0, p1, p2, tmax, t, b0, b1, b2)) return false;
    return tri_leaf_box_inv(p0, p1, p2, rx.o, rx.inv, tmin, tmax);
}
// Candidate primitive in leaf slot `s`: (the f32 prefilter, then) the reference's exact accept test.  Returns true with t.
// `ftmax` is the far end of the caller's CULLING interval (>= best-so-far t / the shadow ray's t_limit): a hit beyond it cannot matter.
// `fetch(o, d)` hands out the ray's f64 origin and direction (prefilter builds without RT_STATE_RAY: read back from the ray's queue).
// skipped (RT_PREFILTER_CHECK only): the prefilter said "skip" — the exact test was run all the same so that the caller can audit the decision.
template <class Fetch>
__device__ __forceinline__ bool slot_test(const DScene& sc, uint32_t s, const RayL& rl, const Fetch& fetch, double tmin, double tmax, float ftmax,
                                          double& t, uint32_t& prim, bool* skipped = nullptr) {
#if RT_PREFILTER
    const float4* w = reinterpret_cast<const float4*>(sc.tri32) + 3 * (size_t)s;
    const float4 w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
    prim = __float_as_uint(w2.y);
    const uint32_t kind = __float_as_uint(w2.z);
    if (kind == 2) return false;     // never hit: a triangle the reference cannot hit (degenerate uv frame ∧ zero normal, hittable.rs:373-378), or a direct shape
    if (kind == 0) {
        const bool pass = tri_prefilter(rl.rp, w0, w1, w2, ftmax);
#if RT_PREFILTER_CHECK
        atomicAdd(&g_prefilter_stats[0], 1ull);
        if (!pass) atomicAdd(&g_prefilter_stats[1], 1ull);
        if (skipped) *skipped = !pass;
#else
        if (!pass) return false;
#endif
    }
    D3 o, d;
#if RT_STATE_RAY
    o = rl.o; d = rl.d;
#else
    fetch(o, d);
#endif
    if (kind == 0) {
        if (!tri_exact(sc.slots + s, o, d, tmin, tmax, t)) return false;
#if RT_PREFILTER_CHECK
        atomicAdd(&g_prefilter_stats[2], 1ull);
#endif
        return true;
    }
    DPrim pr = sc.prims[prim];
    return shape_test(sc, pr, o, d, tmin, tmax, true, t);
#else
    const RayX& rx = rl.rx;
    const double2* q = reinterpret_cast<const double2*>(sc.slots + s);
    double2 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3), q4 = __ldg(q + 4);
    uint2 tag = *reinterpret_cast<uint2*>(&q4.y);
    prim = tag.x;
    if (tag.y == 0) {
        D3 p0 = mk3(q0.x, q0.y, q1.x), p1 = mk3(q1.y, q2.x, q2.y), p2 = mk3(q3.x, q3.y, q4.x);
        double b0, b1, b2;
        if (!tri_test(rx, p0, p1, p2, tmax, t, b0, b1, b2)) return false;
        return tri_leaf_box_inv(p0, p1, p2, rx.o, rx.inv, tmin, tmax);
    }
    if (tag.y == 2) return false;    // never hit: a triangle the reference cannot hit (degenerate uv frame ∧ zero normal, hittable.rs:373-378), or a direct shape
    DPrim pr = sc.prims[prim];
    return shape_test(sc, pr, rx.o, rx.d, tmin, tmax, true, t, &rx.inv);
#endif
}
#if RT_PREFILTER_CHECK
#define RT_AUDIT(skipped) do { if (skipped) atomicAdd(&g_prefilter_stats[3], 1ull); } while (0)
#else
#define RT_AUDIT(skipped) do { } while (0)
#endif

// Direct shapes.  Every scene function of the reference (scenes.rs) sets a handful of parametric shapes — a 20,000-unit floor rect, a light rect,
// the Cornell walls — next to its meshes.  Inside the BVH they sit in leaves right under the root, so nearly every ray ran their exact f64 test
// (slab + rect: ~150 instructions with a division each) from a divergent leaf phase at ~6 of 32 lanes (ncu r02: 8 % of k_extend's warp instructions).
// When a scene has at most MAX_DIRECT such shapes next to at least one triangle, commit leaves them OUT of the tree (DScene::direct) and every ray
// tests them once, right where it is set up — all freshly fetched lanes of the warp together — BEFORE the walk: the closest-hit walk then starts with
// its culling interval already cut at the floor / wall hit, and a shadow ray blocked by one of them never walks at all.  The hit set is unchanged
// (hit(prim) = leafbox_f64 ∧ primtest_f64 does not depend on where a primitive is looked up), ties keep the smaller primitive index as everywhere.
__device__ __forceinline__ void direct_shapes(const DScene& sc, const RayX& rx, const RayF& rf, double tmin, double tmax, bool any, double t_limit, uint32_t skip,
                                              double& best_t, uint32_t& best_prim, bool& occluded, float& ftmax, float ftmin) {
    for (uint32_t k = 0; k < sc.n_direct; k++) {
        const uint32_t prim = sc.direct[k];
        if (any && prim == skip) continue;            // the shadow ray's own target
        // the same conservative f32 slab test the tree's nodes get, on the shape's own box: a shadow ray that leaves the floor towards the lamp
        // does not pay for the floor's f64 test
        float tn;
        if (!box_test(rf, sc.direct_box[k][0], sc.direct_box[k][3], sc.direct_box[k][1], sc.direct_box[k][4], sc.direct_box[k][2], sc.direct_box[k][5], ftmin, ftmax, tn)) continue;
        const DPrim pr = sc.prims[prim];
        double t;
        if (!shape_test(sc, pr, rx.o, rx.d, tmin, tmax, true, t, &rx.inv)) continue;
        if (any) {
            if (prim != skip && (t < t_limit || (t == t_limit && prim < skip))) { occluded = true; return; }
        } else if (best_prim == RT_NONE || t < best_t || (t == best_t && prim < best_prim)) {
            best_t = t; best_prim = prim;
            ftmax = fmaxf(fminf(ftmax, __double2float_ru(t)), ftmin);
        }
    }
}

// closest hit in the reference's sense over [tmin, tmax]
__device__ __forceinline__ Hit trace_closest(const DScene& sc, D3 o, D3 d, double tmin, double tmax) {
    Hit best; best.t = 0.; best.prim = RT_NONE;
    if (sc.n_nodes == 0 && sc.n_direct == 0) return best;
    RayF rf = make_rayf(o, d);
    const RayX rx0 = make_rayx(o, d);
    const RayL rl = make_rayl(rx0);
    float ftmin = f32_down(tmin); ftmin = ftmin - fabsf(ftmin) * 4.8e-7f;
    float ftmax = f32_up(fmin(tmax, 3.0e38));
    { bool occ = false; direct_shapes(sc, rx0, rf, tmin, tmax, false, 0., 0u, best.t, best.prim, occ, ftmax, ftmin); }
    int32_t stack[STACK + 1]; int sp = 0;          // [STACK]: the write-only slot of node4_closest's suppressed pushes
    stack[sp++] = SENTINEL;
    int32_t cur = sc.n_nodes ? 0 : SENTINEL;
#if RT_BVH_WIDTH != 4
    const float4* __restrict__ nodes = reinterpret_cast<const float4*>(sc.nodes);
#endif
    while (cur != SENTINEL) {
        while (cur >= 0 && cur != SENTINEL) {
#if RT_BVH_WIDTH == 4
            cur = node4_closest(sc, cur, rf, ftmin, ftmax, stack, sp);
#else
            const float4* np = nodes + 4 * (size_t)cur;
            float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2);
            int4 ch = __ldg(reinterpret_cast<const int4*>(np + 3));
            float tl, tr;
            bool hl = box_test(rf, n0.x, n0.y, n0.z, n0.w, n2.x, n2.y, ftmin, ftmax, tl);
            bool hr = box_test(rf, n1.x, n1.y, n1.z, n1.w, n2.z, n2.w, ftmin, ftmax, tr);
            if (hl && hr) {
                int32_t near = ch.x, far = ch.y;
                if (tr < tl) { near = ch.y; far = ch.x; }
                if (sp < STACK) stack[sp++] = far;
                cur = near;
            } else if (hl) cur = ch.x;
            else if (hr) cur = ch.y;
            else cur = stack[--sp];
#endif
        }
        if (cur == SENTINEL) break;
        // leaf
        uint32_t v = (uint32_t)~cur, first = v >> 2, count = (v & 3u) + 1u;
        for (uint32_t s = first; s < first + count; s++) {
            double t; uint32_t prim; bool skipped = false;
            if (slot_test(sc, s, rl, [&](D3& fo, D3& fd) { fo = o; fd = d; }, tmin, tmax, ftmax, t, prim, &skipped)) {
                if (best.prim == RT_NONE || t < best.t || (t == best.t && prim < best.prim)) {
                    RT_AUDIT(skipped);
                    best.t = t; best.prim = prim;
                    // Triangles ignore tmin (only t < 1e-4 is rejected, hittable.rs:360), so the best t can lie BELOW tmin; a closer
                    // primitive then still has a box that straddles tmin (its f64 leaf-box test must pass), which is why the
                    // culling interval never shrinks below [ftmin, ftmin]
                    ftmax = fmaxf(fminf(ftmax, f32_up(t)), ftmin);
                }
            }
        }
        cur = stack[--sp];
    }
    return best;
}

// Any-hit: is there a primitive other than `skip` whose reference-exact hit has t < t_limit
// (or t <= t_limit when `tie_blocks`)?  Used for Visibility::unoccluded (hittable.rs:25-39): the light
// is visible iff the closest hit is the light primitive itself.
__device__ __forceinline__ bool trace_any(const DScene& sc, D3 o, D3 d, double tmin, double tmax, double t_limit, uint32_t skip) {
    if (sc.n_nodes == 0 && sc.n_direct == 0) return false;
    RayF rf = make_rayf(o, d);
    const RayX rx0 = make_rayx(o, d);
    const RayL rl = make_rayl(rx0);
    float ftmin = f32_down(tmin); ftmin = ftmin - fabsf(ftmin) * 4.8e-7f;
    float ftmax = f32_up(fmin(fmin(tmax, t_limit), 3.0e38));
    { bool occ = false; double bt = 0.; uint32_t bp = RT_NONE; direct_shapes(sc, rx0, rf, tmin, tmax, true, t_limit, skip, bt, bp, occ, ftmax, ftmin); if (occ) return true; }
    int32_t stack[STACK + 1]; int sp = 0;          // [STACK]: the write-only slot of node4_closest's suppressed pushes
    stack[sp++] = SENTINEL;
    int32_t cur = sc.n_nodes ? 0 : SENTINEL;
#if RT_BVH_WIDTH != 4
    const float4* __restrict__ nodes = reinterpret_cast<const float4*>(sc.nodes);
#endif
    while (cur != SENTINEL) {
        while (cur >= 0 && cur != SENTINEL) {
#if RT_BVH_WIDTH == 4
            cur = node4_any(sc, cur, rf, ftmin, ftmax, stack, sp);
#else
            const float4* np = nodes + 4 * (size_t)cur;
            float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2);
            int4 ch = __ldg(reinterpret_cast<const int4*>(np + 3));
            float tl, tr;
            bool hl = box_test(rf, n0.x, n0.y, n0.z, n0.w, n2.x, n2.y, ftmin, ftmax, tl);
            bool hr = box_test(rf, n1.x, n1.y, n1.z, n1.w, n2.z, n2.w, ftmin, ftmax, tr);
            if (hl && hr) { if (sp < STACK) stack[sp++] = ch.y; cur = ch.x; }
            else if (hl) cur = ch.x;
            else if (hr) cur = ch.y;
            else cur = stack[--sp];
#endif
        }
        if (cur == SENTINEL) break;
        uint32_t v = (uint32_t)~cur, first = v >> 2, count = (v & 3u) + 1u;
        for (uint32_t s = first; s < first + count; s++) {
            double t; uint32_t prim; bool skipped = false;
            if (slot_test(sc, s, rl, [&](D3& fo, D3& fd) { fo = o; fd = d; }, tmin, tmax, ftmax, t, prim, &skipped) && prim != skip) {
                // the closest-hit rule keeps the smaller primitive index on an exact tie
                if (t < t_limit || (t == t_limit && prim < skip)) { RT_AUDIT(skipped); return true; }
            }
        }
        cur = stack[--sp];
    }
    return false;
}

// Visibility::unoccluded (hittable.rs:25-39) for a finite (area) light: true iff the closest hit of the
// shadow ray over (0, INF) is p1's primitive.
__device__ __forceinline__ bool unoccluded(const DScene& sc, D3 p0, D3 p1, uint32_t p1_prim, bool infinite_light) {
    D3 dir = p1 - p0;
    D3 org = p0 + dir * SMALL;
    if (infinite_light) {
        Hit h = trace_closest(sc, org, dir, 0., INF);
        return h.prim == RT_NONE;
    }
    if (p1_prim >= sc.n_prims) return false;
    // exact hit of the target primitive itself (leaf box included)
    DPrim pr = sc.prims[p1_prim];
    double t_l;
    if ((pr.info & 7u) == RT_PRIM_TRIANGLE) {
        Rec r;
        if (!prim_record(sc, p1_prim, org, dir, 0., INF, true, r)) return false;
        t_l = r.t;
    } else {
        if (!shape_test(sc, pr, org, dir, 0., INF, true, t_l)) return false;
    }
    return !trace_any(sc, org, dir, 0., INF, t_l, p1_prim);
}

// ------------------------------------------------------------------------------------------------------------
// Persistent traversal with dynamic ray fetch.
//
// A warp that walks 32 incoherent rays in lock-step idles most of its lanes: rays finish after very different
// numbers of steps (ncu on the first version of k_extend: 7.9 of 32 lanes active on average).  Here every lane owns a
// resumable traversal state; a warp keeps stepping its live rays and, as soon as fewer than REFILL lanes are live,
// the idle lanes fetch fresh rays from a global work counter (one atomic per warp per refill) — the scheme of
// Aila & Laine, "Understanding the efficiency of ray traversal on GPUs", restated for this BVH and the f64 leaf tests.
// A Job supplies: load(i, TState&) -> bool (false = nothing to trace, the job already finished ray i), finish(i, const TState&),
// ray(i, o, d) (the f64 ray of item i again, for the exact test) and tmin(state) / tmax(state).
// The ray's f64 interval [tmin, tmax] is not part of the state: it is a property of the job (extend: (SMALL, INF); connect: (0, INF) for shadow
// rays, (SMALL, INF) for probes; batch entry points: the caller's) and is asked for with job.tmin(state) / job.tmax(state) where the exact test needs it.
struct TState {
    RayL rl; RayF rf;
    float ftmin, ftmax;
    double t_limit;
    uint32_t skip; bool any, occluded;
    Hit best;
    int32_t cur; int sp;
};
__device__ __forceinline__ void tstate_init(const DScene& sc, TState& s, int32_t* stack, D3 o, D3 d, double tmin, double tmax, bool any, double t_limit, uint32_t skip) {
    const RayX rx = make_rayx(o, d);
    s.rl = make_rayl(rx); s.rf = make_rayf(o, d);
    s.any = any; s.t_limit = t_limit; s.skip = skip; s.occluded = false;
    s.ftmin = f32_down(tmin); s.ftmin = s.ftmin - fabsf(s.ftmin) * 4.8e-7f;
    s.ftmax = f32_up(fmin(any ? fmin(tmax, t_limit) : tmax, 3.0e38));
    s.best.t = 0.; s.best.prim = RT_NONE;
    s.sp = 0; stack[s.sp++] = SENTINEL; s.cur = sc.n_nodes ? 0 : SENTINEL;
    if (sc.n_direct) {
        direct_shapes(sc, rx, s.rf, tmin, tmax, any, t_limit, skip, s.best.t, s.best.prim, s.occluded, s.ftmax, s.ftmin);
        if (s.occluded) s.cur = SENTINEL;          // a blocked shadow ray is finished before it walks
    }
}
// one inner-node step: test both children, descend into the nearer hit child, push the other, or pop
__device__ __forceinline__ void node_step(const DScene& sc, TState& s, int32_t* stack) {
#if RT_BVH_WIDTH == 4
    // front-to-back for any-hit rays too: one code path for the warp, and near occluders end a shadow ray sooner
    s.cur = node4_closest(sc, s.cur, s.rf, s.ftmin, s.ftmax, stack, s.sp);
#else
    const float4* np = reinterpret_cast<const float4*>(sc.nodes) + 4 * (size_t)s.cur;
    float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2);
    int4 ch = __ldg(reinterpret_cast<const int4*>(np + 3));
    float tl, tr;
    bool hl = box_test(s.rf, n0.x, n0.y, n0.z, n0.w, n2.x, n2.y, s.ftmin, s.ftmax, tl);
    bool hr = box_test(s.rf, n1.x, n1.y, n1.z, n1.w, n2.z, n2.w, s.ftmin, s.ftmax, tr);
    if (hl && hr) {
        int32_t near = ch.x, far = ch.y;
        if (tr < tl) { near = ch.y; far = ch.x; }
        if (s.sp < STACK) stack[s.sp++] = far;
        s.cur = near;
    } else if (hl) s.cur = ch.x;
    else if (hr) s.cur = ch.y;
    else s.cur = stack[--s.sp];
#endif
}
// one leaf: the reference-exact test of its 1..4 slots, then pop
template <class Fetch>
__device__ __forceinline__ void leaf_step(const DScene& sc, TState& s, int32_t* stack, double tmin, double tmax, const Fetch& fetch) {
    uint32_t v = (uint32_t)~s.cur, first = v >> 2, count = (v & 3u) + 1u;
#if RT_PREFILTER && RT_LEAF_PREFETCH
    // the leaf's 2..4 prefilter records are contiguous (48 B each): pull the rest into L1 while the first one is being tested, so that the
    // slot loop pays one L2 round trip per leaf instead of one per slot
    if (count > 1u) {
        const char* base = reinterpret_cast<const char*>(sc.tri32) + 48 * (size_t)first;
        asm volatile("prefetch.global.L1 [%0];" ::"l"(base + 64));
        if (count > 2u) asm volatile("prefetch.global.L1 [%0];" ::"l"(base + 128));
        if (count > 3u) asm volatile("prefetch.global.L1 [%0];" ::"l"(base + 176));
    }
#endif
    for (uint32_t k = first; k < first + count; k++) {
        double t; uint32_t prim; bool skipped = false;
        if (!slot_test(sc, k, s.rl, fetch, tmin, tmax, s.ftmax, t, prim, &skipped)) continue;
        if (s.any) {
            // the closest-hit rule keeps the smaller primitive index on an exact tie
            if (prim != s.skip && (t < s.t_limit || (t == s.t_limit && prim < s.skip))) { RT_AUDIT(skipped); s.occluded = true; s.cur = SENTINEL; return; }
        } else if (s.best.prim == RT_NONE || t < s.best.t || (t == s.best.t && prim < s.best.prim)) {
            RT_AUDIT(skipped);
            s.best.t = t; s.best.prim = prim;
            s.ftmax = fmaxf(fminf(s.ftmax, f32_up(t)), s.ftmin);     // see trace_closest: best t may lie below tmin for triangles
        }
    }
    s.cur = stack[--s.sp];
}

constexpr int TRACE_BLOCK = 128;            // block size of every persistent traversal kernel
// (A leaf phase compacted over the warp — the (ray, slot) pairs of a warp dealt out one per lane through shared memory — was built and measured in
// round 1: extend 32.9 -> 34.3 ms, i.e. slower; the idle lanes of the exact test come from its early-outs, not from the 1..4 slot loop.  Removed.)

#ifndef RT_REFILL
#define RT_REFILL 16        // refetch when fewer than this many lanes of the warp hold a live ray (BVH4, dragon 64 spp: 16 / 20 / 24 -> extend 31.6 / 31.7 / 32.5 ms)
#endif
#ifndef RT_NODE_MIN
#define RT_NODE_MIN 6       // leave the node phase for the leaf phase when fewer lanes than this still walk inner nodes (2 / 4 / 6 / 8 / 12 / 16 -> extend
                            // 34.1 / 32.2 / 31.6 / 31.6 / 32.2 / 33.5 ms, connect 21.1 / 20.3 / 20.1 / 20.2 / 21.1 / 22.4 ms)
#endif

// Job::shared_leaf_ray: the per-ray f64 constants that only the exact leaf test reads (1/d per axis and the two shear factors of the watertight test: 10
// registers) wait in shared memory between leaves (5 KB per block) instead of occupying registers through the node loop.  For k_connect only: under
// its 96-register cap ptxas had chosen to spill the f32 culling ray instead — 8 local loads in EVERY node step, 47 % of that kernel's local-memory
// sectors (profiles/connect_mem_by_source_r02.txt); connect -1 % (dragon) ... -3.5 % (Cornell + statue).  k_extend does not spill and pays +0.7 % for the
// round trip, so it keeps everything in registers; parking origin and direction as well (22 registers) was slower for both (pass V).
template <class Job>
__device__ __forceinline__ void trace_jobs(const DScene& sc, Job& job, uint32_t n, uint32_t* counter) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    int32_t stack[STACK + 1];                      // [STACK]: the write-only slot of node4_closest's suppressed pushes
    TState st; st.cur = SENTINEL; st.sp = 0;
    constexpr bool PARK = Job::shared_leaf_ray;
    __shared__ double sh_leaf[PARK ? 5 : 1][PARK ? TRACE_BLOCK : 1];
    const int tid = threadIdx.x;
    bool active = false, exhausted = false;
    uint32_t idx = 0;
    if (sc.n_nodes == 0 && sc.n_direct == 0) {      // empty scene: every ray misses
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
            if (job.load(i, st, stack)) { st.best.prim = RT_NONE; st.occluded = false; job.finish(i, st); }
        }
        return;
    }
    for (;;) {
        unsigned idle = __ballot_sync(FULL, !active);
        if (idle && !exhausted) {
            int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if (lane == leader) base = atomicAdd(counter, (uint32_t)__popc(idle));
            base = __shfl_sync(FULL, base, leader);
            if (!active) {
                idx = base + __popc(idle & ((1u << lane) - 1u));
                if (idx < n) active = job.load(idx, st, stack);
                if constexpr (PARK) {
                    if (active) { sh_leaf[0][tid] = st.rl.rx.inv.x; sh_leaf[1][tid] = st.rl.rx.inv.y; sh_leaf[2][tid] = st.rl.rx.inv.z; sh_leaf[3][tid] = st.rl.rx.sx; sh_leaf[4][tid] = st.rl.rx.sy; }
                }
            }
            if (base + __popc(idle) >= n) exhausted = true;
        }
        if (!__any_sync(FULL, active)) { if (exhausted) break; else continue; }
        for (;;) {
            // node phase: all lanes that sit on an inner node step together; lanes already at a leaf wait, but only
            // while at least RT_NODE_MIN lanes are still walking
            for (;;) {
                bool at_node = active && st.cur >= 0 && st.cur != SENTINEL;
                unsigned m = __ballot_sync(FULL, at_node);
                if (m == 0u) break;
                if (__popc(m) < RT_NODE_MIN && __any_sync(FULL, active && st.cur < 0)) break;
                if (at_node) node_step(sc, st, stack);
            }
            // leaf phase
            if (active && st.cur < 0) {
                if constexpr (PARK) {      // live for this leaf only
                    st.rl.rx.inv = mk3(sh_leaf[0][tid], sh_leaf[1][tid], sh_leaf[2][tid]); st.rl.rx.sx = sh_leaf[3][tid]; st.rl.rx.sy = sh_leaf[4][tid];
                }
                leaf_step(sc, st, stack, job.tmin(st), job.tmax(st), [&](D3& fo, D3& fd) { job.ray(idx, fo, fd); });
            }
            if (active && st.cur == SENTINEL) { job.finish(idx, st); active = false; }
            unsigned act = __ballot_sync(FULL, active);
            if (act == 0u) break;
            if (!exhausted && __popc(act) < RT_REFILL) break;
        }
    }
}

}  // namespace rt
