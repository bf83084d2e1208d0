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
        t0 = t0 - fabsf(t0) * 4.8e-7f; t1 = t1 + fabsf(t1) * 4.8e-7f;                                                               \
        const bool h = t0 <= t1;                                                                                                    \
        key[k] = h ? ((__float_as_int(t0) & ~3) | k) : (0x7f800000 | k);                                                            \
        mask |= h ? (1u << k) : 0u;                                                                                                 \
    }
    RT_CHILD(x, 0) RT_CHILD(y, 1) RT_CHILD(z, 2) RT_CHILD(w, 3)
#undef RT_CHILD
    return mask;
}
__device__ __forceinline__ int32_t child_of(const int4& ch, int32_t key) { int k = key & 3; return k == 0 ? ch.x : (k == 1 ? ch.y : (k == 2 ? ch.z : ch.w)); }
__device__ __forceinline__ void sort4(int32_t k[4]) {
#define RT_CSWAP(a, b) { int32_t lo = min(k[a], k[b]), hi = max(k[a], k[b]); k[a] = lo; k[b] = hi; }
    RT_CSWAP(0, 1) RT_CSWAP(2, 3) RT_CSWAP(0, 2) RT_CSWAP(1, 3) RT_CSWAP(1, 2)
#undef RT_CSWAP
}
// closest-hit descent: nearest hit child next, the others pushed far to near
__device__ __forceinline__ int32_t node4_closest(const DScene& sc, int32_t cur, const RayF& rf, float ftmin, float ftmax, int32_t* stack, int& sp) {
    int32_t key[4]; int4 ch;
    const unsigned mask = node4_test(reinterpret_cast<const float4*>(sc.nodes) + 8 * (size_t)cur, rf, ftmin, ftmax, key, ch);
    const int nh = __popc(mask);
    if (nh == 0) return stack[--sp];
    sort4(key);
    if (nh > 3 && sp < STACK) stack[sp++] = child_of(ch, key[3]);
    if (nh > 2 && sp < STACK) stack[sp++] = child_of(ch, key[2]);
    if (nh > 1 && sp < STACK) stack[sp++] = child_of(ch, key[1]);
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


// Candidate primitive in leaf slot `s`: the reference's exact accept test.  Returns true with t.
__device__ __forceinline__ bool slot_test(const DScene& sc, const DSlot* __restrict__ slots, uint32_t s, const RayX& rx, double tmin, double tmax,
                                          double& t, uint32_t& prim) {
    const double2* q = reinterpret_cast<const double2*>(slots + s);
    double2 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3), q4 = __ldg(q + 4);
    uint2 tag = *reinterpret_cast<uint2*>(&q4.y);
    prim = tag.x;
    if (tag.y == 0) {
        D3 p0 = mk3(q0.x, q0.y, q1.x), p1 = mk3(q1.y, q2.x, q2.y), p2 = mk3(q3.x, q3.y, q4.x);
        double b0, b1, b2;
        if (!tri_test(rx, p0, p1, p2, tmax, t, b0, b1, b2)) return false;
        return tri_leaf_box_inv(p0, p1, p2, rx.o, rx.inv, tmin, tmax);
    }
    if (tag.y == 2) return false;    // triangle the reference can never hit (degenerate uv frame ∧ zero normal, hittable.rs:373-378)
    DPrim pr = sc.prims[prim];
    return shape_test(sc, pr, rx.o, rx.d, tmin, tmax, true, t, &rx.inv);
}

// closest hit in the reference's sense over [tmin, tmax]
__device__ __forceinline__ Hit trace_closest(const DScene& sc, D3 o, D3 d, double tmin, double tmax) {
    Hit best; best.t = 0.; best.prim = RT_NONE;
    if (sc.n_nodes == 0) return best;
    RayX rx = make_rayx(o, d);
    RayF rf = make_rayf(o, d);
    float ftmin = f32_down(tmin); ftmin = ftmin - fabsf(ftmin) * 4.8e-7f;
    float ftmax = f32_up(fmin(tmax, 3.0e38));
    int32_t stack[STACK]; int sp = 0;
    stack[sp++] = SENTINEL;
    int32_t cur = 0;
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
            double t; uint32_t prim;
            if (slot_test(sc, sc.slots, s, rx, tmin, tmax, t, prim)) {
                if (best.prim == RT_NONE || t < best.t || (t == best.t && prim < best.prim)) {
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
    if (sc.n_nodes == 0) return false;
    RayX rx = make_rayx(o, d);
    RayF rf = make_rayf(o, d);
    float ftmin = f32_down(tmin); ftmin = ftmin - fabsf(ftmin) * 4.8e-7f;
    float ftmax = f32_up(fmin(fmin(tmax, t_limit), 3.0e38));
    int32_t stack[STACK]; int sp = 0;
    stack[sp++] = SENTINEL;
    int32_t cur = 0;
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
            double t; uint32_t prim;
            if (slot_test(sc, sc.slots, s, rx, tmin, tmax, t, prim) && prim != skip) {
                // the closest-hit rule keeps the smaller primitive index on an exact tie
                if (t < t_limit || (t == t_limit && prim < skip)) return true;
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
// A Job supplies: load(i, TState&) -> bool (false = nothing to trace, the job already finished ray i) and
// finish(i, const TState&).
struct TState {
    RayX rx; RayF rf;
    float ftmin, ftmax;
    double tmin, tmax, t_limit;
    uint32_t skip; bool any, occluded;
    Hit best;
    int32_t cur; int sp;
};
__device__ __forceinline__ void tstate_init(TState& s, int32_t* stack, D3 o, D3 d, double tmin, double tmax, bool any, double t_limit, uint32_t skip) {
    s.rx = make_rayx(o, d); s.rf = make_rayf(o, d);
    s.tmin = tmin; s.tmax = tmax; s.any = any; s.t_limit = t_limit; s.skip = skip; s.occluded = false;
    s.ftmin = f32_down(tmin); s.ftmin = s.ftmin - fabsf(s.ftmin) * 4.8e-7f;
    s.ftmax = f32_up(fmin(any ? fmin(tmax, t_limit) : tmax, 3.0e38));
    s.best.t = 0.; s.best.prim = RT_NONE;
    s.sp = 0; stack[s.sp++] = SENTINEL; s.cur = 0;
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
__device__ __forceinline__ void leaf_step(const DScene& sc, TState& s, int32_t* stack) {
    uint32_t v = (uint32_t)~s.cur, first = v >> 2, count = (v & 3u) + 1u;
    for (uint32_t k = first; k < first + count; k++) {
        double t; uint32_t prim;
        if (!slot_test(sc, sc.slots, k, s.rx, s.tmin, s.tmax, t, prim)) continue;
        if (s.any) {
            // the closest-hit rule keeps the smaller primitive index on an exact tie
            if (prim != s.skip && (t < s.t_limit || (t == s.t_limit && prim < s.skip))) { s.occluded = true; s.cur = SENTINEL; return; }
        } else if (s.best.prim == RT_NONE || t < s.best.t || (t == s.best.t && prim < s.best.prim)) {
            s.best.t = t; s.best.prim = prim;
            s.ftmax = fmaxf(fminf(s.ftmax, f32_up(t)), s.ftmin);     // see trace_closest: best t may lie below tmin for triangles
        }
    }
    s.cur = stack[--s.sp];
}

// Leaf phase, compacted over the warp.  After a node phase a warp holds a handful of lanes that sit on a leaf of 1..4 slots and
// many lanes that do not; running leaf_step per lane leaves most of the warp idle for the most expensive code of the traversal
// (the f64 reference-order tests; ncu: 7.4 of 32 lanes active in it).  Here the warp's (ray, slot) candidate pairs are
// enumerated with a prefix sum and dealt out one per lane, whoever owns the ray: the tester lane reads the owner's ray from
// shared memory (written once when the ray was fetched — it never changes during the walk), runs the exact test and posts
// (t, prim); the owner then folds the results of its own slots.  The fold is a min over (t, prim) for closest-hit rays and an
// "any" for shadow rays, so the order of the slots does not matter and the outcome is identical to leaf_step's.
// MEASURED (dragon, 1024^2, 64 spp): extend 32.9 -> 34.3 ms, connect 21.5 -> 22.0 ms, i.e. slower, so it is OFF by default.  ncu shows why: the
// lanes-per-instruction of the exact tests stays at 7.9 with compaction — the idle lanes come from the early-outs INSIDE the test (edge-function
// reject, then det, then the leaf box), not from the 1..4 slot loop — while the 13 KB of shared memory per block costs 5 points of L1 hit rate.
#ifndef RT_LEAF_COMPACT
#define RT_LEAF_COMPACT 0
#endif
constexpr int TRACE_BLOCK = 128;            // block size of every persistent traversal kernel
struct LeafShared {
    double ray[13][TRACE_BLOCK];            // o.xyz, d.xyz, shear sx sy, 1/d xyz, tmin, tmax of the ray each thread owns
    double res_t[TRACE_BLOCK];
    uint32_t res_prim[TRACE_BLOCK];
    int32_t kz[TRACE_BLOCK];
    uint8_t pair[TRACE_BLOCK];              // per warp: pair p of this round -> owner lane | slot-in-leaf << 5
};
__device__ __forceinline__ void leaf_shared_store(LeafShared& sh, const TState& s) {
    const int i = threadIdx.x;
    sh.ray[0][i] = s.rx.o.x; sh.ray[1][i] = s.rx.o.y; sh.ray[2][i] = s.rx.o.z;
    sh.ray[3][i] = s.rx.d.x; sh.ray[4][i] = s.rx.d.y; sh.ray[5][i] = s.rx.d.z;
    sh.ray[6][i] = s.rx.sx; sh.ray[7][i] = s.rx.sy; sh.ray[8][i] = s.rx.inv.x; sh.ray[11][i] = s.rx.inv.y; sh.ray[12][i] = s.rx.inv.z;
    sh.ray[9][i] = s.tmin; sh.ray[10][i] = s.tmax; sh.kz[i] = s.rx.kz;
}
__device__ __forceinline__ void leaf_phase_compact(const DScene& sc, LeafShared& sh, TState& s, int32_t* stack, bool active) {
    const unsigned FULL = 0xffffffffu;
    const int tid = threadIdx.x, lane = tid & 31, wbase = tid & ~31;
    const bool at_leaf = active && s.cur < 0;
    const uint32_t v = (uint32_t)~s.cur, first = v >> 2;
    const int cnt = at_leaf ? (int)(v & 3u) + 1 : 0;
    // prefix sum of the slot counts from three ballots (count - 1 has two bits)
    const unsigned ml = __ballot_sync(FULL, at_leaf), b0 = __ballot_sync(FULL, at_leaf && (v & 1u)), b1 = __ballot_sync(FULL, at_leaf && (v & 2u));
    const unsigned lt = (1u << lane) - 1u;
    const int excl = __popc(ml & lt) + __popc(b0 & lt) + 2 * __popc(b1 & lt);
    const int total = min(32, __popc(ml) + __popc(b0) + 2 * __popc(b1));
    // one round of at most 32 pairs; a leaf that does not fit keeps its remaining slots for the next leaf phase
    const int take = min(cnt, max(0, 32 - excl));
    for (int k = 0; k < take; k++) sh.pair[wbase + excl + k] = (uint8_t)(lane | (k << 5));
    __syncwarp();
    const bool has = lane < total;
    const int code = has ? sh.pair[tid] : lane;
    const int owner = wbase + (code & 31);
    const uint32_t slot = __shfl_sync(FULL, first, code & 31) + (uint32_t)(code >> 5);
    double t = 0.; uint32_t prim = RT_NONE;
    if (has) {
        RayX rx;
        rx.o = mk3(sh.ray[0][owner], sh.ray[1][owner], sh.ray[2][owner]); rx.d = mk3(sh.ray[3][owner], sh.ray[4][owner], sh.ray[5][owner]);
        rx.sx = sh.ray[6][owner]; rx.sy = sh.ray[7][owner]; rx.inv = mk3(sh.ray[8][owner], sh.ray[11][owner], sh.ray[12][owner]); rx.kz = sh.kz[owner];
        uint32_t p;
        if (slot_test(sc, sc.slots, slot, rx, sh.ray[9][owner], sh.ray[10][owner], t, p)) prim = p;
    }
    sh.res_t[tid] = t; sh.res_prim[tid] = prim;
    __syncwarp();
    for (int k = 0; k < take; k++) {
        const uint32_t hp = sh.res_prim[wbase + excl + k];
        if (hp == RT_NONE) continue;
        const double ht = sh.res_t[wbase + excl + k];
        if (s.any) {
            if (hp != s.skip && (ht < s.t_limit || (ht == s.t_limit && hp < s.skip))) s.occluded = true;
        } else if (s.best.prim == RT_NONE || ht < s.best.t || (ht == s.best.t && hp < s.best.prim)) {
            s.best.t = ht; s.best.prim = hp;
            s.ftmax = fmaxf(fminf(s.ftmax, f32_up(ht)), s.ftmin);
        }
    }
    if (at_leaf) {
        if (s.any && s.occluded) s.cur = SENTINEL;
        else if (take == cnt) s.cur = stack[--s.sp];
        else s.cur = (int32_t)~(((first + (uint32_t)take) << 2) | (uint32_t)(cnt - take - 1));
    }
}

#ifndef RT_REFILL
#define RT_REFILL 16        // refetch when fewer than this many lanes of the warp hold a live ray (BVH4, dragon 64 spp: 16 / 20 / 24 -> extend 31.6 / 31.7 / 32.5 ms)
#endif
#ifndef RT_NODE_MIN
#define RT_NODE_MIN 6       // leave the node phase for the leaf phase when fewer lanes than this still walk inner nodes (2 / 4 / 6 / 8 / 12 / 16 -> extend
                            // 34.1 / 32.2 / 31.6 / 31.6 / 32.2 / 33.5 ms, connect 21.1 / 20.3 / 20.1 / 20.2 / 21.1 / 22.4 ms)
#endif

template <class Job>
__device__ __forceinline__ void trace_jobs(const DScene& sc, Job& job, uint32_t n, uint32_t* counter) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    int32_t stack[STACK];
    TState st; st.cur = SENTINEL; st.sp = 0;
    bool active = false, exhausted = false;
    uint32_t idx = 0;
#if RT_LEAF_COMPACT
    __shared__ LeafShared lsh;
#endif
    if (sc.n_nodes == 0) {      // empty scene: every ray misses
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
#if RT_LEAF_COMPACT
                if (active) leaf_shared_store(lsh, st);
#endif
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
#if RT_LEAF_COMPACT
            leaf_phase_compact(sc, lsh, st, stack, active);
#else
            if (active && st.cur < 0) leaf_step(sc, st, stack);
#endif
            if (active && st.cur == SENTINEL) { job.finish(idx, st); active = false; }
            unsigned act = __ballot_sync(FULL, active);
            if (act == 0u) break;
            if (!exhausted && __popc(act) < RT_REFILL) break;
        }
    }
}

}  // namespace rt
