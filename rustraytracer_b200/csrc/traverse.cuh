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

constexpr int STACK = 64;
constexpr int32_t SENTINEL = 0x7fffffff;

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
        return tri_leaf_box(p0, p1, p2, rx.o, rx.d, tmin, tmax);
    }
    if (tag.y == 2) return false;    // triangle the reference can never hit (degenerate uv frame ∧ zero normal, hittable.rs:373-378)
    DPrim pr = sc.prims[prim];
    return shape_test(sc, pr, rx.o, rx.d, tmin, tmax, true, t);
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
    const float4* __restrict__ nodes = reinterpret_cast<const float4*>(sc.nodes);
    while (cur != SENTINEL) {
        while (cur >= 0 && cur != SENTINEL) {
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
        }
        if (cur == SENTINEL) break;
        // leaf
        uint32_t v = (uint32_t)~cur, first = v >> 2, count = (v & 3u) + 1u;
        for (uint32_t s = first; s < first + count; s++) {
            double t; uint32_t prim;
            if (slot_test(sc, sc.slots, s, rx, tmin, tmax, t, prim)) {
                if (best.prim == RT_NONE || t < best.t || (t == best.t && prim < best.prim)) {
                    best.t = t; best.prim = prim;
                    ftmax = fminf(ftmax, f32_up(t));
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
    const float4* __restrict__ nodes = reinterpret_cast<const float4*>(sc.nodes);
    while (cur != SENTINEL) {
        while (cur >= 0 && cur != SENTINEL) {
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

}  // namespace rt
