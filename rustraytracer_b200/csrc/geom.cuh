// Device geometry: exact (f64, reference operation order) primitive tests and hit-record
// reconstruction.  The library is compiled with -fmad=false so that every a*b+c below is two
// correctly rounded operations, as in the reference's Rust (which never contracts).
// Reference: hittable.rs:292-452 (triangle), intersects.rs:10-258 (rects, sphere),
// hittable.rs:494-508 (slab test), primitive.rs:247-316 (dispatch / FlipFace).
#pragma once
#include <cuda_runtime.h>
#include "rt_internal.h"

namespace rt {

struct D3 { double x, y, z; };
struct D2 { double x, y; };
__device__ __forceinline__ D3 mk3(double x, double y, double z) { D3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ D2 mk2(double x, double y) { D2 r; r.x = x; r.y = y; return r; }
__device__ __forceinline__ D3 operator+(D3 a, D3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ D3 operator-(D3 a, D3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ D3 operator-(D3 a) { return mk3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ D3 operator*(D3 a, double s) { return mk3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ D3 operator*(double s, D3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ D3 operator/(D3 a, double s) { return mk3(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ bool eq3(D3 a, D3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
__device__ __forceinline__ bool is_black(D3 a) { return a.x == 0. && a.y == 0. && a.z == 0.; }
__device__ __forceinline__ D3 cmul(D3 a, D3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ D3 cdiv(D3 a, D3 b) { return mk3(a.x / b.x, a.y / b.y, a.z / b.z); }
__device__ __forceinline__ double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ D3 cross(D3 a, D3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
__device__ __forceinline__ double norm2(D3 a) { return dot(a, a); }
__device__ __forceinline__ double norm(D3 a) { return sqrt(norm2(a)); }
// a / n, bit for bit, for the components of a vector divided by its own norm.  A ZERO numerator takes the slow path of the f64 division (the inline
// fast path only covers numerators of ordinary magnitude; the rest is an out-of-line routine of ~60 instructions the whole warp waits for), and the
// vectors normalised here are mostly axis-aligned — rect normals and tangents, frames built from them: two zero components in three.  ncu r02: that
// routine ran 11 times per shaded vertex and was 19 % of the f64 matte kernel's executed instructions.  +-0 / n = +-0 for any n > 0 (n = 0 or NaN
// still goes through the division and gives the reference's NaN).  The division itself is the same IEEE operation, kept behind an opaque asm so that
// the compiler does not turn the test back into an unconditional quotient plus a select.
// Only in the f64 shade kernels' translation unit (f64 shade -2.3 % dragon / -6.3 % Cornell, all-f64 bench 2,337 -> 2,440 Mrays/s): the f32 kernels
// normalise few f64 vectors, and there the opaque division cost more than the slow path it avoids (shade +2.5 %), so shade_f32.cu keeps the plain quotient.
__device__ __forceinline__ double div_by_norm(double a, double n) {
#ifndef RT_PLAIN_NORMALIZE
    if (a == 0. && n > 0.) return a;
    double q;
    asm volatile("div.rn.f64 %0, %1, %2;" : "=d"(q) : "d"(a), "d"(n));
    return q;
#else
    return a / n;
#endif
}
__device__ __forceinline__ D3 normalize(D3 a) { double n = norm(a); return mk3(div_by_norm(a.x, n), div_by_norm(a.y, n), div_by_norm(a.z, n)); }
__device__ __forceinline__ double comp(D3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
__device__ __forceinline__ double clampd(double x, double lo, double hi) { return fmin(fmax(x, lo), hi); }
__device__ __forceinline__ D3 face_forward(D3 n, D3 v) { return dot(n, v) < 0. ? -n : n; }          // util.rs:578-581
__device__ __forceinline__ void make_coordinate_system(D3 v1, D3& v2, D3& v3) {                     // util.rs:567-576
    if (fabs(v1.x) > fabs(v1.y)) v2 = mk3(-v1.z, 0., v1.x) * (1. / sqrt(v1.x * v1.x + v1.z * v1.z));
    else v2 = mk3(0., v1.z, -v1.y) * (1. / sqrt(v1.y * v1.y + v1.z * v1.z));
    v3 = cross(v1, v2);
}
__device__ __forceinline__ D3 ld3(const double* p) { return mk3(p[0], p[1], p[2]); }

// Per-ray exact constants: permutation and shear of the watertight test depend on the ray only
// (hittable.rs:313-323), so they are computed once per ray instead of once per triangle.
struct RayX {
    D3 o, d;
    int kz;
    double sx, sy;
    D3 inv;                 // 1.0 / d per axis: the `inv_d` of BoundingBox::intersects (hittable.rs:497), computed once per ray; sz = inv[kz]
};
__device__ __forceinline__ RayX make_rayx(D3 o, D3 d) {
    RayX r; r.o = o; r.d = d;
    double ax = fabs(d.x), ay = fabs(d.y), az = fabs(d.z);
    int kz = 0; double best = ax;                       // argmax: first maximum
    if (ay > best) { kz = 1; best = ay; }
    if (az > best) { kz = 2; }
    r.kz = kz;
    int kx = (kz + 1) % 3, ky = (kx + 1) % 3;
    double dx = comp(d, kx), dy = comp(d, ky), dz = comp(d, kz);
    r.sx = -dx / dz; r.sy = -dy / dz;
    r.inv = mk3(1. / d.x, 1. / d.y, 1. / d.z);
    return r;
}
__device__ __forceinline__ D3 permute(D3 v, int kz) {  // (v[kx], v[ky], v[kz])
    return kz == 0 ? mk3(v.y, v.z, v.x) : (kz == 1 ? mk3(v.z, v.x, v.y) : v);
}

// Mesh::intersects_triangle up to the accept/reject decision (hittable.rs:300-362).
// Returns true and t, b0, b1, b2 when the reference would go on to build a record.
__device__ __forceinline__ bool tri_test(const RayX& r, D3 p0, D3 p1, D3 p2, double tmax, double& t, double& b0, double& b1, double& b2) {
    D3 p0t = permute(p0 - r.o, r.kz), p1t = permute(p1 - r.o, r.kz), p2t = permute(p2 - r.o, r.kz);
    p0t.x += r.sx * p0t.z; p0t.y += r.sy * p0t.z;
    p1t.x += r.sx * p1t.z; p1t.y += r.sy * p1t.z;
    p2t.x += r.sx * p2t.z; p2t.y += r.sy * p2t.z;
    double e0 = p1t.x * p2t.y - p1t.y * p2t.x;
    double e1 = p2t.x * p0t.y - p2t.y * p0t.x;
    double e2 = p0t.x * p1t.y - p0t.y * p1t.x;
    if ((e0 < 0. || e1 < 0. || e2 < 0.) && (e0 > 0. || e1 > 0. || e2 > 0.)) return false;
    double det = e0 + e1 + e2;
    if (fabs(det) < SMALL / 10000.) return false;
    const double sz = comp(r.inv, r.kz);                // 1. / d[kz]
    p0t.z *= sz; p1t.z *= sz; p2t.z *= sz;
    double t_scaled = e0 * p0t.z + e1 * p1t.z + e2 * p2t.z;
    if (det < 0. && (t_scaled >= 0. || t_scaled < tmax * det)) return false;
    else if (det > 0. && (t_scaled <= 0. || t_scaled > tmax * det)) return false;
    double inv_det = 1. / det;
    b0 = e0 * inv_det; b1 = e1 * inv_det; b2 = e2 * inv_det;
    t = t_scaled * inv_det;
    if (t < SMALL / 10.) return false;
    return true;
}

// BoundingBox::intersects (hittable.rs:494-508) on the primitive's own f64 leaf box
__device__ __forceinline__ bool slab_test(D3 mn, D3 mx, D3 o, D3 d, double tmin, double tmax) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
        double inv_d = 1.0 / comp(d, a);
        double val1 = (comp(mn, a) - comp(o, a)) * inv_d;
        double val2 = (comp(mx, a) - comp(o, a)) * inv_d;
        double t0 = fmin(val1, val2), t1 = fmax(val1, val2);
        tmin = fmax(tmin, t0); tmax = fmin(tmax, t1);
        if (tmax <= tmin) return false;
    }
    return true;
}
// the same test with the ray's precomputed 1.0 / d (bitwise the same quotients)
__device__ __forceinline__ bool slab_test_inv(D3 mn, D3 mx, D3 o, D3 inv, double tmin, double tmax) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
        double inv_d = comp(inv, a);
        double val1 = (comp(mn, a) - comp(o, a)) * inv_d;
        double val2 = (comp(mx, a) - comp(o, a)) * inv_d;
        double t0 = fmin(val1, val2), t1 = fmax(val1, val2);
        tmin = fmax(tmin, t0); tmax = fmin(tmax, t1);
        if (tmax <= tmin) return false;
    }
    return true;
}
__device__ __forceinline__ bool tri_leaf_box_inv(D3 p0, D3 p1, D3 p2, D3 o, D3 inv, double tmin, double tmax) {
    D3 mn = mk3(fmin(p0.x, fmin(p1.x, p2.x)), fmin(p0.y, fmin(p1.y, p2.y)), fmin(p0.z, fmin(p1.z, p2.z)));
    D3 mx = mk3(fmax(p0.x, fmax(p1.x, p2.x)), fmax(p0.y, fmax(p1.y, p2.y)), fmax(p0.z, fmax(p1.z, p2.z)));
    return slab_test_inv(mn, mx, o, inv, tmin, tmax);
}
__device__ __forceinline__ bool tri_leaf_box(D3 p0, D3 p1, D3 p2, D3 o, D3 d, double tmin, double tmax) {
    D3 mn = mk3(fmin(p0.x, fmin(p1.x, p2.x)), fmin(p0.y, fmin(p1.y, p2.y)), fmin(p0.z, fmin(p1.z, p2.z)));
    D3 mx = mk3(fmax(p0.x, fmax(p1.x, p2.x)), fmax(p0.y, fmax(p1.y, p2.y)), fmax(p0.z, fmax(p1.z, p2.z)));
    return slab_test(mn, mx, o, d, tmin, tmax);
}

// sphere_intersect (intersects.rs:177-205): t only
__device__ __forceinline__ bool sphere_test(D3 center, double r, D3 o, D3 d, double tmin, double tmax, double& t) {
    D3 diff = o - center;
    double a = dot(d, d), b = dot(diff, d), c = dot(diff, diff) - r * r;
    double disc = b * b - a * c;
    if (disc < 0.0) return false;
    double inv_a = 1.0 / a, root = sqrt(disc);
    double ans = (-b - root) * inv_a;
    if (ans < tmax && ans > tmin) { t = ans; return true; }
    ans = (-b + root) * inv_a;
    if (ans < tmax && ans > tmin) { t = ans; return true; }
    return false;
}

__device__ __forceinline__ D3 xf_vector(const double* m, D3 v) {
    return mk3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z, m[8] * v.x + m[9] * v.y + m[10] * v.z);
}
__device__ __forceinline__ D3 xf_point(const double* m, D3 p) {
    double x = m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3];
    double y = m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7];
    double z = m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11];
    double w = m[12] * p.x + m[13] * p.y + m[14] * p.z + m[15];
    return mk3(x / w, y / w, z / w);
}

// x?_rect_intersect (intersects.rs:10-175) up to accept: t, in-plane coordinates, object-space ray
struct RectHit { double t, a, b; D3 to, td; };
__device__ __forceinline__ bool rect_test(uint32_t type, const double* sh, const DXform* xforms, D3 o, D3 d, double t0, double t1, RectHit& h) {
    int kaxis = type == RT_PRIM_XY_RECT ? 2 : (type == RT_PRIM_XZ_RECT ? 1 : 0);
    int aa = kaxis == 0 ? 1 : 0, ba = kaxis == 2 ? 1 : 2;
    D3 to = o, td = d;
    if (sh[5] >= 0.) { const DXform& x = xforms[(uint32_t)sh[5]]; td = xf_vector(x.inv, d); to = xf_point(x.inv, o); }   // Ray::transform geometry.rs:231-235
    double t = (sh[4] - comp(to, kaxis)) / comp(td, kaxis);
    if (t < t0 || t > t1) return false;
    double a = comp(to, aa) + t * comp(td, aa);
    double b = comp(to, ba) + t * comp(td, ba);
    if (a < sh[0] || b < sh[1] || a > sh[2] || b > sh[3]) return false;
    h.t = t; h.a = a; h.b = b; h.to = to; h.td = td;
    return true;
}

// Exact accept test of one non-triangle primitive incl. its f64 leaf box: hit(prim) = leafbox ∧ primtest
__device__ __forceinline__ bool shape_test(const DScene& sc, const DPrim& pr, D3 o, D3 d, double tmin, double tmax, bool leaf_box, double& t, const D3* inv = nullptr) {
    uint32_t type = pr.info & 7u;
    const double* sh = sc.shapes + pr.a;
    if (leaf_box && !(inv ? slab_test_inv(ld3(sh + 6), ld3(sh + 9), o, *inv, tmin, tmax) : slab_test(ld3(sh + 6), ld3(sh + 9), o, d, tmin, tmax))) return false;
    if (type == RT_PRIM_SPHERE) return sphere_test(ld3(sh), sh[3], o, d, tmin, tmax, t);
    RectHit h;
    if (!rect_test(type, sh, sc.xforms, o, d, tmin, tmax, h)) return false;
    t = h.t; return true;
}

// The HitRecord fields the integrator consumes (hittable.rs:51-72)
struct Rec {
    double t; D3 p, n, sn, sdpdu; D2 uv; D3 wo; uint32_t prim, mat; bool front;
};
// HitRecord::new + set_front (hittable.rs:75-117,186-189) for parametric shapes
// PN: only the point and the normal are wanted (prim_hit_pn): the shading tangent is left unnormalised
template <bool PN = false> __device__ __forceinline__ void rec_from_partials(Rec& r, D3 p, D2 uv, D3 dpdu, D3 dpdv, double t, D3 ray_d) {
    D3 n = normalize(cross(dpdu, dpdv));
    r.sn = n; r.sdpdu = PN ? dpdu : normalize(dpdu);
    r.p = p; r.uv = uv; r.t = t; r.wo = -ray_d;
    r.front = dot(ray_d, n) < 0.0;
    r.n = r.front ? n : -n;
}

// RT_SHADE_F32 mode: the same record with the two normalisations and the cross product in f32 (consumed in f32 there)
__device__ __forceinline__ void rec_from_partials_f32(Rec& r, D3 p, D2 uv, D3 dpdu, D3 dpdv, double t, D3 ray_d) {
    const float ux = (float)dpdu.x, uy = (float)dpdu.y, uz = (float)dpdu.z, vx = (float)dpdv.x, vy = (float)dpdv.y, vz = (float)dpdv.z;
    float nx = __fmaf_rn(uy, vz, -(uz * vy)), ny = __fmaf_rn(uz, vx, -(ux * vz)), nz = __fmaf_rn(ux, vy, -(uy * vx));
    const float sn = rsqrtf(__fmaf_rn(nz, nz, __fmaf_rn(ny, ny, nx * nx))), su = rsqrtf(__fmaf_rn(uz, uz, __fmaf_rn(uy, uy, ux * ux)));
    nx *= sn; ny *= sn; nz *= sn;
    r.sn = mk3((double)nx, (double)ny, (double)nz); r.sdpdu = mk3((double)(ux * su), (double)(uy * su), (double)(uz * su));
    r.p = p; r.uv = uv; r.t = t; r.wo = -ray_d;
    r.front = (float)ray_d.x * nx + (float)ray_d.y * ny + (float)ray_d.z * nz < 0.0f;
    r.n = r.front ? r.sn : -r.sn;
}

// Primitive::intersects for ONE primitive with the full record (primitive.rs:247-316): used by shade and
// by the rt_intersect epilogue on the closest primitive only.  Bitwise the same accept arithmetic as the
// traversal, so it cannot disagree with it.  leaf_box = false gives Primitive::intersects_obj semantics
// (primitive.rs:372-425) plus the FlipFace toggle being skipped by the caller.
// CLS: 1 = the primitive is known to be a triangle, 0 = known parametric shape, 2 = decide at run time
// F32FRAME (RT_SHADE_F32 mode, parametric shapes): normal / dpdu normalisation in f32, everything else unchanged
// PN (prim_hit_pn): the caller reads p and n only — uv (two f64 divisions for a rect) and the normalised tangent are not computed
template <int CLS, bool F32FRAME = false, bool PN = false>
static __device__ __noinline__ bool prim_record_t(const DScene& sc, uint32_t prim, D3 o, D3 d, double tmin, double tmax, bool leaf_box, Rec& r) {
    DPrim pr = sc.prims[prim];
    uint32_t type = CLS == 1 ? (uint32_t)RT_PRIM_TRIANGLE : (pr.info & 7u);
    r.prim = prim; r.mat = pr.mat;
    if (CLS != 0 && type == RT_PRIM_TRIANGLE) {
        const uint32_t* ti = sc.tri_index + pr.a;
        uint32_t i1 = ti[0], i2 = ti[1], i3 = ti[2];
        const DMesh ms = sc.meshes[pr.info >> 6];
        D3 p0 = ld3(sc.points + 3 * (ms.p_base + i1)), p1 = ld3(sc.points + 3 * (ms.p_base + i2)), p2 = ld3(sc.points + 3 * (ms.p_base + i3));
        RayX rx = make_rayx(o, d);
        double t, b0, b1, b2;
        if (!tri_test(rx, p0, p1, p2, tmax, t, b0, b1, b2)) return false;
        if (leaf_box && !tri_leaf_box(p0, p1, p2, o, d, tmin, tmax)) return false;
        D2 uv0 = mk2(0., 0.), uv1 = mk2(1., 0.), uv2 = mk2(1., 1.);            // Mesh::get_uv hittable.rs:454-468
        if (pr.info & PI_HAS_UV) {
            const double* u = sc.uvs; uv0 = mk2(u[2 * (ms.uv_base + i1)], u[2 * (ms.uv_base + i1) + 1]);
            uv1 = mk2(u[2 * (ms.uv_base + i2)], u[2 * (ms.uv_base + i2) + 1]); uv2 = mk2(u[2 * (ms.uv_base + i3)], u[2 * (ms.uv_base + i3) + 1]);
        }
        D2 duv02 = mk2(uv0.x - uv2.x, uv0.y - uv2.y), duv12 = mk2(uv1.x - uv2.x, uv1.y - uv2.y);
        D3 dp02 = p0 - p2, dp12 = p1 - p2;
        double determinant = duv02.x * duv12.y - duv02.y * duv12.x;
        D3 dpdu, dpdv;
        if (fabs(determinant) < SMALL / 10000.) {
            D3 n = cross(p2 - p0, p1 - p0);
            if (norm2(n) == 0.) return false;
            make_coordinate_system(n, dpdu, dpdv);
        } else {
            double inv = 1. / determinant;
            dpdu = (duv12.y * dp02 - duv02.y * dp12) * inv;
        }
        r.p = b0 * p0 + b1 * p1 + b2 * p2;
        r.uv = mk2(b0 * uv0.x + b1 * uv1.x + b2 * uv2.x, b0 * uv0.y + b1 * uv1.y + b2 * uv2.y);
        D3 normal;
        if (pr.info & PI_HAS_N) {
            const double* nn = sc.normals;
            normal = b0 * ld3(nn + 3 * (ms.n_base + i1)) + b1 * ld3(nn + 3 * (ms.n_base + i2)) + b2 * ld3(nn + 3 * (ms.n_base + i3));
        } else normal = cross(dp02, dp12);
        D3 gn = normalize(cross(dp02, dp12));
        D3 sn = normalize(normal);
        D3 ss = normalize(dpdu);
        D3 ts = normalize(cross(sn, ss));
        if (norm2(ts) > 0.) ss = normalize(cross(ts, sn));
        else { D3 a, b; make_coordinate_system(sn, a, b); ss = normalize(a); ts = normalize(b); }
        // set_shading_geometry(ss, ts, .., is_auth = true) (hittable.rs:191-210)
        D3 n2 = normalize(cross(ss, ts));
        r.sn = n2;
        gn = face_forward(gn, r.sn);
        r.sdpdu = ss;
        r.front = dot(d, gn) < 0.0;                                              // set_front
        r.n = r.front ? gn : -gn;
        r.t = t; r.wo = -d;
    } else {
        const double* sh = sc.shapes + pr.a;
        if (leaf_box && !slab_test(ld3(sh + 6), ld3(sh + 9), o, d, tmin, tmax)) return false;
        if (type == RT_PRIM_SPHERE) {
            D3 center = ld3(sh); double rad = sh[3], t;
            if (!sphere_test(center, rad, o, d, tmin, tmax, t)) return false;
            // make_sphere_record (intersects.rs:216-258) on the ray translated to the centre
            D3 to = o - center;
            D3 p = to + d * t;
            p = (p * rad) / norm(p);
            if (p.x == 0. && p.y == 0.) p.x = 1e-5 * rad;
            double phi = atan2(p.y, p.x);
            if (phi < 0.) phi = phi + 2. * PI;
            double phi_max = 2. * PI, theta_min = 0., theta_max = PI;
            double u = phi / phi_max;
            double theta = acos(clampd(p.z / rad, -1., 1.));
            double v = (theta - theta_min) / (theta_max - theta_min);
            double z_r = sqrt(p.x * p.x + p.y * p.y);
            double inv_z_r = 1. / z_r;
            double cos_phi = p.x * inv_z_r, sin_phi = p.y * inv_z_r;
            D3 dpdu = mk3(-phi_max * p.y, phi_max * p.x, 0.);
            D3 dpdv = (theta_max - theta_min) * mk3(p.z * cos_phi, p.z * sin_phi, -rad * sin(theta));
            if (F32FRAME) rec_from_partials_f32(r, p, mk2(u, v), dpdu, dpdv, t, d); else rec_from_partials<PN>(r, p, mk2(u, v), dpdu, dpdv, t, d);
            r.p = r.p + center;
        } else {
            RectHit h;
            if (!rect_test(type, sh, sc.xforms, o, d, tmin, tmax, h)) return false;
            int kaxis = type == RT_PRIM_XY_RECT ? 2 : (type == RT_PRIM_XZ_RECT ? 1 : 0);
            int aa = kaxis == 0 ? 1 : 0, ba = kaxis == 2 ? 1 : 2;
            D2 uv = mk2(0., 0.);
            if (!PN) uv = mk2((h.a - sh[0]) / (sh[2] - sh[0]), (h.b - sh[1]) / (sh[3] - sh[1]));
            D3 p = h.to + h.td * h.t;
            D3 dpdu = mk3(aa == 0 ? 1. : 0., aa == 1 ? 1. : 0., 0.);
            D3 dpdv = mk3(0., ba == 1 ? 1. : 0., ba == 2 ? 1. : 0.);
            if (sh[5] >= 0.) { const DXform& x = sc.xforms[(uint32_t)sh[5]]; p = xf_point(x.fwd, p); dpdu = xf_vector(x.fwd, dpdu); dpdv = xf_vector(x.fwd, dpdv); }
            if (F32FRAME) rec_from_partials_f32(r, p, uv, dpdu, dpdv, h.t, d); else rec_from_partials<PN>(r, p, uv, dpdu, dpdv, h.t, d);
        }
    }
    if (leaf_box && (pr.info & PI_FLIP)) r.front = !r.front;                     // FlipFace (primitive.rs:300-310)
    return true;
}

// Primitive::pdf's re-intersection of ONE (light) primitive (primitive.rs:462-473 → intersects_obj, no leaf box, no
// FlipFace toggle): only the hit point and the geometric normal UP TO SIGN are consumed there
// (`dist.magnitude_squared() / (area * |n · -dir|)`), so the triangle arm skips the shading frame.
// F32N (RT_SHADE_F32 kernels, where the pdf this feeds is an f32 MIS weight): the normal is normalised in f32 — the f64 record of a parametric light cost
// two f64 normalisations (a square root and three divisions each, ~280 instructions) per shaded vertex for a value consumed as a float.
template <bool F32N = false>
static __device__ __noinline__ bool prim_hit_pn(const DScene& sc, uint32_t prim, D3 o, D3 d, double tmin, double tmax, D3& p, D3& n) {
    DPrim pr = sc.prims[prim];
    if ((pr.info & 7u) == RT_PRIM_TRIANGLE) {
        const uint32_t* ti = sc.tri_index + pr.a;
        uint32_t i1 = ti[0], i2 = ti[1], i3 = ti[2];
        const DMesh ms = sc.meshes[pr.info >> 6];
        D3 p0 = ld3(sc.points + 3 * (ms.p_base + i1)), p1 = ld3(sc.points + 3 * (ms.p_base + i2)), p2 = ld3(sc.points + 3 * (ms.p_base + i3));
        RayX rx = make_rayx(o, d);
        double t, b0, b1, b2;
        if (!tri_test(rx, p0, p1, p2, tmax, t, b0, b1, b2)) return false;
        if (pr.info & PI_HAS_UV) {                                   // degenerate uv frame ∧ zero normal → None (hittable.rs:366-378)
            const double* u = sc.uvs;
            D2 uv0 = mk2(u[2 * (ms.uv_base + i1)], u[2 * (ms.uv_base + i1) + 1]), uv1 = mk2(u[2 * (ms.uv_base + i2)], u[2 * (ms.uv_base + i2) + 1]);
            D2 uv2 = mk2(u[2 * (ms.uv_base + i3)], u[2 * (ms.uv_base + i3) + 1]);
            double determinant = (uv0.x - uv2.x) * (uv1.y - uv2.y) - (uv0.y - uv2.y) * (uv1.x - uv2.x);
            if (fabs(determinant) < SMALL / 10000. && norm2(cross(p2 - p0, p1 - p0)) == 0.) return false;
        }
        p = b0 * p0 + b1 * p1 + b2 * p2;
        if constexpr (F32N) {
            const D3 e1 = p0 - p2, e2 = p1 - p2;
            const float ax = (float)e1.x, ay = (float)e1.y, az = (float)e1.z, bx = (float)e2.x, by = (float)e2.y, bz = (float)e2.z;
            float nx = __fmaf_rn(ay, bz, -(az * by)), ny = __fmaf_rn(az, bx, -(ax * bz)), nz = __fmaf_rn(ax, by, -(ay * bx));
            const float s = rsqrtf(__fmaf_rn(nz, nz, __fmaf_rn(ny, ny, nx * nx)));
            n = mk3((double)(nx * s), (double)(ny * s), (double)(nz * s));
        } else n = normalize(cross(p0 - p2, p1 - p2));
        return true;
    }
    Rec r;
    if (!prim_record_t<0, F32N, true>(sc, prim, o, d, tmin, tmax, false, r)) return false;
    p = r.p; n = r.n;
    return true;
}
static __device__ __forceinline__ bool prim_record(const DScene& sc, uint32_t prim, D3 o, D3 d, double tmin, double tmax, bool leaf_box, Rec& r) {
    return prim_record_t<2>(sc, prim, o, d, tmin, tmax, leaf_box, r);
}

}  // namespace rt
