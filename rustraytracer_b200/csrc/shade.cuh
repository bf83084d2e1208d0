// Device shading: textures, materials → BSDF, BxDF lobes, Trowbridge–Reitz microfacets, area lights.
// Reference operation order throughout; every function cites the reference item it replaces.  Quirks of
// SURVEY.md §8.1 are kept on purpose: image parity is against what the reference computes.
//
// Everything radiometric is templated on the real type R:
//   R = double : the reference's own arithmetic (compiled with -fmad=false) — films match the f64 oracle per pixel;
//   R = float  : the production path — same formulas in f32 (half the registers, a third of the code, cheap
//                transcendentals); geometry (hit points, normals, ray origins, light sample points) stays f64
//                in both, so ray/primitive decisions remain bit-exact.
// Heavy functions are __noinline__: the fully inlined kernel was 41 K SASS instructions and spent 81 % of its
// stall samples on instruction-cache misses (profiles/shade_r01_notes.md).
#pragma once
#include "geom.cuh"

namespace rt {

constexpr uint32_t BSDF_REFLECTION = 1, BSDF_TRANSMISSION = 2, BSDF_DIFFUSE = 4, BSDF_GLOSSY = 8, BSDF_SPECULAR = 16, BSDF_ALL = 31;  // consts.rs:14-20

// ---- counter-based RNG replacing thread_rng()/StdRng::from_entropy() (sampler.rs:318,334-337; util.rs:32-34):
// Inlining level of the shading code (measured r01, dragon 64 spp, shade ms: 0 = everything heavy is a call 40.8; 1 = BSDF code of the
// Lambert / Oren-Nayar materials inlined 33.6; 2 = BSDF + lobe code of every material inlined 30.1; 3 / 4 see profiles/shade_r01_notes.md)
#ifndef SHADE_INLINE_SMALL
#define SHADE_INLINE_SMALL 2
#endif
#define RT_HELPER __noinline__          // microfacet / Fresnel helpers: inlining them too is slower (31.7 vs 30.2 ms)
#if SHADE_INLINE_SMALL >= 3
#define RT_HELPER2 __forceinline__      // functions that return through references (record, texture, area sample)
#else
#define RT_HELPER2 __noinline__
#endif
#if SHADE_INLINE_SMALL >= 4
#define RT_HELPER3 __forceinline__      // the Philox block
#else
#define RT_HELPER3 __noinline__
#endif
// Philox4x32-10, key = seed, counter = (pixel, sample, dim >> 1, 'RTCX').
// Each Philox block yields two uniforms: dimension d uses block d >> 1, words (0,1) when d is even, (2,3) when odd.
// The block computed for an even dimension is kept for the odd one that follows.
struct Rng { uint32_t k0, k1, pixel, sample, dim; uint32_t w2, w3; };
// The block function is pure (registers in, registers out); the bookkeeping around it is always inlined, so that an Rng lives in the caller's
// registers.  (As one out-of-line function taking Rng& it forced the generator into local memory: five local loads and up to three stores per draw,
// about a third of the matte shade kernel's local-memory sectors — profiles/shade_matte_mem_by_source_r02.txt.)
static __device__ RT_HELPER3 uint4 philox4x32_10(uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1, uint32_t c2) {
    uint32_t c3 = 0x52544358u;
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}
__device__ __forceinline__ double rng_next(Rng& g) {
    const uint32_t d = g.dim++;
    uint32_t a, b;
    if ((d & 1u) && g.w2 != 0xffffffffu) {                 // cached second half (0xffffffff marks "none"; a genuine all-ones word just recomputes)
        a = g.w2; b = g.w3; g.w2 = 0xffffffffu;
    } else {
        const uint4 r = philox4x32_10(g.k0, g.k1, g.pixel, g.sample, d >> 1);
        if (d & 1u) { a = r.z; b = r.w; }
        else { g.w2 = r.z; g.w3 = r.w; a = r.x; b = r.y; }
    }
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) / 9007199254740992.0;
}
// n dimensions whose values nobody reads: the counter moves on, the block function is not run.  Later draws are keyed by their absolute dimension, so
// they are the same numbers either way.
__device__ __forceinline__ void rng_skip(Rng& g, uint32_t n) { g.dim += n; g.w2 = 0xffffffffu; }
template <class R> __device__ __forceinline__ R to_unit(double u);
template <> __device__ __forceinline__ double to_unit<double>(double u) { return u; }
template <> __device__ __forceinline__ float to_unit<float>(double u) { return fminf((float)u, 0.99999994f); }   // keep [0,1)
template <class R> __device__ __forceinline__ R rng_real(Rng& g) { return to_unit<R>(rng_next(g)); }

// ---- small vector algebra in R
template <class R> struct V3 { R x, y, z; };
template <class R> struct V2 { R x, y; };
template <class R> __device__ __forceinline__ V3<R> v3(R x, R y, R z) { V3<R> r; r.x = x; r.y = y; r.z = z; return r; }
template <class R> __device__ __forceinline__ V3<R> cast3(D3 a) { return v3<R>((R)a.x, (R)a.y, (R)a.z); }
template <class R> __device__ __forceinline__ D3 dbl3(V3<R> a) { return mk3((double)a.x, (double)a.y, (double)a.z); }
template <class R> __device__ __forceinline__ V3<R> operator+(V3<R> a, V3<R> b) { return v3<R>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <class R> __device__ __forceinline__ V3<R> operator-(V3<R> a, V3<R> b) { return v3<R>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <class R> __device__ __forceinline__ V3<R> operator-(V3<R> a) { return v3<R>(-a.x, -a.y, -a.z); }
template <class R> __device__ __forceinline__ V3<R> operator*(V3<R> a, R s) { return v3<R>(a.x * s, a.y * s, a.z * s); }
template <class R> __device__ __forceinline__ V3<R> operator*(R s, V3<R> a) { return v3<R>(s * a.x, s * a.y, s * a.z); }
template <class R> __device__ __forceinline__ bool is_black(V3<R> a) { return a.x == R(0) && a.y == R(0) && a.z == R(0); }
template <class R> __device__ __forceinline__ V3<R> cmul(V3<R> a, V3<R> b) { return v3<R>(a.x * b.x, a.y * b.y, a.z * b.z); }
template <class R> __device__ __forceinline__ V3<R> cdiv(V3<R> a, V3<R> b) { return v3<R>(a.x / b.x, a.y / b.y, a.z / b.z); }
template <class R> __device__ __forceinline__ R dot(V3<R> a, V3<R> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <class R> __device__ __forceinline__ V3<R> cross(V3<R> a, V3<R> b) { return v3<R>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
// f32 (production) mode: the library is compiled with -fmad=false for the sake of the f64 expressions, so the f32 vector
// algebra asks for its fused multiply-adds explicitly (fewer instructions, one rounding less; R = double is untouched)
template <> __device__ __forceinline__ float dot<float>(V3<float> a, V3<float> b) { return __fmaf_rn(a.z, b.z, __fmaf_rn(a.y, b.y, a.x * b.x)); }
template <> __device__ __forceinline__ V3<float> cross<float>(V3<float> a, V3<float> b) {
    return v3<float>(__fmaf_rn(a.y, b.z, -(a.z * b.y)), __fmaf_rn(a.z, b.x, -(a.x * b.z)), __fmaf_rn(a.x, b.y, -(a.y * b.x)));
}
template <class R> __device__ __forceinline__ R norm(V3<R> a) { return sqrt(dot(a, a)); }
template <class R> __device__ __forceinline__ V3<R> normalize(V3<R> a) {
    R n = norm(a);
    if constexpr (sizeof(R) == 8) return v3<R>(div_by_norm(a.x, n), div_by_norm(a.y, n), div_by_norm(a.z, n));      // geom.cuh: zero numerators skip the f64 slow path
    else return v3<R>(a.x / n, a.y / n, a.z / n);
}
template <class R> __device__ __forceinline__ R clampr(R x, R lo, R hi) { return fmin(fmax(x, lo), hi); }
template <class R> __device__ __forceinline__ V3<R> face_forward(V3<R> n, V3<R> v) { return dot(n, v) < R(0) ? -n : n; }
template <class R> __device__ __forceinline__ bool finite3(V3<R> a) { return isfinite(a.x) && isfinite(a.y) && isfinite(a.z); }
template <class R> __device__ __forceinline__ R pi() { return (R)PI; }
template <class R> __device__ __forceinline__ R inv_pi() { return (R)(1.0 / PI); }          // consts.rs:41

// ---- RT_SHADE_F32 mode, triangle hits: Mesh::intersects_triangle's record (hittable.rs:362-449) with the same f64 accept
// arithmetic, hit point and uv as geom.cuh:prim_record_t<1> — the next ray starts at that point, textures are looked up at that
// uv — but the shading frame (six normalisations, four cross products: a third of that function's instructions as f64
// division / square-root sequences) in f32.  Its consumers in this mode are f32 anyway: the Bsdf frame and Light::l's sign test.
__device__ __forceinline__ V3<float> nrm3f(V3<float> a) { float s = rsqrtf(__fmaf_rn(a.z, a.z, __fmaf_rn(a.y, a.y, a.x * a.x))); V3<float> r; r.x = a.x * s; r.y = a.y * s; r.z = a.z * s; return r; }
__device__ __forceinline__ V3<float> crs3f(V3<float> a, V3<float> b) {
    V3<float> r; r.x = __fmaf_rn(a.y, b.z, -(a.z * b.y)); r.y = __fmaf_rn(a.z, b.x, -(a.x * b.z)); r.z = __fmaf_rn(a.x, b.y, -(a.y * b.x)); return r;
}
__device__ __forceinline__ V3<float> f3(D3 a) { V3<float> r; r.x = (float)a.x; r.y = (float)a.y; r.z = (float)a.z; return r; }
__device__ __forceinline__ D3 d3(V3<float> a) { return mk3((double)a.x, (double)a.y, (double)a.z); }
static __device__ RT_HELPER2 bool prim_record_tri_f32(const DScene& sc, uint32_t prim, D3 o, D3 d, double tmin, double tmax, Rec& r) {
    const DPrim pr = sc.prims[prim];
    r.prim = prim; r.mat = pr.mat;
    const uint32_t* ti = sc.tri_index + pr.a;
    const uint32_t i1 = ti[0], i2 = ti[1], i3 = ti[2];
    const DMesh ms = sc.meshes[pr.info >> 6];
    const D3 p0 = ld3(sc.points + 3 * (ms.p_base + i1)), p1 = ld3(sc.points + 3 * (ms.p_base + i2)), p2 = ld3(sc.points + 3 * (ms.p_base + i3));
    const RayX rx = make_rayx(o, d);
    double t, b0, b1, b2;
    if (!tri_test(rx, p0, p1, p2, tmax, t, b0, b1, b2)) return false;
    // no leaf-box test here: extend has already accepted this primitive with the reference's full predicate (traverse.cuh)
    D2 uv0 = mk2(0., 0.), uv1 = mk2(1., 0.), uv2 = mk2(1., 1.);
    if (pr.info & PI_HAS_UV) {
        const double* u = sc.uvs; uv0 = mk2(u[2 * (ms.uv_base + i1)], u[2 * (ms.uv_base + i1) + 1]);
        uv1 = mk2(u[2 * (ms.uv_base + i2)], u[2 * (ms.uv_base + i2) + 1]); uv2 = mk2(u[2 * (ms.uv_base + i3)], u[2 * (ms.uv_base + i3) + 1]);
    }
    const D2 duv02 = mk2(uv0.x - uv2.x, uv0.y - uv2.y), duv12 = mk2(uv1.x - uv2.x, uv1.y - uv2.y);
    const D3 dp02 = p0 - p2, dp12 = p1 - p2;
    const double determinant = duv02.x * duv12.y - duv02.y * duv12.x;
    V3<float> dpdu;
    if (fabs(determinant) < SMALL / 10000.) {
        D3 n = cross(p2 - p0, p1 - p0);
        if (norm2(n) == 0.) return false;                    // a hit/miss decision: stays f64
        V3<float> nf = f3(n);                                // util::make_coordinate_system (util.rs:567-576), first vector only
        if (fabsf(nf.x) > fabsf(nf.y)) { float s = rsqrtf(nf.x * nf.x + nf.z * nf.z); dpdu.x = -nf.z * s; dpdu.y = 0.f; dpdu.z = nf.x * s; }
        else { float s = rsqrtf(nf.y * nf.y + nf.z * nf.z); dpdu.x = 0.f; dpdu.y = nf.z * s; dpdu.z = -nf.y * s; }
    } else {
        const double inv = 1. / determinant;
        dpdu = f3((duv12.y * dp02 - duv02.y * dp12) * inv);
    }
    r.p = b0 * p0 + b1 * p1 + b2 * p2;
    r.uv = mk2(b0 * uv0.x + b1 * uv1.x + b2 * uv2.x, b0 * uv0.y + b1 * uv1.y + b2 * uv2.y);
    const V3<float> e02 = f3(dp02), e12 = f3(dp12);
    V3<float> normal;
    if (pr.info & PI_HAS_N) {
        const double* nn = sc.normals;
        normal = f3(b0 * ld3(nn + 3 * (ms.n_base + i1)) + b1 * ld3(nn + 3 * (ms.n_base + i2)) + b2 * ld3(nn + 3 * (ms.n_base + i3)));
    } else normal = crs3f(e02, e12);
    V3<float> gn = nrm3f(crs3f(e02, e12));
    const V3<float> sn = nrm3f(normal);
    V3<float> ss = nrm3f(dpdu);
    V3<float> ts = nrm3f(crs3f(sn, ss));
    if (ts.x * ts.x + ts.y * ts.y + ts.z * ts.z > 0.f) ss = nrm3f(crs3f(ts, sn));
    else {                                                   // degenerate frame: make_coordinate_system(sn)
        V3<float> a;
        if (fabsf(sn.x) > fabsf(sn.y)) { float s = rsqrtf(sn.x * sn.x + sn.z * sn.z); a.x = -sn.z * s; a.y = 0.f; a.z = sn.x * s; }
        else { float s = rsqrtf(sn.y * sn.y + sn.z * sn.z); a.x = 0.f; a.y = sn.z * s; a.z = -sn.y * s; }
        ss = nrm3f(a); ts = nrm3f(crs3f(sn, a));
    }
    const V3<float> n2 = nrm3f(crs3f(ss, ts));               // set_shading_geometry(ss, ts, .., is_auth = true) (hittable.rs:191-210)
    if (gn.x * n2.x + gn.y * n2.y + gn.z * n2.z < 0.f) { gn.x = -gn.x; gn.y = -gn.y; gn.z = -gn.z; }
    const V3<float> df = f3(d);
    r.front = df.x * gn.x + df.y * gn.y + df.z * gn.z < 0.f;   // set_front
    r.sn = d3(n2); r.sdpdu = d3(ss);
    r.n = r.front ? d3(gn) : -d3(gn);
    r.t = t; r.wo = -d;
    if (pr.info & PI_FLIP) r.front = !r.front;               // FlipFace (primitive.rs:300-310)
    return true;
}

// ---- bxdf.rs:12-68
template <class R> __device__ __forceinline__ R cos_sq_theta(V3<R> v) { return v.z * v.z; }
template <class R> __device__ __forceinline__ R abs_cos_theta(V3<R> v) { return fabs(v.z); }
template <class R> __device__ __forceinline__ R sin_sq_theta(V3<R> v) { return fmax(R(0), R(1) - cos_sq_theta(v)); }
template <class R> __device__ __forceinline__ R sin_theta(V3<R> v) { return sqrt(sin_sq_theta(v)); }
template <class R> __device__ __forceinline__ R tan_theta(V3<R> v) { return sin_theta(v) / v.z; }
template <class R> __device__ __forceinline__ R tan_sq_theta(V3<R> v) { return sin_sq_theta(v) / cos_sq_theta(v); }
template <class R> __device__ __forceinline__ R cos_phi(V3<R> v) { R s = sin_theta(v); return s == R(0) ? R(1) : clampr(v.x / s, R(-1), R(1)); }
template <class R> __device__ __forceinline__ R sin_phi(V3<R> v) { R s = sin_theta(v); return s == R(0) ? R(1) : clampr(v.y / s, R(-1), R(1)); }
__device__ __forceinline__ bool matches_flags(uint32_t flag, uint32_t other) { return (flag & other) == flag; }
template <class R> __device__ __forceinline__ bool same_hemisphere(V3<R> v, V3<R> w) { return v.z * w.z > R(0); }              // util.rs:591-593
template <class R> __device__ __forceinline__ V3<R> reflect(V3<R> v, V3<R> n) { R s = R(2) * dot(v, n); return -v + n * s; }   // util.rs:203-206
template <class R> static __device__ RT_HELPER bool refract(V3<R> vec, V3<R> n, R eta, V3<R>& out) {                        // util.rs:376-385
    R cos_theta_i = dot(n, vec) / norm(vec);
    R sin2_theta_i = fmax(R(0), R(1) - cos_theta_i * cos_theta_i);
    R sin2_theta_t = eta * eta * sin2_theta_i;
    if (sin2_theta_t >= R(1)) return false;
    R cos_theta_t = sqrt(R(1) - sin2_theta_t);
    out = eta * (-vec) + (eta * cos_theta_i - cos_theta_t) * n;
    return true;
}
template <class R> static __device__ RT_HELPER R fr_dielectric(R cos_theta_i, R eta_i, R eta_t) {                           // bxdf.rs:113-136
    cos_theta_i = clampr(cos_theta_i, R(-1), R(1));
    R index_i = eta_i, index_t = eta_t;
    if (cos_theta_i < R(0)) { index_i = eta_t; index_t = eta_i; cos_theta_i = fabs(cos_theta_i); }
    R sin_theta_i = sqrt(fmax(R(0), R(1) - cos_theta_i * cos_theta_i));
    R sin_theta_t = index_i / index_t * sin_theta_i;
    R cos_theta_t = sqrt(fmax(R(0), R(1) - sin_theta_t * sin_theta_t));
    if (sin_theta_t >= R(1)) return R(1);
    R r_parl = ((index_t * cos_theta_i) - (index_i * cos_theta_t)) / ((index_t * cos_theta_i) + (index_i * cos_theta_t));
    R r_perp = ((index_i * cos_theta_i) - (index_t * cos_theta_t)) / ((index_i * cos_theta_i) + (index_t * cos_theta_t));
    return (r_parl * r_parl + r_perp * r_perp) / R(2);
}
template <class R> static __device__ RT_HELPER V3<R> fr_conductor(R cos_theta_i, V3<R> eta, V3<R> eta_k) {                  // bxdf.rs:141-170
    cos_theta_i = clampr(cos_theta_i, R(-1), R(1));
    R c2 = cos_theta_i * cos_theta_i, s2 = R(1) - c2;
    V3<R> eta2 = cmul(eta, eta), etak2 = cmul(eta_k, eta_k);
    V3<R> t0 = (eta2 - etak2) - v3<R>(s2, s2, s2);
    V3<R> a2pb2 = cmul(t0, t0) + cmul(eta2, etak2) * R(4);
    a2pb2 = v3<R>(sqrt(a2pb2.x), sqrt(a2pb2.y), sqrt(a2pb2.z));
    V3<R> t1 = a2pb2 + v3<R>(c2, c2, c2);
    V3<R> a = (a2pb2 + t0) * R(0.5);
    // f32 (production) mode only: with k = 0 and eta^2 < sin^2 the exact value is 0, but sqrt(t0^2) from the fast f32 square root can come out an ulp
    // below |t0| and the next sqrt would take the root of -1e-9: clamp.  The f64 kernels keep the reference's expression as it is (bxdf.rs:141-170).
    if constexpr (sizeof(R) == 4) a = v3<R>(fmaxf(a.x, 0.f), fmaxf(a.y, 0.f), fmaxf(a.z, 0.f));
    a = v3<R>(sqrt(a.x), sqrt(a.y), sqrt(a.z));
    V3<R> t2 = a * (R(2) * cos_theta_i);
    V3<R> rs = cdiv(t1 - t2, t1 + t2);
    V3<R> t3 = a2pb2 * c2 + v3<R>(s2 * s2, s2 * s2, s2 * s2);
    V3<R> t4 = t2 * s2;
    V3<R> rp = cmul(rs, cdiv(t3 - t4, t3 + t4));
    return (rp + rs) * R(0.5);
}

// ---- lobes (bxdf.rs:213-300, in-scope subset) kept in registers/local memory: no heap, no Vec<Box<..>>
enum { F_DIELECTRIC = 0, F_CONDUCTOR = 1, F_NOOP = 2, F_DISNEY = 3 };
// every lobe a Material can build (bxdf.rs:213-300); ScaledBxdf and FresnelBlend have constructors but no call site in the reference
enum { L_LAMBERT = 0, L_OREN_NAYAR = 1, L_SPEC_R = 2, L_SPEC_T = 3, L_FRESNEL_SPEC = 4, L_MICRO_R = 5, L_MICRO_T = 6,
       L_DISNEY_DIFFUSE = 7, L_DISNEY_FAKESS = 8, L_DISNEY_RETRO = 9, L_DISNEY_SHEEN = 10, L_DISNEY_CLEARCOAT = 11, L_LAMBERT_T = 12 };
// Compile-time lobe/fresnel masks: the shade kernel is specialised per material type (the bins extend fills), and each
// specialisation only keeps the arms its material can build — small kernels that fit the instruction cache instead of
// one 200 KB kernel (profiles/shade_r01_notes.md).  KM_ALL keeps every arm.
constexpr uint32_t KM_F_SHIFT = 16;
__host__ __device__ constexpr uint32_t km(int k) { return 1u << k; }
__host__ __device__ constexpr uint32_t kmf(int f) { return 1u << (KM_F_SHIFT + f); }
constexpr uint32_t KM_ALL = 0xffffffffu;
// A material whose lobes are only Lambert / Oren-Nayar (Matte, Light) gets its BSDF code inlined into the shade kernel: the code is
// small, and as calls the Bsdf / Lobe structs go through local memory on every call (ncu r01: 14 % of the matte kernel's
// instructions were local loads / stores).  The heavy materials keep the calls: inlined, their kernels overflow the instruction cache.
__host__ __device__ constexpr bool km_small(uint32_t KM) { return SHADE_INLINE_SMALL >= 2 || (SHADE_INLINE_SMALL == 1 && (KM & 0xffffu & ~(km(0) | km(1))) == 0u); }
#define KM_HAS(k) ((KM & km(k)) != 0u)
#define KM_HASF(f) ((KM & kmf(f)) != 0u)
template <class R> struct Lobe {
    int kind; uint32_t type;
    V3<R> color;              // color / r
    V3<R> aux;                // FresnelSpecular t ; conductor k
    V3<R> eta3;               // conductor eta
    R a, b;                   // OrenNayer A,B ; dielectric fresnel eta_i, eta_t ; DisneyFresnel metallic, eta (r0 in eta3) ;
                              // DisneyFakeSS / DisneyRetro roughness (a) ; DisneyClearCoat weight (a), gloss (b)
    int fresnel;
    bool mfd_empty, mfd_disney; R alpha_x, alpha_y;     // mfd_disney: MicrofacetDistribution::DisneyMicrofacet — TrowbridgeReitz with the separable G
    R eta_a, eta_b;
};
// at most 2 lobes for Matte .. Mirror; Material::Disney builds up to 8 (material.rs:245-394) and is shaded by the f64 kernels only
template <class R> struct Bsdf { V3<R> ns, ng, ss, ts; int n; uint32_t mask; bool empty; Lobe<R> lobes[sizeof(R) == 8 ? 8 : 2]; };
// The f32 (production) BSDF never indexes `lobes` with a run-time value: a material's lobes have FIXED slots (Plastic: 0 = Lambert, 1 = microfacet, ...) and
// `mask` says which are present, so walking the present slots in slot order visits exactly the lobes, in exactly the order, that the reference's
// `bxdfs` vector holds (a lobe that is left out there is a cleared bit here).  With static indices and no address handed to an out-of-line function
// the whole Bsdf lives in registers and the constant fields (kind, type, fresnel) fold away; as an array in local memory it was the largest part of the
// shade kernels' local-memory sectors, which outnumbered their global ones (profiles/shade_*_mem_by_source_r02.txt).  Only a Material::Disney BSDF (up to
// 8 lobes, f64 kernels) keeps the dense `n`-long array.
// Static slots apply whenever the kernel's compile-time lobe mask KM rules out Material::Disney (the only material with more than two lobes): every f32
// kernel, and the f64 kernels of Matte .. Mirror.
template <uint32_t KM, class R> __host__ __device__ constexpr bool lobes_static() {
    return sizeof(R) == 4 || (KM & (km(L_DISNEY_DIFFUSE) | km(L_DISNEY_FAKESS) | km(L_DISNEY_RETRO) | km(L_DISNEY_SHEEN) | km(L_DISNEY_CLEARCOAT) | km(L_LAMBERT_T) | kmf(F_DISNEY))) == 0u;
}
template <uint32_t KM, class R> __device__ __forceinline__ Lobe<R>& add_lobe(Bsdf<R>& b, int slot) {
    if constexpr (lobes_static<KM, R>()) { b.mask |= 1u << slot; b.n++; return b.lobes[slot]; }
    else { b.mask |= 1u << b.n; return b.lobes[b.n++]; }       // dense append; `slot` is ignored
}
template <uint32_t KM, class R, class F> __device__ __forceinline__ void for_lobes(const Bsdf<R>& b, F&& fn) {          // fn(lobe, slot), present lobes in order
    if constexpr (lobes_static<KM, R>()) { if (b.mask & 1u) fn(b.lobes[0], 0); if (b.mask & 2u) fn(b.lobes[1], 1); }
    else { for (int i = 0; i < b.n; i++) fn(b.lobes[i], i); }
}
template <uint32_t KM, class R, class F> __device__ __forceinline__ void with_lobe(const Bsdf<R>& b, int slot, F&& fn) {   // fn(lobe) for one slot chosen at run time
    if constexpr (lobes_static<KM, R>()) { if (slot == 0) fn(b.lobes[0]); else fn(b.lobes[1]); }
    else fn(b.lobes[slot]);
}

// helpers of the Disney lobes (bxdf.rs:73-108, util.rs:173-183)
template <class R> __device__ __forceinline__ R lerp_r(R t, R mn, R mx) { return (R(1) - t) * mn + t * mx; }
template <class R> __device__ __forceinline__ R schlick_weight(R u) { const R m = clampr(R(1) - u, R(0), R(1)); const R m2 = m * m; return m2 * m2 * m; }
template <class R> __device__ __forceinline__ V3<R> fr_schlick(V3<R> r0, R cos_theta) { const R w = schlick_weight<R>(cos_theta); return r0 + w * (v3<R>(R(1), R(1), R(1)) - r0); }
template <class R> static __device__ RT_HELPER R gtr1(R cos_theta, R a) {
    if (cos_theta > R(1)) return inv_pi<R>();
    const R a2 = a * a;
    return (a2 - R(1)) / (pi<R>() * log(a2) * (R(1) + (a2 - R(1)) * cos_theta * cos_theta));
}
template <class R> __device__ __forceinline__ R smithg_ggx(R cos_theta, R alpha) { const R a2 = alpha * alpha, c2 = cos_theta * cos_theta; return R(2) / (R(1) + sqrt(a2 + (R(1) - a2) * c2)); }
// Fresnel::evaluate (bxdf.rs:190-211); dielectric arm passes (|cos|, eta_t, eta_i)
template <class R, uint32_t KM> __device__ __forceinline__ V3<R> fresnel_eval(const Lobe<R>& l, R cos_theta_i) {
    if (KM_HASF(F_DIELECTRIC) && l.fresnel == F_DIELECTRIC) { R v = fr_dielectric<R>(fabs(cos_theta_i), l.b, l.a); return v3<R>(v, v, v); }
    if (KM_HASF(F_CONDUCTOR) && l.fresnel == F_CONDUCTOR) return fr_conductor<R>(fabs(cos_theta_i), l.eta3, l.aux);
    if (KM_HASF(F_DISNEY) && l.fresnel == F_DISNEY) {            // bxdf.rs:199-208: no |cos| here; r0 = eta3, metallic = a, eta = b
        const R mn = fr_dielectric<R>(cos_theta_i, R(1), l.b);
        const V3<R> mx = fr_schlick<R>(l.eta3, cos_theta_i);
        return v3<R>(lerp_r<R>(l.a, mn, mx.x), lerp_r<R>(l.a, mn, mx.y), lerp_r<R>(l.a, mn, mx.z));
    }
    return v3<R>(R(1), R(1), R(1));
}
// ---- microfacet.rs, TrowbridgeReitz{sample_visible_area:true} / Empty
// the distribution's four fields travel BY VALUE into the out-of-line helpers: a `const Lobe&` parameter would pin the caller's lobe in local memory
template <class R> struct Mfd { R alpha_x, alpha_y; bool mfd_empty, mfd_disney; };
template <class R> __device__ __forceinline__ Mfd<R> mfd_of(const Lobe<R>& l) { Mfd<R> m; m.alpha_x = l.alpha_x; m.alpha_y = l.alpha_y; m.mfd_empty = l.mfd_empty; m.mfd_disney = l.mfd_disney; return m; }
template <class R> static __device__ RT_HELPER R mfd_d(const Mfd<R> m, V3<R> wh) {                                        // microfacet.rs:53-68
    if (m.mfd_empty) return R(0);
    R t2 = tan_sq_theta(wh);
    if (isinf(t2)) return R(0);
    R cos4 = cos_sq_theta(wh) * cos_sq_theta(wh);
    R cp = cos_phi(wh), sp = sin_phi(wh);
    R e = ((cp * cp) / (m.alpha_x * m.alpha_x) + (sp * sp) / (m.alpha_y * m.alpha_y)) * t2;
    return R(1) / (pi<R>() * m.alpha_x * m.alpha_y * cos4 * (R(1) + e) * (R(1) + e));
}
template <class R> static __device__ RT_HELPER R mfd_lambda(const Mfd<R> m, V3<R> w) {                                    // microfacet.rs:109-123
    if (m.mfd_empty) return R(0);
    R att = fabs(tan_theta(w));
    if (isinf(att)) return R(0);
    R cp = cos_phi(w), sp = sin_phi(w);
    R alpha = sqrt((cp * cp) * m.alpha_x * m.alpha_x + (sp * sp) * m.alpha_y * m.alpha_y);
    R a2t2 = (alpha * att) * (alpha * att);
    return (R(-1) + sqrt(R(1) + a2t2)) / R(2);
}
template <class R> __device__ __forceinline__ R mfd_g1(const Mfd<R> m, V3<R> w) { return R(1) / (R(1) + mfd_lambda(m, w)); }
template <class R> __device__ __forceinline__ R mfd_g(const Mfd<R> m, V3<R> wo, V3<R> wi) {       // microfacet.rs:146-157
    if (sizeof(R) == 8 && m.mfd_disney) return mfd_g1(m, wo) * mfd_g1(m, wi);                       // DisneyMicrofacet: separable
    return R(1) / (R(1) + mfd_lambda(m, wo) + mfd_lambda(m, wi));
}
template <class R> __device__ __forceinline__ R mfd_pdf(const Mfd<R> m, V3<R> wo, V3<R> wh) {                                // microfacet.rs:159-172
    if (!m.mfd_empty) return mfd_d(m, wh) * mfd_g1(m, wo) * fabs(dot(wo, wh)) / abs_cos_theta(wo);
    return mfd_d(m, wh) * abs_cos_theta(wh);
}
// microfacet.rs:442-446 — evaluated in f64 in both modes (material constants, once per vertex)
static __device__ __noinline__ double tr_roughness_to_alpha(double roughness) {
    roughness = fmax(roughness, 1e-5);
    double x = log(roughness);
    return 1.62142 + 0.819955 * x + 0.1734 * x * x + 0.0171201 * x * x * x + 0.000640711 * x * x * x * x;
}
template <class R> static __device__ RT_HELPER V2<R> tr_sample_11(R cos_t, R u1, R u2) {                                    // microfacet.rs:469-512
    V2<R> out;
    if (cos_t > R(0.9999)) { R r = sqrt(u1 / (R(1) - u1)), phi = R(2) * pi<R>() * u2; out.x = r * cos(phi); out.y = r * sin(phi); return out; }
    R sin_t = fmax(R(0), sqrt(R(1) - cos_t * cos_t));
    R tan_t = sin_t / cos_t;
    R a = R(1) / tan_t;
    R g1 = R(2) / (R(1) + sqrt(R(1) + R(1) / (a * a)));
    a = R(2) * u1 / g1 - R(1);
    R tmp = R(1) / (a * a - R(1));
    if (tmp > R(1e10)) tmp = R(1e10);
    R b = tan_t;
    R d = sqrt(fmax(R(0), b * b * tmp * tmp - (a * a - b * b) * tmp));
    R sx1 = b * tmp - d, sx2 = b * tmp + d;
    R slope_x = (a < R(0) || sx2 > R(1) / tan_t) ? sx1 : sx2;
    R s, nu2;
    if (u2 > R(0.5)) { s = R(1); nu2 = R(2) * (u2 - R(0.5)); } else { s = R(-1); nu2 = R(2) * (R(0.5) - u2); }
    R z = (nu2 * (nu2 * (nu2 * R(0.27385) - R(0.73369)) + R(0.46341))) / (nu2 * (nu2 * (nu2 * R(0.093073) + R(0.309420)) - R(1)) + R(0.597999));
    out.x = slope_x; out.y = s * z * sqrt(R(1) + slope_x * slope_x);
    return out;
}
template <class R> static __device__ RT_HELPER V3<R> mfd_sample_wh(const Mfd<R> m, V3<R> wo, V2<R> u) {                   // microfacet.rs:240-283, 448-467
    if (m.mfd_empty) return normalize(v3<R>(R(0), R(0), R(0)));
    bool flip = wo.z < R(0);
    V3<R> wi = flip ? -wo : wo;
    V3<R> wi_s = normalize(v3<R>(m.alpha_x * wi.x, m.alpha_y * wi.y, wi.z));
    V2<R> slope = tr_sample_11<R>(wi_s.z, u.x, u.y);
    R sp = sin_phi(wi_s), cp = cos_phi(wi_s);
    R tmp = cp * slope.x - sp * slope.y;
    R sy = sp * slope.x + cp * slope.y;
    R sx = tmp;
    sx = m.alpha_x * sx; sy = m.alpha_y * sy;
    V3<R> wh = normalize(v3<R>(-sx, -sy, R(1)));
    return flip ? -wh : wh;
}

// Bxdf::f (bxdf.rs:328-527)
template <class R, uint32_t KM> __device__ __forceinline__ V3<R> lobe_f_impl(const Lobe<R>& l, V3<R> wo, V3<R> wi) {
    const V3<R> zero = v3<R>(R(0), R(0), R(0));
    if (KM_HAS(L_LAMBERT) && l.kind == L_LAMBERT) return l.color * inv_pi<R>();
    {
        if (KM_HAS(L_OREN_NAYAR) && l.kind == L_OREN_NAYAR) {
            R sti = sin_theta(wi), sto = sin_theta(wo), max_cos = R(0);
            if (sti > R(SMALL) && sto > R(SMALL)) { R d_cos = cos_phi(wi) * cos_phi(wo) + sin_phi(wi) * sin_phi(wo); max_cos = fmax(R(0), d_cos); }
            R sin_alpha, tan_beta;
            if (abs_cos_theta(wi) > abs_cos_theta(wo)) { sin_alpha = sto; tan_beta = sti / abs_cos_theta(wi); }
            else { sin_alpha = sti; tan_beta = sto / abs_cos_theta(wi); }                   // |cos θi| in both arms (§8.1 #11)
            return l.color * (inv_pi<R>() * (l.a + l.b * max_cos * sin_alpha * tan_beta));
        }
        if (KM_HAS(L_MICRO_R) && l.kind == L_MICRO_R) {
            R cto = abs_cos_theta(wo), cti = abs_cos_theta(wi);
            V3<R> wh = wi + wo;
            if (cti == R(0) || cto == R(0)) return zero;
            if (is_black(wh)) return zero;
            wh = normalize(wh);
            V3<R> f = fresnel_eval<R, KM>(l, dot(wi, face_forward(wh, v3<R>(R(0), R(0), R(1)))));
            V3<R> comp1 = l.color * mfd_d(mfd_of(l), wh) * mfd_g(mfd_of(l), wo, wi);
            return cmul(comp1, f * (R(1) / (R(4) * cti * cto)));
        }
        if (KM_HAS(L_MICRO_T) && l.kind == L_MICRO_T) {
            if (same_hemisphere(wo, wi)) return zero;
            R cto = wo.z, cti = wi.z;
            if (cti == R(0) || cto == R(0)) return zero;
            R eta = wo.z > R(0) ? l.eta_b / l.eta_a : l.eta_a / l.eta_b;
            V3<R> wh = normalize(wo + wi * eta);
            if (wh.z < R(0)) wh = -wh;
            if (dot(wo, wh) * dot(wi, wh) > R(0)) return zero;
            V3<R> f = fresnel_eval<R, KM>(l, dot(wo, wh));
            R sqrt_denom = dot(wo, wh) + eta * dot(wi, wh);
            R factor = R(1) / eta;                                                           // mode == RADIANCE
            V3<R> c = cmul(v3<R>(R(1), R(1), R(1)) - f, l.color);
            return c * fabs(mfd_d(mfd_of(l), wh) * mfd_g(mfd_of(l), wo, wi) * eta * eta * fabs(dot(wi, wh)) * fabs(dot(wo, wh)) * factor * factor /
                            (cti * cto * sqrt_denom * sqrt_denom));
        }
        if (KM_HAS(L_LAMBERT_T) && l.kind == L_LAMBERT_T) return l.color * inv_pi<R>();                    // bxdf.rs:337
        if (KM_HAS(L_DISNEY_DIFFUSE) && l.kind == L_DISNEY_DIFFUSE) {                                      // bxdf.rs:465-470
            const R fo = schlick_weight<R>(abs_cos_theta(wo)), fi = schlick_weight<R>(abs_cos_theta(wi));
            return (inv_pi<R>() * (R(1) - fo / R(2)) * (R(1) - fi / R(2))) * l.color;
        }
        if ((KM_HAS(L_DISNEY_FAKESS) && l.kind == L_DISNEY_FAKESS) || (KM_HAS(L_DISNEY_RETRO) && l.kind == L_DISNEY_RETRO) ||
            (KM_HAS(L_DISNEY_SHEEN) && l.kind == L_DISNEY_SHEEN) || (KM_HAS(L_DISNEY_CLEARCOAT) && l.kind == L_DISNEY_CLEARCOAT)) {
            V3<R> wh = wi + wo;
            if (is_black(wh)) return zero;
            wh = normalize(wh);
            const R cos_theta_d = dot(wi, wh);
            if (l.kind == L_DISNEY_SHEEN) return schlick_weight<R>(cos_theta_d) * l.color;                  // bxdf.rs:505-513
            if (l.kind == L_DISNEY_CLEARCOAT) {                                                            // bxdf.rs:514-526
                const R dr = gtr1<R>(abs_cos_theta(wh), l.b);
                const V3<R> fr = fr_schlick<R>(v3<R>(R(1), R(1), R(1)) * R(0.04), dot(wo, wh));
                const R gr = smithg_ggx<R>(abs_cos_theta(wo), R(0.25)) * smithg_ggx<R>(abs_cos_theta(wi), R(0.25));
                return (l.a / R(4) * gr * dr) * fr;
            }
            const R fo = schlick_weight<R>(abs_cos_theta(wo)), fi = schlick_weight<R>(abs_cos_theta(wi));
            if (l.kind == L_DISNEY_FAKESS) {                                                               // bxdf.rs:471-489
                const R fss90 = cos_theta_d * cos_theta_d * l.a;
                const R fss = lerp_r<R>(fo, R(1), fss90) * lerp_r<R>(fi, R(1), fss90);
                const R ss = R(1.25) * (fss * (R(1) / (abs_cos_theta(wo) + abs_cos_theta(wi)) - R(0.5)) + R(0.5));
                return (inv_pi<R>() * ss) * l.color;
            }
            const R rr = R(2) * l.a * cos_theta_d * cos_theta_d;                                            // DisneyRetro bxdf.rs:490-504
            return (inv_pi<R>() * rr * (fo + fi + fo * fi * (rr - R(1)))) * l.color;
        }
        return zero;
    }
}
template <class R, uint32_t KM> static __device__ __noinline__ V3<R> lobe_f_call(const Lobe<R>& l, V3<R> wo, V3<R> wi) { return lobe_f_impl<R, KM>(l, wo, wi); }
template <class R, uint32_t KM> __device__ __forceinline__ V3<R> lobe_f(const Lobe<R>& l, V3<R> wo, V3<R> wi) { if constexpr (km_small(KM)) { return lobe_f_impl<R, KM>(l, wo, wi); } else { return lobe_f_call<R, KM>(l, wo, wi); } }
template <class R> __device__ __forceinline__ R default_pdf(V3<R> wo, V3<R> wi) { return same_hemisphere(wo, wi) ? abs_cos_theta(wi) * inv_pi<R>() : R(0); }   // bxdf.rs:829-835
// Bxdf::pdf (bxdf.rs:721-813)
template <class R, uint32_t KM> __device__ __forceinline__ R lobe_pdf_impl(const Lobe<R>& l, V3<R> wo, V3<R> wi) {
    {
        if ((KM_HAS(L_LAMBERT) && l.kind == L_LAMBERT) || (KM_HAS(L_OREN_NAYAR) && l.kind == L_OREN_NAYAR) || (KM_HAS(L_SPEC_R) && l.kind == L_SPEC_R)) return default_pdf(wo, wi);
        if ((KM_HAS(L_DISNEY_DIFFUSE) && l.kind == L_DISNEY_DIFFUSE) || (KM_HAS(L_DISNEY_FAKESS) && l.kind == L_DISNEY_FAKESS) ||
            (KM_HAS(L_DISNEY_RETRO) && l.kind == L_DISNEY_RETRO) || (KM_HAS(L_DISNEY_SHEEN) && l.kind == L_DISNEY_SHEEN)) return default_pdf(wo, wi);   // bxdf.rs:792-795
        if (KM_HAS(L_LAMBERT_T) && l.kind == L_LAMBERT_T) return !same_hemisphere(wo, wi) ? abs_cos_theta(wi) * inv_pi<R>() : R(0);                   // bxdf.rs:726-732
        if (KM_HAS(L_DISNEY_CLEARCOAT) && l.kind == L_DISNEY_CLEARCOAT) {                                                                              // bxdf.rs:796-811
            if (!same_hemisphere(wo, wi)) return R(0);
            V3<R> wh = wi + wo;
            if (is_black(wh)) return R(0);
            wh = normalize(wh);
            return gtr1<R>(abs_cos_theta(wh), l.b) * abs_cos_theta(wh) / (R(4) * dot(wo, wh));
        }
        if (KM_HAS(L_MICRO_R) && l.kind == L_MICRO_R) {
            if (!same_hemisphere(wo, wi)) return R(0);
            V3<R> wh = normalize(wo + wi);
            return mfd_pdf(mfd_of(l), wo, wh) / (R(4) * dot(wo, wh));
        }
        if (KM_HAS(L_MICRO_T) && l.kind == L_MICRO_T) {
            if (same_hemisphere(wo, wi)) return R(0);
            R eta = wo.z > R(0) ? l.eta_b / l.eta_a : l.eta_a / l.eta_b;
            V3<R> wh = normalize(wo + wi * eta);
            if (dot(wo, wh) * dot(wi, wh) > R(0)) return R(0);
            R sqrt_denom = dot(wo, wh) + eta * dot(wi, wh);
            R dwh_dwi = fabs(eta * eta * dot(wi, wh)) / (sqrt_denom * sqrt_denom);
            return mfd_pdf(mfd_of(l), wo, wh) * dwh_dwi;
        }
        return R(0);     // SpecularTransmission: 0 ; FresnelSpecular: "unimplemented" → 0 (bxdf.rs:777-791)
    }
}
template <class R, uint32_t KM> static __device__ __noinline__ R lobe_pdf_call(const Lobe<R>& l, V3<R> wo, V3<R> wi) { return lobe_pdf_impl<R, KM>(l, wo, wi); }
template <class R, uint32_t KM> __device__ __forceinline__ R lobe_pdf(const Lobe<R>& l, V3<R> wo, V3<R> wi) { if constexpr (km_small(KM)) { return lobe_pdf_impl<R, KM>(l, wo, wi); } else { return lobe_pdf_call<R, KM>(l, wo, wi); } }
// util::rand_cosine_dir (util.rs:127-148)
template <class R> static __device__ RT_HELPER V3<R> cosine_dir_of(R r1, R r2) {
    R u1 = R(2) * r1 - R(1), u2 = R(2) * r2 - R(1);
    if (u1 == R(0) && u2 == R(0)) return v3<R>(R(0), R(0), R(1));
    R theta, r;
    if (fabs(u1) > fabs(u2)) { r = u1; theta = pi<R>() / R(4) * (u2 / u1); }
    else { r = u2; theta = pi<R>() / R(2) - pi<R>() / R(4) * (u1 / u2); }
    R x = r * cos(theta), y = r * sin(theta);
    return v3<R>(x, y, sqrt(fmax(R(0), R(1) - x * x - y * y)));
}
template <class R> __device__ __forceinline__ V3<R> rand_cosine_dir(Rng& g) { const R r1 = rng_real<R>(g), r2 = rng_real<R>(g); return cosine_dir_of<R>(r1, r2); }   // the draws stay in the caller: no Rng&
// Bxdf::sample_f (bxdf.rs:532-719)
template <class R, uint32_t KM> __device__ __forceinline__ void lobe_sample_f_impl(const Lobe<R>& l, V3<R> wo, V2<R> sample, Rng& g, V3<R>& color, V3<R>& wi, R& pdf) {
    const V3<R> zero = v3<R>(R(0), R(0), R(0)), one = v3<R>(R(1), R(1), R(1));
    color = zero; wi = zero; pdf = R(0);
    {
        if (KM_HAS(L_LAMBERT_T) && l.kind == L_LAMBERT_T) {            // bxdf.rs:709-717: util::cosine_sample_hemisphere of the SUPPLIED sample (util.rs:69-97)
            const R ux = R(2) * sample.x - R(1), uy = R(2) * sample.y - R(1);
            R dx = R(0), dy = R(0);
            if (!(ux == R(0) && uy == R(0))) {
                R theta, r;
                if (fabs(ux) > fabs(uy)) { r = ux; theta = pi<R>() / R(4) * (uy / ux); }
                else { r = uy; theta = pi<R>() / R(2) - pi<R>() / R(4) * (ux / uy); }
                dx = r * cos(theta); dy = r * sin(theta);
            }
            wi = v3<R>(dx, dy, sqrt(fmax(R(0), R(1) - dx * dx - dy * dy)));
            if (wi.z > R(0)) wi.z = -wi.z;
            pdf = lobe_pdf<R, KM>(l, wo, wi); color = lobe_f<R, KM>(l, wo, wi); return;
        }
        if (KM_HAS(L_DISNEY_CLEARCOAT) && l.kind == L_DISNEY_CLEARCOAT) {       // bxdf.rs:689-708
            if (wo.z == R(0)) return;
            const R alpha2 = l.b * l.b;
            const R cos_t = sqrt(fmax(R(0), (R(1) - pow(alpha2, R(1) - sample.x)) / (R(1) - alpha2)));
            const R sin_t = sqrt(fmax(R(0), R(1) - cos_t * cos_t));
            const R phi = R(2) * pi<R>() * sample.y;
            V3<R> wh = v3<R>(sin_t * cos(phi), sin_t * sin(phi), cos_t);         // util::make_spherical
            if (!same_hemisphere(wo, wh)) wh = -wh;
            const V3<R> w = reflect(wo, normalize(wh));
            if (!same_hemisphere(wo, w)) return;
            wi = w; color = lobe_f<R, KM>(l, wo, wi); pdf = lobe_pdf<R, KM>(l, wo, wi); return;
        }
        if ((KM_HAS(L_LAMBERT) && l.kind == L_LAMBERT) || (KM_HAS(L_OREN_NAYAR) && l.kind == L_OREN_NAYAR) ||
            (KM_HAS(L_DISNEY_DIFFUSE) && l.kind == L_DISNEY_DIFFUSE) || (KM_HAS(L_DISNEY_FAKESS) && l.kind == L_DISNEY_FAKESS) ||
            (KM_HAS(L_DISNEY_RETRO) && l.kind == L_DISNEY_RETRO) || (KM_HAS(L_DISNEY_SHEEN) && l.kind == L_DISNEY_SHEEN)) {     // default_sample_f (bxdf.rs:815-827): ignores `sample`
            wi = rand_cosine_dir<R>(g);
            if (wo.z < R(0)) wi.z *= R(-1);
            pdf = lobe_pdf<R, KM>(l, wo, wi); color = lobe_f<R, KM>(l, wo, wi); return;
        }
        if (KM_HAS(L_SPEC_R) && l.kind == L_SPEC_R) {
            wi = v3<R>(-wo.x, -wo.y, wo.z);
            color = cmul(l.color, fresnel_eval<R, KM>(l, wi.z)); pdf = R(1); return;
        }
        if (KM_HAS(L_SPEC_T) && l.kind == L_SPEC_T) {
            bool entering = wo.z > R(0);
            R eta_i = entering ? l.eta_a : l.eta_b, eta_t = entering ? l.eta_b : l.eta_a;
            V3<R> dir;
            if (!refract<R>(wo, face_forward(v3<R>(R(0), R(0), R(1)), wo), eta_i / eta_t, dir)) return;
            V3<R> ft = cmul(one - fresnel_eval<R, KM>(l, dir.z), l.color);
            ft = ft * ((eta_i * eta_i) / (eta_t * eta_t));
            color = ft; wi = dir; pdf = R(1); return;
        }
        if (KM_HAS(L_MICRO_R) && l.kind == L_MICRO_R) {
            if (wo.z == R(0)) return;
            V3<R> wh = mfd_sample_wh<R>(mfd_of(l), wo, sample);   // the wo·wh<0 guard has no `return` (bxdf.rs:598-600)
            V3<R> w = reflect(wo, wh);
            if (!same_hemisphere(wo, w)) return;
            pdf = mfd_pdf(mfd_of(l), wo, wh) / (R(4) * dot(wo, wh));
            wi = w; color = lobe_f<R, KM>(l, wo, wi); return;
        }
        if (KM_HAS(L_MICRO_T) && l.kind == L_MICRO_T) {
            if (wo.z == R(0)) return;
            V3<R> wh = mfd_sample_wh<R>(mfd_of(l), wo, sample);
            if (dot(wo, wh) < R(0)) return;
            R eta = wo.z > R(0) ? l.eta_a / l.eta_b : l.eta_b / l.eta_a;
            V3<R> w;
            if (!refract<R>(wo, wh, eta, w)) return;
            wi = w; pdf = lobe_pdf<R, KM>(l, wo, wi); color = lobe_f<R, KM>(l, wo, wi); return;
        }
        if (KM_HAS(L_FRESNEL_SPEC) && l.kind == L_FRESNEL_SPEC) {
            R f = fr_dielectric<R>(wo.z / norm(wo), l.eta_a, l.eta_b);
            if (sample.x < f) { wi = v3<R>(-wo.x, -wo.y, wo.z); pdf = f; color = l.color * f; return; }
            bool entering = wo.z > R(0);
            R eta_i = entering ? l.eta_a : l.eta_b, eta_t = entering ? l.eta_b : l.eta_a;
            V3<R> dir;
            if (!refract<R>(wo, face_forward(v3<R>(R(0), R(0), R(1)), wo), eta_i / eta_t, dir)) return;
            V3<R> ft = l.aux * (R(1) - f);
            ft = ft * ((eta_i * eta_i) / (eta_t * eta_t));
            color = ft; wi = dir; pdf = R(1) - f; return;
        }
    }
}
template <class R, uint32_t KM> static __device__ __noinline__ void lobe_sample_f_call(const Lobe<R>& l, V3<R> wo, V2<R> sample, Rng& g, V3<R>& color, V3<R>& wi, R& pdf) { lobe_sample_f_impl<R, KM>(l, wo, sample, g, color, wi, pdf); }
template <class R, uint32_t KM> __device__ __forceinline__ void lobe_sample_f(const Lobe<R>& l, V3<R> wo, V2<R> sample, Rng& g, V3<R>& color, V3<R>& wi, R& pdf) { if constexpr (km_small(KM)) { lobe_sample_f_impl<R, KM>(l, wo, sample, g, color, wi, pdf); } else { lobe_sample_f_call<R, KM>(l, wo, sample, g, color, wi, pdf); } }

// ---- bsdf.rs
template <class R> __device__ __forceinline__ V3<R> world_to_local(const Bsdf<R>& b, V3<R> v) { return v3<R>(dot(v, b.ss), dot(v, b.ts), dot(v, b.ns)); }
template <class R> __device__ __forceinline__ V3<R> local_to_world(const Bsdf<R>& b, V3<R> v) {
    return v3<R>(b.ss.x * v.x + b.ts.x * v.y + b.ns.x * v.z, b.ss.y * v.x + b.ts.y * v.y + b.ns.y * v.z, b.ss.z * v.x + b.ts.z * v.y + b.ns.z * v.z);
}
template <> __device__ __forceinline__ V3<float> local_to_world<float>(const Bsdf<float>& b, V3<float> v) {
    return v3<float>(__fmaf_rn(b.ns.x, v.z, __fmaf_rn(b.ts.x, v.y, b.ss.x * v.x)), __fmaf_rn(b.ns.y, v.z, __fmaf_rn(b.ts.y, v.y, b.ss.y * v.x)),
                     __fmaf_rn(b.ns.z, v.z, __fmaf_rn(b.ts.z, v.y, b.ss.z * v.x)));
}
template <uint32_t KM, class R> __device__ __forceinline__ uint32_t num_components(const Bsdf<R>& b, uint32_t t) { uint32_t c = 0; for_lobes<KM>(b, [&](const Lobe<R>& l, int) { if (matches_flags(l.type, t)) c++; }); return c; }
// Bsdf::f (bsdf.rs:83-98), `a && b || c` precedence kept (§8.1 #6)
template <class R, uint32_t KM> __device__ __forceinline__ V3<R> bsdf_f_impl(const Bsdf<R>& b, V3<R> wow, V3<R> wiw, uint32_t flags) {
    V3<R> wi = world_to_local(b, wiw), wo = world_to_local(b, wow);
    bool refl = dot(wiw, b.ng) * dot(wow, b.ng) > R(0);
    V3<R> f = v3<R>(R(0), R(0), R(0));
    for_lobes<KM>(b, [&](const Lobe<R>& l, int) {
        const uint32_t ty = l.type;
        if ((matches_flags(ty, flags) && (refl && (ty & BSDF_REFLECTION) > 0)) || (!refl && (ty & BSDF_TRANSMISSION) > 0)) f = f + lobe_f<R, KM>(l, wo, wi);
    });
    return f;
}
template <class R, uint32_t KM> static __device__ __noinline__ V3<R> bsdf_f_call(const Bsdf<R>& b, V3<R> wow, V3<R> wiw, uint32_t flags) { return bsdf_f_impl<R, KM>(b, wow, wiw, flags); }
template <class R, uint32_t KM> __device__ __forceinline__ V3<R> bsdf_f(const Bsdf<R>& b, V3<R> wow, V3<R> wiw, uint32_t flags) { if constexpr (km_small(KM)) { return bsdf_f_impl<R, KM>(b, wow, wiw, flags); } else { return bsdf_f_call<R, KM>(b, wow, wiw, flags); } }
// Bsdf::pdf (bsdf.rs:166-189): sums, does not average (§8.1 #8); wo is NOT normalised here
template <class R, uint32_t KM> __device__ __forceinline__ R bsdf_pdf_impl(const Bsdf<R>& b, V3<R> wow, V3<R> wiw, uint32_t flags) {
    if (b.n == 0) return R(0);
    V3<R> wo = world_to_local(b, wow), wi = world_to_local(b, wiw);
    if (wo.z == R(0)) return R(0);
    R pdf = R(0); int matching = 0;
    for_lobes<KM>(b, [&](const Lobe<R>& l, int) { if (matches_flags(l.type, flags)) { matching++; pdf += lobe_pdf<R, KM>(l, wo, wi); } });
    return matching > 0 ? pdf : R(0);
}
template <class R, uint32_t KM> static __device__ __noinline__ R bsdf_pdf_call(const Bsdf<R>& b, V3<R> wow, V3<R> wiw, uint32_t flags) { return bsdf_pdf_impl<R, KM>(b, wow, wiw, flags); }
template <class R, uint32_t KM> __device__ __forceinline__ R bsdf_pdf(const Bsdf<R>& b, V3<R> wow, V3<R> wiw, uint32_t flags) { if constexpr (km_small(KM)) { return bsdf_pdf_impl<R, KM>(b, wow, wiw, flags); } else { return bsdf_pdf_call<R, KM>(b, wow, wiw, flags); } }
// Bsdf::sample_f (bsdf.rs:102-164)
template <class R, uint32_t KM> __device__ __forceinline__ void bsdf_sample_f_impl(const Bsdf<R>& b, V3<R> wow, V2<R> sample, uint32_t bxdf_type, Rng& g, V3<R>& color, V3<R>& wiw, R& pdf, uint32_t& sampled) {
    const V3<R> zero = v3<R>(R(0), R(0), R(0));
    color = zero; wiw = zero; pdf = R(0); sampled = 0;
    uint32_t matching = num_components<KM>(b, bxdf_type);
    if (matching == 0) return;
    uint32_t comp = min((uint32_t)floor(sample.x * (R)matching), matching - 1);
    uint32_t count = comp; int used = -1;
    for_lobes<KM>(b, [&](const Lobe<R>& l, int i) { if (used < 0 && matches_flags(l.type, bxdf_type)) { if (count == 0) used = i; else count--; } });
    if (used < 0) used = 0;                        // cannot happen (comp < matching); the reference's `used` starts at 0
    V3<R> wo = normalize(world_to_local(b, wow));
    if (wo.z == R(0)) return;
    V3<R> c, wi; R p; uint32_t bx_type = 0;
    with_lobe<KM>(b, used, [&](const Lobe<R>& bx) { bx_type = bx.type; lobe_sample_f<R, KM>(bx, wo, sample, g, c, wi, p); });
    if (p == R(0)) return;
    V3<R> w = local_to_world(b, wi);
    bool spec = (bx_type & BSDF_SPECULAR) != 0;
    if (!spec && matching > 1)
        for_lobes<KM>(b, [&](const Lobe<R>& l, int i) { if (i != used && matches_flags(l.type, bxdf_type)) p += lobe_pdf<R, KM>(l, wo, wi); });
    if (matching > 1) p = p / (R)matching;
    if (!spec) {
        bool refl = dot(w, b.ng) * dot(wow, b.ng) > R(0);
        c = zero;
        for_lobes<KM>(b, [&](const Lobe<R>& l, int) {
            const uint32_t ty = l.type;
            if (matches_flags(ty, bxdf_type) && ((refl && (ty & BSDF_REFLECTION) > 0) || (!refl && (ty & BSDF_TRANSMISSION) > 0))) c = c + lobe_f<R, KM>(l, wo, wi);
        });
    }
    color = c; wiw = w; pdf = p; sampled = bx_type;
}
template <class R, uint32_t KM> static __device__ __noinline__ void bsdf_sample_f_call(const Bsdf<R>& b, V3<R> wow, V2<R> sample, uint32_t bxdf_type, Rng& g, V3<R>& color, V3<R>& wiw, R& pdf, uint32_t& sampled) { bsdf_sample_f_impl<R, KM>(b, wow, sample, bxdf_type, g, color, wiw, pdf, sampled); }
template <class R, uint32_t KM> __device__ __forceinline__ void bsdf_sample_f(const Bsdf<R>& b, V3<R> wow, V2<R> sample, uint32_t bxdf_type, Rng& g, V3<R>& color, V3<R>& wiw, R& pdf, uint32_t& sampled) { if constexpr (km_small(KM)) { bsdf_sample_f_impl<R, KM>(b, wow, sample, bxdf_type, g, color, wiw, pdf, sampled); } else { bsdf_sample_f_call<R, KM>(b, wow, sample, bxdf_type, g, color, wiw, pdf, sampled); } }

// ---- material.rs
// Texture::value (material.rs:542-589): SolidColor, Checkered (ids may nest), Hdr, Image.  Always f64: the checker's
// sin(frequency*u*2π) argument reaches ~3e4 rad in plastic_dragon and decides which colour a point gets.
__device__ __forceinline__ uint32_t sat_u32(double x) { return !(x > 0.) ? 0u : (x >= 4294967295. ? 4294967295u : (uint32_t)x); }   // Rust `f64 as u32`
static __device__ RT_HELPER2 D3 texel_value(const DScene& sc, const rt_texture& t, double u, double v) {
    const DImage im = sc.images[t.image];
    if (t.type == RT_TEX_HDR) {              // nearest texel, RGBE8 re-quantised at upload, (c+0.5)*2^(e-128)/256 (material.rs:570-587)
        uint32_t x = sat_u32(round((1. - u) * (double)im.w)) % im.w, y = sat_u32(round(v * (double)im.h)) % im.h;
        uint32_t q = __ldg(sc.texels + im.offset + (size_t)y * im.w + x);
        double scale = __hiloint2double((int)(((q >> 24) - 128u + 1023u) << 20), 0);     // 2^(e-128), exact
        return mk3(((double)(q & 255u) + 0.5) * scale / 256., ((double)((q >> 8) & 255u) + 0.5) * scale / 256., ((double)((q >> 16) & 255u) + 0.5) * scale / 256.);
    }
    uint32_t x = sat_u32(round(u * (double)im.w)) % im.w, y = sat_u32(round((1. - v) * (double)im.h)) % im.h;      // material.rs:601-611
    uint32_t q = __ldg(sc.texels + im.offset + (size_t)y * im.w + x);
    return mk3((double)(q & 255u) / 255., (double)((q >> 8) & 255u) / 255., (double)((q >> 16) & 255u) / 255.);
}
// Sign of sin(x) without evaluating it: k = round(x/π), r = x − kπ with π split into three doubles (Cody–Waite; the
// reduction error is ~|k|·1e-33).  Returns +1 / −1, or 0 when r is too close to a zero of sin (or x too large) to be sure —
// the caller then evaluates the real expression.  The checker texture only needs `sin(a)·sin(b) < 0` (material.rs:556-561),
// and two f64 sin() were 13 % of the floor's shade kernel (ncu r01).
__device__ __forceinline__ int sin_sign(double x) {
    if (!(fabs(x) < 1e9)) return 0;
    const double k = rint(x * 0.31830988618379067154);
    double r = fma(-k, 3.141592653589793116, x);
    r = fma(-k, 1.2246467991473532072e-16, r);
    r = fma(-k, -2.9947698097183396659e-33, r);
    if (!(fabs(r) > 1e-9)) return 0;
    const bool odd = ((long long)k) & 1ll;
    return ((r < 0.) != odd) ? -1 : 1;
}
static __device__ RT_HELPER2 D3 tex_value(const DScene& sc, uint32_t id, double u, double v) {
    for (int depth = 0; depth < 8; depth++) {
        const rt_texture& t = sc.textures[id];
        if (t.type == RT_TEX_SOLID) return mk3(t.color[0], t.color[1], t.color[2]);
        if (t.type != RT_TEX_CHECKERED) return texel_value(sc, t, u, v);
        const double a = t.frequency * u * 2. * PI, b = t.frequency * v * 2. * PI;
        const int sa = sin_sign(a), sb = sin_sign(b);
        bool neg;
        if (sa != 0 && sb != 0) neg = sa != sb;               // both sines are safely non-zero: their product is negative iff the signs differ
        else { double mult = sin(a) * sin(b); neg = mult < 0.; }
        id = neg ? t.even : t.odd;
    }
    return mk3(0., 0., 0.);
}
// Texture::Perlin (material.rs:566-569) over perlin::Perlin::{get_noise, turb} (perlin.rs:36-100).  `x.floor() as usize` saturates:
// negative coordinates index lattice cell 0 while the fractional part still wraps into [0, 1).
__device__ __forceinline__ uint32_t sat_index(double x) { return !(x > 0.) ? 0u : (x >= 1.8446744073709552e19 ? 0xffffffffu : (uint32_t)((unsigned long long)x & 0xffffffffull)); }
static __device__ __noinline__ double perlin_noise(const double* __restrict__ pl, D3 p) {
    const double scale = pl[0];
    const double* px = pl + 1; const double* py = pl + 257; const double* pz = pl + 513; const double* vec = pl + 769;
    const double x = p.x * scale, y = p.y * scale, z = p.z * scale;
    const uint32_t i0 = sat_index(floor(x)), j0 = sat_index(floor(y)), k0 = sat_index(floor(z));
    double u = x - trunc(x), v = y - trunc(y), w = z - trunc(z);
    if (u < 0.) u += 1.;
    if (v < 0.) v += 1.;
    if (w < 0.) w += 1.;
    u = u * u * (3. - 2. * u); v = v * v * (3. - 2. * v); w = w * w * (3. - 2. * w);
    double accum = 0.;
    // the reference fills c[di + 2 dj + 4 dk] and then reads c[i + 2 j + 4 k] with weight (u - i, v - j, w - k) built from the SMOOTHED u, v, w
    for (int i = 0; i < 2; i++)
        for (int j = 0; j < 2; j++)
            for (int k = 0; k < 2; k++) {
                const uint32_t index = (uint32_t)px[(i0 + i) & 255u] ^ (uint32_t)py[(j0 + j) & 255u] ^ (uint32_t)pz[(k0 + k) & 255u];
                const double cx = vec[3 * index], cy = vec[3 * index + 1], cz = vec[3 * index + 2];
                const double wx = u - (double)i, wy = v - (double)j, wz = w - (double)k;
                const double i_part = (double)i * u + (double)(1 - i) * (1. - u);
                const double j_part = (double)j * v + (double)(1 - j) * (1. - v);
                const double k_part = (double)k * w + (double)(1 - k) * (1. - w);
                accum += i_part * j_part * k_part * (cx * wx + cy * wy + cz * wz);
            }
    return accum;
}
static __device__ __noinline__ D3 perlin_value(const DScene& sc, const rt_texture& t, D3 p) {
    const double* pl = sc.perlin + (size_t)t.image * PERLIN_STRIDE;
    double accum = 0., weight = 1.; D3 q = p;
    for (int d = 0; d < 7; d++) { accum += weight * perlin_noise(pl, q); weight *= 0.5; q = mk3(q.x * 2., q.y * 2., q.z * 2.); }
    const double c = 0.5 * (1. + sin(pl[0] * p.z + 10. * fabs(accum)));
    return mk3(c, c, c);
}
// Texture::value with the hit point: only scenes that hold a Texture::Perlin take this path (DScene::has_perlin, uniform per launch)
static __device__ __noinline__ D3 tex_value_point(const DScene& sc, uint32_t id, double u, double v, D3 p) {
    for (int depth = 0; depth < 8; depth++) {
        const rt_texture& t = sc.textures[id];
        if (t.type == RT_TEX_PERLIN) return perlin_value(sc, t, p);
        if (t.type != RT_TEX_CHECKERED) return tex_value(sc, id, u, v);
        const double mult = sin(t.frequency * u * 2. * PI) * sin(t.frequency * v * 2. * PI);
        id = mult < 0. ? t.even : t.odd;
    }
    return mk3(0., 0., 0.);
}
// POINT = false (the f32 production kernels): the Perlin arm is not compiled in — present, it cost the dragon scene 2 % of its shade time
// (30.29 vs 29.67 ms at 64 spp, r01) — and rt_render shades scenes that hold a Texture::Perlin with the f64 kernels instead.
template <bool POINT> __device__ __forceinline__ D3 tex_value_at(const DScene& sc, uint32_t id, double u, double v, const D3& p) {
    if constexpr (POINT) { if (sc.has_perlin) return tex_value_point(sc, id, u, v, p); }
    return tex_value(sc, id, u, v);
}
template <class R> __device__ __forceinline__ void lobe_init(Lobe<R>& l, int kind, uint32_t type, D3 color) {
    const V3<R> zero = v3<R>(R(0), R(0), R(0));
    l.kind = kind; l.type = type; l.color = cast3<R>(color); l.aux = zero; l.eta3 = zero; l.a = R(0); l.b = R(0);
    l.fresnel = F_NOOP; l.mfd_empty = true; l.mfd_disney = false; l.alpha_x = R(0); l.alpha_y = R(0); l.eta_a = R(1); l.eta_b = R(1);
}
template <class R> __device__ __forceinline__ void bsdf_new(Bsdf<R>& b, const Rec& r) {                 // bsdf.rs:26-35
    b.ns = cast3<R>(r.sn); b.ng = cast3<R>(r.n); b.ss = cast3<R>(r.sdpdu); b.empty = false;
    if constexpr (sizeof(R) == 4) b.ts = nrm3f(crs3f(b.ns, b.ss));        // f32 mode: the frame is consumed in f32
    else b.ts = cast3<R>(normalize(cross(r.sn, r.sdpdu)));
}
// Material::compute_scattering (material.rs:80-244), mode = RADIANCE, allow_lobes = true
// lobe / fresnel arms a material type can build (material.rs:80-244); MAT < 0 = any
template <int MAT> __host__ __device__ constexpr uint32_t mat_km() {
    return MAT == RT_MAT_MATTE ? (km(L_LAMBERT) | km(L_OREN_NAYAR))
         : MAT == RT_MAT_LIGHT ? 0u
         : MAT == RT_MAT_PLASTIC ? (km(L_LAMBERT) | km(L_MICRO_R) | kmf(F_DIELECTRIC))
         : MAT == RT_MAT_GLASS ? (km(L_FRESNEL_SPEC) | km(L_MICRO_R) | km(L_MICRO_T) | kmf(F_DIELECTRIC))
         : MAT == RT_MAT_METAL ? (km(L_MICRO_R) | kmf(F_CONDUCTOR))
         : MAT == RT_MAT_MIRROR ? (km(L_SPEC_R) | kmf(F_NOOP))
         : MAT == RT_MAT_DISNEY ? (km(L_DISNEY_DIFFUSE) | km(L_DISNEY_FAKESS) | km(L_DISNEY_RETRO) | km(L_DISNEY_SHEEN) | km(L_DISNEY_CLEARCOAT) | km(L_LAMBERT_T) |
                                   km(L_SPEC_T) | km(L_MICRO_R) | km(L_MICRO_T) | kmf(F_DISNEY) | kmf(F_DIELECTRIC))
         : KM_ALL;
}
// Materials whose BSDF holds at most ONE lobe and whose lobe ignores the 2-D sample handed to Bsdf::sample_f: Matte (Lambert / Oren-Nayar sample through
// default_sample_f = rand_cosine_dir's own two draws, bxdf.rs:815-827; the component pick floor(u.x * 1) is 0) and Mirror (SpecularReflection, bxdf.rs:571-578).
// The shade kernels of these materials advance the dimension counter over the two sample pairs of a vertex without running Philox for them.
template <int MAT> __host__ __device__ constexpr bool mat_ignores_sample2d() { return MAT == RT_MAT_MATTE || MAT == RT_MAT_MIRROR; }
template <class R, int MAT> __device__ __forceinline__ void compute_scattering_impl(const DScene& sc, const Rec& r, Bsdf<R>& b) {
    V3<R> e = cast3<R>(normalize(mk3(1., 1., 1.)));    // Bsdf::empty frame (bsdf.rs:37-47)
    b.ns = e; b.ng = e; b.ss = e; b.ts = e; b.n = 0; b.mask = 0u; b.empty = true;
    const rt_material& m = sc.materials[r.mat];
    double u = r.uv.x, v = r.uv.y;
    const uint32_t mt = MAT >= 0 ? (uint32_t)MAT : m.type;
    {
        if (mt == RT_MAT_MATTE) {
            double sig = clampd(m.f[0], 0., 90.);
            D3 color = tex_value_at<sizeof(R) == 8>(sc, m.tex[0], u, v, r.p);
            if (!is_black(color)) {
                bsdf_new(b, r);
                Lobe<R>& l = add_lobe<mat_km<MAT>()>(b, 0);
                if (sig == 0.) lobe_init(l, L_LAMBERT, BSDF_REFLECTION | BSDF_DIFFUSE, color);
                else {
                    lobe_init(l, L_OREN_NAYAR, BSDF_REFLECTION | BSDF_DIFFUSE, color);     // bxdf.rs:909-921
                    double sr = sig * PI / 180.0, s2 = sr * sr;
                    l.a = (R)(1. - (s2 / (2. * (s2 + 0.33)))); l.b = (R)(0.45 * s2 / (s2 + 0.09));
                }
            }
            return;
        }
        if (mt == RT_MAT_LIGHT) return;
        if (mt == RT_MAT_PLASTIC) {
            D3 color = tex_value_at<sizeof(R) == 8>(sc, m.tex[0], u, v, r.p);
            if (!is_black(color)) { bsdf_new(b, r); lobe_init(add_lobe<mat_km<MAT>()>(b, 0), L_LAMBERT, BSDF_REFLECTION | BSDF_DIFFUSE, color); }
            D3 spec = tex_value_at<sizeof(R) == 8>(sc, m.tex[1], u, v, r.p);
            if (!is_black(spec)) {
                if (b.empty) bsdf_new(b, r);
                double rough = sc.mat_alpha[2 * r.mat];                                     // constant per material: remapped at commit
                if (rough != rough) { rough = m.f[0]; if (m.remap_roughness) rough = tr_roughness_to_alpha(rough); }
                Lobe<R>& l = add_lobe<mat_km<MAT>()>(b, 1);
                lobe_init(l, L_MICRO_R, BSDF_REFLECTION | BSDF_GLOSSY, spec);
                l.fresnel = F_DIELECTRIC; l.a = R(1.5); l.b = R(1);                        // eta_i, eta_t
                l.mfd_empty = false; l.alpha_x = (R)fmax(rough, 1e-3); l.alpha_y = (R)fmax(rough, 1e-3);
            }
            return;
        }
        if (mt == RT_MAT_GLASS) {
            double eta = m.f[2], urough = m.f[0], vrough = m.f[1];
            D3 rr = tex_value_at<sizeof(R) == 8>(sc, m.tex[0], u, v, r.p), t = tex_value_at<sizeof(R) == 8>(sc, m.tex[1], u, v, r.p);
            bsdf_new(b, r);
            if (is_black(rr) && is_black(t)) return;
            bool is_specular = urough == 0. && vrough == 0.;
            if (is_specular) {                                                             // allow_lobes = true
                Lobe<R>& l = add_lobe<mat_km<MAT>()>(b, 0);
                lobe_init(l, L_FRESNEL_SPEC, BSDF_TRANSMISSION | BSDF_REFLECTION | BSDF_SPECULAR, rr);
                l.aux = cast3<R>(t); l.eta_a = (R)eta; l.eta_b = R(1);
                return;
            }
            if (sc.mat_alpha[2 * r.mat] == sc.mat_alpha[2 * r.mat]) { urough = sc.mat_alpha[2 * r.mat]; vrough = sc.mat_alpha[2 * r.mat + 1]; }
            else if (m.remap_roughness) { urough = tr_roughness_to_alpha(urough); vrough = tr_roughness_to_alpha(vrough); }
            if (!is_black(rr)) {
                Lobe<R>& l = add_lobe<mat_km<MAT>()>(b, 0);
                lobe_init(l, L_MICRO_R, BSDF_REFLECTION | BSDF_GLOSSY, rr);
                l.fresnel = F_DIELECTRIC; l.a = (R)eta; l.b = R(1);
                l.mfd_empty = false; l.alpha_x = (R)fmax(urough, 1e-3); l.alpha_y = (R)fmax(vrough, 1e-3);
            }
            if (!is_black(t)) {
                Lobe<R>& l = add_lobe<mat_km<MAT>()>(b, 1);
                lobe_init(l, L_MICRO_T, BSDF_TRANSMISSION | BSDF_GLOSSY, t);
                l.fresnel = F_DIELECTRIC; l.a = (R)eta; l.b = R(1); l.eta_a = (R)eta; l.eta_b = R(1);
                l.mfd_empty = false; l.alpha_x = (R)fmax(urough, 1e-3); l.alpha_y = (R)fmax(vrough, 1e-3);
            }
            return;
        }
        if (mt == RT_MAT_METAL) {
            bsdf_new(b, r);
            D3 u_rough = tex_value_at<sizeof(R) == 8>(sc, m.tex[3] == RT_NONE ? m.tex[2] : m.tex[3], u, v, r.p);
            D3 v_rough = tex_value_at<sizeof(R) == 8>(sc, m.tex[4] == RT_NONE ? m.tex[2] : m.tex[4], u, v, r.p);
            const double au = sc.mat_alpha[2 * r.mat], av = sc.mat_alpha[2 * r.mat + 1];
            double ur = au == au ? au : (m.remap_roughness ? tr_roughness_to_alpha(u_rough.x) : u_rough.x);
            double vr = av == av ? av : (m.remap_roughness ? tr_roughness_to_alpha(v_rough.x) : v_rough.x);
            Lobe<R>& l = add_lobe<mat_km<MAT>()>(b, 0);
            lobe_init(l, L_MICRO_R, BSDF_REFLECTION | BSDF_GLOSSY, mk3(1., 1., 1.));
            l.fresnel = F_CONDUCTOR; l.eta3 = cast3<R>(tex_value_at<sizeof(R) == 8>(sc, m.tex[0], u, v, r.p)); l.aux = cast3<R>(tex_value_at<sizeof(R) == 8>(sc, m.tex[1], u, v, r.p));
            l.mfd_empty = false; l.alpha_x = (R)fmax(ur, 1e-3); l.alpha_y = (R)fmax(vr, 1e-3);
            return;
        }
        if (mt == RT_MAT_MIRROR) {
            bsdf_new(b, r);
            D3 color = tex_value_at<sizeof(R) == 8>(sc, m.tex[0], u, v, r.p);
            if (!is_black(color)) lobe_init(add_lobe<mat_km<MAT>()>(b, 0), L_SPEC_R, BSDF_REFLECTION | BSDF_SPECULAR, color);
            return;
        }
        if constexpr (sizeof(R) == 8) {
        if (mt == RT_MAT_DISNEY) {
            // material.rs:245-394.  The arm never calls Bsdf::new: the lobes go into the record's Bsdf::empty(), whose frame is
            // ns = ng = ss = ts = normalize((1,1,1)) (bsdf.rs:37-47) — `b` holds exactly that frame at this point.  Kept as the reference has it.
            const rt_disney d = sc.disneys[m.tex[0]];
            auto tex1 = [&](uint32_t id) { return tex_value_at<true>(sc, id, u, v, r.p).x; };
            const D3 color = tex_value_at<true>(sc, d.color, u, v, r.p);
            const double metallic_weight = tex1(d.metallic), e = tex1(d.eta), strans = tex1(d.spec_trans);
            const double diffuse_weight = (1. - metallic_weight) * (1. - strans);
            const double dt = tex1(d.diff_trans) / 2.;
            const double rough = tex1(d.roughness);
            const double lum = 0.2126 * color.x + 0.7152 * color.y + 0.0722 * color.z;       // util::color_to_luminance
            const D3 white = mk3(1., 1., 1.);
            const D3 c_tint = lum > 0. ? (1. / lum) * color : white;
            const double sheen_weight = tex1(d.sheen);
            auto lerp_d = [](double t, double a, double bb) { return (1. - t) * a + t * bb; };
            auto lerp_v = [&](double t, D3 a, D3 bb) { return mk3(lerp_d(t, a.x, bb.x), lerp_d(t, a.y, bb.y), lerp_d(t, a.z, bb.z)); };
            D3 c_sheen = mk3(0., 0., 0.);
            if (sheen_weight > 0.) c_sheen = lerp_v(tex1(d.sheen_tint), white, c_tint);
            b.empty = false;
            if (diffuse_weight > 0.) {
                if (d.thin) {
                    const double flat = tex1(d.flatness);
                    lobe_init(add_lobe<mat_km<MAT>()>(b, 0), L_DISNEY_DIFFUSE, BSDF_DIFFUSE | BSDF_REFLECTION, (diffuse_weight * (1. - flat) * (1. - dt)) * color);
                    Lobe<R>& l = add_lobe<mat_km<MAT>()>(b, 0);
                    lobe_init(l, L_DISNEY_FAKESS, BSDF_DIFFUSE | BSDF_REFLECTION, (diffuse_weight * flat * (1. - dt)) * color); l.a = rough;
                } else {
                    const D3 sd = tex_value_at<true>(sc, d.scatter_distance, u, v, r.p);
                    if (is_black(sd)) lobe_init(add_lobe<mat_km<MAT>()>(b, 0), L_DISNEY_DIFFUSE, BSDF_DIFFUSE | BSDF_REFLECTION, diffuse_weight * color);
                    else {                               // "use bssrdf instead; right now, that doesn't exist, so just specular transmission" (material.rs:300-307)
                        Lobe<R>& l = add_lobe<mat_km<MAT>()>(b, 0);
                        lobe_init(l, L_SPEC_T, BSDF_TRANSMISSION | BSDF_SPECULAR, white);
                        l.fresnel = F_DIELECTRIC; l.a = 1.; l.b = e; l.eta_a = 1.; l.eta_b = e;
                    }
                }
                { Lobe<R>& l = add_lobe<mat_km<MAT>()>(b, 0); lobe_init(l, L_DISNEY_RETRO, BSDF_DIFFUSE | BSDF_REFLECTION, diffuse_weight * color); l.a = rough; }
                if (sheen_weight > 0.) lobe_init(add_lobe<mat_km<MAT>()>(b, 0), L_DISNEY_SHEEN, BSDF_DIFFUSE | BSDF_REFLECTION, (diffuse_weight * sheen_weight) * c_sheen);
            }
            const double ani = tex1(d.anisotropic);
            const double aspect = sqrt(1. - ani * 0.9);
            const double ax = fmax(0.001, rough * rough / aspect), ay = fmax(0.001, rough * rough * aspect);
            const double spec_tint = tex1(d.specular_tint);
            const double r0e = (e - 1.) * (e - 1.) / ((e + 1.) * (e + 1.));                  // eta_to_r0
            const D3 c_spec_0 = lerp_v(metallic_weight, r0e * lerp_v(spec_tint, white, c_tint), color);
            {   // MicrofacetReflection(white, DisneyFresnel{r0, metallic, eta}, DisneyMicrofacet(ax, ay)); make_disney clamps alpha at 1e-3
                Lobe<R>& l = add_lobe<mat_km<MAT>()>(b, 0);
                lobe_init(l, L_MICRO_R, BSDF_REFLECTION | BSDF_GLOSSY, white);
                l.fresnel = F_DISNEY; l.eta3 = cast3<R>(c_spec_0); l.a = metallic_weight; l.b = e;
                l.mfd_empty = false; l.mfd_disney = true; l.alpha_x = fmax(ax, 1e-3); l.alpha_y = fmax(ay, 1e-3);
            }
            const double cc = tex1(d.clearcoat);
            if (cc > 0.) { Lobe<R>& l = add_lobe<mat_km<MAT>()>(b, 0); lobe_init(l, L_DISNEY_CLEARCOAT, BSDF_GLOSSY | BSDF_REFLECTION, mk3(0., 0., 0.)); l.a = cc; l.b = lerp_d(tex1(d.clearcoat_gloss), 0.1, 0.001); }
            if (strans > 0.) {
                const D3 t = mk3(strans * sqrt(color.x), strans * sqrt(color.y), strans * sqrt(color.z));
                Lobe<R>& l = add_lobe<mat_km<MAT>()>(b, 0);
                lobe_init(l, L_MICRO_T, BSDF_TRANSMISSION | BSDF_GLOSSY, t);
                l.fresnel = F_DIELECTRIC; l.a = 1.; l.b = e; l.eta_a = 1.; l.eta_b = e; l.mfd_empty = false;
                if (d.thin) {                            // "approximate with scaled distribution" (Burley 2015): a plain TrowbridgeReitz
                    const double rscaled = (0.65 * e - 0.35) * rough;
                    l.alpha_x = fmax(fmax(0.001, rscaled * rscaled / aspect), 1e-3); l.alpha_y = fmax(fmax(0.001, rscaled * rscaled * aspect), 1e-3);
                } else { l.mfd_disney = true; l.alpha_x = fmax(ax, 1e-3); l.alpha_y = fmax(ay, 1e-3); }
            }
            if (d.thin) lobe_init(add_lobe<mat_km<MAT>()>(b, 0), L_LAMBERT_T, BSDF_TRANSMISSION | BSDF_DIFFUSE, dt * color);
            return;
        }
        }
    }
}
template <class R, int MAT> static __device__ __noinline__ void compute_scattering_call(const DScene& sc, const Rec& r, Bsdf<R>& b) { compute_scattering_impl<R, MAT>(sc, r, b); }
template <class R, int MAT> __device__ __forceinline__ void compute_scattering(const DScene& sc, const Rec& r, Bsdf<R>& b) { if constexpr (km_small(mat_km<MAT>())) { compute_scattering_impl<R, MAT>(sc, r, b); } else { compute_scattering_call<R, MAT>(sc, r, b); } }

// ---- primitive.rs sampling / light.rs Diffuse — f64 in both modes (the sampled point is a shadow-ray endpoint)
// Primitive::sample_area (primitive.rs:478-539): point and normal on the light's primitive
static __device__ RT_HELPER2 void sample_area(const DScene& sc, uint32_t prim, D2 u, D3& p, D3& n) {
    DPrim pr = sc.prims[prim];
    uint32_t type = pr.info & 7u;
    if (type == RT_PRIM_TRIANGLE) {
        const uint32_t* ti = sc.tri_index + pr.a;
        const DMesh ms = sc.meshes[pr.info >> 6];
        D3 p0 = ld3(sc.points + 3 * (ms.p_base + ti[0])), p1 = ld3(sc.points + 3 * (ms.p_base + ti[1])), p2 = ld3(sc.points + 3 * (ms.p_base + ti[2]));
        double s0 = sqrt(u.x); double b0 = 1. - s0, b1 = u.y * s0;                         // util.rs:62-65
        p = mk3(0., 0., 0.) + (b0 * p0 + b1 * p1 + (1. - b0 - b1) * p2);
        if (pr.info & PI_HAS_N) {
            const double* nn = sc.normals;
            n = normalize(b0 * ld3(nn + 3 * (ms.n_base + ti[0])) + b1 * ld3(nn + 3 * (ms.n_base + ti[1])) + (1. - b0 - b1) * ld3(nn + 3 * (ms.n_base + ti[2])));
        } else n = normalize(cross(p1 - p0, p2 - p0));
    } else {
        const double* sh = sc.shapes + pr.a;
        if (type == RT_PRIM_SPHERE) {
            double z = 1. - 2. * u.x, r = sqrt(fmax(0., 1. - z * z)), phi = 2. * PI * u.y;   // util.rs:51-56
            p = mk3(0., 0., 0.) + sh[3] * mk3(r * cos(phi), r * sin(phi), z);              // around the ORIGIN (primitive.rs:482)
            n = normalize(p - mk3(0., 0., 0.));
        } else {
            double a = sh[0] + u.x * (sh[2] - sh[0]), bb = sh[1] + u.y * (sh[3] - sh[1]), k = sh[4];
            if (type == RT_PRIM_XY_RECT) { p = mk3(a, bb, k); n = normalize(mk3(0., 0., 1.)); }
            else if (type == RT_PRIM_XZ_RECT) { p = mk3(a, k, bb); n = normalize(mk3(0., 1., 0.)); }
            else { p = mk3(k, a, bb); n = normalize(mk3(1., 0., 0.)); }
        }
    }
    if (pr.info & PI_FLIP) n = -n;
}

// ---- light.rs Infinite + distribution.rs — f64 in both modes (directions and pdfs of light samples)
constexpr double INV_PI_C = 1.0 / PI, INV_2PI_C = 1.0 / (2.0 * PI);           // consts.rs:41-42
// find_interval (distribution.rs:150-166) with pred = |i| cdf[i] <= u
__device__ __forceinline__ uint32_t find_interval_cdf(const double* __restrict__ cdf, uint32_t size, double u) {
    uint32_t first = 0, len = size;
    while (len > 0) {
        uint32_t half = len >> 1, middle = first + half;
        if (__ldg(cdf + middle) <= u) { first = middle + 1; len = len - half - 1; } else len = half;
    }
    double c = clampd((double)first - 1., 0., (double)size - 2.);            // cdf[0] = 0 <= u, so first >= 1
    return (uint32_t)c;
}
// Distribution1D::sample_continuous (distribution.rs:64-78): cdf[0..n], func[i] = func[i * 1], → x, pdf, offset
__device__ __forceinline__ void dist1d_sample(const double* __restrict__ cdf, const double* __restrict__ func, uint32_t n, double func_int, double u, double& x, double& pdf, uint32_t& offset) {
    offset = find_interval_cdf(cdf, n + 1, u);
    double c0 = __ldg(cdf + offset), c1 = __ldg(cdf + offset + 1);
    double du = u - c0;
    if (c1 - c0 > 0.) du = du / (c1 - c0);
    pdf = func_int > 0. ? __ldg(func + offset) / func_int : 0.;
    x = ((double)offset + du) / (double)n;
}
__device__ __forceinline__ D3 light_to_world(const DScene& sc, const DLight& l, D3 v) { return l.xform == RT_NONE ? v : xf_vector(sc.xforms[l.xform].fwd, v); }
__device__ __forceinline__ D3 light_to_obj(const DScene& sc, const DLight& l, D3 v) { return l.xform == RT_NONE ? v : xf_vector(sc.xforms[l.xform].inv, v); }
__device__ __forceinline__ double spherical_theta(D3 v) { return acos(clampd(v.y, -1., 1.)); }                 // util.rs:153-155
__device__ __forceinline__ double spherical_phi(D3 v) { double p = atan2(v.z, v.x); return p < 0. ? p + 2. * PI : p; }   // util.rs:160-167
// Light::sample_li, Infinite arm (light.rs:204-245): → wi (normalised), pdf, colour, shadow-ray end point
static __device__ __noinline__ void infinite_sample_li(const DScene& sc, const DLight& l, D3 p, D2 u, D3& wi, double& pdf, D3& color, D3& p1) {
    // Distribution2D::sample_continuous (distribution.rs:135-139): marginal with u[1], then row v with u[0]
    double d1, pdf1, d0, pdf0; uint32_t v, off;
    dist1d_sample(l.marg_cdf, l.row_int, l.nv, l.marg_int, u.y, d1, pdf1, v);
    dist1d_sample(l.dist_cdf + (size_t)v * (l.nu + 1), l.dist_f + v, l.nu, __ldg(l.row_int + v), u.x, d0, pdf0, off);   // row v's func is f[v..v+nu)
    double map_pdf = pdf1 * pdf0;
    wi = mk3(0., 0., 0.); color = mk3(0., 0., 0.); p1 = p; pdf = 0.;
    if (map_pdf == 0.) return;
    double theta = d1 * PI, phi = d0 * 2. * PI;
    double cos_theta = cos(theta), sin_theta = sin(theta), cos_phi = cos(phi), sin_phi = sin(phi);
    D3 w = light_to_world(sc, l, mk3(sin_theta * cos_phi, sin_theta * sin_phi, cos_theta));      // Z-up here, Y-up in le/pdf_li (§8.1 #14)
    pdf = map_pdf / (2. * PI * PI * sin_theta);
    if (sin_theta == 0.) pdf = 0.;
    p1 = p + w * (2. * WORLD_RADIUS);
    color = tex_value(sc, l.text_id, d0, d1);
    wi = normalize(w);
}
// Light::sample_li, Point / Spot / Distant arms (light.rs:142-175): pdf = 1, the shadow ray ends at the light position (Distant:
// 2 x 1e10 along dir).  The visibility target is primitive 0 (HitRecord::make_basic), as the reference has it.
static __device__ __noinline__ void delta_sample_li(const DScene& sc, const DLight& l, D3 p, D3& wi, double& pdf, D3& color, D3& p1) {
    const D3 lp = ld3(l.pos), c = ld3(l.color);
    pdf = 1.;
    if (l.type == RT_LIGHT_DISTANT) { p1 = p + lp * (2. * DISTANT_WORLD_RADIUS); color = c; wi = normalize(lp); return; }
    wi = normalize(lp - p); p1 = lp;
    double scale = 1. / norm2(lp - p);
    if (l.type == RT_LIGHT_SPOT) {
        // falloff (light.rs:15-26)
        const D3 wl = normalize(light_to_obj(sc, l, -wi));
        const double cos_theta = wl.z;
        double fo;
        if (cos_theta < l.cos_width) fo = 0.;
        else if (cos_theta > l.cos_falloff) fo = 1.;
        else { const double delta = (cos_theta - l.cos_width) / (l.cos_falloff - l.cos_width); fo = (delta * delta) * (delta * delta); }
        scale = fo / norm2(lp - p);
    }
    color = c * scale;
}
// Light::pdf_li, Infinite arm (light.rs:285-294) — transforms with to_world, as the reference does
static __device__ __noinline__ double infinite_pdf_li(const DScene& sc, const DLight& l, D3 wi) {
    D3 w = light_to_world(sc, l, wi);
    double theta = spherical_theta(w), phi = spherical_phi(w);
    double sin_theta = sin(theta);
    if (sin_theta == 0.) return 0.;
    // Distribution2D::pdf (distribution.rs:141-147)
    double pu = phi * INV_2PI_C, pv = theta * INV_PI_C;
    uint32_t iu = min(sat_u32(pu * (double)l.nu), l.nu - 1u), iv = min(sat_u32(pv * (double)l.nv), l.nv - 1u);
    return (__ldg(l.dist_f + iv + iu) / l.marg_int) / (2. * PI * PI * sin_theta);
}
// Light::le (light.rs:499-513): radiance along a ray that leaves the scene
static __device__ __noinline__ D3 infinite_le(const DScene& sc, const DLight& l, D3 ray_d) {
    D3 w = normalize(light_to_obj(sc, l, ray_d));
    return tex_value(sc, l.text_id, spherical_phi(w) * INV_2PI_C, spherical_theta(w) * INV_PI_C);
}

}  // namespace rt
