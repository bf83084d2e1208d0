// Device shading: textures, materials → BSDF, BxDF lobes, Trowbridge–Reitz microfacets, area lights.
// f64, reference operation order (compiled with -fmad=false).  Each function cites the reference item
// it replaces.  Quirks of SURVEY.md §8.1 are kept on purpose: image parity is against what the
// reference computes.
#pragma once
#include "geom.cuh"

namespace rt {

constexpr uint32_t BSDF_REFLECTION = 1, BSDF_TRANSMISSION = 2, BSDF_DIFFUSE = 4, BSDF_GLOSSY = 8, BSDF_SPECULAR = 16, BSDF_ALL = 31;  // consts.rs:14-20
constexpr double INV_PI = 1.0 / PI;                                                                                                  // consts.rs:41

// ---- counter-based RNG replacing thread_rng()/StdRng::from_entropy() (sampler.rs:318,334-337; util.rs:32-34):
// Philox4x32-10, key = seed, counter = (pixel, sample, dim, 'RTCX'); one uniform per counter.
struct Rng { uint32_t k0, k1, pixel, sample, dim; };
static __device__ __noinline__ double rng_next(Rng& g) {
    uint32_t c0 = g.pixel, c1 = g.sample, c2 = g.dim++, c3 = 0x52544358u, k0 = g.k0, k1 = g.k1;
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return ((double)(c0 >> 5) * 67108864.0 + (double)(c1 >> 6)) / 9007199254740992.0;
}

// ---- bxdf.rs:12-68
__device__ __forceinline__ double cos_sq_theta(D3 v) { return v.z * v.z; }
__device__ __forceinline__ double abs_cos_theta(D3 v) { return fabs(v.z); }
__device__ __forceinline__ double sin_sq_theta(D3 v) { return fmax(0., 1. - cos_sq_theta(v)); }
__device__ __forceinline__ double sin_theta(D3 v) { return sqrt(sin_sq_theta(v)); }
__device__ __forceinline__ double tan_theta(D3 v) { return sin_theta(v) / v.z; }
__device__ __forceinline__ double tan_sq_theta(D3 v) { return sin_sq_theta(v) / cos_sq_theta(v); }
__device__ __forceinline__ double cos_phi(D3 v) { double s = sin_theta(v); return s == 0. ? 1. : clampd(v.x / s, -1., 1.); }
__device__ __forceinline__ double sin_phi(D3 v) { double s = sin_theta(v); return s == 0. ? 1. : clampd(v.y / s, -1., 1.); }
__device__ __forceinline__ bool matches_flags(uint32_t flag, uint32_t other) { return (flag & other) == flag; }
__device__ __forceinline__ bool same_hemisphere(D3 v, D3 w) { return v.z * w.z > 0.; }                 // util.rs:591-593
__device__ __forceinline__ D3 reflect(D3 v, D3 n) { double s = 2. * dot(v, n); return -v + n * s; }    // util.rs:203-206
static __device__ __noinline__ bool refract(D3 vec, D3 n, double eta, D3& out) {                           // util.rs:376-385
    double cos_theta_i = dot(n, vec) / norm(vec);
    double sin2_theta_i = fmax(0., 1. - cos_theta_i * cos_theta_i);
    double sin2_theta_t = eta * eta * sin2_theta_i;
    if (sin2_theta_t >= 1.) return false;
    double cos_theta_t = sqrt(1. - sin2_theta_t);
    out = eta * (-vec) + (eta * cos_theta_i - cos_theta_t) * n;
    return true;
}
static __device__ __noinline__ double fr_dielectric(double cos_theta_i, double eta_i, double eta_t) {               // bxdf.rs:113-136
    cos_theta_i = clampd(cos_theta_i, -1., 1.);
    double index_i = eta_i, index_t = eta_t;
    if (cos_theta_i < 0.) { index_i = eta_t; index_t = eta_i; cos_theta_i = fabs(cos_theta_i); }
    double sin_theta_i = sqrt(fmax(0., 1. - cos_theta_i * cos_theta_i));
    double sin_theta_t = index_i / index_t * sin_theta_i;
    double cos_theta_t = sqrt(fmax(0., 1. - sin_theta_t * sin_theta_t));
    if (sin_theta_t >= 1.) return 1.;
    double r_parl = ((index_t * cos_theta_i) - (index_i * cos_theta_t)) / ((index_t * cos_theta_i) + (index_i * cos_theta_t));
    double r_perp = ((index_i * cos_theta_i) - (index_t * cos_theta_t)) / ((index_i * cos_theta_i) + (index_t * cos_theta_t));
    return (r_parl * r_parl + r_perp * r_perp) / 2.;
}
static __device__ __noinline__ D3 fr_conductor(double cos_theta_i, D3 eta, D3 eta_k) {                              // bxdf.rs:141-170
    cos_theta_i = clampd(cos_theta_i, -1., 1.);
    double c2 = cos_theta_i * cos_theta_i, s2 = 1. - c2;
    D3 eta2 = cmul(eta, eta), etak2 = cmul(eta_k, eta_k);
    D3 t0 = (eta2 - etak2) - mk3(s2, s2, s2);
    D3 a2pb2 = cmul(t0, t0) + cmul(eta2, etak2) * 4.;
    a2pb2 = mk3(sqrt(a2pb2.x), sqrt(a2pb2.y), sqrt(a2pb2.z));
    D3 t1 = a2pb2 + mk3(c2, c2, c2);
    D3 a = (a2pb2 + t0) * 0.5;
    a = mk3(sqrt(a.x), sqrt(a.y), sqrt(a.z));
    D3 t2 = a * (2. * cos_theta_i);
    D3 rs = cdiv(t1 - t2, t1 + t2);
    D3 t3 = a2pb2 * c2 + mk3(s2 * s2, s2 * s2, s2 * s2);
    D3 t4 = t2 * s2;
    D3 rp = cmul(rs, cdiv(t3 - t4, t3 + t4));
    return (rp + rs) * 0.5;
}

// ---- lobes (bxdf.rs:213-300, in-scope subset) kept in registers/local memory: no heap, no Vec<Box<..>>
enum { F_DIELECTRIC = 0, F_CONDUCTOR = 1, F_NOOP = 2 };
enum { L_LAMBERT = 0, L_OREN_NAYAR = 1, L_SPEC_R = 2, L_SPEC_T = 3, L_FRESNEL_SPEC = 4, L_MICRO_R = 5, L_MICRO_T = 6 };
struct Lobe {
    int kind; uint32_t type;
    D3 color;                 // color / r
    D3 aux;                   // FresnelSpecular t ; conductor k
    D3 eta3;                  // conductor eta
    double a, b;              // OrenNayer A,B ; or dielectric fresnel eta_i, eta_t
    int fresnel;
    bool mfd_empty; double alpha_x, alpha_y;
    double eta_a, eta_b;
};
struct Bsdf { D3 ns, ng, ss, ts; int n; bool empty; Lobe lobes[2]; };

// Fresnel::evaluate (bxdf.rs:190-211); dielectric arm passes (|cos|, eta_t, eta_i)
__device__ __forceinline__ D3 fresnel_eval(const Lobe& l, double cos_theta_i) {
    if (l.fresnel == F_DIELECTRIC) { double v = fr_dielectric(fabs(cos_theta_i), l.b, l.a); return mk3(v, v, v); }
    if (l.fresnel == F_CONDUCTOR) return fr_conductor(fabs(cos_theta_i), l.eta3, l.aux);
    return mk3(1., 1., 1.);
}
// ---- microfacet.rs, TrowbridgeReitz{sample_visible_area:true} / Empty
static __device__ __noinline__ double mfd_d(const Lobe& m, D3 wh) {                                                 // microfacet.rs:53-68
    if (m.mfd_empty) return 0.;
    double t2 = tan_sq_theta(wh);
    if (isinf(t2)) return 0.;
    double cos4 = cos_sq_theta(wh) * cos_sq_theta(wh);
    double cp = cos_phi(wh), sp = sin_phi(wh);
    double e = ((cp * cp) / (m.alpha_x * m.alpha_x) + (sp * sp) / (m.alpha_y * m.alpha_y)) * t2;
    return 1. / (PI * m.alpha_x * m.alpha_y * cos4 * (1. + e) * (1. + e));
}
static __device__ __noinline__ double mfd_lambda(const Lobe& m, D3 w) {                                             // microfacet.rs:109-123
    if (m.mfd_empty) return 0.;
    double att = fabs(tan_theta(w));
    if (isinf(att)) return 0.;
    double cp = cos_phi(w), sp = sin_phi(w);
    double alpha = sqrt((cp * cp) * m.alpha_x * m.alpha_x + (sp * sp) * m.alpha_y * m.alpha_y);
    double a2t2 = (alpha * att) * (alpha * att);
    return (-1. + sqrt(1. + a2t2)) / 2.;
}
__device__ __forceinline__ double mfd_g1(const Lobe& m, D3 w) { return 1. / (1. + mfd_lambda(m, w)); }
__device__ __forceinline__ double mfd_g(const Lobe& m, D3 wo, D3 wi) { return 1. / (1. + mfd_lambda(m, wo) + mfd_lambda(m, wi)); }
__device__ __forceinline__ double mfd_pdf(const Lobe& m, D3 wo, D3 wh) {                               // microfacet.rs:159-172
    if (!m.mfd_empty) return mfd_d(m, wh) * mfd_g1(m, wo) * fabs(dot(wo, wh)) / abs_cos_theta(wo);
    return mfd_d(m, wh) * abs_cos_theta(wh);
}
static __device__ __noinline__ double tr_roughness_to_alpha(double roughness) {                            // microfacet.rs:442-446
    roughness = fmax(roughness, 1e-5);
    double x = log(roughness);
    return 1.62142 + 0.819955 * x + 0.1734 * x * x + 0.0171201 * x * x * x + 0.000640711 * x * x * x * x;
}
static __device__ __noinline__ D2 tr_sample_11(double cos_t, double u1, double u2) {                                // microfacet.rs:469-512
    if (cos_t > 0.9999) { double r = sqrt(u1 / (1. - u1)), phi = 2. * PI * u2; return mk2(r * cos(phi), r * sin(phi)); }
    double sin_t = fmax(0., sqrt(1. - cos_t * cos_t));
    double tan_t = sin_t / cos_t;
    double a = 1. / tan_t;
    double g1 = 2. / (1. + sqrt(1. + 1. / (a * a)));
    a = 2. * u1 / g1 - 1.;
    double tmp = 1. / (a * a - 1.);
    if (tmp > 1e10) tmp = 1e10;
    double b = tan_t;
    double d = sqrt(fmax(0., b * b * tmp * tmp - (a * a - b * b) * tmp));
    double sx1 = b * tmp - d, sx2 = b * tmp + d;
    double slope_x = (a < 0. || sx2 > 1. / tan_t) ? sx1 : sx2;
    double s, nu2;
    if (u2 > 0.5) { s = 1.; nu2 = 2. * (u2 - 0.5); } else { s = -1.; nu2 = 2. * (0.5 - u2); }
    double z = (nu2 * (nu2 * (nu2 * 0.27385 - 0.73369) + 0.46341)) / (nu2 * (nu2 * (nu2 * 0.093073 + 0.309420) - 1.) + 0.597999);
    return mk2(slope_x, s * z * sqrt(1. + slope_x * slope_x));
}
static __device__ __noinline__ D3 mfd_sample_wh(const Lobe& m, D3 wo, D2 u) {                                       // microfacet.rs:240-283, 448-467
    if (m.mfd_empty) return normalize(mk3(0., 0., 0.));
    bool flip = wo.z < 0.;
    D3 wi = flip ? -wo : wo;
    D3 wi_s = normalize(mk3(m.alpha_x * wi.x, m.alpha_y * wi.y, wi.z));
    D2 slope = tr_sample_11(wi_s.z, u.x, u.y);
    double sp = sin_phi(wi_s), cp = cos_phi(wi_s);
    double tmp = cp * slope.x - sp * slope.y;
    double sy = sp * slope.x + cp * slope.y;
    double sx = tmp;
    sx = m.alpha_x * sx; sy = m.alpha_y * sy;
    D3 wh = normalize(mk3(-sx, -sy, 1.));
    return flip ? -wh : wh;
}

// Bxdf::f (bxdf.rs:328-527)
static __device__ __noinline__ D3 lobe_f(const Lobe& l, D3 wo, D3 wi) {
    switch (l.kind) {
        case L_LAMBERT: return l.color * INV_PI;
        case L_OREN_NAYAR: {
            double sti = sin_theta(wi), sto = sin_theta(wo), max_cos = 0.;
            if (sti > SMALL && sto > SMALL) { double d_cos = cos_phi(wi) * cos_phi(wo) + sin_phi(wi) * sin_phi(wo); max_cos = fmax(0., d_cos); }
            double sin_alpha, tan_beta;
            if (abs_cos_theta(wi) > abs_cos_theta(wo)) { sin_alpha = sto; tan_beta = sti / abs_cos_theta(wi); }
            else { sin_alpha = sti; tan_beta = sto / abs_cos_theta(wi); }
            return l.color * (INV_PI * (l.a + l.b * max_cos * sin_alpha * tan_beta));
        }
        case L_MICRO_R: {
            double cto = abs_cos_theta(wo), cti = abs_cos_theta(wi);
            D3 wh = wi + wo;
            if (cti == 0. || cto == 0.) return mk3(0., 0., 0.);
            if (is_black(wh)) return mk3(0., 0., 0.);
            wh = normalize(wh);
            D3 f = fresnel_eval(l, dot(wi, face_forward(wh, mk3(0., 0., 1.))));
            D3 comp1 = l.color * mfd_d(l, wh) * mfd_g(l, wo, wi);
            return cmul(comp1, f * (1. / (4. * cti * cto)));
        }
        case L_MICRO_T: {
            if (same_hemisphere(wo, wi)) return mk3(0., 0., 0.);
            double cto = wo.z, cti = wi.z;
            if (cti == 0. || cto == 0.) return mk3(0., 0., 0.);
            double eta = wo.z > 0. ? l.eta_b / l.eta_a : l.eta_a / l.eta_b;
            D3 wh = normalize(wo + wi * eta);
            if (wh.z < 0.) wh = -wh;
            if (dot(wo, wh) * dot(wi, wh) > 0.) return mk3(0., 0., 0.);
            D3 f = fresnel_eval(l, dot(wo, wh));
            double sqrt_denom = dot(wo, wh) + eta * dot(wi, wh);
            double factor = 1. / eta;                                           // mode == RADIANCE
            D3 c = cmul(mk3(1., 1., 1.) - f, l.color);
            return c * fabs(mfd_d(l, wh) * mfd_g(l, wo, wi) * eta * eta * fabs(dot(wi, wh)) * fabs(dot(wo, wh)) * factor * factor /
                            (cti * cto * sqrt_denom * sqrt_denom));
        }
        default: return mk3(0., 0., 0.);
    }
}
__device__ __forceinline__ double default_pdf(D3 wo, D3 wi) { return same_hemisphere(wo, wi) ? abs_cos_theta(wi) * INV_PI : 0.; }   // bxdf.rs:829-835
// Bxdf::pdf (bxdf.rs:721-813)
static __device__ __noinline__ double lobe_pdf(const Lobe& l, D3 wo, D3 wi) {
    switch (l.kind) {
        case L_LAMBERT: case L_OREN_NAYAR: case L_SPEC_R: return default_pdf(wo, wi);
        case L_MICRO_R: {
            if (!same_hemisphere(wo, wi)) return 0.;
            D3 wh = normalize(wo + wi);
            return mfd_pdf(l, wo, wh) / (4. * dot(wo, wh));
        }
        case L_MICRO_T: {
            if (same_hemisphere(wo, wi)) return 0.;
            double eta = wo.z > 0. ? l.eta_b / l.eta_a : l.eta_a / l.eta_b;
            D3 wh = normalize(wo + wi * eta);
            if (dot(wo, wh) * dot(wi, wh) > 0.) return 0.;
            double sqrt_denom = dot(wo, wh) + eta * dot(wi, wh);
            double dwh_dwi = fabs(eta * eta * dot(wi, wh)) / (sqrt_denom * sqrt_denom);
            return mfd_pdf(l, wo, wh) * dwh_dwi;
        }
        default: return 0.;       // SpecularTransmission: 0 ; FresnelSpecular: "unimplemented" → 0 (bxdf.rs:777-791)
    }
}
// util::rand_cosine_dir (util.rs:127-148)
static __device__ __noinline__ D3 rand_cosine_dir(Rng& g) {
    double r1 = rng_next(g), r2 = rng_next(g);
    double u1 = 2. * r1 - 1., u2 = 2. * r2 - 1.;
    if (u1 == 0. && u2 == 0.) return mk3(0., 0., 1.);
    double theta, r;
    if (fabs(u1) > fabs(u2)) { r = u1; theta = PI / 4. * (u2 / u1); }
    else { r = u2; theta = PI / 2. - PI / 4. * (u1 / u2); }
    double x = r * cos(theta), y = r * sin(theta);
    return mk3(x, y, sqrt(fmax(0., 1. - x * x - y * y)));
}
// Bxdf::sample_f (bxdf.rs:532-719)
static __device__ __noinline__ void lobe_sample_f(const Lobe& l, D3 wo, D2 sample, Rng& g, D3& color, D3& wi, double& pdf) {
    color = mk3(0., 0., 0.); wi = mk3(0., 0., 0.); pdf = 0.;
    switch (l.kind) {
        case L_LAMBERT: case L_OREN_NAYAR: {             // default_sample_f (bxdf.rs:815-827): ignores `sample`
            wi = rand_cosine_dir(g);
            if (wo.z < 0.) wi.z *= -1.;
            pdf = lobe_pdf(l, wo, wi); color = lobe_f(l, wo, wi); return;
        }
        case L_SPEC_R: {
            wi = mk3(-wo.x, -wo.y, wo.z);
            color = cmul(l.color, fresnel_eval(l, wi.z)); pdf = 1.; return;
        }
        case L_SPEC_T: {
            bool entering = wo.z > 0.;
            double eta_i = entering ? l.eta_a : l.eta_b, eta_t = entering ? l.eta_b : l.eta_a;
            D3 dir;
            if (!refract(wo, face_forward(mk3(0., 0., 1.), wo), eta_i / eta_t, dir)) return;
            D3 ft = cmul(mk3(1., 1., 1.) - fresnel_eval(l, dir.z), l.color);
            ft = ft * ((eta_i * eta_i) / (eta_t * eta_t));
            color = ft; wi = dir; pdf = 1.; return;
        }
        case L_MICRO_R: {
            if (wo.z == 0.) return;
            D3 wh = mfd_sample_wh(l, wo, sample);       // the wo·wh<0 guard has no `return` (bxdf.rs:598-600)
            D3 w = reflect(wo, wh);
            if (!same_hemisphere(wo, w)) return;
            pdf = mfd_pdf(l, wo, wh) / (4. * dot(wo, wh));
            wi = w; color = lobe_f(l, wo, wi); return;
        }
        case L_MICRO_T: {
            if (wo.z == 0.) return;
            D3 wh = mfd_sample_wh(l, wo, sample);
            if (dot(wo, wh) < 0.) return;
            double eta = wo.z > 0. ? l.eta_a / l.eta_b : l.eta_b / l.eta_a;
            D3 w;
            if (!refract(wo, wh, eta, w)) return;
            wi = w; pdf = lobe_pdf(l, wo, wi); color = lobe_f(l, wo, wi); return;
        }
        case L_FRESNEL_SPEC: {
            double f = fr_dielectric(wo.z / norm(wo), l.eta_a, l.eta_b);
            if (sample.x < f) { wi = mk3(-wo.x, -wo.y, wo.z); pdf = f; color = l.color * f; return; }
            bool entering = wo.z > 0.;
            double eta_i = entering ? l.eta_a : l.eta_b, eta_t = entering ? l.eta_b : l.eta_a;
            D3 dir;
            if (!refract(wo, face_forward(mk3(0., 0., 1.), wo), eta_i / eta_t, dir)) return;
            D3 ft = l.aux * (1. - f);
            ft = ft * ((eta_i * eta_i) / (eta_t * eta_t));
            color = ft; wi = dir; pdf = 1. - f; return;
        }
    }
}

// ---- bsdf.rs
__device__ __forceinline__ D3 world_to_local(const Bsdf& b, D3 v) { return mk3(dot(v, b.ss), dot(v, b.ts), dot(v, b.ns)); }
__device__ __forceinline__ D3 local_to_world(const Bsdf& b, D3 v) {
    return mk3(b.ss.x * v.x + b.ts.x * v.y + b.ns.x * v.z, b.ss.y * v.x + b.ts.y * v.y + b.ns.y * v.z, b.ss.z * v.x + b.ts.z * v.y + b.ns.z * v.z);
}
__device__ __forceinline__ uint32_t num_components(const Bsdf& b, uint32_t t) { uint32_t c = 0; for (int i = 0; i < b.n; i++) if (matches_flags(b.lobes[i].type, t)) c++; return c; }
// Bsdf::f (bsdf.rs:83-98), `a && b || c` precedence kept
static __device__ __noinline__ D3 bsdf_f(const Bsdf& b, D3 wow, D3 wiw, uint32_t flags) {
    D3 wi = world_to_local(b, wiw), wo = world_to_local(b, wow);
    bool refl = dot(wiw, b.ng) * dot(wow, b.ng) > 0.;
    D3 f = mk3(0., 0., 0.);
    for (int i = 0; i < b.n; i++) {
        uint32_t ty = b.lobes[i].type;
        if ((matches_flags(ty, flags) && (refl && (ty & BSDF_REFLECTION) > 0)) || (!refl && (ty & BSDF_TRANSMISSION) > 0)) f = f + lobe_f(b.lobes[i], wo, wi);
    }
    return f;
}
// Bsdf::pdf (bsdf.rs:166-189)
static __device__ __noinline__ double bsdf_pdf(const Bsdf& b, D3 wow, D3 wiw, uint32_t flags) {
    if (b.n == 0) return 0.;
    D3 wo = world_to_local(b, wow), wi = world_to_local(b, wiw);
    if (wo.z == 0.) return 0.;
    double pdf = 0.; int matching = 0;
    for (int i = 0; i < b.n; i++) if (matches_flags(b.lobes[i].type, flags)) { matching++; pdf += lobe_pdf(b.lobes[i], wo, wi); }
    return matching > 0 ? pdf : 0.;
}
// Bsdf::sample_f (bsdf.rs:102-164)
static __device__ __noinline__ void bsdf_sample_f(const Bsdf& b, D3 wow, D2 sample, uint32_t bxdf_type, Rng& g, D3& color, D3& wiw, double& pdf, uint32_t& sampled) {
    color = mk3(0., 0., 0.); wiw = mk3(0., 0., 0.); pdf = 0.; sampled = 0;
    uint32_t matching = num_components(b, bxdf_type);
    if (matching == 0) return;
    uint32_t comp = min((uint32_t)floor(sample.x * (double)matching), matching - 1);
    uint32_t count = comp; int used = 0;
    for (int i = 0; i < b.n; i++) if (matches_flags(b.lobes[i].type, bxdf_type)) { if (count == 0) { used = i; break; } count--; }
    const Lobe& bx = b.lobes[used];
    D3 wo = normalize(world_to_local(b, wow));
    if (wo.z == 0.) return;
    D3 c, wi; double p;
    lobe_sample_f(bx, wo, sample, g, c, wi, p);
    if (p == 0.) return;
    D3 w = local_to_world(b, wi);
    bool spec = (bx.type & BSDF_SPECULAR) != 0;
    if (!spec && matching > 1)
        for (int i = 0; i < b.n; i++) if (i != used && matches_flags(b.lobes[i].type, bxdf_type)) p += lobe_pdf(b.lobes[i], wo, wi);
    if (matching > 1) p = p / (double)matching;
    if (!spec) {
        bool refl = dot(w, b.ng) * dot(wow, b.ng) > 0.;
        c = mk3(0., 0., 0.);
        for (int i = 0; i < b.n; i++) {
            uint32_t ty = b.lobes[i].type;
            if (matches_flags(ty, bxdf_type) && ((refl && (ty & BSDF_REFLECTION) > 0) || (!refl && (ty & BSDF_TRANSMISSION) > 0))) c = c + lobe_f(b.lobes[i], wo, wi);
        }
    }
    color = c; wiw = w; pdf = p; sampled = bx.type;
}

// ---- material.rs
// Texture::value (material.rs:542-589): SolidColor, Checkered (ids may nest)
static __device__ __noinline__ D3 tex_value(const DScene& sc, uint32_t id, double u, double v) {
    for (int depth = 0; depth < 8; depth++) {
        const rt_texture& t = sc.textures[id];
        if (t.type != RT_TEX_CHECKERED) return mk3(t.color[0], t.color[1], t.color[2]);
        double mult = sin(t.frequency * u * 2. * PI) * sin(t.frequency * v * 2. * PI);
        id = mult < 0. ? t.even : t.odd;
    }
    return mk3(0., 0., 0.);
}
__device__ __forceinline__ void lobe_init(Lobe& l, int kind, uint32_t type, D3 color) {
    l.kind = kind; l.type = type; l.color = color; l.aux = mk3(0., 0., 0.); l.eta3 = mk3(0., 0., 0.); l.a = 0.; l.b = 0.;
    l.fresnel = F_NOOP; l.mfd_empty = true; l.alpha_x = 0.; l.alpha_y = 0.; l.eta_a = 1.; l.eta_b = 1.;
}
__device__ __forceinline__ void bsdf_new(Bsdf& b, const Rec& r) {                                       // bsdf.rs:26-35
    b.ns = r.sn; b.ng = r.n; b.ss = r.sdpdu; b.ts = normalize(cross(r.sn, r.sdpdu)); b.empty = false;
}
// Material::compute_scattering (material.rs:80-244), mode = RADIANCE, allow_lobes = true
static __device__ __noinline__ void compute_scattering(const DScene& sc, const Rec& r, Bsdf& b) {
    double n1 = 1. / sqrt(3.);
    (void)n1;
    D3 e = normalize(mk3(1., 1., 1.));                 // Bsdf::empty frame (bsdf.rs:37-47)
    b.ns = e; b.ng = e; b.ss = e; b.ts = e; b.n = 0; b.empty = true;
    const rt_material& m = sc.materials[r.mat];
    double u = r.uv.x, v = r.uv.y;
    switch (m.type) {
        case RT_MAT_MATTE: {
            double sig = clampd(m.f[0], 0., 90.);
            D3 color = tex_value(sc, m.tex[0], u, v);
            if (!is_black(color)) {
                bsdf_new(b, r);
                Lobe& l = b.lobes[b.n++];
                if (sig == 0.) lobe_init(l, L_LAMBERT, BSDF_REFLECTION | BSDF_DIFFUSE, color);
                else {
                    lobe_init(l, L_OREN_NAYAR, BSDF_REFLECTION | BSDF_DIFFUSE, color);    // bxdf.rs:909-921
                    double sr = sig * PI / 180.0, s2 = sr * sr;
                    l.a = 1. - (s2 / (2. * (s2 + 0.33))); l.b = 0.45 * s2 / (s2 + 0.09);
                }
            }
            return;
        }
        case RT_MAT_LIGHT: return;
        case RT_MAT_PLASTIC: {
            D3 color = tex_value(sc, m.tex[0], u, v);
            if (!is_black(color)) { bsdf_new(b, r); lobe_init(b.lobes[b.n++], L_LAMBERT, BSDF_REFLECTION | BSDF_DIFFUSE, color); }
            D3 spec = tex_value(sc, m.tex[1], u, v);
            if (!is_black(spec)) {
                if (b.empty) bsdf_new(b, r);
                double rough = m.f[0];
                if (m.remap_roughness) rough = tr_roughness_to_alpha(rough);
                Lobe& l = b.lobes[b.n++];
                lobe_init(l, L_MICRO_R, BSDF_REFLECTION | BSDF_GLOSSY, spec);
                l.fresnel = F_DIELECTRIC; l.a = 1.5; l.b = 1.;                           // eta_i, eta_t
                l.mfd_empty = false; l.alpha_x = fmax(rough, 1e-3); l.alpha_y = fmax(rough, 1e-3);
            }
            return;
        }
        case RT_MAT_GLASS: {
            double eta = m.f[2], urough = m.f[0], vrough = m.f[1];
            D3 rr = tex_value(sc, m.tex[0], u, v), t = tex_value(sc, m.tex[1], u, v);
            bsdf_new(b, r);
            if (is_black(rr) && is_black(t)) return;
            bool is_specular = urough == 0. && vrough == 0.;
            if (is_specular) {                                                            // allow_lobes = true
                Lobe& l = b.lobes[b.n++];
                lobe_init(l, L_FRESNEL_SPEC, BSDF_TRANSMISSION | BSDF_REFLECTION | BSDF_SPECULAR, rr);
                l.aux = t; l.eta_a = eta; l.eta_b = 1.;
                return;
            }
            if (m.remap_roughness) { urough = tr_roughness_to_alpha(urough); vrough = tr_roughness_to_alpha(vrough); }
            if (!is_black(rr)) {
                Lobe& l = b.lobes[b.n++];
                lobe_init(l, L_MICRO_R, BSDF_REFLECTION | BSDF_GLOSSY, rr);
                l.fresnel = F_DIELECTRIC; l.a = eta; l.b = 1.;
                l.mfd_empty = false; l.alpha_x = fmax(urough, 1e-3); l.alpha_y = fmax(vrough, 1e-3);
            }
            if (!is_black(t)) {
                Lobe& l = b.lobes[b.n++];
                lobe_init(l, L_MICRO_T, BSDF_TRANSMISSION | BSDF_GLOSSY, t);
                l.fresnel = F_DIELECTRIC; l.a = eta; l.b = 1.; l.eta_a = eta; l.eta_b = 1.;
                l.mfd_empty = false; l.alpha_x = fmax(urough, 1e-3); l.alpha_y = fmax(vrough, 1e-3);
            }
            return;
        }
        case RT_MAT_METAL: {
            bsdf_new(b, r);
            D3 u_rough = tex_value(sc, m.tex[3] == RT_NONE ? m.tex[2] : m.tex[3], u, v);
            D3 v_rough = tex_value(sc, m.tex[4] == RT_NONE ? m.tex[2] : m.tex[4], u, v);
            double ur = m.remap_roughness ? tr_roughness_to_alpha(u_rough.x) : u_rough.x;
            double vr = m.remap_roughness ? tr_roughness_to_alpha(v_rough.x) : v_rough.x;
            Lobe& l = b.lobes[b.n++];
            lobe_init(l, L_MICRO_R, BSDF_REFLECTION | BSDF_GLOSSY, mk3(1., 1., 1.));
            l.fresnel = F_CONDUCTOR; l.eta3 = tex_value(sc, m.tex[0], u, v); l.aux = tex_value(sc, m.tex[1], u, v);
            l.mfd_empty = false; l.alpha_x = fmax(ur, 1e-3); l.alpha_y = fmax(vr, 1e-3);
            return;
        }
        case RT_MAT_MIRROR: {
            bsdf_new(b, r);
            D3 color = tex_value(sc, m.tex[0], u, v);
            if (!is_black(color)) lobe_init(b.lobes[b.n++], L_SPEC_R, BSDF_REFLECTION | BSDF_SPECULAR, color);
            return;
        }
    }
}

// ---- primitive.rs sampling / light.rs Diffuse
// Primitive::sample_area (primitive.rs:478-539): point and normal on the light's primitive
static __device__ __noinline__ void sample_area(const DScene& sc, uint32_t prim, D2 u, D3& p, D3& n) {
    DPrim pr = sc.prims[prim];
    uint32_t type = pr.info & 7u;
    if (type == RT_PRIM_TRIANGLE) {
        const uint32_t* ti = sc.tri_index + pr.a;
        const DMesh ms = sc.meshes[pr.info >> 6];
        D3 p0 = ld3(sc.points + 3 * (ms.p_base + ti[0])), p1 = ld3(sc.points + 3 * (ms.p_base + ti[1])), p2 = ld3(sc.points + 3 * (ms.p_base + ti[2]));
        double s0 = sqrt(u.x); double b0 = 1. - s0, b1 = u.y * s0;                        // util.rs:62-65
        p = mk3(0., 0., 0.) + (b0 * p0 + b1 * p1 + (1. - b0 - b1) * p2);
        if (pr.info & PI_HAS_N) {
            const double* nn = sc.normals;
            n = normalize(b0 * ld3(nn + 3 * (ms.n_base + ti[0])) + b1 * ld3(nn + 3 * (ms.n_base + ti[1])) + (1. - b0 - b1) * ld3(nn + 3 * (ms.n_base + ti[2])));
        } else n = normalize(cross(p1 - p0, p2 - p0));
    } else {
        const double* sh = sc.shapes + pr.a;
        if (type == RT_PRIM_SPHERE) {
            double z = 1. - 2. * u.x, r = sqrt(fmax(0., 1. - z * z)), phi = 2. * PI * u.y;  // util.rs:51-56
            p = mk3(0., 0., 0.) + sh[3] * mk3(r * cos(phi), r * sin(phi), z);             // around the ORIGIN (primitive.rs:482)
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

}  // namespace rt
