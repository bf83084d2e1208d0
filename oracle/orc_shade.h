// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.h).  PARITY UNPINNED.
// L1/L2 of the reference: textures, materials, BSDF/BxDF lobes, microfacets, area lights,
// the MIS path integrator.  Quirks listed in SURVEY.md §8.1 are reproduced on purpose.
#pragma once
#include "orc_geom.h"

namespace orc {

// ------------------------------------------------------------------ bxdf.rs:12-68 local-frame helpers
inline double cos_theta(const V3& v) { return v.z; }
inline double cos_sq_theta(const V3& v) { return v.z * v.z; }
inline double abs_cos_theta(const V3& v) { return std::fabs(v.z); }
inline double sin_sq_theta(const V3& v) { return rmax(0., 1. - cos_sq_theta(v)); }
inline double sin_theta(const V3& v) { return std::sqrt(sin_sq_theta(v)); }
inline double tan_theta(const V3& v) { return sin_theta(v) / cos_theta(v); }
inline double tan_sq_theta(const V3& v) { return sin_sq_theta(v) / cos_sq_theta(v); }
inline double cos_phi(const V3& v) { double s = sin_theta(v); return s == 0. ? 1. : clampd(v.x / s, -1., 1.); }
inline double sin_phi(const V3& v) { double s = sin_theta(v); return s == 0. ? 1. : clampd(v.y / s, -1., 1.); }
inline double cos_sq_phi(const V3& v) { return cos_phi(v) * cos_phi(v); }
inline double sin_sq_phi(const V3& v) { return sin_phi(v) * sin_phi(v); }
inline bool matches_flags(uint8_t flag, uint8_t other) { return (flag & other) == flag; }
inline bool same_hemisphere(const V3& v, const V3& w) { return v.z * w.z > 0.; }           // util.rs:591-593
inline V3 reflect(const V3& v, const V3& n) { double s = 2. * dot(v, n); return -v + n * s; }  // util.rs:203-206
// util.rs:376-385
inline bool refract(const V3& vec, const V3& n, double eta, V3& out) {
    double cos_theta_i = dot(n, vec) / norm(vec);
    double sin2_theta_i = rmax(0., 1. - cos_theta_i * cos_theta_i);
    double sin2_theta_t = eta * eta * sin2_theta_i;
    if (sin2_theta_t >= 1.) return false;
    double cos_theta_t = std::sqrt(1. - sin2_theta_t);
    out = eta * (-vec) + (eta * cos_theta_i - cos_theta_t) * V3(n.x, n.y, n.z);
    return true;
}

// bxdf.rs:113-136
inline double fr_dielectric(double cos_theta_i, double eta_i, double eta_t) {
    cos_theta_i = clampd(cos_theta_i, -1., 1.);
    double index_i = eta_i, index_t = eta_t;
    if (cos_theta_i < 0.) { index_i = eta_t; index_t = eta_i; cos_theta_i = std::fabs(cos_theta_i); }
    double sin_theta_i = std::sqrt(rmax(0., 1. - cos_theta_i * cos_theta_i));
    double sin_theta_t = index_i / index_t * sin_theta_i;
    double cos_theta_t = std::sqrt(rmax(0., 1. - sin_theta_t * sin_theta_t));
    if (sin_theta_t >= 1.) return 1.;
    double r_parl = ((index_t * cos_theta_i) - (index_i * cos_theta_t)) / ((index_t * cos_theta_i) + (index_i * cos_theta_t));
    double r_perp = ((index_i * cos_theta_i) - (index_t * cos_theta_t)) / ((index_i * cos_theta_i) + (index_t * cos_theta_t));
    return (r_parl * r_parl + r_perp * r_perp) / 2.;
}
// bxdf.rs:141-170
inline V3 fr_conductor(double cos_theta_i, const V3& eta, const V3& eta_k) {
    cos_theta_i = clampd(cos_theta_i, -1., 1.);
    double c2 = cos_theta_i * cos_theta_i, s2 = 1. - c2;
    V3 eta2 = cmul(eta, eta), etak2 = cmul(eta_k, eta_k);
    V3 t0 = (eta2 - etak2) - V3(s2, s2, s2);
    V3 a2pb2 = cmul(t0, t0) + cmul(eta2, etak2) * 4.;
    a2pb2 = V3(std::sqrt(a2pb2.x), std::sqrt(a2pb2.y), std::sqrt(a2pb2.z));
    V3 t1 = a2pb2 + V3(c2, c2, c2);
    V3 a = (a2pb2 + t0) * 0.5;
    a = V3(std::sqrt(a.x), std::sqrt(a.y), std::sqrt(a.z));
    V3 t2 = a * (2. * cos_theta_i);
    V3 rs = cdiv(t1 - t2, t1 + t2);
    V3 t3 = a2pb2 * c2 + V3(s2 * s2, s2 * s2, s2 * s2);
    V3 t4 = t2 * s2;
    V3 rp = cmul(rs, cdiv(t3 - t4, t3 + t4));
    return (rp + rs) * 0.5;
}
inline double color_to_luminance_(const V3& v) { return 0.2126 * v.x + 0.7152 * v.y + 0.0722 * v.z; }   // util.rs:169-171
// bxdf.rs:73-108, util.rs:173-183: helpers of the Disney lobes
inline double lerp(double t, double mn, double mx) { return (1. - t) * mn + t * mx; }
inline V3 lerp_v(double t, const V3& mn, const V3& mx) { return V3(lerp(t, mn.x, mx.x), lerp(t, mn.y, mx.y), lerp(t, mn.z, mx.z)); }
inline double schlick_weight(double u) { double m = clampd(1. - u, 0., 1.); double m2 = m * m; return m2 * m2 * m; }
inline V3 fr_schlick(const V3& r0, double cos_theta) { double w = schlick_weight(cos_theta); return r0 + w * (white() - r0); }
inline double gtr1(double cos_theta, double a) {
    if (cos_theta > 1.) return INV_PI;
    double a2 = a * a;
    return (a2 - 1.) / (PI * std::log(a2) * (1. + (a2 - 1.) * cos_theta * cos_theta));
}
inline double smithg_ggx(double cos_theta, double alpha) { double a2 = alpha * alpha, c2 = cos_theta * cos_theta; return 2. / (1. + std::sqrt(a2 + (1. - a2) * c2)); }
inline double eta_to_r0(double eta) { return (eta - 1.) * (eta - 1.) / ((eta + 1.) * (eta + 1.)); }
// Fresnel::evaluate (bxdf.rs:190-211): the dielectric arm passes (|cos|, eta_t, eta_i) — swapped
inline V3 fresnel_evaluate(const Fresnel& f, double cos_theta_i) {
    switch (f.kind) {
        case F_DIELECTRIC: { double v = fr_dielectric(std::fabs(cos_theta_i), f.eta_t, f.eta_i); return V3(v, v, v); }
        case F_CONDUCTOR: return fr_conductor(std::fabs(cos_theta_i), f.eta, f.k);
        case F_DISNEY: {                                                       // bxdf.rs:199-208: no |cos| here
            double mn = fr_dielectric(cos_theta_i, 1., f.eta_t);
            V3 mx = fr_schlick(f.eta, cos_theta_i);
            return V3(lerp(f.eta_i, mn, mx.x), lerp(f.eta_i, mn, mx.y), lerp(f.eta_i, mn, mx.z));
        }
        default: return white();
    }
}

// ------------------------------------------------------------------ microfacet.rs (TrowbridgeReitz, Empty)
inline double mfd_d(const Mfd& m, const V3& wh) {                       // microfacet.rs:53-68
    if (m.empty) return 0.;
    double t2 = tan_sq_theta(wh);
    if (t2 == INFINITY) return 0.;
    double cos4 = cos_sq_theta(wh) * cos_sq_theta(wh);
    double e = (cos_sq_phi(wh) / (m.alpha_x * m.alpha_x) + sin_sq_phi(wh) / (m.alpha_y * m.alpha_y)) * t2;
    return 1. / (PI * m.alpha_x * m.alpha_y * cos4 * (1. + e) * (1. + e));
}
inline double mfd_lambda(const Mfd& m, const V3& w) {                   // microfacet.rs:109-123
    if (m.empty) return 0.;
    double att = std::fabs(tan_theta(w));
    if (att == INFINITY) return 0.;
    double alpha = std::sqrt(cos_sq_phi(w) * m.alpha_x * m.alpha_x + sin_sq_phi(w) * m.alpha_y * m.alpha_y);
    double a2t2 = (alpha * att) * (alpha * att);
    return (-1. + std::sqrt(1. + a2t2)) / 2.;
}
inline double mfd_g1(const Mfd& m, const V3& w) { return 1. / (1. + mfd_lambda(m, w)); }                 // :143-145
inline double mfd_g(const Mfd& m, const V3& wo, const V3& wi) {          // :146-157: DisneyMicrofacet uses the separable form
    if (m.disney) return mfd_g1(m, wo) * mfd_g1(m, wi);
    return 1. / (1. + mfd_lambda(m, wo) + mfd_lambda(m, wi));
}
inline double mfd_pdf(const Mfd& m, const V3& wo, const V3& wh) {       // microfacet.rs:159-172
    if (!m.empty) return mfd_d(m, wh) * mfd_g1(m, wo) * std::fabs(dot(wo, wh)) / abs_cos_theta(wo);
    return mfd_d(m, wh) * abs_cos_theta(wh);
}
inline double tr_roughness_to_alpha(double roughness) {                 // microfacet.rs:442-446
    roughness = rmax(roughness, 1e-5);
    double x = std::log(roughness);
    return 1.62142 + 0.819955 * x + 0.1734 * x * x + 0.0171201 * x * x * x + 0.000640711 * x * x * x * x;
}
inline Mfd make_trowbridge_reitz(double ax, double ay) { Mfd m; m.empty = false; m.alpha_x = rmax(ax, 1e-3); m.alpha_y = rmax(ay, 1e-3); return m; }  // :340-348
inline V2 tr_sample_11(double cos_t, double u1, double u2) {            // microfacet.rs:469-512
    if (cos_t > 0.9999) {
        double r = std::sqrt(u1 / (1. - u1)), phi = 2. * PI * u2;
        return V2(r * std::cos(phi), r * std::sin(phi));
    }
    double sin_t = rmax(0., std::sqrt(1. - cos_t * cos_t));
    double tan_t = sin_t / cos_t;
    double a = 1. / tan_t;
    double g1 = 2. / (1. + std::sqrt(1. + 1. / (a * a)));
    a = 2. * u1 / g1 - 1.;
    double tmp = 1. / (a * a - 1.);
    if (tmp > 1e10) tmp = 1e10;
    double b = tan_t;
    double d = std::sqrt(rmax(0., b * b * tmp * tmp - (a * a - b * b) * tmp));
    double sx1 = b * tmp - d, sx2 = b * tmp + d;
    double slope_x = (a < 0. || sx2 > 1. / tan_t) ? sx1 : sx2;
    double s, nu2;
    if (u2 > 0.5) { s = 1.; nu2 = 2. * (u2 - 0.5); } else { s = -1.; nu2 = 2. * (0.5 - u2); }
    double z = (nu2 * (nu2 * (nu2 * 0.27385 - 0.73369) + 0.46341)) / (nu2 * (nu2 * (nu2 * 0.093073 + 0.309420) - 1.) + 0.597999);
    double slope_y = s * z * std::sqrt(1. + slope_x * slope_x);
    return V2(slope_x, slope_y);
}
inline V3 tr_sample(const V3& wi, double ax, double ay, double u1, double u2) {  // microfacet.rs:448-467
    V3 wi_s = normalize(V3(ax * wi.x, ay * wi.y, wi.z));
    V2 slope = tr_sample_11(cos_theta(wi_s), u1, u2);
    double sx = slope.x, sy = slope.y;
    double sp = sin_phi(wi_s), cp = cos_phi(wi_s);
    double tmp = cp * sx - sp * sy;
    sy = sp * sx + cp * sy;
    sx = tmp;
    sx = ax * sx; sy = ay * sy;
    return normalize(V3(-sx, -sy, 1.));
}
inline V3 mfd_sample_wh(const Mfd& m, const V3& wo, const V2& u) {     // microfacet.rs:240-283
    if (m.empty) return normalize(black());
    bool flip = wo.z < 0.;
    V3 fwo = flip ? -wo : wo;
    V3 wh = tr_sample(fwo, m.alpha_x, m.alpha_y, u.x, u.y);
    return flip ? -wh : wh;
}

// ------------------------------------------------------------------ bxdf.rs lobes
inline Lobe make_lambert(const V3& c) { Lobe l; l.kind = L_LAMBERT_R; l.type = BSDF_REFLECTION | BSDF_DIFFUSE; l.color = c; return l; }   // :904-907
inline Lobe make_oren_nayar(const V3& c, double sigma) {                // bxdf.rs:909-921
    Lobe l; l.kind = L_OREN_NAYAR; l.type = BSDF_REFLECTION | BSDF_DIFFUSE; l.color = c;
    double sr = sigma * PI / 180.0, s2 = sr * sr;
    l.a = 1. - (s2 / (2. * (s2 + 0.33))); l.b = 0.45 * s2 / (s2 + 0.09);
    return l;
}
inline Lobe make_specular_reflection(const V3& c, const Fresnel& f) { Lobe l; l.kind = L_SPEC_R; l.type = BSDF_REFLECTION | BSDF_SPECULAR; l.color = c; l.fresnel = f; return l; }  // :878-884
inline Lobe make_specular_transmission(const V3& c, double ea, double eb, uint8_t mode) {  // :886-902
    Lobe l; l.kind = L_SPEC_T; l.type = BSDF_TRANSMISSION | BSDF_SPECULAR; l.color = c; l.eta_a = ea; l.eta_b = eb; l.mode = mode;
    l.fresnel.kind = F_DIELECTRIC; l.fresnel.eta_i = ea; l.fresnel.eta_t = eb; return l;
}
inline Lobe make_microfacet_reflection(const V3& c, const Fresnel& f, const Mfd& m) { Lobe l; l.kind = L_MICRO_R; l.type = BSDF_REFLECTION | BSDF_GLOSSY; l.color = c; l.fresnel = f; l.mfd = m; return l; }  // :923-935
inline Lobe make_microfacet_transmission(const V3& c, const Mfd& m, double ea, double eb, uint8_t mode) {  // :937-958
    Lobe l; l.kind = L_MICRO_T; l.type = BSDF_TRANSMISSION | BSDF_GLOSSY; l.color = c; l.mfd = m; l.eta_a = ea; l.eta_b = eb; l.mode = mode;
    l.fresnel.kind = F_DIELECTRIC; l.fresnel.eta_i = ea; l.fresnel.eta_t = eb; return l;
}
inline Mfd make_disney_mfd(double ax, double ay) { Mfd m = make_trowbridge_reitz(ax, ay); m.disney = true; return m; }       // microfacet.rs:349-357
// bxdf.rs:992-1030, 237
inline Lobe make_disney_diffuse(const V3& c) { Lobe l; l.kind = L_DISNEY_DIFFUSE; l.type = BSDF_DIFFUSE | BSDF_REFLECTION; l.color = c; return l; }
inline Lobe make_disney_fake_ss(const V3& c, double rough) { Lobe l; l.kind = L_DISNEY_FAKESS; l.type = BSDF_DIFFUSE | BSDF_REFLECTION; l.color = c; l.a = rough; return l; }
inline Lobe make_disney_retro(const V3& c, double rough) { Lobe l; l.kind = L_DISNEY_RETRO; l.type = BSDF_DIFFUSE | BSDF_REFLECTION; l.color = c; l.a = rough; return l; }
inline Lobe make_disney_sheen(const V3& c) { Lobe l; l.kind = L_DISNEY_SHEEN; l.type = BSDF_DIFFUSE | BSDF_REFLECTION; l.color = c; return l; }
inline Lobe make_disney_clearcoat(double weight, double gloss) { Lobe l; l.kind = L_DISNEY_CLEARCOAT; l.type = BSDF_GLOSSY | BSDF_REFLECTION; l.a = weight; l.b = gloss; return l; }
inline Lobe make_lambertian_transmission(const V3& c) { Lobe l; l.kind = L_LAMBERT_T; l.type = BSDF_TRANSMISSION | BSDF_DIFFUSE; l.color = c; return l; }
inline Lobe make_fresnel_specular(const V3& r, const V3& t, double ea, double eb, uint8_t mode) {  // :974-990
    Lobe l; l.kind = L_FRESNEL_SPEC; l.type = BSDF_TRANSMISSION | BSDF_REFLECTION | BSDF_SPECULAR; l.color = r; l.t = t; l.eta_a = ea; l.eta_b = eb; l.mode = mode; return l;
}

// Bxdf::f (bxdf.rs:328-527)
inline V3 lobe_f(const Lobe& l, const V3& wo, const V3& wi) {
    switch (l.kind) {
        case L_LAMBERT_R: return l.color * INV_PI;
        case L_OREN_NAYAR: {
            double sti = sin_theta(wi), sto = sin_theta(wo), max_cos = 0.;
            if (sti > SMALL && sto > SMALL) {
                double d_cos = cos_phi(wi) * cos_phi(wo) + sin_phi(wi) * sin_phi(wo);
                max_cos = rmax(0., d_cos);
            }
            double sin_alpha, tan_beta;
            if (abs_cos_theta(wi) > abs_cos_theta(wo)) { sin_alpha = sto; tan_beta = sti / abs_cos_theta(wi); }
            else { sin_alpha = sti; tan_beta = sto / abs_cos_theta(wi); }      // |cos θi| in both arms (§8.1 #11)
            return l.color * (INV_PI * (l.a + l.b * max_cos * sin_alpha * tan_beta));
        }
        case L_MICRO_R: {
            double cto = abs_cos_theta(wo), cti = abs_cos_theta(wi);
            V3 wh = wi + wo;
            if (cti == 0. || cto == 0.) return black();
            if (wh == black()) return black();
            wh = normalize(wh);
            V3 f = fresnel_evaluate(l.fresnel, dot(wi, face_forward(wh, V3(0., 0., 1.))));
            V3 comp1 = l.color * mfd_d(l.mfd, wh) * mfd_g(l.mfd, wo, wi);
            return cmul(comp1, f * (1. / (4. * cti * cto)));
        }
        case L_MICRO_T: {
            if (same_hemisphere(wo, wi)) return black();
            double cto = cos_theta(wo), cti = cos_theta(wi);
            if (cti == 0. || cto == 0.) return black();
            double eta = cos_theta(wo) > 0. ? l.eta_b / l.eta_a : l.eta_a / l.eta_b;
            V3 wh = normalize(wo + wi * eta);
            if (wh.z < 0.) wh = -wh;
            if (dot(wo, wh) * dot(wi, wh) > 0.) return black();
            V3 f = fresnel_evaluate(l.fresnel, dot(wo, wh));
            double sqrt_denom = dot(wo, wh) + eta * dot(wi, wh);
            double factor = l.mode == RADIANCE ? 1. / eta : 1.;
            V3 c = cmul(white() - f, l.color);
            return c * std::fabs(mfd_d(l.mfd, wh) * mfd_g(l.mfd, wo, wi) * eta * eta * std::fabs(dot(wi, wh)) * std::fabs(dot(wo, wh)) * factor * factor /
                                 (cti * cto * sqrt_denom * sqrt_denom));
        }
        case L_LAMBERT_T: return l.color * INV_PI;                             // bxdf.rs:337
        case L_DISNEY_DIFFUSE: {                                               // bxdf.rs:465-470
            double fo = schlick_weight(abs_cos_theta(wo)), fi = schlick_weight(abs_cos_theta(wi));
            return (INV_PI * (1. - fo / 2.) * (1. - fi / 2.)) * l.color;
        }
        case L_DISNEY_FAKESS: {                                                // bxdf.rs:471-489
            V3 wh = wi + wo;
            if (wh == black()) return black();
            wh = normalize(wh);
            double cos_theta_d = dot(wi, wh);
            double fss90 = cos_theta_d * cos_theta_d * l.a;
            double fo = schlick_weight(abs_cos_theta(wo)), fi = schlick_weight(abs_cos_theta(wi));
            double fss = lerp(fo, 1., fss90) * lerp(fi, 1., fss90);
            double ss = 1.25 * (fss * (1. / (abs_cos_theta(wo) + abs_cos_theta(wi)) - 0.5) + 0.5);
            return (INV_PI * ss) * l.color;
        }
        case L_DISNEY_RETRO: {                                                 // bxdf.rs:490-504
            V3 wh = wi + wo;
            if (wh == black()) return black();
            wh = normalize(wh);
            double cos_theta_d = dot(wi, wh);
            double fo = schlick_weight(abs_cos_theta(wo)), fi = schlick_weight(abs_cos_theta(wi));
            double rr = 2. * l.a * cos_theta_d * cos_theta_d;
            return (INV_PI * rr * (fo + fi + fo * fi * (rr - 1.))) * l.color;
        }
        case L_DISNEY_SHEEN: {                                                 // bxdf.rs:505-513
            V3 wh = wi + wo;
            if (wh == black()) return black();
            wh = normalize(wh);
            return schlick_weight(dot(wi, wh)) * l.color;
        }
        case L_DISNEY_CLEARCOAT: {                                             // bxdf.rs:514-526
            V3 wh = wi + wo;
            if (wh == black()) return black();
            wh = normalize(wh);
            double dr = gtr1(abs_cos_theta(wh), l.b);
            V3 fr = fr_schlick(white() * 0.04, dot(wo, wh));
            double gr = smithg_ggx(abs_cos_theta(wo), 0.25) * smithg_ggx(abs_cos_theta(wi), 0.25);
            return (l.a / 4. * gr * dr) * fr;
        }
        default: return black();  // SpecularReflection / SpecularTransmission / FresnelSpecular
    }
}
inline double default_pdf(const V3& wo, const V3& wi) { return same_hemisphere(wo, wi) ? abs_cos_theta(wi) * INV_PI : 0.; }  // bxdf.rs:829-835
// Bxdf::pdf (bxdf.rs:721-813)
inline double lobe_pdf(const Lobe& l, const V3& wo, const V3& wi) {
    switch (l.kind) {
        case L_LAMBERT_R: case L_OREN_NAYAR: case L_SPEC_R: return default_pdf(wo, wi);
        case L_DISNEY_DIFFUSE: case L_DISNEY_FAKESS: case L_DISNEY_RETRO: case L_DISNEY_SHEEN: return default_pdf(wo, wi);   // bxdf.rs:792-795
        case L_LAMBERT_T: return !same_hemisphere(wo, wi) ? abs_cos_theta(wi) * INV_PI : 0.;                                // bxdf.rs:726-732
        case L_DISNEY_CLEARCOAT: {                                                                                           // bxdf.rs:796-811
            if (!same_hemisphere(wo, wi)) return 0.;
            V3 wh = wi + wo;
            if (wh == black()) return 0.;
            wh = normalize(wh);
            double dr = gtr1(abs_cos_theta(wh), l.b);
            return dr * abs_cos_theta(wh) / (4. * dot(wo, wh));
        }
        case L_SPEC_T: return 0.;
        case L_MICRO_R: {
            if (!same_hemisphere(wo, wi)) return 0.;
            V3 wh = normalize(wo + wi);
            return mfd_pdf(l.mfd, wo, wh) / (4. * dot(wo, wh));
        }
        case L_MICRO_T: {
            if (same_hemisphere(wo, wi)) return 0.;
            double eta = cos_theta(wo) > 0. ? l.eta_b / l.eta_a : l.eta_a / l.eta_b;
            V3 wh = normalize(wo + wi * eta);
            if (dot(wo, wh) * dot(wi, wh) > 0.) return 0.;
            double sqrt_denom = dot(wo, wh) + eta * dot(wi, wh);
            double dwh_dwi = std::fabs(eta * eta * dot(wi, wh)) / (sqrt_denom * sqrt_denom);
            return mfd_pdf(l.mfd, wo, wh) * dwh_dwi;
        }
        default: return 0.;  // FresnelSpecular: "unimplemented", returns 0 (bxdf.rs:777-791)
    }
}
// util::rand_cosine_dir (util.rs:127-148): two fresh uniforms, NOT the supplied sample (§8.1 #9)
inline V3 rand_cosine_dir(Rng& rng) {
    double r1 = rng.next(), r2 = rng.next();
    double u1 = 2. * r1 - 1., u2 = 2. * r2 - 1.;
    if (u1 == 0. && u2 == 0.) return V3(0., 0., 1.);
    double theta, r;
    if (std::fabs(u1) > std::fabs(u2)) { r = u1; theta = PI / 4. * (u2 / u1); }
    else { r = u2; theta = PI / 2. - PI / 4. * (u1 / u2); }
    double x = r * std::cos(theta), y = r * std::sin(theta);
    double z = std::sqrt(rmax(0., 1. - x * x - y * y));
    return V3(x, y, z);
}
// Bxdf::sample_f (bxdf.rs:532-719) → (color, wi, pdf)
inline void lobe_sample_f(const Lobe& l, const V3& wo, const V2& sample, Rng& rng, V3& color, V3& wi, double& pdf) {
    color = black(); wi = black(); pdf = 0.;
    switch (l.kind) {
        case L_LAMBERT_T: {                                                    // bxdf.rs:709-717: uses the SUPPLIED sample (concentric map)
            V2 uo(2. * sample.x - 1., 2. * sample.y - 1.);                      // util::concentric_sample_disk (util.rs:79-97)
            V2 d(0., 0.);
            if (!(uo.x == 0. && uo.y == 0.)) {
                double theta, r;
                if (std::fabs(uo.x) > std::fabs(uo.y)) { r = uo.x; theta = PI / 4. * (uo.y / uo.x); }
                else { r = uo.y; theta = PI / 2. - PI / 4. * (uo.x / uo.y); }
                d = V2(r * std::cos(theta), r * std::sin(theta));
            }
            wi = V3(d.x, d.y, std::sqrt(rmax(0., 1. - d.x * d.x - d.y * d.y)));
            if (wi.z > 0.) wi.z = -wi.z;
            pdf = lobe_pdf(l, wo, wi); color = lobe_f(l, wo, wi); return;
        }
        case L_DISNEY_CLEARCOAT: {                                             // bxdf.rs:689-708
            if (wo.z == 0.) return;
            double alpha2 = l.b * l.b;
            double cos_t = std::sqrt(rmax(0., (1. - std::pow(alpha2, 1. - sample.x)) / (1. - alpha2)));
            double sin_t = std::sqrt(rmax(0., 1. - cos_t * cos_t));
            double phi = 2. * PI * sample.y;
            V3 wh(sin_t * std::cos(phi), sin_t * std::sin(phi), cos_t);         // util::make_spherical
            if (!same_hemisphere(wo, wh)) wh = -wh;
            V3 w = reflect(wo, normalize(wh));
            if (!same_hemisphere(wo, w)) return;
            wi = w; color = lobe_f(l, wo, wi); pdf = lobe_pdf(l, wo, wi); return;
        }
        case L_LAMBERT_R: case L_OREN_NAYAR:                                    // default_sample_f bxdf.rs:815-827
        case L_DISNEY_DIFFUSE: case L_DISNEY_FAKESS: case L_DISNEY_RETRO: case L_DISNEY_SHEEN: {
            wi = rand_cosine_dir(rng);
            if (wo.z < 0.) wi.z *= -1.;
            pdf = lobe_pdf(l, wo, wi); color = lobe_f(l, wo, wi); return;
        }
        case L_SPEC_R: {
            wi = V3(-wo.x, -wo.y, wo.z);
            color = cmul(l.color, fresnel_evaluate(l.fresnel, cos_theta(wi))); pdf = 1.; return;
        }
        case L_SPEC_T: {
            bool entering = cos_theta(wo) > 0.;
            double eta_i = entering ? l.eta_a : l.eta_b, eta_t = entering ? l.eta_b : l.eta_a;
            V3 dir;
            if (!refract(wo, face_forward(V3(0., 0., 1.), wo), eta_i / eta_t, dir)) { wi = black(); return; }
            V3 ft = cmul(white() - fresnel_evaluate(l.fresnel, cos_theta(dir)), l.color);
            if (l.mode == RADIANCE) ft = ft * ((eta_i * eta_i) / (eta_t * eta_t));
            color = ft; wi = dir; pdf = 1.; return;
        }
        case L_MICRO_R: {
            if (wo.z == 0.) return;
            V3 wh = mfd_sample_wh(l.mfd, wo, sample);
            // the `wo·wh < 0` guard has no `return` in the reference (bxdf.rs:598-600, §8.1 #10)
            V3 w = reflect(wo, wh);
            if (!same_hemisphere(wo, w)) return;
            pdf = mfd_pdf(l.mfd, wo, wh) / (4. * dot(wo, wh));
            wi = w; color = lobe_f(l, wo, wi); return;
        }
        case L_MICRO_T: {
            if (wo.z == 0.) return;
            V3 wh = mfd_sample_wh(l.mfd, wo, sample);
            if (dot(wo, wh) < 0.) return;
            double eta = cos_theta(wo) > 0. ? l.eta_a / l.eta_b : l.eta_b / l.eta_a;
            V3 w;
            if (!refract(wo, wh, eta, w)) return;
            wi = w; pdf = lobe_pdf(l, wo, wi); color = lobe_f(l, wo, wi); return;
        }
        case L_FRESNEL_SPEC: {
            double f = fr_dielectric(cos_theta(wo) / norm(wo), l.eta_a, l.eta_b);
            if (sample.x < f) { wi = V3(-wo.x, -wo.y, wo.z); pdf = f; color = l.color * f; return; }
            bool entering = cos_theta(wo) > 0.;
            double eta_i = entering ? l.eta_a : l.eta_b, eta_t = entering ? l.eta_b : l.eta_a;
            V3 dir;
            if (!refract(wo, face_forward(V3(0., 0., 1.), wo), eta_i / eta_t, dir)) return;
            V3 ft = l.t * (1. - f);
            if (l.mode == RADIANCE) ft = ft * ((eta_i * eta_i) / (eta_t * eta_t));
            color = ft; wi = dir; pdf = 1. - f; return;
        }
    }
}

// ------------------------------------------------------------------ bsdf.rs
inline Bsdf bsdf_new(const HitRecord& r, double eta) {                  // bsdf.rs:26-35
    Bsdf b; b.eta = eta; b.ns = r.shading.n; b.ng = r.n; b.ss = r.shading.dpdu;
    b.ts = normalize(cross(r.shading.n, r.shading.dpdu)); b.n_lobes = 0;
    return b;
}
inline bool bsdf_is_empty(const Bsdf& b) { return b.eta == -100.0; }
inline void bsdf_add(Bsdf& b, const Lobe& l) { b.lobes[b.n_lobes++] = l; }
inline uint32_t bsdf_num_components(const Bsdf& b, uint8_t t) { uint32_t c = 0; for (int i = 0; i < b.n_lobes; i++) if (matches_flags(b.lobes[i].type, t)) c++; return c; }
inline V3 world_to_local(const Bsdf& b, const V3& v) { return V3(dot(v, b.ss), dot(v, b.ts), dot(v, b.ns)); }  // :67-69
inline V3 local_to_world(const Bsdf& b, const V3& v) {                  // bsdf.rs:71-77
    return V3(b.ss.x * v.x + b.ts.x * v.y + b.ns.x * v.z, b.ss.y * v.x + b.ts.y * v.y + b.ns.y * v.z, b.ss.z * v.x + b.ts.z * v.y + b.ns.z * v.z);
}
// Bsdf::f (bsdf.rs:83-98) — `a && b || c` precedence kept (§8.1 #6)
inline V3 bsdf_f(const Bsdf& b, const V3& wow, const V3& wiw, uint8_t flags) {
    V3 wi = world_to_local(b, wiw), wo = world_to_local(b, wow);
    bool refl = dot(wiw, b.ng) * dot(wow, b.ng) > 0.;
    V3 f = black();
    for (int i = 0; i < b.n_lobes; i++) {
        const Lobe& l = b.lobes[i];
        if ((matches_flags(l.type, flags) && (refl && (l.type & BSDF_REFLECTION) > 0)) || (!refl && (l.type & BSDF_TRANSMISSION) > 0))
            f = f + lobe_f(l, wo, wi);
    }
    return f;
}
// Bsdf::sample_f (bsdf.rs:102-164) → (color, wi world, pdf, sampled type)
inline void bsdf_sample_f(const Bsdf& b, const V3& wow, const V2& sample, uint8_t bxdf_type, Rng& rng, V3& color, V3& wiw, double& pdf, uint8_t& sampled) {
    color = black(); wiw = black(); pdf = 0.; sampled = 0;
    uint32_t matching = bsdf_num_components(b, bxdf_type);
    if (matching == 0) return;
    uint32_t comp = std::min((uint32_t)std::floor(sample.x * (double)matching), matching - 1);
    uint32_t count = comp; int used = -1;
    for (int i = 0; i < b.n_lobes; i++) if (matches_flags(b.lobes[i].type, bxdf_type)) { if (count == 0) { used = i; break; } count--; }
    const Lobe& bx = b.lobes[used];
    uint8_t st = bx.type;
    V3 wo = normalize(world_to_local(b, wow));
    if (wo.z == 0.) return;
    V3 c, wi; double p;
    lobe_sample_f(bx, wo, sample, rng, c, wi, p);
    if (p == 0.) return;
    V3 w = local_to_world(b, wi);
    if ((bx.type & BSDF_SPECULAR) == 0 && matching > 1)
        for (int i = 0; i < b.n_lobes; i++) if (i != used && matches_flags(b.lobes[i].type, bxdf_type)) p += lobe_pdf(b.lobes[i], wo, wi);
    if (matching > 1) p = p / (double)matching;
    if ((bx.type & BSDF_SPECULAR) == 0) {
        bool refl = dot(w, b.ng) * dot(wow, b.ng) > 0.;
        c = V3(0., 0., 0.);
        for (int i = 0; i < b.n_lobes; i++) {
            const Lobe& l = b.lobes[i];
            if (matches_flags(l.type, bxdf_type) && ((refl && (l.type & BSDF_REFLECTION) > 0) || (!refl && (l.type & BSDF_TRANSMISSION) > 0)))
                c = c + lobe_f(l, wo, wi);
        }
    }
    color = c; wiw = w; pdf = p; sampled = st;
}
// Bsdf::pdf (bsdf.rs:166-189): sums, does not average (§8.1 #8); wo is NOT normalised here
inline double bsdf_pdf(const Bsdf& b, const V3& wow, const V3& wiw, uint8_t flags) {
    uint32_t n = bsdf_num_components(b, BSDF_ALL);
    if (n == 0) return 0.;
    V3 wo = world_to_local(b, wow), wi = world_to_local(b, wiw);
    if (wo.z == 0.) return 0.;
    double pdf = 0.; int matching = 0;
    for (uint32_t i = 0; i < n; i++) if (matches_flags(b.lobes[i].type, flags)) { matching++; pdf += lobe_pdf(b.lobes[i], wo, wi); }
    return matching > 0 ? pdf : 0.;
}

// ------------------------------------------------------------------ material.rs
// Rust `f64 as u32` / `as usize` / `f32 as u8`: saturating, NaN -> 0
inline uint32_t as_u32(double x) { if (!(x > 0.)) return 0u; if (x >= 4294967295.) return 4294967295u; return (uint32_t)x; }
inline size_t as_usize(double x) { if (!(x > 0.)) return 0; if (x >= 18446744073709551615.) return (size_t)-1; return (size_t)x; }
inline uint8_t as_u8(float x) { if (!(x > 0.f)) return 0; if (x >= 255.f) return 255; return (uint8_t)x; }
// image 0.23.12 (crates.io, not vendored) src/codecs/hdr/decoder.rs `to_rgbe8`, called by material.rs:578
struct Rgbe8 { uint8_t c[3]; uint8_t e; };
inline Rgbe8 to_rgbe8(const float pix[3]) {
    float mx = std::fmax(pix[0], std::fmax(pix[1], pix[2]));
    Rgbe8 o{{0, 0, 0}, 0};
    if (mx <= 0.0f) return o;
    int exp = (int)std::floor(std::log2(mx)) + 1;
    float mul = std::pow(2.0f, (float)exp);                 // f32::powi(2.0, exp): exact for a power of two
    for (int k = 0; k < 3; k++) o.c[k] = as_u8(std::trunc(pix[k] / mul * 256.0f));
    o.e = (uint8_t)(exp + 128);
    return o;
}
// perlin::Perlin::get_noise (perlin.rs:36-85).  `x.floor() as usize` saturates (negative -> 0); the gradient weights use the SMOOTHED u, v, w
inline double perlin_noise(const double* pl, const V3& p) {
    const double scale = pl[0];
    const double *px = pl + 1, *py = pl + 257, *pz = pl + 513, *vec = pl + 769;
    double x = p.x * scale, y = p.y * scale, z = p.z * scale;
    auto as_usize = [](double f) -> uint64_t { if (!(f > 0.)) return 0; if (f >= 1.8446744073709552e19) return ~0ull; return (uint64_t)f; };
    uint64_t i0 = as_usize(std::floor(x)), j0 = as_usize(std::floor(y)), k0 = as_usize(std::floor(z));
    double u = x - std::trunc(x), v = y - std::trunc(y), w = z - std::trunc(z);      // f64::fract
    if (u < 0.) u += 1.;
    if (v < 0.) v += 1.;
    if (w < 0.) w += 1.;
    V3 c[8];
    for (int di = 0; di < 2; di++) for (int dj = 0; dj < 2; dj++) for (int dk = 0; dk < 2; dk++) {
        uint32_t index = (uint32_t)px[(i0 + di) & 255] ^ (uint32_t)py[(j0 + dj) & 255] ^ (uint32_t)pz[(k0 + dk) & 255];
        c[di + 2 * dj + 4 * dk] = V3(vec[3 * index], vec[3 * index + 1], vec[3 * index + 2]);
    }
    u = u * u * (3. - 2. * u); v = v * v * (3. - 2. * v); w = w * w * (3. - 2. * w);
    double accum = 0.;
    for (int i = 0; i < 2; i++) for (int j = 0; j < 2; j++) for (int k = 0; k < 2; k++) {
        V3 weight(u - i, v - j, w - k);
        double i_part = i * u + (1 - i) * (1. - u), j_part = j * v + (1 - j) * (1. - v), k_part = k * w + (1 - k) * (1. - w);
        accum += i_part * j_part * k_part * dot(c[i + 2 * j + 4 * k], weight);
    }
    return accum;
}
// perlin::Perlin::turb (perlin.rs:87-100)
inline double perlin_turb(const double* pl, const V3& p, int depth) {
    double accum = 0., weight = 1.; V3 q = p;
    for (int d = 0; d < depth; d++) { accum += weight * perlin_noise(pl, q); weight *= 0.5; q = V3(q.x * 2., q.y * 2., q.z * 2.); }
    return std::fabs(accum);
}
// Texture::value / get_value (material.rs:542-589): SolidColor, Checkered, Perlin, Hdr; image_value (material.rs:601-611)
inline V3 tex_value(const Scene& s, uint32_t id, double u, double v, const V3& p = V3(0, 0, 0)) {
    const rt_texture& t = s.textures[id];
    if (t.type == RT_TEX_CHECKERED) {
        double mult = std::sin(t.frequency * u * 2. * PI) * std::sin(t.frequency * v * 2. * PI);
        return mult < 0. ? tex_value(s, t.even, u, v, p) : tex_value(s, t.odd, u, v, p);
    }
    if (t.type == RT_TEX_PERLIN) {                                            // material.rs:566-569
        const double* pl = s.perlin.data() + (size_t)t.image * 1537;
        double noise = perlin_turb(pl, p, 7);
        double c = 0.5 * (1. + std::sin(pl[0] * p.z + 10. * noise));
        return V3(c, c, c);
    }
    if (t.type == RT_TEX_HDR) {
        const Image& im = s.images[t.image];
        uint32_t width = im.w, height = im.h;
        uint32_t x = as_u32(std::round((1. - u) * (double)width));
        uint32_t y = as_u32(std::round(v * (double)height));
        x = x % width; y = y % height;
        Rgbe8 q = to_rgbe8(&im.f[3 * ((size_t)y * width + x)]);
        double sc = std::pow(2., (double)((int)q.e - 128));  // (2f64).powi(e - 128)
        return V3(((double)q.c[0] + 0.5) * sc / 256., ((double)q.c[1] + 0.5) * sc / 256., ((double)q.c[2] + 0.5) * sc / 256.);
    }
    if (t.type == RT_TEX_IMAGE) {
        const Image& im = s.images[t.image];
        uint32_t x = as_u32(std::round(u * (double)im.w));
        uint32_t y = as_u32(std::round((1. - v) * (double)im.h));
        x = x % im.w; y = y % im.h;
        const uint8_t* px = &im.b[3 * ((size_t)y * im.w + x)];
        return V3((double)px[0] / 255., (double)px[1] / 255., (double)px[2] / 255.);
    }
    return V3(t.color[0], t.color[1], t.color[2]);
}
// Texture::get_size (material.rs:591-599)
inline void tex_size(const Scene& s, uint32_t id, uint32_t& w, uint32_t& h) {
    const rt_texture& t = s.textures[id];
    if (t.type == RT_TEX_HDR || t.type == RT_TEX_IMAGE) { w = s.images[t.image].w; h = s.images[t.image].h; }
    else if (t.type == RT_TEX_SOLID) { w = h = 1; }
    else { w = h = 0; }
}
// Material::compute_scattering (material.rs:80-244), mode = RADIANCE, allow_lobes = true in the path integrator
inline void compute_scattering(const Scene& s, HitRecord& r, uint8_t mode, bool allow_lobes) {
    const rt_material& m = s.materials[r.mat_index];
    double u = r.uv.x, v = r.uv.y;
    switch (m.type) {
        case RT_MAT_MATTE: {
            double sig = clampd(m.f[0], 0., 90.);
            V3 color = tex_value(s, m.tex[0], u, v, r.p);
            if (color != black()) {
                r.bsdf = bsdf_new(r, 1.);
                if (sig == 0.) bsdf_add(r.bsdf, make_lambert(color)); else bsdf_add(r.bsdf, make_oren_nayar(color, sig));
            }
            return;
        }
        case RT_MAT_LIGHT: return;
        case RT_MAT_PLASTIC: {
            V3 color = tex_value(s, m.tex[0], u, v, r.p);
            if (color != black()) { r.bsdf = bsdf_new(r, 1.); bsdf_add(r.bsdf, make_lambert(color)); }
            V3 spec = tex_value(s, m.tex[1], u, v, r.p);
            if (spec != black()) {
                if (bsdf_is_empty(r.bsdf)) r.bsdf = bsdf_new(r, 1.);
                Fresnel fr; fr.kind = F_DIELECTRIC; fr.eta_i = 1.5; fr.eta_t = 1.;
                double rough = m.f[0];
                if (m.remap_roughness) rough = tr_roughness_to_alpha(rough);
                bsdf_add(r.bsdf, make_microfacet_reflection(spec, fr, make_trowbridge_reitz(rough, rough)));
            }
            return;
        }
        case RT_MAT_GLASS: {
            double eta = m.f[2], urough = m.f[0], vrough = m.f[1];
            V3 rr = tex_value(s, m.tex[0], u, v, r.p), t = tex_value(s, m.tex[1], u, v, r.p);
            r.bsdf = bsdf_new(r, eta);
            if (rr == black() && t == black()) return;
            bool is_specular = urough == 0. && vrough == 0.;
            if (is_specular && allow_lobes) { bsdf_add(r.bsdf, make_fresnel_specular(rr, t, eta, 1., mode)); return; }
            if (m.remap_roughness) { urough = tr_roughness_to_alpha(urough); vrough = tr_roughness_to_alpha(vrough); }
            Mfd mfd = is_specular ? Mfd() : make_trowbridge_reitz(urough, vrough);
            if (!(rr == black())) {
                Fresnel fr; fr.kind = F_DIELECTRIC; fr.eta_i = eta; fr.eta_t = 1.;
                if (is_specular) bsdf_add(r.bsdf, make_specular_reflection(rr, fr)); else bsdf_add(r.bsdf, make_microfacet_reflection(rr, fr, mfd));
            }
            if (!(t == black())) {
                if (is_specular) bsdf_add(r.bsdf, make_specular_transmission(t, eta, 1., mode)); else bsdf_add(r.bsdf, make_microfacet_transmission(t, mfd, eta, 1., mode));
            }
            return;
        }
        case RT_MAT_METAL: {
            r.bsdf = bsdf_new(r, 1.);
            V3 u_rough = tex_value(s, m.tex[3] == RT_NONE ? m.tex[2] : m.tex[3], u, v, r.p);
            V3 v_rough = tex_value(s, m.tex[4] == RT_NONE ? m.tex[2] : m.tex[4], u, v, r.p);
            double ur = m.remap_roughness ? tr_roughness_to_alpha(u_rough.x) : u_rough.x;
            double vr = m.remap_roughness ? tr_roughness_to_alpha(v_rough.x) : v_rough.x;
            Fresnel fr; fr.kind = F_CONDUCTOR; fr.eta = tex_value(s, m.tex[0], u, v, r.p); fr.k = tex_value(s, m.tex[1], u, v, r.p);
            bsdf_add(r.bsdf, make_microfacet_reflection(white(), fr, make_trowbridge_reitz(ur, vr)));
            return;
        }
        case RT_MAT_MIRROR: {
            r.bsdf = bsdf_new(r, 1.);
            V3 color = tex_value(s, m.tex[0], u, v, r.p);
            if (color != black()) { Fresnel fr; fr.kind = F_NOOP; bsdf_add(r.bsdf, make_specular_reflection(color, fr)); }
            return;
        }
        case RT_MAT_DISNEY: {
            // material.rs:245-394.  This arm never calls Bsdf::new: the lobes are added to the record's Bsdf::empty(), whose frame is
            // ns = ng = ss = ts = normalize((1,1,1)) (bsdf.rs:37-47) — kept as it is.
            const rt_disney& d = s.disneys[m.tex[0]];
            const V3& p = r.p;
            V3 color = tex_value(s, d.color, u, v, p);
            double metallic_weight = tex_value(s, d.metallic, u, v, p).x;
            double e = tex_value(s, d.eta, u, v, p).x;
            double strans = tex_value(s, d.spec_trans, u, v, p).x;
            double diffuse_weight = (1. - metallic_weight) * (1. - strans);
            double dt = tex_value(s, d.diff_trans, u, v, p).x / 2.;
            double rough = tex_value(s, d.roughness, u, v, p).x;
            double lum = color_to_luminance_(color);
            V3 c_tint = lum > 0. ? (1. / lum) * color : white();
            double sheen_weight = tex_value(s, d.sheen, u, v, p).x;
            V3 c_sheen = black();
            if (sheen_weight > 0.) { double sheen_tint = tex_value(s, d.sheen_tint, u, v, p).x; c_sheen = lerp_v(sheen_tint, white(), c_tint); }
            if (diffuse_weight > 0.) {
                if (d.thin) {
                    double flat = tex_value(s, d.flatness, u, v, p).x;
                    bsdf_add(r.bsdf, make_disney_diffuse((diffuse_weight * (1. - flat) * (1. - dt)) * color));
                    bsdf_add(r.bsdf, make_disney_fake_ss((diffuse_weight * flat * (1. - dt)) * color, rough));
                } else {
                    V3 sd = tex_value(s, d.scatter_distance, u, v, p);
                    if (sd == black()) bsdf_add(r.bsdf, make_disney_diffuse(diffuse_weight * color));
                    else bsdf_add(r.bsdf, make_specular_transmission(white(), 1., e, mode));
                }
                bsdf_add(r.bsdf, make_disney_retro(diffuse_weight * color, rough));
                if (sheen_weight > 0.) bsdf_add(r.bsdf, make_disney_sheen((diffuse_weight * sheen_weight) * c_sheen));
            }
            double ani = tex_value(s, d.anisotropic, u, v, p).x;
            double aspect = std::sqrt(1. - ani * 0.9);
            double ax = rmax(0.001, rough * rough / aspect), ay = rmax(0.001, rough * rough * aspect);
            Mfd dist = make_disney_mfd(ax, ay);
            double spec_tint = tex_value(s, d.specular_tint, u, v, p).x;
            V3 c_spec_0 = lerp_v(metallic_weight, eta_to_r0(e) * lerp_v(spec_tint, white(), c_tint), color);
            Fresnel fr; fr.kind = F_DISNEY; fr.eta = c_spec_0; fr.eta_i = metallic_weight; fr.eta_t = e;
            bsdf_add(r.bsdf, make_microfacet_reflection(white(), fr, dist));
            double cc = tex_value(s, d.clearcoat, u, v, p).x;
            if (cc > 0.) { double gloss = tex_value(s, d.clearcoat_gloss, u, v, p).x; bsdf_add(r.bsdf, make_disney_clearcoat(cc, lerp(gloss, 0.1, 0.001))); }
            if (strans > 0.) {
                V3 t(strans * std::sqrt(color.x), strans * std::sqrt(color.y), strans * std::sqrt(color.z));
                if (d.thin) {
                    double rscaled = (0.65 * e - 0.35) * rough;
                    double sax = rmax(0.001, rscaled * rscaled / aspect), say = rmax(0.001, rscaled * rscaled * aspect);
                    bsdf_add(r.bsdf, make_microfacet_transmission(t, make_trowbridge_reitz(sax, say), 1., e, mode));
                } else bsdf_add(r.bsdf, make_microfacet_transmission(t, dist, 1., e, mode));
            }
            if (d.thin) bsdf_add(r.bsdf, make_lambertian_transmission(dt * color));
            return;
        }
    }
}

// ------------------------------------------------------------------ primitive.rs sampling, light.rs (Diffuse)
inline V3 uniform_sample_sphere(const V2& u) {                          // util.rs:51-56
    double z = 1. - 2. * u.x, r = std::sqrt(rmax(0., 1. - z * z)), phi = 2. * PI * u.y;
    return V3(r * std::cos(phi), r * std::sin(phi), z);
}
inline V2 uniform_sample_triangle(const V2& u) { double s0 = std::sqrt(u.x); return V2(1. - s0, u.y * s0); }  // util.rs:62-65
// Primitive::sample_area (primitive.rs:478-539) → (record, pdf)
inline void sample_area(const Scene& s, const rt_primitive& pr, const V2& u, HitRecord& rec, double& pdf) {
    switch (pr.type) {
        case RT_PRIM_SPHERE: {
            V3 p_obj = V3(0, 0, 0) + pr.v[3] * uniform_sample_sphere(u);   // around the ORIGIN, not the centre (§8.1 #13)
            rec = record_normal(p_obj, normalize(p_obj - V3(0, 0, 0)));
            break;
        }
        case RT_PRIM_TRIANGLE: {
            const Mesh& m = s.meshes[pr.mesh_index];
            V2 b = uniform_sample_triangle(u);
            V3 p0 = m.p[m.ind[pr.ind]], p1 = m.p[m.ind[pr.ind + 1]], p2 = m.p[m.ind[pr.ind + 2]];
            V3 p = V3(0, 0, 0) + (b.x * p0 + b.y * p1 + (1. - b.x - b.y) * p2);
            rec = record_normal(p, normalize(V3(1., 1., 1.)));
            if (!m.n.empty()) rec.n = normalize(b.x * m.n[m.ind[pr.ind]] + b.y * m.n[m.ind[pr.ind + 1]] + (1. - b.x - b.y) * m.n[m.ind[pr.ind + 2]]);
            else rec.n = normalize(cross(p1 - p0, p2 - p0));
            break;
        }
        case RT_PRIM_XY_RECT: rec = record_normal(V3(pr.v[0] + u.x * (pr.v[2] - pr.v[0]), pr.v[1] + u.y * (pr.v[3] - pr.v[1]), pr.v[4]), normalize(V3(0., 0., 1.))); break;
        case RT_PRIM_XZ_RECT: rec = record_normal(V3(pr.v[0] + u.x * (pr.v[2] - pr.v[0]), pr.v[4], pr.v[1] + u.y * (pr.v[3] - pr.v[1])), normalize(V3(0., 1., 0.))); break;
        default:              rec = record_normal(V3(pr.v[4], pr.v[0] + u.x * (pr.v[2] - pr.v[0]), pr.v[1] + u.y * (pr.v[3] - pr.v[1])), normalize(V3(1., 0., 0.))); break;
    }
    pdf = 1. / prim_area(s, pr);
    if (pr.flip_face) rec.n = -rec.n;                                     // primitive.rs:533-537
}
// Primitive::sample (primitive.rs:445-456) → (record, pdf, wi unnormalised)
inline void prim_sample(const Scene& s, const rt_primitive& pr, const HitRecord& in, const V2& u, HitRecord& rec, double& pdf, V3& wi) {
    sample_area(s, pr, u, rec, pdf);
    wi = rec.p - in.p;
    if (norm2(wi) == 0.) pdf = 0.;
    else { V3 wn = normalize(wi); pdf = pdf * norm2(in.p - rec.p) / std::fabs(dot(rec.n, -wn)); }
}
// Primitive::pdf (primitive.rs:462-473)
inline double prim_pdf(const Scene& s, const rt_primitive& pr, const HitRecord& rec, const V3& dir) {
    Ray ray = spawn_ray(rec, dir);
    HitRecord nr;
    if (!intersects_obj(s, pr, ray, 0., INF, nr)) return 0.;
    V3 dist = rec.p - nr.p;
    return norm2(dist) / (prim_area(s, pr) * std::fabs(dot(nr.n, -dir)));
}
// Light::l (light.rs:475-496)
inline V3 light_l(const Light& l, const HitRecord& rec, const V3& w) {
    if (l.d.type != RT_LIGHT_DIFFUSE) return black();
    if (dot(rec.n, w) > 0. || l.d.two_sided) return V3(l.d.color[0], l.d.color[1], l.d.color[2]);
    return black();
}
// HitRecord::le (hittable.rs:215-225)
inline V3 record_le(const Scene& s, const HitRecord& rec, const V3& w) {
    uint32_t li = s.prims[rec.prim_index].light_index;
    if (li == RT_NONE) return black();
    return light_l(s.lights[li], rec, w);
}
// ------------------------------------------------------------------ distribution.rs
// Distribution1D::make_distribution (distribution.rs:28-55)
inline Distribution1D make_distribution(const double* f, size_t n) {
    Distribution1D d; d.func.assign(f, f + n); d.cdf.assign(n + 1, 0.);
    d.cdf[0] = 0.;
    for (size_t i = 1; i < n + 1; i++) d.cdf[i] = d.cdf[i - 1] + d.func[i - 1] / (double)n;
    d.func_int = d.cdf[n];
    if (d.func_int == 0.) { for (size_t i = 1; i < n + 1; i++) d.cdf[i] = (double)i / (double)n; }
    else { for (size_t i = 1; i < n + 1; i++) d.cdf[i] = d.cdf[i] / d.func_int; }
    return d;
}
// find_interval (distribution.rs:150-166)
template <class Pred> inline size_t find_interval(size_t size, Pred pred) {
    size_t first = 0, len = size;
    while (len > 0) {
        size_t half = len >> 1, middle = first + half;
        if (pred(middle)) { first = middle + 1; len = len - half - 1; } else len = half;
    }
    return as_usize(clampd((double)(first - 1), 0., (double)(size - 2)));   // `first - 1` would wrap at 0; cdf[0] <= u keeps first >= 1
}
// Distribution1D::sample_continuous (distribution.rs:64-78) → (x, pdf, offset)
inline void dist1_sample_continuous(const Distribution1D& d, double u, double& x, double& pdf, size_t& offset) {
    offset = find_interval(d.cdf.size(), [&](size_t i) { return d.cdf[i] <= u; });
    double du = u - d.cdf[offset];
    if (d.cdf[offset + 1] - d.cdf[offset] > 0.) du = du / (d.cdf[offset + 1] - d.cdf[offset]);
    pdf = d.func_int > 0. ? d.func[offset] / d.func_int : 0.;
    x = ((double)offset + du) / (double)d.count();
}
// Distribution2D::make_distribution_2d (distribution.rs:117-130): row v is built from the SLICE f[v..v+nu]
inline Distribution2D make_distribution_2d(const std::vector<double>& f, size_t nu, size_t nv) {
    Distribution2D d;
    for (size_t v = 0; v < nv; v++) d.p_cond_v.push_back(make_distribution(&f[v], nu));
    std::vector<double> marginal;
    for (size_t v = 0; v < nv; v++) marginal.push_back(d.p_cond_v[v].func_int);
    d.p_marginal = make_distribution(marginal.data(), nv);
    return d;
}
// Distribution2D::sample_continuous (distribution.rs:135-139)
inline void dist2_sample_continuous(const Distribution2D& d, const V2& u, V2& out, double& pdf) {
    double d1, pdf1, d0, pdf0; size_t v, off;
    dist1_sample_continuous(d.p_marginal, u.y, d1, pdf1, v);
    dist1_sample_continuous(d.p_cond_v[v], u.x, d0, pdf0, off);
    out = V2(d0, d1); pdf = pdf1 * pdf0;
}
// Distribution2D::pdf (distribution.rs:141-147)
inline double dist2_pdf(const Distribution2D& d, const V2& p) {
    size_t iu = as_usize(p.x * (double)d.p_cond_v[0].count()); iu = std::min(std::max(iu, (size_t)0), d.p_cond_v[0].count() - 1);
    size_t iv = as_usize(p.y * (double)d.p_marginal.count()); iv = std::min(std::max(iv, (size_t)0), d.p_marginal.count() - 1);
    return d.p_cond_v[iv].func[iu] / d.p_marginal.func_int;
}
// ------------------------------------------------------------------ light.rs (Infinite)
constexpr double WORLD_RADIUS = 10000.;                                    // light.rs:636
inline double color_to_luminance(const V3& v) { return 0.2126 * v.x + 0.7152 * v.y + 0.0722 * v.z; }   // util.rs:169-171
inline double spherical_theta(const V3& v) { return std::acos(clampd(v.y, -1., 1.)); }                  // util.rs:153-155
inline double spherical_phi(const V3& v) { double p = std::atan2(v.z, v.x); return p < 0. ? p + 2. * PI : p; }   // util.rs:160-167
inline V3 light_to_world(const Scene& s, const Light& l, const V3& v) { return l.d.transform == RT_NONE ? v : xf_vector(s.xforms[l.d.transform].fwd, v); }
inline V3 light_to_obj(const Scene& s, const Light& l, const V3& v) { return l.d.transform == RT_NONE ? v : xf_vector(s.xforms[l.d.transform].inv, v); }
// Light::make_infinite_light (light.rs:608-638)
inline bool make_infinite_light(const Scene& s, Light& l) {
    uint32_t tw, th; tex_size(s, l.d.text_id, tw, th);
    size_t width = (size_t)tw * 2, height = (size_t)th * 2;
    if (width == 0 || height == 0) return false;
    std::vector<double> img(width * height, 0.);
    for (size_t v = 0; v < height; v++) {
        double vp = ((double)v + 0.5) / (double)height;
        double sin_theta = std::sin(PI * ((double)v + 0.5) / (double)height);
        for (size_t u = 0; u < width; u++) {
            double up = (double)u / (double)width;
            V3 color = tex_value(s, l.d.text_id, up, vp);
            img[u + v * width] = color_to_luminance(color);
            img[u + v * width] = img[u + v * width] * sin_theta;
        }
    }
    l.dist = make_distribution_2d(img, width, height);
    return true;
}
// Light::le (light.rs:499-513)
inline V3 light_le(const Scene& s, const Light& l, const Ray& ray) {
    if (l.d.type != RT_LIGHT_INFINITE) return black();
    V3 w = normalize(light_to_obj(s, l, ray.d));
    V2 uv(spherical_phi(w) * INV_2PI, spherical_theta(w) * INV_PI);
    return tex_value(s, l.d.text_id, uv.x, uv.y);
}

// Light::is_delta_light (light.rs:124-127): Point / Spot carry DELTA_POSITION, Distant DELTA_DIRECTION
inline bool is_delta_light(const Light& l) { return l.d.type == RT_LIGHT_POINT || l.d.type == RT_LIGHT_SPOT || l.d.type == RT_LIGHT_DISTANT; }
constexpr double DISTANT_WORLD_RADIUS = 1e10;                              // light.rs:579
// make_point_light / make_spot_light (light.rs:522-560): position = to_world x origin; the spot cosines are taken of
// (angle x 180 / PI) — the reference converts radians to degrees and then calls cos() on the result; kept as is
inline void make_delta_light(const Scene& s, Light& l) {
    if (l.d.type == RT_LIGHT_DISTANT) return;
    l.pos = l.d.transform == RT_NONE ? V3(0, 0, 0) : xf_point(s.xforms[l.d.transform].fwd, V3(0, 0, 0));
    if (l.d.type == RT_LIGHT_SPOT) { l.cos_width = std::cos(l.d.param[0] * 180. * INV_PI); l.cos_falloff = std::cos(l.d.param[1] * 180. * INV_PI); }
}
// falloff (light.rs:15-26)
inline double spot_falloff(const Scene& s, const Light& l, const V3& w) {
    V3 wl = normalize(light_to_obj(s, l, w));
    double cos_theta = wl.z;
    if (cos_theta < l.cos_width) return 0.;
    else if (cos_theta > l.cos_falloff) return 1.;
    double delta = (cos_theta - l.cos_width) / (l.cos_falloff - l.cos_width);
    return (delta * delta) * (delta * delta);
}

// Light::sample_li (light.rs:136-247) → wi, pdf, color, visibility target.  The delta lights' target record is
// HitRecord::make_basic: prim_index 0, so Visibility::unoccluded (hittable.rs:25-39) lets their light through only where the
// closest hit of the unbounded shadow ray is primitive 0 — the reference's behaviour, restated as it is.
struct LiSample { V3 wi; double pdf; V3 color; V3 p1; uint32_t p1_prim; double p1_t; };
inline LiSample sample_li(const Scene& s, const Light& l, const HitRecord& rec, const V2& u) {
    LiSample o;
    if (l.d.type == RT_LIGHT_POINT || l.d.type == RT_LIGHT_SPOT) {          // light.rs:142-162
        o.wi = normalize(l.pos - rec.p); o.pdf = 1.;
        o.p1 = l.pos; o.p1_prim = 0; o.p1_t = rec.t;
        V3 c(l.d.color[0], l.d.color[1], l.d.color[2]);
        double scale = 1. / norm2(l.pos - rec.p);
        if (l.d.type == RT_LIGHT_SPOT) scale = spot_falloff(s, l, -o.wi) / norm2(l.pos - rec.p);
        o.color = c * scale;
        return o;
    }
    if (l.d.type == RT_LIGHT_DISTANT) {                                     // light.rs:163-175
        V3 dir(l.d.param[0], l.d.param[1], l.d.param[2]);
        o.pdf = 1.;
        o.p1 = rec.p + dir * (2. * DISTANT_WORLD_RADIUS); o.p1_prim = 0; o.p1_t = rec.t;
        o.color = V3(l.d.color[0], l.d.color[1], l.d.color[2]);
        o.wi = normalize(dir);
        return o;
    }
    if (l.d.type == RT_LIGHT_INFINITE) {                                   // light.rs:204-245
        V2 uv; double map_pdf;
        dist2_sample_continuous(l.dist, u, uv, map_pdf);
        if (map_pdf == 0.) { o.wi = normalize(black()); o.pdf = 0.; o.color = black(); o.p1 = V3(0, 0, 0); o.p1_prim = 0; o.p1_t = 1.; return o; }
        double theta = uv.y * PI, phi = uv.x * 2. * PI;
        double cos_theta = std::cos(theta), sin_theta = std::sin(theta), cos_phi = std::cos(phi), sin_phi = std::sin(phi);
        V3 v(sin_theta * cos_phi, sin_theta * sin_phi, cos_theta);
        V3 wi = light_to_world(s, l, v);
        double pdf = map_pdf / (2. * PI * PI * sin_theta);
        if (sin_theta == 0.) pdf = 0.;
        o.p1 = rec.p + wi * (2. * WORLD_RADIUS); o.p1_prim = 0; o.p1_t = rec.t;     // HitRecord::make_basic: prim_index 0
        o.color = tex_value(s, l.d.text_id, uv.x, uv.y);
        o.wi = normalize(wi); o.pdf = pdf;
        return o;
    }
    const rt_primitive& pr = s.prims[l.d.prim_index];
    HitRecord nr; double pdf; V3 wi;
    prim_sample(s, pr, rec, u, nr, pdf, wi);
    nr.prim_index = l.d.prim_index;
    if (pdf == 0. || norm2(rec.p - nr.p) == 0.) {
        o.wi = normalize(black()); o.pdf = 0.; o.color = V3(l.d.color[0], l.d.color[1], l.d.color[2]);
        o.p1 = V3(0, 0, 0); o.p1_prim = 0; o.p1_t = 1.;
        return o;
    }
    o.wi = normalize(nr.p - rec.p);
    o.pdf = pdf;
    o.color = light_l(l, nr, -o.wi);
    o.p1 = nr.p; o.p1_prim = nr.prim_index; o.p1_t = nr.t;
    return o;
}
// Light::pdf_li, Diffuse arm (light.rs:284)
inline double pdf_li(const Scene& s, const Light& l, const HitRecord& rec, const V3& wi) {
    if (is_delta_light(l)) return 0.;                                       // light.rs:281-283
    if (l.d.type == RT_LIGHT_INFINITE) {                                   // light.rs:285-294
        V3 w = light_to_world(s, l, wi);
        double theta = spherical_theta(w), phi = spherical_phi(w);
        double sin_theta = std::sin(theta);
        if (sin_theta == 0.) return 0.;
        return dist2_pdf(l.dist, V2(phi * INV_2PI, theta * INV_PI)) / (2. * PI * PI * sin_theta);
    }
    return prim_pdf(s, s.prims[l.d.prim_index], rec, wi);
}

// ------------------------------------------------------------------ integrator.rs (Path)
struct RayCounts {
    uint64_t extend = 0, shadow = 0, probe = 0;
    // optional: algorithmic work of the extend rays under the ordered "reference traversal" of SURVEY.md §8(d)
    bool count_ordered = false; TravStats ordered;
};
inline double power_heuristic(int nf, double f_pdf, int ng, double g_pdf) {     // integrator.rs:655-659
    double f = nf * f_pdf, g = ng * g_pdf;
    return (f * f) / (f * f + g * g);
}
// estimate_direct (integrator.rs:559-634), specular = false
inline V3 estimate_direct(const Scene& s, const HitRecord& rec, const V2& u_scatter, uint32_t light_num, const V2& u_light, Rng& rng, RayCounts& rc) {
    const Light& light = s.lights[light_num];
    uint8_t flags = BSDF_ALL - BSDF_SPECULAR;
    V3 ld = black();
    LiSample ls = sample_li(s, light, rec, u_light);
    double light_pdf = ls.pdf; V3 color = ls.color;
    if (light_pdf > 0. && color != black()) {
        V3 f = bsdf_f(rec.bsdf, rec.wo, ls.wi, flags) * std::fabs(dot(ls.wi, rec.shading.n));
        double scattering_pdf = bsdf_pdf(rec.bsdf, rec.wo, ls.wi, flags);
        if (f != black()) {
            rc.shadow++;
            if (!unoccluded(s, rec.p, ls.p1, ls.p1_prim, ls.p1_t, light.d.type == RT_LIGHT_INFINITE)) color = black();
            if (color != black()) {
                if (is_delta_light(light)) ld = ld + cmul(f, color) * (1. / light_pdf);   // integrator.rs:587-588
                else {
                    double weight = power_heuristic(1, light_pdf, 1, scattering_pdf);
                    ld = ld + cmul(f, color) * (weight / light_pdf);
                }
            }
        }
    }
    if (!is_delta_light(light)) {                                            // integrator.rs:601
        V3 f, wi; double scattering_pdf; uint8_t st;
        bsdf_sample_f(rec.bsdf, rec.wo, u_scatter, flags, rng, f, wi, scattering_pdf, st);
        f = f * std::fabs(dot(wi, rec.shading.n));
        bool sampled_specular = (st & BSDF_SPECULAR) != 0;
        if (f != black() && scattering_pdf > 0.) {
            double weight = 1.;
            if (!sampled_specular) {
                light_pdf = pdf_li(s, light, rec, wi);
                if (light_pdf == 0.) return ld;
                weight = power_heuristic(1, scattering_pdf, 1, light_pdf);
            }
            Ray nray = spawn_ray(rec, wi);
            HitRecord nrec; rc.probe++;
            V3 c = black();
            if (get_intersection(s, nray, nrec)) {
                uint32_t li = s.prims[nrec.prim_index].light_index;
                if (li != RT_NONE && s.lights[li].eq_class == light.eq_class && s.lights[li].eq_class < RT_NONE / 2) c = record_le(s, nrec, -wi);      // `lights[index].eq(light)`: field equality, not identity
            } else c = light_le(s, light, nray);
            if (c != black()) ld = ld + cmul(f, c) * (weight / scattering_pdf);
        }
    }
    return ld;
}
// uniform_sample_one_light (integrator.rs:530-557)
inline V3 uniform_sample_one_light(const Scene& s, const HitRecord& rec, Rng& rng, RayCounts& rc) {
    size_t n = s.lights.size();
    if (n == 0) return black();
    size_t light_num = std::min(n - 1, (size_t)(rng.next() * (double)n));
    V2 u_light; u_light.x = rng.next(); u_light.y = rng.next();
    V2 u_scatter; u_scatter.x = rng.next(); u_scatter.y = rng.next();
    return estimate_direct(s, rec, u_scatter, (uint32_t)light_num, u_light, rng, rc) * (double)n;
}
// PathIntegrator::li (integrator.rs:375-445)
inline V3 path_li(const Scene& s, Ray ray, uint32_t max_depth, bool invisible_light, Rng& rng, RayCounts& rc) {
    V3 beta = white(), l = black();
    bool specular_bounce = false;
    uint32_t bounces = 0;
    for (;;) {
        HitRecord rec; rc.extend++;
        bool is_some = get_intersection(s, ray, rec);
        if (rc.count_ordered) { bool h2; HitRecord r2; closest_ordered(s, ray, SMALL, INF, h2, r2, rc.ordered); }
        if (!is_some) rec = record_basic(V3(0, 0, 0), 0.);
        if (bounces == 0 || specular_bounce) {
            if (is_some) {
                if (s.prims[rec.prim_index].light_index != RT_NONE) {
                    if (invisible_light) { ray = spawn_ray(rec, ray.d); continue; }
                    l = l + cmul(record_le(s, rec, -ray.d), beta);
                }
            } else { for (const Light& light : s.lights) l = l + cmul(light_le(s, light, ray), beta); }
        }
        if (!is_some || bounces >= max_depth) break;
        compute_scattering(s, rec, RADIANCE, true);
        l = l + cmul(uniform_sample_one_light(s, rec, rng, rc), beta);
        V3 wo = -ray.d;
        V2 u; u.x = rng.next(); u.y = rng.next();
        V3 f, wi; double pdf; uint8_t flags;
        bsdf_sample_f(rec.bsdf, wo, u, BSDF_ALL, rng, f, wi, pdf, flags);
        if (f == black() || pdf == 0.) break;
        beta = cmul(beta, f) * (std::fabs(dot(wi, rec.shading.n)) / pdf);
        specular_bounce = (flags & BSDF_SPECULAR) != 0;
        ray = spawn_ray(rec, wi);
        if (bounces > 3) {
            double q = rmax(0.05, 1. - rmax(beta.x, rmax(beta.y, beta.z)));
            if (rng.next() < q) break;
            beta = beta * (1. / (1. - q));
        }
        bounces++;
    }
    return l;
}

// WhittedIntegrator::li (integrator.rs:131-165) and DirectLightingIntegrator::li (integrator.rs:257-297) with specular_reflect /
// specular_transmit (integrator.rs:464-504) and uniform_sample_all_lights (integrator.rs:506-528: ONE (u_light, u_scatter) pair per
// light, summed, not divided by the sample count — the loop over `samples` is commented out in the reference).
// kind: 1 = Direct{UniformOne}, 2 = Direct{UniformAll}, 3 = Whitted (rtcore.h RT_INTEGRATOR_*; 4 = Basic is basic_li below).
// Random numbers: the reference threads one sampler through the recursion depth-first; here both branch samples are drawn at the
// vertex (reflect, then transmit) and each child continues from the vertex's dimension counter — the rule the wavefront kernel
// follows, where the two children are independent queue entries.  Both are streams of iid uniforms.
inline V3 sampler_li(const Scene& s, const Ray& ray, uint32_t depth, uint32_t kind, Rng& rng, RayCounts& rc) {
    HitRecord rec; rc.extend++;
    if (!get_intersection(s, ray, rec)) {
        V3 a = black();
        for (const Light& light : s.lights) a = a + light_le(s, light, ray);
        return a;
    }
    V3 ans = record_le(s, rec, rec.wo);
    const V3 n = rec.shading.n, wo = rec.wo;
    compute_scattering(s, rec, RADIANCE, true);
    if (kind == 3) {
        for (const Light& light : s.lights) {
            V2 u; u.x = rng.next(); u.y = rng.next();
            LiSample ls = sample_li(s, light, rec, u);
            if (ls.color == black() || ls.pdf == 0.) continue;
            V3 f = bsdf_f(rec.bsdf, wo, ls.wi, BSDF_ALL);
            if (f != black()) {
                rc.shadow++;
                if (unoccluded(s, rec.p, ls.p1, ls.p1_prim, ls.p1_t, light.d.type == RT_LIGHT_INFINITE)) ans = ans + cmul(f, ls.color) * (std::fabs(dot(ls.wi, n)) / ls.pdf);
            }
        }
    } else if (!s.lights.empty()) {
        if (kind == 2) {
            for (size_t j = 0; j < s.lights.size(); j++) {
                V2 u_light; u_light.x = rng.next(); u_light.y = rng.next();
                V2 u_scatter; u_scatter.x = rng.next(); u_scatter.y = rng.next();
                ans = ans + estimate_direct(s, rec, u_scatter, (uint32_t)j, u_light, rng, rc);
            }
        } else ans = ans + uniform_sample_one_light(s, rec, rng, rc);
    }
    if (depth - 1 > 0) {
        V2 ur; ur.x = rng.next(); ur.y = rng.next();
        V2 ut; ut.x = rng.next(); ut.y = rng.next();
        const uint8_t want[2] = {(uint8_t)(BSDF_REFLECTION | BSDF_SPECULAR), (uint8_t)(BSDF_TRANSMISSION | BSDF_SPECULAR)};
        const V2 us[2] = {ur, ut};
        for (int b = 0; b < 2; b++) {
            V3 f, wi; double pdf; uint8_t st;
            Rng scratch = rng;            // specular lobes draw nothing
            bsdf_sample_f(rec.bsdf, wo, us[b], want[b], scratch, f, wi, pdf, st);
            if (pdf > 0. && f != black() && std::fabs(dot(n, wi)) != 0.) {
                Rng child = rng;
                V3 li = sampler_li(s, spawn_ray(rec, wi), depth - 1, kind, child, rc);
                ans = ans + cmul(f, li) * (std::fabs(dot(n, wi)) / pdf);
            }
        }
    }
    return ans;
}

// BasicIntegrator → geometry::cast_ray (geometry.rs:242-277): no light sampling; a hit light returns its emission and ends the walk,
// anything else is continued through one BSDF sample over ALL lobes, weighted by color / pdf and by |ray.dir . ns| — the INCOMING,
// unnormalised direction (so camera-ray lengths scale the image; kept, it is what the reference computes); a miss returns
// sum(light.le) / 4 pi; `depth` counts vertices.
inline V3 basic_li(const Scene& s, const Ray& ray, uint32_t depth, Rng& rng, RayCounts& rc) {
    if (depth == 0) return black();
    HitRecord rec; rc.extend++;
    if (!get_intersection(s, ray, rec)) {
        V3 l = black();
        for (const Light& light : s.lights) l = l + light_le(s, light, ray) * (1. / (4. * PI));
        return l;
    }
    if (s.prims[rec.prim_index].light_index != RT_NONE) return record_le(s, rec, -ray.d);
    compute_scattering(s, rec, RADIANCE, true);
    V2 u; u.x = rng.next(); u.y = rng.next();
    V3 f, wi; double pdf; uint8_t st;
    bsdf_sample_f(rec.bsdf, -ray.d, u, BSDF_ALL, rng, f, wi, pdf, st);
    if (pdf == 0.) return black();
    V3 incoming = basic_li(s, Ray{rec.p, wi, ray.time}, depth - 1, rng, rc);
    return cmul(incoming, f * (1. / pdf)) * std::fabs(dot(ray.d, rec.shading.n));
}

// Camera::get_ray (geometry.rs:177-190) with util::rand_in_disk (util.rs:105-113) and rand_range
inline Ray camera_get_ray(const rt_camera& c, double u, double v, Rng& rng) {
    V3 in_disk;
    for (;;) { double x = rng.next(), y = rng.next(); if (x * x + y * y < 1.) { in_disk = V3(x, y, 0.); break; } }
    in_disk = in_disk * c.lens_radius;
    V3 cu(c.u[0], c.u[1], c.u[2]), cv(c.v[0], c.v[1], c.v[2]);
    V3 offset = cu * in_disk.x + cv * in_disk.y;
    V3 ulc(c.upper_left_corner[0], c.upper_left_corner[1], c.upper_left_corner[2]);
    V3 ho(c.horizontal_offset[0], c.horizontal_offset[1], c.horizontal_offset[2]);
    V3 vo(c.vertical_offset[0], c.vertical_offset[1], c.vertical_offset[2]);
    V3 origin(c.origin[0], c.origin[1], c.origin[2]);
    V3 to = ulc + ho * u - vo * v;
    V3 dir = to - origin;
    double time = c.t0 + (c.t1 - c.t0) * rng.next();
    return Ray{origin + offset, dir - offset, time};
}
// PathIntegrator::single_sample (integrator.rs:357-369) + Samplers::get_camera_sample (sampler.rs:606-613)
inline V3 single_sample(const Scene& s, const rt_camera& cam, uint32_t width, uint32_t height, uint32_t px, uint32_t py,
                        uint32_t sample, uint64_t seed, uint32_t max_depth, bool invisible_light, RayCounts& rc, uint32_t kind = 0) {
    Rng rng(seed, py * width + px, sample);
    double ox = rng.next(), oy = rng.next();
    double fx = (double)px + ox, fy = (double)py + oy;
    rng.next();              // time
    rng.next(); rng.next();  // lens
    if (kind == 4) {      // BasicIntegrator::single_sample (integrator.rs:196-204) divides by width - 1 / height - 1
        Ray ray = camera_get_ray(cam, fx / (double)(width - 1), fy / (double)(height - 1), rng);
        return basic_li(s, ray, max_depth, rng, rc);
    }
    Ray ray = camera_get_ray(cam, fx / (double)width, fy / (double)height, rng);
    if (kind != 0) return sampler_li(s, ray, max_depth, kind, rng, rc);
    return path_li(s, ray, max_depth, invisible_light, rng, rc);
}

// Camera::new_motion_blur (geometry.rs:133-175)
inline rt_camera camera_new(const V3& from, const V3& to, const V3& up, double aspect, double vfov, double aperture, double focus, double t0, double t1) {
    V3 w = normalize(to - from);
    V3 u = -normalize(cross(up, w));
    V3 v = -normalize(cross(w, u));
    double theta = vfov * PI / 180., h = std::tan(theta / 2.);
    double vh = 2. * h, vw = vh * aspect;
    V3 ho = u * (vw * focus), vo = v * (vh * focus);
    V3 ulc = from - ho * 0.5 + vo * 0.5 + w * focus;
    rt_camera c;
    auto put = [](double* d, const V3& a) { d[0] = a.x; d[1] = a.y; d[2] = a.z; };
    put(c.origin, from); put(c.upper_left_corner, ulc); put(c.horizontal_offset, ho); put(c.vertical_offset, vo);
    put(c.u, u); put(c.v, v); put(c.w, w);
    c.lens_radius = aperture / 2.; c.t0 = t0; c.t1 = t1;
    return c;
}

// util::point_to_color (util.rs:441-471) with aces_comp (util.rs:400-408)
inline double aces_comp(double x) {
    x = x * 0.6;
    double a = 2.51, b = 0.03, c = 2.43, d = 0.59, e = 0.14;
    return clampd((x * (a * x + b)) / (x * (c * x + d) + e), 0., 1.);
}
inline uint8_t to_u8(double v) { double r = std::round(v); if (!(r > 0.)) return 0; if (r > 255.) return 255; return (uint8_t)r; }  // Rust `as u8` saturates, NaN → 0
inline void point_to_color(const double rgbn[4], uint8_t out[3]) {
    double scale = 1. / rgbn[3];
    for (int i = 0; i < 3; i++) out[i] = to_u8(std::pow(aces_comp(rgbn[i] * scale), 1. / GAMMA) * 256.);
}

}  // namespace orc
