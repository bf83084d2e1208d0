// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.h).  PARITY UNPINNED.
// L0 of the reference: rays, boxes, primitives, intersection, hit record, BVH.
#pragma once
#include <thread>
#include "orc_math.h"
#include "../include/rtcore.h"

namespace orc {

struct Ray { V3 o, d; double time; };                       // geometry.rs:211-236
inline V3 ray_at(const Ray& r, double t) { return r.o + r.d * t; }  // geometry.rs:227

struct Box { V3 mn, mx; };                                  // hittable.rs:485-488

// BoundingBox::intersects (hittable.rs:494-508)
inline bool box_intersects(const Box& b, const Ray& ray, double tmin, double tmax) {
    for (int a = 0; a < 3; a++) {
        double inv_d = 1.0 / ray.d[a];
        double val1 = (b.mn[a] - ray.o[a]) * inv_d;
        double val2 = (b.mx[a] - ray.o[a]) * inv_d;
        double t0 = rmin(val1, val2);
        double t1 = rmax(val1, val2);
        tmin = rmax(tmin, t0);
        tmax = rmin(tmax, t1);
        if (tmax <= tmin) return false;
    }
    return true;
}
inline Box box_union(const Box& a, const Box& b) {         // hittable.rs:551-563
    Box r;
    r.mn = V3(rmin(a.mn.x, b.mn.x), rmin(a.mn.y, b.mn.y), rmin(a.mn.z, b.mn.z));
    r.mx = V3(rmax(a.mx.x, b.mx.x), rmax(a.mx.y, b.mx.y), rmax(a.mx.z, b.mx.z));
    return r;
}

struct Mesh {                                               // hittable.rs:242-250
    std::vector<uint32_t> ind;
    std::vector<V3> p, n;
    std::vector<V2> uv;
};

struct Shading { V3 n, dpdu, dpdv; };                       // hittable.rs:43-49 (dndu/dndv never read)

// bxdf.rs:172-188.  DisneyFresnel { r0, metallic, eta }: r0 in `eta`, metallic in `eta_i`, eta in `eta_t`
enum FresnelKind { F_DIELECTRIC = 0, F_CONDUCTOR = 1, F_NOOP = 2, F_DISNEY = 3 };
struct Fresnel { int kind = F_NOOP; double eta_i = 1, eta_t = 1; V3 eta, k; };
// microfacet.rs:11-28: TrowbridgeReitz{sample_visible_area: true}, DisneyMicrofacet{sample_visible_area: true} (same D, lambda and sampling,
// separable G: microfacet.rs:146-155) and Empty are what the materials build (material.rs:129,166,168,228,334,378); Beckmann is never constructed
struct Mfd { bool empty = true; double alpha_x = 0, alpha_y = 0; bool disney = false; };
// bxdf.rs:213-300: every lobe a Material can build.  ScaledBxdf and FresnelBlend have constructors (bxdf.rs:877,960) but no call site.
enum LobeKind { L_LAMBERT_R = 0, L_OREN_NAYAR = 1, L_SPEC_R = 2, L_SPEC_T = 3, L_FRESNEL_SPEC = 4, L_MICRO_R = 5, L_MICRO_T = 6,
                L_DISNEY_DIFFUSE = 7, L_DISNEY_FAKESS = 8, L_DISNEY_RETRO = 9, L_DISNEY_SHEEN = 10, L_DISNEY_CLEARCOAT = 11, L_LAMBERT_T = 12 };
struct Lobe {
    int kind = L_LAMBERT_R;
    uint8_t type = 0;        // bxdf_type flag bits
    V3 color;                // color / r
    V3 t;                    // FresnelSpecular t
    double a = 0, b = 0;     // OrenNayer A, B ; DisneyFakeSS / DisneyRetro roughness (a) ; DisneyClearCoat weight (a), gloss (b)
    Fresnel fresnel;
    Mfd mfd;
    double eta_a = 1, eta_b = 1;
    uint8_t mode = RADIANCE;
};
struct Bsdf {                                               // bsdf.rs:13-20
    double eta = -100.0;                                    // EMPTY_ETA bsdf.rs:11
    V3 ns, ng, ss, ts;
    int n_lobes = 0;
    Lobe lobes[8];           // Vec<Box<Bxdf>>: at most 2 for Matte .. Mirror, at most 8 for Disney (material.rs:245-394)
};

struct HitRecord {                                          // hittable.rs:51-72
    double t = 0;
    V3 n, p;
    bool front = false;
    V2 uv;
    uint32_t mat_index = 0, prim_index = 0;
    V3 dpdu, dpdv;
    Shading shading;
    V3 wo;
    Bsdf bsdf;
};

// util.rs:567-576
inline void make_coordinate_system(const V3& v1, V3& v2, V3& v3) {
    if (std::fabs(v1.x) > std::fabs(v1.y)) v2 = V3(-v1.z, 0., v1.x) * (1. / std::sqrt(v1.x * v1.x + v1.z * v1.z));
    else v2 = V3(0., v1.z, -v1.y) * (1. / std::sqrt(v1.y * v1.y + v1.z * v1.z));
    v3 = cross(v1, v2);
}
// util.rs:578-581
inline V3 face_forward(const V3& n, const V3& v) { return dot(n, v) < 0. ? -n : n; }

// Bsdf::empty (bsdf.rs:37-47)
inline Bsdf bsdf_empty() {
    Bsdf b; V3 n = normalize(white());
    b.eta = -100.0; b.ns = n; b.ng = n; b.ss = n; b.ts = n; b.n_lobes = 0;
    return b;
}

// HitRecord::new (hittable.rs:75-117)
inline HitRecord record_new(const V3& p, const V2& uv, const V3& wo, const V3& dpdu, const V3& dpdv, double t,
                            uint32_t prim_index, uint32_t mat_index) {
    HitRecord r;
    V3 n = normalize(cross(dpdu, dpdv));
    r.shading.n = n; r.shading.dpdu = normalize(dpdu); r.shading.dpdv = normalize(dpdv);
    r.p = p; r.n = n; r.t = t; r.front = false; r.uv = uv; r.mat_index = mat_index;
    r.dpdu = dpdu; r.dpdv = dpdv; r.wo = wo; r.prim_index = prim_index;
    r.bsdf = bsdf_empty();
    return r;
}
// HitRecord::make_basic (hittable.rs:119-151): unit vectors are normalize(0) = NaN
inline HitRecord record_basic(const V3& p, double t) {
    HitRecord r; V3 ub = normalize(black());
    r.shading.n = ub; r.shading.dpdu = ub; r.shading.dpdv = ub; r.p = p; r.n = ub; r.t = t;
    r.bsdf = bsdf_empty();
    return r;
}
// HitRecord::make_normal (hittable.rs:153-184)
inline HitRecord record_normal(const V3& p, const V3& n) {
    HitRecord r;
    r.shading.n = n; r.shading.dpdu = n; r.shading.dpdv = n; r.p = p; r.n = n; r.t = 0;
    r.bsdf = bsdf_empty();
    return r;
}
// HitRecord::set_front (hittable.rs:186-189)
inline void set_front(HitRecord& r, const Ray& ray) {
    r.front = dot(ray.d, r.n) < 0.0;
    r.n = r.front ? r.n : -r.n;
}
// HitRecord::set_shading_geometry (hittable.rs:191-210)
inline void set_shading_geometry(HitRecord& r, const V3& dpdus, const V3& dpdvs, bool is_auth) {
    V3 n = normalize(cross(dpdus, dpdvs));
    r.shading.n = n;
    if (is_auth) r.n = face_forward(r.n, r.shading.n);
    else r.shading.n = face_forward(r.shading.n, r.n);
    r.shading.dpdu = dpdus; r.shading.dpdv = dpdvs;
}
// HitRecord::spawn_ray (hittable.rs:230-232)
inline Ray spawn_ray(const HitRecord& r, const V3& d) { return Ray{r.p, d, r.t}; }

// Mesh::get_uv (hittable.rs:454-468)
inline void mesh_get_uv(const Mesh& m, uint32_t ind, V2 uvs[3]) {
    if (m.uv.empty()) { uvs[0] = V2(0., 0.); uvs[1] = V2(1., 0.); uvs[2] = V2(1., 1.); }
    else { uvs[0] = m.uv[m.ind[ind]]; uvs[1] = m.uv[m.ind[ind + 1]]; uvs[2] = m.uv[m.ind[ind + 2]]; }
}

// Mesh::intersects_triangle (hittable.rs:292-452).  `tmin` is ignored by the reference.
inline bool intersects_triangle(const Mesh& mesh, const Ray& ray, uint32_t ind, uint32_t mat_index, double /*tmin*/,
                                double tmax, HitRecord& out) {
    uint32_t ind1 = mesh.ind[ind], ind2 = mesh.ind[ind + 1], ind3 = mesh.ind[ind + 2];
    V3 p0 = mesh.p[ind1], p1 = mesh.p[ind2], p2 = mesh.p[ind3];
    V3 dir = ray.d;
    V3 p0t = p0 - ray.o, p1t = p1 - ray.o, p2t = p2 - ray.o;
    // dir.abs().argmax(): first maximum
    double ax = std::fabs(dir.x), ay = std::fabs(dir.y), az = std::fabs(dir.z);
    int kz = 0; double best = ax;
    if (ay > best) { kz = 1; best = ay; }
    if (az > best) { kz = 2; best = az; }
    int kx = (kz + 1) % 3, ky = (kx + 1) % 3;
    V3 d(dir[kx], dir[ky], dir[kz]);
    p0t = V3(p0t[kx], p0t[ky], p0t[kz]);
    p1t = V3(p1t[kx], p1t[ky], p1t[kz]);
    p2t = V3(p2t[kx], p2t[ky], p2t[kz]);
    double s_x = -d.x / d.z, s_y = -d.y / d.z, s_z = 1. / d.z;
    p0t.x += s_x * p0t.z; p0t.y += s_y * p0t.z;
    p1t.x += s_x * p1t.z; p1t.y += s_y * p1t.z;
    p2t.x += s_x * p2t.z; p2t.y += s_y * p2t.z;
    double e0 = p1t.x * p2t.y - p1t.y * p2t.x;
    double e1 = p2t.x * p0t.y - p2t.y * p0t.x;
    double e2 = p0t.x * p1t.y - p0t.y * p1t.x;
    if ((e0 < 0. || e1 < 0. || e2 < 0.) && (e0 > 0. || e1 > 0. || e2 > 0.)) return false;
    double det = e0 + e1 + e2;
    if (std::fabs(det) < SMALL / 10000.) return false;
    p0t.z *= s_z; p1t.z *= s_z; p2t.z *= s_z;
    double t_scaled = e0 * p0t.z + e1 * p1t.z + e2 * p2t.z;
    if (det < 0. && (t_scaled >= 0. || t_scaled < tmax * det)) return false;
    else if (det > 0. && (t_scaled <= 0. || t_scaled > tmax * det)) return false;
    double inv_det = 1. / det;
    double b0 = e0 * inv_det, b1 = e1 * inv_det, b2 = e2 * inv_det;
    double t = t_scaled * inv_det;
    if (t < SMALL / 10.) return false;
    V3 dpdu, dpdv;
    V2 uvs[3]; mesh_get_uv(mesh, ind, uvs);
    V2 duv02(uvs[0].x - uvs[2].x, uvs[0].y - uvs[2].y), duv12(uvs[1].x - uvs[2].x, uvs[1].y - uvs[2].y);
    V3 dp02 = p0 - p2, dp12 = p1 - p2;
    double determinant = duv02.x * duv12.y - duv02.y * duv12.x;
    if (std::fabs(determinant) < SMALL / 10000.) {
        V3 n = cross(p2 - p0, p1 - p0);
        if (norm2(n) == 0.) return false;  // degenerate triangle
        make_coordinate_system(n, dpdu, dpdv);
    } else {
        double inv = 1. / determinant;
        dpdu = (duv12.y * dp02 - duv02.y * dp12) * inv;
        dpdv = ((-duv12.x) * dp02 + duv02.x * dp12) * inv;
    }
    V3 p_hit = b0 * p0 + b1 * p1 + b2 * p2;
    V2 uv_hit(b0 * uvs[0].x + b1 * uvs[1].x + b2 * uvs[2].x, b0 * uvs[0].y + b1 * uvs[1].y + b2 * uvs[2].y);
    V3 normal = mesh.n.empty() ? cross(dp02, dp12) : (b0 * mesh.n[ind1] + b1 * mesh.n[ind2] + b2 * mesh.n[ind3]);
    HitRecord record = record_new(p_hit, uv_hit, -ray.d, dpdu, dpdv, t, 0, mat_index);
    record.n = normalize(cross(dp02, dp12));
    record.shading.n = normalize(normal);
    V3 ss = normalize(dpdu);
    V3 ts = normalize(cross(record.shading.n, ss));
    if (norm2(ts) > 0.) {
        ss = normalize(cross(ts, record.shading.n));
    } else {
        V3 a, b; make_coordinate_system(record.shading.n, a, b);
        ss = normalize(a); ts = normalize(b);
    }
    set_shading_geometry(record, ss, ts, true);
    set_front(record, ray);
    record.uv = uv_hit;
    out = record;
    return true;
}

// make_sphere_record (intersects.rs:216-258); `ray` is already translated to the sphere centre
inline HitRecord make_sphere_record(double t, double r, uint32_t mat_index, const Ray& ray) {
    V3 p = ray_at(ray, t);
    p = (p * r) / norm(p);
    if (p.x == 0. && p.y == 0.) p.x = 1e-5 * r;
    double phi = std::atan2(p.y, p.x);
    if (phi < 0.) phi = phi + 2. * PI;
    double phi_max = 2. * PI, theta_min = 0., theta_max = PI;
    double u = phi / phi_max;
    double theta = std::acos(clampd(p.z / r, -1., 1.));
    double v = (theta - theta_min) / (theta_max - theta_min);
    double z_r = std::sqrt(p.x * p.x + p.y * p.y);
    double inv_z_r = 1. / z_r;
    double cos_phi = p.x * inv_z_r, sin_phi = p.y * inv_z_r;
    V3 dpdu(-phi_max * p.y, phi_max * p.x, 0.);
    V3 dpdv = (theta_max - theta_min) * V3(p.z * cos_phi, p.z * sin_phi, -r * std::sin(theta));
    HitRecord record = record_new(p, V2(u, v), -ray.d, dpdu, dpdv, t, 0, mat_index);
    set_front(record, ray);
    return record;
}
// sphere_intersect (intersects.rs:177-213)
inline bool sphere_intersect(const V3& center, double r, uint32_t mat_index, const Ray& ray, double tmin, double tmax,
                             HitRecord& out) {
    V3 diff = ray.o - center;
    double a = dot(ray.d, ray.d), b = dot(diff, ray.d), c = dot(diff, diff) - r * r;
    double disc = b * b - a * c;
    if (disc < 0.0) return false;
    double inv_a = 1.0 / a, root = std::sqrt(disc);
    double ans = (-b - root) * inv_a, t;
    if (ans < tmax && ans > tmin) t = ans;
    else {
        ans = (-b + root) * inv_a;
        if (ans < tmax && ans > tmin) t = ans; else return false;
    }
    Ray tr{ray.o - center, ray.d, ray.time};
    out = make_sphere_record(t, r, mat_index, tr);
    out.p = out.p + center;
    return true;
}

struct Xform { M4 fwd, inv; };

// Ray::transform (geometry.rs:231-235)
inline Ray ray_transform(const Ray& r, const Xform& x) {
    return Ray{xf_point(x.inv, r.o), xf_vector(x.inv, r.d), r.time};
}
// xy/xz/yz_rect_intersect (intersects.rs:10-175); axis = 2 (xy), 1 (xz), 0 (yz): the constant axis
inline bool rect_intersect(int kaxis, double a0, double b0, double a1, double b1, double k, uint32_t mat_index,
                           const Xform* xf, const Ray& ray, double t0, double t1, HitRecord& out) {
    Ray tr = xf ? ray_transform(ray, *xf) : ray;
    int aa = (kaxis == 0) ? 1 : 0;                  // first in-plane axis
    int ba = (kaxis == 2) ? 1 : 2;                  // second in-plane axis
    double t = (k - tr.o[kaxis]) / tr.d[kaxis];
    if (t < t0 || t > t1) return false;
    double a = tr.o[aa] + t * tr.d[aa];
    double b = tr.o[ba] + t * tr.d[ba];
    if (a < a0 || b < b0 || a > a1 || b > b1) return false;
    V2 uv((a - a0) / (a1 - a0), (b - b0) / (b1 - b0));
    V3 p = ray_at(tr, t);
    V3 dpdu(0, 0, 0), dpdv(0, 0, 0);
    dpdu.at(aa) = 1.; dpdv.at(ba) = 1.;
    if (xf) { p = xf_point(xf->fwd, p); dpdu = xf_vector(xf->fwd, dpdu); dpdv = xf_vector(xf->fwd, dpdv); }
    out = record_new(p, uv, -ray.d, dpdu, dpdv, t, 0, mat_index);
    set_front(out, ray);
    return true;
}

struct Scene;  // below

struct Texture { rt_texture d; };
struct Material { rt_material d; };
// distribution.rs:4-8, 93-96
struct Distribution1D { std::vector<double> func, cdf; double func_int = 0.; size_t count() const { return func.size(); } };
struct Distribution2D { std::vector<Distribution1D> p_cond_v; Distribution1D p_marginal; };
// pos / cos_width / cos_falloff: what make_point_light / make_spot_light derive at construction (light.rs:522-560)
// eq_class: the smallest index of a light that is field-equal to this one — `#[derive(PartialEq)] enum Light` (light.rs:28) as integrator.rs:621 uses it
struct Light { rt_light d; double area; Distribution2D dist; V3 pos; double cos_width = 0., cos_falloff = 0.; uint32_t eq_class = 0; };
// Texture::Hdr { meta, data: Vec<Rgb<f32>> } / Texture::Image { img: RgbImage } (material.rs:521-539)
struct Image { uint32_t w = 0, h = 0; bool hdr = false; std::vector<float> f; std::vector<uint8_t> b; };

struct BvhNode { Box box; int32_t left, right; uint32_t prim; };  // left<0 ⇒ leaf

struct Scene {
    std::vector<Mesh> meshes;
    std::vector<rt_primitive> prims;
    std::vector<Box> boxes;
    std::vector<Xform> xforms;
    std::vector<rt_texture> textures;
    std::vector<rt_material> materials;
    std::vector<rt_disney> disneys;  // Material::Disney records (material.rs:55-73), indexed by the material's tex[0]
    std::vector<Light> lights;
    std::vector<Image> images;
    std::vector<double> perlin;      // 1 + 768 + 768 doubles per Texture::Perlin: scale, perm_x | perm_y | perm_z, 256 gradient vectors
    std::vector<BvhNode> nodes;   // faithful tree (hittable.rs:637-723), root = 0
    int32_t root = -1;
    // ordered tree statistics: counted by closest_ordered()
};

// util::get_new_box (util.rs:493-517)
inline Box get_new_box(const Box& b, const M4& t) {
    V3 mn(INF, INF, INF), mx(-INF, -INF, -INF);
    for (int i = 0; i < 2; i++) for (int j = 0; j < 2; j++) for (int k = 0; k < 2; k++) {
        V3 p(i == 0 ? b.mn.x : b.mx.x, j == 0 ? b.mn.y : b.mx.y, k == 0 ? b.mn.z : b.mx.z);
        V3 np = xf_point(t, p);
        mn.x = rmin(mn.x, np.x); mn.y = rmin(mn.y, np.y); mn.z = rmin(mn.z, np.z);
        mx.x = rmax(mx.x, np.x); mx.y = rmax(mx.y, np.y); mx.z = rmax(mx.z, np.z);
    }
    return Box{mn, mx};
}
// bounding boxes of the constructors (primitive.rs:64-233) and generate_triangles (hittable.rs:257-288)
inline Box prim_box(const Scene& s, const rt_primitive& pr) {
    Box b;
    switch (pr.type) {
        case RT_PRIM_SPHERE: {
            V3 c(pr.v[0], pr.v[1], pr.v[2]); V3 rv(pr.v[3], pr.v[3], pr.v[3]);
            b.mn = c - rv; b.mx = c + rv; return b;
        }
        case RT_PRIM_TRIANGLE: {
            const Mesh& m = s.meshes[pr.mesh_index];
            V3 v1 = m.p[m.ind[pr.ind]], v2 = m.p[m.ind[pr.ind + 1]], v3 = m.p[m.ind[pr.ind + 2]];
            b.mn = V3(rmin(v1.x, rmin(v2.x, v3.x)), rmin(v1.y, rmin(v2.y, v3.y)), rmin(v1.z, rmin(v2.z, v3.z)));
            b.mx = V3(rmax(v1.x, rmax(v2.x, v3.x)), rmax(v1.y, rmax(v2.y, v3.y)), rmax(v1.z, rmax(v2.z, v3.z)));
            return b;
        }
        case RT_PRIM_XY_RECT: b.mn = V3(pr.v[0], pr.v[1], pr.v[4] - SMALL); b.mx = V3(pr.v[2], pr.v[3], pr.v[4] + SMALL); break;
        case RT_PRIM_XZ_RECT: b.mn = V3(pr.v[0], pr.v[4] - SMALL, pr.v[1]); b.mx = V3(pr.v[2], pr.v[4] + SMALL, pr.v[3]); break;
        default:              b.mn = V3(pr.v[4] - SMALL, pr.v[0], pr.v[1]); b.mx = V3(pr.v[4] + SMALL, pr.v[2], pr.v[3]); break;
    }
    if (pr.transform != RT_NONE) b = get_new_box(b, s.xforms[pr.transform].fwd);
    return b;
}

// Primitive::intersects_obj (primitive.rs:372-425): FlipFace delegates, no `front` toggle, prim_index left 0
inline bool intersects_obj(const Scene& s, const rt_primitive& pr, const Ray& ray, double tmin, double tmax, HitRecord& out) {
    switch (pr.type) {
        case RT_PRIM_SPHERE: return sphere_intersect(V3(pr.v[0], pr.v[1], pr.v[2]), pr.v[3], pr.mat_index, ray, tmin, tmax, out);
        case RT_PRIM_TRIANGLE: return intersects_triangle(s.meshes[pr.mesh_index], ray, pr.ind, pr.mat_index, tmin, tmax, out);
        default: {
            int kaxis = pr.type == RT_PRIM_XY_RECT ? 2 : (pr.type == RT_PRIM_XZ_RECT ? 1 : 0);
            const Xform* xf = pr.transform != RT_NONE ? &s.xforms[pr.transform] : nullptr;
            return rect_intersect(kaxis, pr.v[0], pr.v[1], pr.v[2], pr.v[3], pr.v[4], pr.mat_index, xf, ray, tmin, tmax, out);
        }
    }
}
// Primitive::intersects (primitive.rs:247-316)
inline bool prim_intersects(const Scene& s, uint32_t index, const Ray& ray, double tmin, double tmax, HitRecord& out) {
    const rt_primitive& pr = s.prims[index];
    if (!intersects_obj(s, pr, ray, tmin, tmax, out)) return false;
    if (pr.flip_face) out.front = !out.front;
    out.prim_index = index;
    return true;
}
// Primitive::area (primitive.rs:339-359) — sphere area is 2*PI*r in the reference
inline double prim_area(const Scene& s, const rt_primitive& pr) {
    switch (pr.type) {
        case RT_PRIM_SPHERE: return 2. * PI * pr.v[3];
        case RT_PRIM_TRIANGLE: {
            const Mesh& m = s.meshes[pr.mesh_index];
            V3 p0 = m.p[m.ind[pr.ind]], p1 = m.p[m.ind[pr.ind + 1]], p2 = m.p[m.ind[pr.ind + 2]];
            return 0.5 * norm(cross(p1 - p0, p2 - p0));
        }
        default: return (pr.v[2] - pr.v[0]) * (pr.v[3] - pr.v[1]);
    }
}

// ---- hit-set oracle: brute force over all primitives; hit(prim) = leafbox ∧ primtest
// (SURVEY.md §8a row 4).  Returns the min-t record (ties: last prim wins here; `ties` holds all
// prims with t == t_min so that tree-dependent outcomes can be accepted).
inline bool closest_brute(const Scene& s, const Ray& ray, double tmin, double tmax, HitRecord& best,
                          std::vector<uint32_t>* ties) {
    bool any = false; HitRecord rec;
    if (ties) ties->clear();
    for (uint32_t i = 0; i < s.prims.size(); i++) {
        if (!box_intersects(s.boxes[i], ray, tmin, tmax)) continue;
        if (!prim_intersects(s, i, ray, tmin, tmax, rec)) continue;
        if (!any || rec.t < best.t) { best = rec; any = true; if (ties) { ties->clear(); ties->push_back(i); } }
        else if (rec.t == best.t) { best = rec; if (ties) ties->push_back(i); }
    }
    return any;
}

// ---- faithful tree: BvhNode::new (hittable.rs:637-723) with a seeded axis choice
struct BuildRng { uint64_t s; double next() { s = s * 6364136223846793005ULL + 1442695040888963407ULL; return (double)(s >> 11) / 9007199254740992.0; } };
inline int32_t bvh_build(Scene& s, std::vector<uint32_t>& idx, size_t start, size_t end, BuildRng& rng) {
    double r = rng.next();
    int axis = r < 1. / 3. ? 0 : (r < 2. / 3. ? 1 : 2);
    size_t n = end - start;
    auto leaf = [&](uint32_t prim) { BvhNode nd; nd.box = s.boxes[prim]; nd.left = -1; nd.right = -1; nd.prim = prim; s.nodes.push_back(nd); return (int32_t)s.nodes.size() - 1; };
    auto less = [&](uint32_t a, uint32_t b) { return s.boxes[a].mn[axis] < s.boxes[b].mn[axis]; };
    if (n == 1) return leaf(idx[start]);
    int32_t l, rr;
    if (n == 2) {
        // handle_two (hittable.rs:725-752): `curr` becomes the RIGHT leaf, `next` the LEFT leaf
        // box_compare never returns Equal: a.min < b.min ? Less : Greater (util.rs:531-554)
        size_t curr, next;
        if (less(idx[start], idx[start + 1])) { curr = start; next = start + 1; }    // comp != Greater
        else { curr = start + 1; next = start; }
        l = leaf(idx[next]); rr = leaf(idx[curr]);
    } else {
        std::stable_sort(idx.begin() + start, idx.begin() + end, less);
        size_t mid = start + n / 2;
        l = bvh_build(s, idx, start, mid, rng);
        rr = bvh_build(s, idx, mid, end, rng);
    }
    BvhNode nd; nd.box = box_union(s.nodes[l].box, s.nodes[rr].box); nd.left = l; nd.right = rr; nd.prim = RT_NONE;
    s.nodes.push_back(nd);
    return (int32_t)s.nodes.size() - 1;
}

// BvhNode::intersects (hittable.rs:591-634): both children, unchanged [tmin,tmax], left iff l.t < r.t
struct TravStats { uint64_t box_tests = 0, prim_tests = 0; };
inline bool bvh_intersects(const Scene& s, int32_t node, const Ray& ray, double tmin, double tmax, HitRecord& out, TravStats* st) {
    const BvhNode& nd = s.nodes[node];
    if (st) st->box_tests++;
    if (!box_intersects(nd.box, ray, tmin, tmax)) return false;
    if (nd.left < 0) { if (st) st->prim_tests++; return prim_intersects(s, nd.prim, ray, tmin, tmax, out); }
    HitRecord l, r;
    bool hl = bvh_intersects(s, nd.left, ray, tmin, tmax, l, st);
    bool hr = bvh_intersects(s, nd.right, ray, tmin, tmax, r, st);
    if (hl && hr) { out = (l.t < r.t) ? l : r; return true; }
    if (hl) { out = l; return true; }
    if (hr) { out = r; return true; }
    return false;
}
// geometry::get_intersection (geometry.rs:238-240) over the faithful tree
inline bool get_intersection(const Scene& s, const Ray& ray, HitRecord& out, TravStats* st = nullptr) {
    if (s.root < 0) return false;
    return bvh_intersects(s, s.root, ray, SMALL, INF, out, st);
}
inline bool intersect_range(const Scene& s, const Ray& ray, double tmin, double tmax, HitRecord& out, TravStats* st = nullptr) {
    if (s.root < 0) return false;
    return bvh_intersects(s, s.root, ray, tmin, tmax, out, st);
}

// ---- hit-set oracle over a tree (SURVEY.md §8c "a BVH-accelerated variant of the same predicate"): every node whose f64 box passes
// BoundingBox::intersects with the UNCHANGED [tmin, tmax] is visited (a leaf box that passes implies every ancestor passes — monotonic IEEE
// rounding, §8a row 4 — so no primitive of the hit set is missed), leaves run the reference's exact predicate, and the result is what
// closest_brute returns: the min-t record and the tie set {prim : t == t_min}.  The tree only decides which boxes are looked at, not the
// answer; bvh_build_fast builds one quickly for the 1 M - 50 M triangle soups (median split with nth_element on the widest axis of the
// box minima, one primitive per leaf, subtrees in parallel), where the faithful builder's stable_sort per level takes minutes.
inline bool closest_hitset(const Scene& s, const Ray& ray, double tmin, double tmax, HitRecord& best, std::vector<uint32_t>* ties) {
    bool any = false; HitRecord rec;
    if (ties) ties->clear();
    if (s.root < 0) return false;
    std::vector<int32_t> stack; stack.reserve(256); stack.push_back(s.root);
    while (!stack.empty()) {
        const BvhNode& nd = s.nodes[stack.back()]; stack.pop_back();
        if (!box_intersects(nd.box, ray, tmin, tmax)) continue;
        if (nd.left >= 0) { stack.push_back(nd.left); stack.push_back(nd.right); continue; }
        if (!prim_intersects(s, nd.prim, ray, tmin, tmax, rec)) continue;
        if (!any || rec.t < best.t) { best = rec; any = true; if (ties) { ties->clear(); ties->push_back(nd.prim); } }
        else if (rec.t == best.t) { if (ties) ties->push_back(nd.prim); if (nd.prim > best.prim_index) best = rec; }
    }
    if (ties) std::sort(ties->begin(), ties->end());
    return any;
}
inline int32_t bvh_build_fast_rec(Scene& s, uint32_t* idx, size_t n, int32_t base, int par_depth);
inline void bvh_build_fast(Scene& s, int threads) {
    s.nodes.clear(); s.root = -1;
    const size_t n = s.prims.size();
    if (n == 0) return;
    s.nodes.resize(2 * n - 1);                       // subtree over k primitives occupies 2k-1 consecutive nodes, root first
    std::vector<uint32_t> idx(n);
    for (size_t i = 0; i < n; i++) idx[i] = (uint32_t)i;
    int depth = 0; while ((1 << depth) < threads) depth++;
    s.root = bvh_build_fast_rec(s, idx.data(), n, 0, depth + 1);
}
inline int32_t bvh_build_fast_rec(Scene& s, uint32_t* idx, size_t n, int32_t base, int par_depth) {
    BvhNode& nd = s.nodes[base];
    if (n == 1) { nd.box = s.boxes[idx[0]]; nd.left = -1; nd.right = -1; nd.prim = idx[0]; return base; }
    double lo[3] = {1e308, 1e308, 1e308}, hi[3] = {-1e308, -1e308, -1e308};
    for (size_t i = 0; i < n; i++) for (int a = 0; a < 3; a++) { lo[a] = rmin(lo[a], s.boxes[idx[i]].mn[a]); hi[a] = rmax(hi[a], s.boxes[idx[i]].mn[a]); }
    int axis = 0; for (int a = 1; a < 3; a++) if (hi[a] - lo[a] > hi[axis] - lo[axis]) axis = a;
    const size_t mid = n / 2;
    std::nth_element(idx, idx + mid, idx + n, [&](uint32_t a, uint32_t b) { return s.boxes[a].mn[axis] < s.boxes[b].mn[axis]; });
    const int32_t lbase = base + 1, rbase = base + 1 + (int32_t)(2 * mid - 1);
    int32_t l, r;
    if (par_depth > 0 && n > 100000) {
        std::thread th([&]() { l = bvh_build_fast_rec(s, idx, mid, lbase, par_depth - 1); });
        r = bvh_build_fast_rec(s, idx + mid, n - mid, rbase, par_depth - 1);
        th.join();
    } else {
        l = bvh_build_fast_rec(s, idx, mid, lbase, 0);
        r = bvh_build_fast_rec(s, idx + mid, n - mid, rbase, 0);
    }
    nd.box = box_union(s.nodes[l].box, s.nodes[r].box); nd.left = l; nd.right = r; nd.prim = RT_NONE;
    return base;
}

// The "reference traversal" of SURVEY.md §8(d) used ONLY to count algorithmic work per ray:
// same binary tree and primitive order, nearest-child-first, tmax shrinking, one prim per leaf.
inline double box_entry(const Box& b, const Ray& ray, double tmin, double tmax) {
    for (int a = 0; a < 3; a++) {
        double inv_d = 1.0 / ray.d[a];
        double v1 = (b.mn[a] - ray.o[a]) * inv_d, v2 = (b.mx[a] - ray.o[a]) * inv_d;
        tmin = rmax(tmin, rmin(v1, v2)); tmax = rmin(tmax, rmax(v1, v2));
        if (tmax <= tmin) return INF * 10;  // +inf ⇒ miss
    }
    return tmin;
}
inline void closest_ordered(const Scene& s, const Ray& ray, double tmin, double tmax, bool& hit, HitRecord& best, TravStats& st) {
    hit = false;
    if (s.root < 0) return;
    int32_t stack[128]; int sp = 0; stack[sp++] = s.root;
    st.box_tests++;
    if (!box_intersects(s.nodes[s.root].box, ray, tmin, tmax)) return;
    while (sp > 0) {
        const BvhNode& nd = s.nodes[stack[--sp]];
        if (nd.left < 0) {
            st.prim_tests++;
            HitRecord rec;
            if (prim_intersects(s, nd.prim, ray, tmin, tmax, rec) && (!hit || rec.t < best.t)) { best = rec; hit = true; tmax = rec.t; }
            continue;
        }
        st.box_tests += 2;
        double tl = box_entry(s.nodes[nd.left].box, ray, tmin, tmax), tr = box_entry(s.nodes[nd.right].box, ray, tmin, tmax);
        bool hl = tl < INF, hr = tr < INF;
        if (hl && hr) { if (tl <= tr) { stack[sp++] = nd.right; stack[sp++] = nd.left; } else { stack[sp++] = nd.left; stack[sp++] = nd.right; } }
        else if (hl) stack[sp++] = nd.left;
        else if (hr) stack[sp++] = nd.right;
    }
}

// Visibility::unoccluded (hittable.rs:25-39)
inline bool unoccluded(const Scene& s, const V3& p0, const V3& p1, uint32_t p1_prim, double p1_t, bool infinite_light) {
    V3 dir = p1 - p0;
    Ray ray{p0 + dir * SMALL, dir, p1_t};
    HitRecord rec;
    bool hit = s.root >= 0 && bvh_intersects(s, s.root, ray, 0., INF, rec, nullptr);
    if (!hit) return infinite_light;
    return rec.prim_index == p1_prim && !infinite_light;
}

}  // namespace orc
