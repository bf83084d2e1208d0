// Host half of the C ABI: scene tables, primitive boxes, commit (BVH + SoA flatten).
// No CUDA in this file.  Reference items are cited per function.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <cstring>
#include <mutex>
#include <thread>
#include "scene.h"
#include "camera_host.h"

namespace rt {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }

static inline double rmin(double a, double b) { return std::fmin(a, b); }
static inline double rmax(double a, double b) { return std::fmax(a, b); }

// closed-form 4x4 inverse by cofactors (what nalgebra's 4x4 `try_inverse` does); row-major in/out
bool invert4(const double m[16], double out[16]) {
    double inv[16];
    inv[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
    inv[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
    inv[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
    inv[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
    inv[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
    inv[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
    inv[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
    inv[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
    inv[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
    inv[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
    inv[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
    inv[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
    inv[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
    inv[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
    inv[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
    inv[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
    double det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
    if (det == 0.0) return false;
    double inv_det = 1.0 / det;
    for (int i = 0; i < 16; i++) out[i] = inv[i] * inv_det;
    return true;
}

static void xf_point(const double* m, const double p[3], double o[3]) {   // Projective3::transform_point
    double x = m[0] * p[0] + m[1] * p[1] + m[2] * p[2] + m[3];
    double y = m[4] * p[0] + m[5] * p[1] + m[6] * p[2] + m[7];
    double z = m[8] * p[0] + m[9] * p[1] + m[10] * p[2] + m[11];
    double w = m[12] * p[0] + m[13] * p[1] + m[14] * p[2] + m[15];
    o[0] = x / w; o[1] = y / w; o[2] = z / w;
}

// Bounding boxes exactly as the reference constructors make them: sphere c±r (primitive.rs:64-77),
// rects padded ±SMALL on the constant axis (primitive.rs:93-94,111-112,...), transformed rects through
// util::get_new_box (util.rs:493-517), triangles min/max of the vertices (hittable.rs:267-276).
Box64 prim_box(const rt_scene& s, const HostPrim& p) {
    Box64 b;
    if (p.type == RT_PRIM_SPHERE) {
        const double* v = &s.shapes[p.shape];
        for (int k = 0; k < 3; k++) { b.mn[k] = v[k] - v[3]; b.mx[k] = v[k] + v[3]; }
        return b;
    }
    if (p.type == RT_PRIM_TRIANGLE) {
        const HostMesh& m = s.meshes[p.mesh];
        const double* v1 = &m.p[3 * (size_t)m.ind[p.ind]]; const double* v2 = &m.p[3 * (size_t)m.ind[p.ind + 1]]; const double* v3 = &m.p[3 * (size_t)m.ind[p.ind + 2]];
        for (int k = 0; k < 3; k++) { b.mn[k] = rmin(v1[k], rmin(v2[k], v3[k])); b.mx[k] = rmax(v1[k], rmax(v2[k], v3[k])); }
        return b;
    }
    const double* v = &s.shapes[p.shape];
    double a0 = v[0], b0 = v[1], a1 = v[2], b1 = v[3], k = v[4];
    if (p.type == RT_PRIM_XY_RECT) { b.mn[0] = a0; b.mn[1] = b0; b.mn[2] = k - SMALL; b.mx[0] = a1; b.mx[1] = b1; b.mx[2] = k + SMALL; }
    else if (p.type == RT_PRIM_XZ_RECT) { b.mn[0] = a0; b.mn[1] = k - SMALL; b.mn[2] = b0; b.mx[0] = a1; b.mx[1] = k + SMALL; b.mx[2] = b1; }
    else { b.mn[0] = k - SMALL; b.mn[1] = a0; b.mn[2] = b0; b.mx[0] = k + SMALL; b.mx[1] = a1; b.mx[2] = b1; }
    if (p.xform != RT_NONE) {
        const double* t = s.xforms[p.xform].fwd;
        Box64 nb; for (int a = 0; a < 3; a++) { nb.mn[a] = INF; nb.mx[a] = -INF; }
        for (int i = 0; i < 2; i++) for (int j = 0; j < 2; j++) for (int l = 0; l < 2; l++) {
            double pt[3] = {i == 0 ? b.mn[0] : b.mx[0], j == 0 ? b.mn[1] : b.mx[1], l == 0 ? b.mn[2] : b.mx[2]}, np[3];
            xf_point(t, pt, np);
            for (int a = 0; a < 3; a++) { nb.mn[a] = rmin(nb.mn[a], np[a]); nb.mx[a] = rmax(nb.mx[a], np[a]); }
        }
        b = nb;
    }
    return b;
}

// Primitive::area (primitive.rs:339-359); the sphere's 2*PI*r is the reference's value
double prim_area(const rt_scene& s, const HostPrim& p) {
    if (p.type == RT_PRIM_SPHERE) return 2. * PI * s.shapes[p.shape + 3];
    if (p.type == RT_PRIM_TRIANGLE) {
        const HostMesh& m = s.meshes[p.mesh];
        const double* p0 = &m.p[3 * (size_t)m.ind[p.ind]]; const double* p1 = &m.p[3 * (size_t)m.ind[p.ind + 1]]; const double* p2 = &m.p[3 * (size_t)m.ind[p.ind + 2]];
        double v1[3] = {p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2]}, v2[3] = {p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2]};
        double c[3] = {v1[1] * v2[2] - v1[2] * v2[1], v1[2] * v2[0] - v1[0] * v2[2], v1[0] * v2[1] - v1[1] * v2[0]};
        return 0.5 * std::sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
    }
    const double* v = &s.shapes[p.shape];
    return (v[2] - v[0]) * (v[3] - v[1]);
}

// ---- textures on the host: the same lookups the device performs (shade.cuh tex_value), needed at commit by
// Light::make_infinite_light (light.rs:608-638), which evaluates the environment texture into its sampling table.
static inline uint32_t sat_u32(double x) {                    // Rust `f64 as u32`: saturating, NaN -> 0
    if (!(x > 0.)) return 0u;
    if (x >= 4294967295.) return 4294967295u;
    return (uint32_t)x;
}
static void texel_rgb(const rt_scene& s, const rt_texture& t, double u, double v, double out[3]) {
    const DImage& im = s.images[t.image];
    if (t.type == RT_TEX_HDR) {                               // material.rs:570-587
        uint32_t x = sat_u32(std::round((1. - u) * (double)im.w)) % im.w, y = sat_u32(std::round(v * (double)im.h)) % im.h;
        uint32_t q = s.texels[im.offset + (size_t)y * im.w + x];
        double sc = std::ldexp(1.0, (int)(q >> 24) - 128);    // 2f64.powi(e - 128)
        for (int k = 0; k < 3; k++) out[k] = ((double)((q >> (8 * k)) & 255u) + 0.5) * sc / 256.;
    } else {                                                  // material.rs:601-611
        uint32_t x = sat_u32(std::round(u * (double)im.w)) % im.w, y = sat_u32(std::round((1. - v) * (double)im.h)) % im.h;
        uint32_t q = s.texels[im.offset + (size_t)y * im.w + x];
        for (int k = 0; k < 3; k++) out[k] = (double)((q >> (8 * k)) & 255u) / 255.;
    }
}
static void host_tex_value(const rt_scene& s, uint32_t id, double u, double v, double out[3]) {   // material.rs:542-589
    for (int depth = 0; depth < 8; depth++) {
        const rt_texture& t = s.textures[id];
        if (t.type == RT_TEX_SOLID) { out[0] = t.color[0]; out[1] = t.color[1]; out[2] = t.color[2]; return; }
        if (t.type == RT_TEX_HDR || t.type == RT_TEX_IMAGE) { texel_rgb(s, t, u, v, out); return; }
        double mult = std::sin(t.frequency * u * 2. * PI) * std::sin(t.frequency * v * 2. * PI);
        id = mult < 0. ? t.even : t.odd;
    }
    out[0] = out[1] = out[2] = 0.;
}
// Distribution1D::make_distribution (distribution.rs:28-55) of f[0..n): writes the normalised cdf[0..n], returns func_int
static double make_distribution_1d(const double* f, size_t n, double* cdf) {
    cdf[0] = 0.;
    for (size_t i = 1; i < n + 1; i++) cdf[i] = cdf[i - 1] + f[i - 1] / (double)n;
    double func_int = cdf[n];
    if (func_int == 0.) { for (size_t i = 1; i < n + 1; i++) cdf[i] = (double)i / (double)n; }
    else { for (size_t i = 1; i < n + 1; i++) cdf[i] = cdf[i] / func_int; }
    return func_int;
}
// Light::make_infinite_light (light.rs:608-638) + Distribution2D::make_distribution_2d (distribution.rs:117-130).
// The pointers of `d` receive OFFSETS into `dist` (in doubles, +1 so that 0 stays "null"); upload_scene rebases them.
static rt_status build_infinite_light(const rt_scene& s, const rt_light& l, DLight& d, std::vector<double>& dist) {
    if (l.text_id >= s.textures.size()) { set_error("commit: infinite light texture out of range"); return RT_ERR_INVALID; }
    if (l.transform != RT_NONE && l.transform >= s.xforms.size()) { set_error("commit: infinite light transform out of range"); return RT_ERR_INVALID; }
    const rt_texture& t = s.textures[l.text_id];
    // Texture::get_size (material.rs:591-599): Image/Hdr -> (w,h); SolidColor -> (1,1); Checkered -> (0,0)
    uint32_t tw = 0, th = 0;
    if (t.type == RT_TEX_HDR || t.type == RT_TEX_IMAGE) { tw = s.images[t.image].w; th = s.images[t.image].h; }
    else if (t.type == RT_TEX_SOLID) { tw = th = 1; }
    if (tw == 0 || th == 0) { set_error("commit: infinite light needs a sized texture (the reference's table would be empty)"); return RT_ERR_INVALID; }
    size_t nu = (size_t)tw * 2, nv = (size_t)th * 2;
    d.text_id = l.text_id; d.xform = l.transform; d.nu = (uint32_t)nu; d.nv = (uint32_t)nv;
    size_t base = dist.size();
    dist.resize(base + nu * nv + nv * (nu + 1) + nv + (nv + 1));
    double* f = &dist[base]; double* cdf = f + nu * nv; double* row_int = cdf + nv * (nu + 1); double* mcdf = row_int + nv;
    for (size_t v = 0; v < nv; v++) {
        double vp = ((double)v + 0.5) / (double)nv;
        double sin_theta = std::sin(PI * ((double)v + 0.5) / (double)nv);
        for (size_t u = 0; u < nu; u++) {
            double up = (double)u / (double)nu, c[3];
            host_tex_value(s, l.text_id, up, vp, c);
            double lum = 0.2126 * c[0] + 0.7152 * c[1] + 0.0722 * c[2];       // util::color_to_luminance (util.rs:169-171)
            f[u + v * nu] = lum * sin_theta;
        }
    }
    for (size_t v = 0; v < nv; v++) row_int[v] = make_distribution_1d(f + v, nu, cdf + v * (nu + 1));   // `&f[v..(v + nu)]` (distribution.rs:120)
    d.marg_int = make_distribution_1d(row_int, nv, mcdf);
    d.dist_f = (const double*)(uintptr_t)(base + 1); d.dist_cdf = (const double*)(uintptr_t)(base + nu * nv + 1);
    d.row_int = (const double*)(uintptr_t)(base + nu * nv + nv * (nu + 1) + 1); d.marg_cdf = (const double*)(uintptr_t)(base + nu * nv + nv * (nu + 1) + nv + 1);
    return RT_OK;
}

rt_status build_host_image(const rt_scene& s, HostImage& img, bool device_build) {
    size_t n = s.prims.size();
    // validate cross indices
    for (size_t i = 0; i < n; i++) {
        const HostPrim& p = s.prims[i];
        if (p.mat >= s.materials.size()) { set_error("primitive " + std::to_string(i) + ": mat_index out of range"); return RT_ERR_INVALID; }
        if (p.light != RT_NONE && p.light >= s.lights.size()) { set_error("primitive: light_index out of range"); return RT_ERR_INVALID; }
    }
    for (auto& m : s.materials) if (m.type == RT_MAT_DISNEY) {
        if (m.tex[0] >= s.disneys.size()) { set_error("Disney material: record index out of range (rt_scene_add_disney first)"); return RT_ERR_INVALID; }
        const uint32_t* ids = &s.disneys[m.tex[0]].color;       // the 14 texture ids are the record's first 14 words
        for (int k = 0; k < 14; k++) if (ids[k] >= s.textures.size()) { set_error("Disney material: texture id out of range"); return RT_ERR_INVALID; }
    }
    for (auto& m : s.materials) for (int k = 0; k < 5; k++) {
        if (m.type == RT_MAT_DISNEY) break;
        int used = m.type == RT_MAT_METAL ? 5 : (m.type == RT_MAT_PLASTIC || m.type == RT_MAT_GLASS ? 2 : 1);
        if (k < used && m.tex[k] != RT_NONE && m.tex[k] >= s.textures.size()) { set_error("material: texture id out of range"); return RT_ERR_INVALID; }
        if (k < used && m.tex[k] == RT_NONE && !(m.type == RT_MAT_METAL && k >= 3)) { set_error("material: missing texture id"); return RT_ERR_INVALID; }
    }
    for (auto& t : s.textures) if (t.type == RT_TEX_CHECKERED && (t.odd >= s.textures.size() || t.even >= s.textures.size())) { set_error("checkered texture: id out of range"); return RT_ERR_INVALID; }
    for (auto& l : s.lights) if (l.type == RT_LIGHT_DIFFUSE && l.prim_index >= n) { set_error("light: prim_index out of range"); return RT_ERR_INVALID; }
    for (auto& t : s.textures) if ((t.type == RT_TEX_HDR || t.type == RT_TEX_IMAGE) && t.image >= s.images.size()) { set_error("texture: image index out of range"); return RT_ERR_INVALID; }
    // meshes → element offsets into the concatenated f64 / index arrays
    img.meshes.resize(s.meshes.size());
    std::vector<uint64_t> idx_base(s.meshes.size());
    uint64_t pb = 0, nb = 0, ub = 0, ib = 0;
    for (size_t m = 0; m < s.meshes.size(); m++) {
        img.meshes[m] = DMesh{pb, nb, ub};
        idx_base[m] = ib;
        pb += s.meshes[m].p.size() / 3; nb += s.meshes[m].n.size() / 3; ub += s.meshes[m].uv.size() / 2; ib += s.meshes[m].ind.size();
    }
    // the arrays themselves go to the device mesh by mesh (device.cu:up_meshes)
    // per-primitive records and boxes
    img.shapes = s.shapes;
    img.prims.resize(n);
    img.device_build = device_build && n >= 8;
    img.n_slots = n;
    {   // direct shapes: few parametric shapes next to at least one triangle (RT_DIRECT=0 keeps them in the tree, for A/B runs)
        size_t n_shapes = 0, n_tris = 0;
        for (size_t i = 0; i < n; i++) (s.prims[i].type == RT_PRIM_TRIANGLE ? n_tris : n_shapes)++;
        const char* e = getenv("RT_DIRECT");
        if (n_shapes >= 1 && n_shapes <= MAX_DIRECT && n_tris >= 8 && !(e && e[0] == '0'))
            for (size_t i = 0; i < n; i++) if (s.prims[i].type != RT_PRIM_TRIANGLE) img.direct.push_back((uint32_t)i);
    }
    const bool use_direct = !img.direct.empty();
    std::vector<Box64> boxes(img.device_build ? 0 : n);
    std::mutex other_mu;
    int threads = (int)host_threads();
    {
        std::vector<std::thread> pool; size_t chunk = (n + threads - 1) / std::max(threads, 1);
        for (int t = 0; t < threads; t++) pool.emplace_back([&, t]() {
            size_t b = t * chunk, e = std::min(n, b + chunk);
            for (size_t i = b; i < e; i++) {
                const HostPrim& p = s.prims[i];
                DPrim d; d.info = p.type | (p.flip ? PI_FLIP : 0); d.mat = p.mat; d.light = p.light; d.a = 0;
                if (p.type == RT_PRIM_TRIANGLE) {
                    const HostMesh& m = s.meshes[p.mesh];
                    if (!m.n.empty()) d.info |= PI_HAS_N;
                    if (!m.uv.empty()) d.info |= PI_HAS_UV;
                    d.info |= p.mesh << 6;
                    d.a = (uint32_t)(idx_base[p.mesh] + p.ind);
                } else d.a = p.shape;
                img.prims[i] = d;
                if (img.device_build && p.type == RT_PRIM_TRIANGLE) continue;      // box and slot record: device.cu:k_fill_tri_slots
                const Box64 bx = prim_box(s, p);
                if (p.type != RT_PRIM_TRIANGLE) for (int k = 0; k < 3; k++) { img.shapes[p.shape + 6 + k] = bx.mn[k]; img.shapes[p.shape + 9 + k] = bx.mx[k]; }
                if (!img.device_build) boxes[i] = bx;
                else {
                    // a direct shape keeps its slot (slots are indexed by primitive there) but can never be reached: inverted box, slot kind 2
                    Box64 tb = bx;
                    if (use_direct && p.type != RT_PRIM_TRIANGLE) for (int k = 0; k < 3; k++) { tb.mn[k] = 1e300; tb.mx[k] = -1e300; }
                    std::lock_guard<std::mutex> g(other_mu); img.other_idx.push_back((uint32_t)i); img.other_boxes.push_back(tb);
                }
            }
        });
        for (auto& th : pool) th.join();
    }
    if (ib >= (1ull << 32)) { set_error("index buffer exceeds 2^32 entries"); return RT_ERR_INVALID; }
    BuiltBvh bvh;
    std::vector<uint32_t> tree_prims;              // host build with direct shapes: the primitives handed to the builder (the triangles), in scene order
    if (!img.device_build) {
        if (use_direct) {
            std::vector<Box64> tb; tb.reserve(n - img.direct.size()); tree_prims.reserve(n - img.direct.size());
            for (size_t i = 0; i < n; i++) if (s.prims[i].type == RT_PRIM_TRIANGLE) { tree_prims.push_back((uint32_t)i); tb.push_back(boxes[i]); }
            boxes.swap(tb);
            img.n_slots = tree_prims.size();
        }
        const size_t n_tree = boxes.size();
        auto t0 = std::chrono::steady_clock::now();
        rt_status bst = build_bvh(boxes, bvh, threads);
        if (bst != RT_OK) return bst;
        if (bvh.order.size() != n_tree) { set_error("commit: BVH build returned an incomplete primitive order"); return RT_ERR_INVALID; }
        img.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        img.max_stack = bvh.max_stack;
        if (bvh.max_stack > TRAVERSAL_STACK) { set_error("commit: BVH needs a deeper traversal stack than the kernels have (" + std::to_string(bvh.max_stack) + " > " + std::to_string(TRAVERSAL_STACK) + ")"); return RT_ERR_INVALID; }
        if (getenv("RT_TRACE_BUILD")) fprintf(stderr, "bvh: %zu nodes of %zu bytes, %zu slots, worst-case stack %d\n", bvh.nodes.size(), sizeof(DNodeT), n, bvh.max_stack);
        img.nodes = std::move(bvh.nodes);
    }
    const size_t n_host_slots = img.device_build ? 0 : img.n_slots;
    img.slots.resize(n_host_slots);
    auto fill_slots = [&](size_t sb, size_t se) {
    for (size_t sidx = sb; sidx < se; sidx++) {
        uint32_t pi = tree_prims.empty() ? bvh.order[sidx] : tree_prims[bvh.order[sidx]];
        const HostPrim& p = s.prims[pi];
        DSlot sl; memset(&sl, 0, sizeof(sl));
        sl.prim = pi;
        if (p.type == RT_PRIM_TRIANGLE) {
            const HostMesh& m = s.meshes[p.mesh];
            for (int v = 0; v < 3; v++) for (int k = 0; k < 3; k++) sl.v[3 * v + k] = m.p[3 * (size_t)m.ind[p.ind + v] + k];
            sl.kind = 0;
            if (!m.uv.empty()) {
                // hittable.rs:366-378: a triangle whose uv determinant is below 1e-7 AND whose geometric
                // normal (p2-p0)x(p1-p0) is exactly zero always returns None
                const double* u0 = &m.uv[2 * (size_t)m.ind[p.ind]]; const double* u1 = &m.uv[2 * (size_t)m.ind[p.ind + 1]]; const double* u2 = &m.uv[2 * (size_t)m.ind[p.ind + 2]];
                double d02x = u0[0] - u2[0], d02y = u0[1] - u2[1], d12x = u1[0] - u2[0], d12y = u1[1] - u2[1];
                double determinant = d02x * d12y - d02y * d12x;
                if (std::fabs(determinant) < SMALL / 10000.) {
                    double a[3] = {sl.v[6] - sl.v[0], sl.v[7] - sl.v[1], sl.v[8] - sl.v[2]}, b[3] = {sl.v[3] - sl.v[0], sl.v[4] - sl.v[1], sl.v[5] - sl.v[2]};
                    double c[3] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
                    if (c[0] * c[0] + c[1] * c[1] + c[2] * c[2] == 0.) sl.kind = 2;
                }
            }
        } else sl.kind = 1;
        img.slots[sidx] = sl;
    }
    };
    if (!img.device_build) {
        const size_t ns = n_host_slots;
        std::vector<std::thread> pool; size_t chunk = (ns + threads - 1) / std::max(threads, 1);
        for (int t = 0; t < threads; t++) { size_t b = std::min(ns, t * chunk), e = std::min(ns, b + chunk); if (e > b) pool.emplace_back(fill_slots, b, e); }
        for (auto& th : pool) th.join();
    }
    // per-material microfacet alphas that do not depend on the hit point (material.rs:110-117,170-173,205-214)
    auto remap_alpha = [](double roughness) {          // trowbridge_reitz_roughness_to_alpha (microfacet.rs:442-446)
        roughness = std::fmax(roughness, 1e-5);
        double x = std::log(roughness);
        return 1.62142 + 0.819955 * x + 0.1734 * x * x + 0.0171201 * x * x * x + 0.000640711 * x * x * x * x;
    };
    img.mat_alpha.assign(2 * s.materials.size(), std::nan(""));
    for (size_t i = 0; i < s.materials.size(); i++) {
        const rt_material& m = s.materials[i];
        double a[2] = {std::nan(""), std::nan("")};
        if (m.type == RT_MAT_PLASTIC) { a[0] = a[1] = m.remap_roughness ? remap_alpha(m.f[0]) : m.f[0]; }
        else if (m.type == RT_MAT_GLASS) { a[0] = m.remap_roughness ? remap_alpha(m.f[0]) : m.f[0]; a[1] = m.remap_roughness ? remap_alpha(m.f[1]) : m.f[1]; }
        else if (m.type == RT_MAT_METAL) {
            for (int k = 0; k < 2; k++) {
                uint32_t tid = m.tex[3 + k] == RT_NONE ? m.tex[2] : m.tex[3 + k];
                if (tid < s.textures.size() && s.textures[tid].type == RT_TEX_SOLID) a[k] = m.remap_roughness ? remap_alpha(s.textures[tid].color[0]) : s.textures[tid].color[0];
            }
        }
        img.mat_alpha[2 * i] = a[0]; img.mat_alpha[2 * i + 1] = a[1];
    }
    img.lights.resize(s.lights.size());
    for (size_t i = 0; i < s.lights.size(); i++) {
        DLight d; memset(&d, 0, sizeof(d));
        const rt_light& l = s.lights[i];
        d.type = l.type;
        if (l.type == RT_LIGHT_DIFFUSE) {
            if (l.prim_index >= s.prims.size()) { set_error("commit: light primitive out of range"); return RT_ERR_INVALID; }
            d.prim = l.prim_index; d.two_sided = l.two_sided;
            for (int k = 0; k < 3; k++) d.color[k] = l.color[k];
            d.area = prim_area(s, s.prims[d.prim]);          // light.rs:603
        } else if (l.type == RT_LIGHT_POINT || l.type == RT_LIGHT_SPOT || l.type == RT_LIGHT_DISTANT) {
            // make_point_light / make_spot_light / make_distant_light (light.rs:522-582)
            if (l.transform != RT_NONE && l.transform >= s.xforms.size()) { set_error("commit: light transform out of range"); return RT_ERR_INVALID; }
            d.prim = 0; d.xform = l.transform;               // HitRecord::make_basic: the visibility target is primitive 0 (hittable.rs:138)
            for (int k = 0; k < 3; k++) d.color[k] = l.color[k];
            if (l.type == RT_LIGHT_DISTANT) { for (int k = 0; k < 3; k++) d.pos[k] = l.param[k]; }
            else if (l.transform != RT_NONE) {
                const double* m = s.xforms[l.transform].fwd;  // to_world.transform_point(origin): translation column over w
                for (int k = 0; k < 3; k++) d.pos[k] = m[4 * k + 3] / m[15];
            }
            if (l.type == RT_LIGHT_SPOT) {
                // light.rs:548-549 takes cos(angle * 180 / PI) of the radian arguments; restated as it is
                d.cos_width = std::cos(l.param[0] * 180. * (1. / M_PI)); d.cos_falloff = std::cos(l.param[1] * 180. * (1. / M_PI));
            }
        } else {
            rt_status st = build_infinite_light(s, l, d, img.dist);
            if (st != RT_OK) return st;
        }
        img.lights[i] = d;
    }
    // `lights[index].eq(light)` (integrator.rs:621) is the derived PartialEq of `enum Light` (light.rs:28): field equality, so two entries that
    // describe the same light (same primitive, colour, ...) count as one.  A NaN field makes a light unequal to everything, itself included.
    auto same = [](const rt_light& x, const rt_light& y) {
        if (x.type != y.type || x.prim_index != y.prim_index || x.two_sided != y.two_sided || x.text_id != y.text_id || x.transform != y.transform) return false;
        for (int k = 0; k < 3; k++) if (!(x.color[k] == y.color[k]) || !(x.param[k] == y.param[k])) return false;
        return true;
    };
    for (size_t i = 0; i < s.lights.size(); i++) {
        img.lights[i].eq_class = (uint32_t)i;
        for (size_t j = 0; j < i; j++) if (same(s.lights[j], s.lights[i])) { img.lights[i].eq_class = img.lights[j].eq_class; break; }
    }
    return RT_OK;
}

}  // namespace rt

using namespace rt;

extern "C" {

const char* rt_last_error_string(void) { return g_error.c_str(); }
const char* rt_version_string(void) { return "rtcore 0.1 (sm_100a)"; }
uint32_t rt_round_spp(uint32_t spp) { uint32_t p = 1; while (p < spp) p <<= 1; return p; }   // sampler.rs:637-640

// Camera::new_motion_blur (geometry.rs:133-175)
rt_status rt_camera_look_at(const double from[3], const double to[3], const double up[3], double aspect, double vfov,
                            double aperture, double focus, double t0, double t1, rt_camera* out) {
    if (!from || !to || !up || !out) { set_error("rt_camera_look_at: null argument"); return RT_ERR_INVALID; }
    camera_look_at(from, to, up, aspect, vfov, aperture, focus, t0, t1, out);
    return RT_OK;
}

rt_status rt_scene_create(rt_scene** out) {
    if (!out) { set_error("rt_scene_create: null out"); return RT_ERR_INVALID; }
    *out = new rt_scene(); return RT_OK;
}
rt_status rt_scene_destroy(rt_scene* s) {
    if (!s) return RT_OK;
    free_wavefront(*s); free_scene_device(*s); delete s; return RT_OK;
}
rt_status rt_scene_add_mesh(rt_scene* s, uint64_t n_points, const double* p, uint64_t n_normals, const double* n, uint64_t n_uvs,
                            const double* uv, uint64_t n_ind, const uint32_t* ind, uint32_t* out) {
    if (!s || (n_points && !p) || (n_normals && !n) || (n_uvs && !uv) || (n_ind && !ind)) { set_error("rt_scene_add_mesh: null argument"); return RT_ERR_INVALID; }
    if (n_ind % 3) { set_error("rt_scene_add_mesh: index count not a multiple of 3"); return RT_ERR_INVALID; }
    if (n_normals && n_normals != n_points) { set_error("rt_scene_add_mesh: normals must be per point (single-index mesh)"); return RT_ERR_INVALID; }
    if (n_uvs && n_uvs != n_points) { set_error("rt_scene_add_mesh: uvs must be per point (single-index mesh)"); return RT_ERR_INVALID; }
    {   // validation and the copies run on all host threads (3.6 GB of points at 50 M soup triangles)
        std::atomic<int> bad{0};
        par_for(n_ind, (size_t)1 << 20, [&](size_t b, size_t e) { for (size_t i = b; i < e; i++) if (ind[i] >= n_points) { bad.store(1); break; } });
        if (bad.load()) { set_error("rt_scene_add_mesh: index out of range"); return RT_ERR_INVALID; }
    }
    HostMesh m; par_assign(m.p, p, 3 * n_points); if (n_normals) par_assign(m.n, n, 3 * n_normals); if (n_uvs) par_assign(m.uv, uv, 2 * n_uvs); par_assign(m.ind, ind, n_ind);
    s->meshes.push_back(std::move(m)); s->committed = false;
    if (out) *out = (uint32_t)s->meshes.size() - 1;
    return RT_OK;
}
rt_status rt_scene_add_transform(rt_scene* s, const double m[16], uint32_t* out) {
    if (!s || !m) { set_error("rt_scene_add_transform: null argument"); return RT_ERR_INVALID; }
    DXform x; memcpy(x.fwd, m, sizeof(double) * 16);
    if (!invert4(x.fwd, x.inv)) { set_error("rt_scene_add_transform: singular matrix"); return RT_ERR_INVALID; }
    s->xforms.push_back(x); s->committed = false;
    if (out) *out = (uint32_t)s->xforms.size() - 1;
    return RT_OK;
}
rt_status rt_scene_add_primitives(rt_scene* s, uint64_t n, const rt_primitive* prims, uint32_t* first) {
    if (!s || (n && !prims)) { set_error("rt_scene_add_primitives: null argument"); return RT_ERR_INVALID; }
    for (uint64_t i = 0; i < n; i++) {
        const rt_primitive& p = prims[i];
        if (p.type > RT_PRIM_YZ_RECT) { set_error("rt_scene_add_primitives: unknown type"); return RT_ERR_INVALID; }
        if (p.type == RT_PRIM_TRIANGLE && (p.mesh_index >= s->meshes.size() || (uint64_t)p.ind + 3 > s->meshes[p.mesh_index].ind.size())) { set_error("rt_scene_add_primitives: triangle out of range"); return RT_ERR_INVALID; }
        if (p.type >= RT_PRIM_XY_RECT && p.transform != RT_NONE && p.transform >= s->xforms.size()) { set_error("rt_scene_add_primitives: transform out of range"); return RT_ERR_INVALID; }
    }
    if (first) *first = (uint32_t)s->prims.size();
    for (uint64_t i = 0; i < n; i++) {
        const rt_primitive& p = prims[i];
        HostPrim h{}; h.type = (uint8_t)p.type; h.flip = p.flip_face ? 1 : 0; h.mat = p.mat_index; h.light = p.light_index;
        h.mesh = p.mesh_index; h.ind = p.ind; h.xform = p.type >= RT_PRIM_XY_RECT ? p.transform : RT_NONE; h.shape = 0;
        if (p.type != RT_PRIM_TRIANGLE) {
            h.shape = (uint32_t)s->shapes.size();
            for (int k = 0; k < 5; k++) s->shapes.push_back(p.v[k]);
            s->shapes.push_back(h.xform == RT_NONE ? -1.0 : (double)h.xform);
            for (int k = 0; k < 6; k++) s->shapes.push_back(0.0);     // f64 leaf box, filled at commit
        }
        s->prims.push_back(h);
    }
    s->committed = false;
    return RT_OK;
}
// Mesh::generate_triangles (hittable.rs:257-288)
rt_status rt_scene_add_mesh_triangles(rt_scene* s, uint32_t mesh, uint32_t mat, uint32_t* first) {
    if (!s || mesh >= s->meshes.size()) { set_error("rt_scene_add_mesh_triangles: bad mesh"); return RT_ERR_INVALID; }
    if (first) *first = (uint32_t)s->prims.size();
    const size_t nt = s->meshes[mesh].ind.size() / 3, base = s->prims.size();
    s->prims.resize(base + nt);                            // HVec: not zeroed first; filled by the threads below
    HostPrim* out = s->prims.data() + base;
    par_for(nt, (size_t)1 << 18, [=](size_t b, size_t e) {
        for (size_t t = b; t < e; t++) {
            HostPrim h{}; h.type = RT_PRIM_TRIANGLE; h.flip = 0; h.mat = mat; h.light = RT_NONE; h.mesh = mesh; h.ind = (uint32_t)(3 * t); h.xform = RT_NONE; h.shape = 0;
            out[t] = h;
        }
    });
    s->committed = false;
    return RT_OK;
}
rt_status rt_scene_set_light_index(rt_scene* s, uint32_t prim, uint32_t light) {
    if (!s || prim >= s->prims.size()) { set_error("rt_scene_set_light_index: bad primitive"); return RT_ERR_INVALID; }
    s->prims[prim].light = light; s->committed = false; return RT_OK;
}
rt_status rt_scene_set_textures(rt_scene* s, uint64_t n, const rt_texture* t) {
    if (!s || (n && !t)) { set_error("rt_scene_set_textures: null argument"); return RT_ERR_INVALID; }
    for (uint64_t i = 0; i < n; i++) {
        if (t[i].type > RT_TEX_PERLIN) { set_error("rt_scene_set_textures: unsupported texture type"); return RT_ERR_INVALID; }
        if (t[i].type == RT_TEX_PERLIN && (size_t)t[i].image >= s->perlin.size() / PERLIN_STRIDE) { set_error("rt_scene_set_textures: perlin index out of range (rt_scene_add_perlin first)"); return RT_ERR_INVALID; }
        if (t[i].type == RT_TEX_CHECKERED && (t[i].odd >= n || t[i].even >= n)) { set_error("rt_scene_set_textures: checkered texture id out of range"); return RT_ERR_INVALID; }
        if ((t[i].type == RT_TEX_HDR || t[i].type == RT_TEX_IMAGE) && t[i].image >= s->images.size()) { set_error("rt_scene_set_textures: image index out of range (add images first)"); return RT_ERR_INVALID; }
    }
    s->textures.assign(t, t + n); s->committed = false; return RT_OK;
}
// Texture::new_hdr (material.rs:626-642).  `data: Vec<Rgb<f32>>` is what image 0.23's HdrDecoder::read_image_hdr returns;
// Texture::get_value re-quantises the texel with image::hdr::to_rgbe8 on every lookup (material.rs:578) — a pure function
// of the texel, so it is applied once here (image 0.23.12, src/codecs/hdr/decoder.rs `to_rgbe8`: mx = max(r,g,b);
// mx <= 0 -> (0,0,0,0); else exp = floor(log2(mx)) + 1, c = trunc(v / 2^exp * 256) as u8, e = exp + 128).
rt_status rt_scene_add_image_hdr(rt_scene* s, uint32_t w, uint32_t h, const float* rgb, uint32_t* out) {
    if (!s || !rgb || !w || !h) { set_error("rt_scene_add_image_hdr: bad argument"); return RT_ERR_INVALID; }
    DImage im; im.offset = s->texels.size(); im.w = w; im.h = h;
    s->texels.resize(im.offset + (size_t)w * h);
    for (size_t i = 0; i < (size_t)w * h; i++) {
        const float* p = rgb + 3 * i;
        float mx = std::fmax(p[0], std::fmax(p[1], p[2]));
        uint32_t q = 0;
        if (mx > 0.0f) {
            int e = (int)std::floor(std::log2(mx)) + 1;
            float mul = std::ldexp(1.0f, e);                       // f32::powi(2.0, exp)
            for (int k = 0; k < 3; k++) {
                float c = std::trunc(p[k] / mul * 256.0f);
                uint32_t cu = !(c > 0.0f) ? 0u : (c >= 255.0f ? 255u : (uint32_t)c);   // `as u8` saturates
                q |= cu << (8 * k);
            }
            q |= ((uint32_t)(e + 128) & 255u) << 24;             // `(exp + 128) as u8` wraps
        }
        s->texels[im.offset + i] = q;
    }
    s->images.push_back(im); s->committed = false;
    if (out) *out = (uint32_t)s->images.size() - 1;
    return RT_OK;
}
// Texture::new_texture (material.rs:613-616)
rt_status rt_scene_add_image_rgb8(rt_scene* s, uint32_t w, uint32_t h, const uint8_t* rgb, uint32_t* out) {
    if (!s || !rgb || !w || !h) { set_error("rt_scene_add_image_rgb8: bad argument"); return RT_ERR_INVALID; }
    DImage im; im.offset = s->texels.size(); im.w = w; im.h = h;
    s->texels.resize(im.offset + (size_t)w * h);
    for (size_t i = 0; i < (size_t)w * h; i++) s->texels[im.offset + i] = (uint32_t)rgb[3 * i] | ((uint32_t)rgb[3 * i + 1] << 8) | ((uint32_t)rgb[3 * i + 2] << 16);
    s->images.push_back(im); s->committed = false;
    if (out) *out = (uint32_t)s->images.size() - 1;
    return RT_OK;
}
// Texture::new_perlin → Perlin::new(256, scale) (material.rs:628-631, perlin.rs:17-34)
rt_status rt_scene_add_perlin(rt_scene* s, double scale, const uint32_t* perm, const double* vecs, uint32_t* out) {
    if (!s || !perm || !vecs) { set_error("rt_scene_add_perlin: null argument"); return RT_ERR_INVALID; }
    for (int i = 0; i < 768; i++) if (perm[i] > 255u) { set_error("rt_scene_add_perlin: permutation value > 255"); return RT_ERR_INVALID; }
    const size_t base = s->perlin.size();
    s->perlin.resize(base + PERLIN_STRIDE);
    s->perlin[base] = scale;
    for (int i = 0; i < 768; i++) { s->perlin[base + 1 + i] = (double)perm[i]; s->perlin[base + 769 + i] = vecs[i]; }
    s->committed = false;
    if (out) *out = (uint32_t)(base / PERLIN_STRIDE);
    return RT_OK;
}
// Material::make_disney (material.rs:480-516)
rt_status rt_scene_add_disney(rt_scene* s, const rt_disney* d, uint32_t* out) {
    if (!s || !d) { set_error("rt_scene_add_disney: null argument"); return RT_ERR_INVALID; }
    s->disneys.push_back(*d); s->committed = false;
    if (out) *out = (uint32_t)s->disneys.size() - 1;
    return RT_OK;
}
rt_status rt_scene_set_materials(rt_scene* s, uint64_t n, const rt_material* m) {
    if (!s || (n && !m)) { set_error("rt_scene_set_materials: null argument"); return RT_ERR_INVALID; }
    for (uint64_t i = 0; i < n; i++) if (m[i].type > RT_MAT_DISNEY) { set_error("rt_scene_set_materials: unsupported material type"); return RT_ERR_INVALID; }
    s->materials.assign(m, m + n); s->committed = false; return RT_OK;
}
rt_status rt_scene_set_lights(rt_scene* s, uint64_t n, const rt_light* l) {
    if (!s || (n && !l)) { set_error("rt_scene_set_lights: null argument"); return RT_ERR_INVALID; }
    for (uint64_t i = 0; i < n; i++) if (l[i].type > RT_LIGHT_DISTANT) { set_error("rt_scene_set_lights: unsupported light type"); return RT_ERR_INVALID; }
    s->lights.assign(l, l + n); s->committed = false; return RT_OK;
}
rt_status rt_scene_commit(rt_scene* s) {
    if (!s) { set_error("rt_scene_commit: null scene"); return RT_ERR_INVALID; }
    auto t0 = std::chrono::steady_clock::now();
    bool device_build = s->build_mode == RT_BUILD_DEVICE_LBVH;
    if (const char* e = getenv("RT_BUILD")) { if (!strcmp(e, "lbvh")) device_build = true; else if (!strcmp(e, "sah")) device_build = false; }
    HostImage img;
    rt_status st = build_host_image(*s, img, device_build);
    if (st != RT_OK) return st;
    free_wavefront(*s); free_scene_device(*s);
    st = upload_scene(*s, img);
    if (st != RT_OK) return st;
    if (img.device_build && img.max_stack > TRAVERSAL_STACK) {
        // an LBVH over badly clustered Morton codes can be deeper than the traversal stack: take the host build instead
        if (getenv("RT_TRACE_BUILD")) fprintf(stderr, "bvh: device LBVH needs a stack of %d > %d, falling back to the host build\n", img.max_stack, TRAVERSAL_STACK);
        free_scene_device(*s);
        img = HostImage();
        if ((st = build_host_image(*s, img, false)) != RT_OK) return st;
        if ((st = upload_scene(*s, img)) != RT_OK) return st;
    }
    s->build_stats.mode = img.device_build ? RT_BUILD_DEVICE_LBVH : RT_BUILD_HOST_SAH;
    s->build_stats.max_stack = (uint32_t)img.max_stack; s->build_stats.nodes = s->d.n_nodes; s->build_stats.build_ms = img.build_ms;
    s->build_stats.commit_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    s->bin_mask = 0;
    for (auto& p : s->prims) s->bin_mask |= 1u << (s->materials[p.mat].type * 2 + (p.type == RT_PRIM_TRIANGLE ? 1 : 0));
    s->committed = true;
    return RT_OK;
}
rt_status rt_scene_set_build_mode(rt_scene* s, uint32_t mode) {
    if (!s || mode > RT_BUILD_DEVICE_LBVH) { set_error("rt_scene_set_build_mode: bad argument"); return RT_ERR_INVALID; }
    s->build_mode = mode; return RT_OK;
}
rt_status rt_scene_build_stats(const rt_scene* s, rt_build_stats* out) {
    if (!s || !out) { set_error("rt_scene_build_stats: null argument"); return RT_ERR_INVALID; }
    if (!s->committed) { set_error("rt_scene_build_stats: scene not committed"); return RT_ERR_STATE; }
    *out = s->build_stats; return RT_OK;
}
rt_status rt_scene_num_primitives(const rt_scene* s, uint64_t* out) {
    if (!s || !out) { set_error("rt_scene_num_primitives: null argument"); return RT_ERR_INVALID; }
    *out = s->prims.size(); return RT_OK;
}
rt_status rt_scene_device_bytes(const rt_scene* s, uint64_t* nodes, uint64_t* leaves, uint64_t* tables) {
    if (!s) { set_error("rt_scene_device_bytes: null scene"); return RT_ERR_INVALID; }
    if (nodes) *nodes = s->bytes_nodes; if (leaves) *leaves = s->bytes_slots; if (tables) *tables = s->bytes_tables;
    return RT_OK;
}
rt_status rt_set_wavefront_size(rt_scene* s, uint32_t paths) {
    if (!s || paths < 1024) { set_error("rt_set_wavefront_size: need >= 1024 paths"); return RT_ERR_INVALID; }
    free_wavefront(*s); s->wavefront_size = paths; return RT_OK;
}
rt_status rt_render_cancel(rt_scene* s) { if (!s) return RT_ERR_INVALID; s->cancel.store(1); return RT_OK; }
rt_status rt_get_counters(const rt_scene* s, rt_counters* out) {
    if (!s || !out) { set_error("rt_get_counters: null argument"); return RT_ERR_INVALID; }
    *out = s->counters; return RT_OK;
}

}  // extern "C"
