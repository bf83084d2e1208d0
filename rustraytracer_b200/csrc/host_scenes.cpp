// Host scene layer: the reference's scene functions (scenes.rs), OBJ ingest (parser.rs), and the
// deterministic stand-ins / generators of SURVEY.md §8d.  Not on the hot path.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <atomic>
#include <cstdint>
#include <thread>
#include <mutex>
#include <string>
#include <tuple>
#include <type_traits>
#include <vector>
#include "rt_internal.h"
#include "camera_host.h"
#include "../../include/rthost.h"

using rt::PI;
using rt::SMALL;

// librthost.so keeps its own last-error string (rth_last_error_string): it does not link against librtcore.so
static thread_local std::string g_host_error;
static void host_error(const std::string& msg) { g_host_error = msg; }

struct rth_scene {
    std::vector<rt::HostMesh> meshes;
    // primitives in order; a triangle run of a mesh is kept implicit
    struct Item { bool run; rt_primitive prim; uint32_t mesh, mat; };
    std::vector<Item> items;
    std::vector<std::pair<uint64_t, uint32_t>> light_sets;   // (prim index, light index) applied after expansion
    std::vector<double> xforms;                               // 16 each
    std::vector<rt_texture> textures;
    std::vector<rt_material> materials;
    std::vector<rt_disney> disneys;                           // Material::Disney records (material tex[0] = index)
    std::vector<rt_light> lights;
    struct Image { uint32_t w = 0, h = 0; bool hdr = false; std::vector<float> f; std::vector<uint8_t> b; };   // Texture::Hdr data / Texture::Image img
    std::vector<Image> images;
    rt_camera camera{};
    std::string png;
    bool standin = false;

    uint64_t n_prims() const { uint64_t n = 0; for (auto& it : items) n += it.run ? meshes[it.mesh].ind.size() / 3 : 1; return n; }
};

namespace {

// ---- small constructors mirroring the reference's
uint32_t push_solid(rth_scene& s, double r, double g, double b) {                  // Texture::new_solid_color
    rt_texture t{}; t.type = RT_TEX_SOLID; t.color[0] = r; t.color[1] = g; t.color[2] = b; s.textures.push_back(t); return (uint32_t)s.textures.size() - 1;
}
uint32_t push_checkered(rth_scene& s, uint32_t even, uint32_t odd, double freq) {  // Texture::new_checkered(even, odd, f) material.rs:622-628
    rt_texture t{}; t.type = RT_TEX_CHECKERED; t.even = even; t.odd = odd; t.frequency = freq; s.textures.push_back(t); return (uint32_t)s.textures.size() - 1;
}
rt_material mat_base(uint32_t type) { rt_material m{}; m.type = type; for (int k = 0; k < 5; k++) m.tex[k] = RT_NONE; return m; }
uint32_t push_matte(rth_scene& s, uint32_t kd, double sigma) { rt_material m = mat_base(RT_MAT_MATTE); m.tex[0] = kd; m.f[0] = sigma; s.materials.push_back(m); return (uint32_t)s.materials.size() - 1; }
uint32_t push_light_mat(rth_scene& s, uint32_t tex) { rt_material m = mat_base(RT_MAT_LIGHT); m.tex[0] = tex; s.materials.push_back(m); return (uint32_t)s.materials.size() - 1; }
uint32_t push_plastic(rth_scene& s, uint32_t kd, uint32_t ks, double rough, bool remap) { rt_material m = mat_base(RT_MAT_PLASTIC); m.tex[0] = kd; m.tex[1] = ks; m.f[0] = rough; m.remap_roughness = remap; s.materials.push_back(m); return (uint32_t)s.materials.size() - 1; }
uint32_t push_glass(rth_scene& s, uint32_t kr, uint32_t kt, double ur, double vr, double index, bool remap) { rt_material m = mat_base(RT_MAT_GLASS); m.tex[0] = kr; m.tex[1] = kt; m.f[0] = ur; m.f[1] = vr; m.f[2] = index; m.remap_roughness = remap; s.materials.push_back(m); return (uint32_t)s.materials.size() - 1; }
// Material::make_metal(eta_id, k_id, u_r_id, v_r_id, r_id, bump, remap) (material.rs:452-470)
uint32_t push_metal(rth_scene& s, uint32_t eta, uint32_t k, uint32_t ur, uint32_t vr, uint32_t r, bool remap) { rt_material m = mat_base(RT_MAT_METAL); m.tex[0] = eta; m.tex[1] = k; m.tex[2] = r; m.tex[3] = ur; m.tex[4] = vr; m.remap_roughness = remap; s.materials.push_back(m); return (uint32_t)s.materials.size() - 1; }
uint32_t push_mirror(rth_scene& s, uint32_t c) { rt_material m = mat_base(RT_MAT_MIRROR); m.tex[0] = c; s.materials.push_back(m); return (uint32_t)s.materials.size() - 1; }

rt_primitive prim_base(uint32_t type, uint32_t mat) { rt_primitive p{}; p.type = type; p.mat_index = mat; p.light_index = RT_NONE; p.transform = RT_NONE; return p; }
rt_primitive make_sphere(double cx, double cy, double cz, double r, uint32_t mat) { rt_primitive p = prim_base(RT_PRIM_SPHERE, mat); p.v[0] = cx; p.v[1] = cy; p.v[2] = cz; p.v[3] = r; return p; }
rt_primitive make_rect(uint32_t type, double a0, double b0, double a1, double b1, double k, uint32_t mat, uint32_t xf = RT_NONE) {
    rt_primitive p = prim_base(type, mat); p.v[0] = a0; p.v[1] = b0; p.v[2] = a1; p.v[3] = b1; p.v[4] = k; p.transform = xf; return p;
}
rt_primitive flip(rt_primitive p) { p.flip_face = 1; return p; }
uint64_t push_prim(rth_scene& s, const rt_primitive& p) { uint64_t i = s.n_prims(); rth_scene::Item it{}; it.run = false; it.prim = p; s.items.push_back(it); return i; }
uint64_t push_mesh_triangles(rth_scene& s, uint32_t mesh, uint32_t mat) { uint64_t i = s.n_prims(); rth_scene::Item it{}; it.run = true; it.mesh = mesh; it.mat = mat; s.items.push_back(it); return i; }
uint32_t push_diffuse_light(rth_scene& s, uint64_t prim, double r, double g, double b, bool two_sided) {   // Light::make_diffuse_light light.rs:585-606
    rt_light l{}; l.type = RT_LIGHT_DIFFUSE; l.prim_index = (uint32_t)prim; l.two_sided = two_sided; l.color[0] = r; l.color[1] = g; l.color[2] = b;
    s.lights.push_back(l); return (uint32_t)s.lights.size() - 1;
}
uint32_t push_xform(rth_scene& s, const double m[16]) { s.xforms.insert(s.xforms.end(), m, m + 16); return (uint32_t)(s.xforms.size() / 16) - 1; }

void camera(rth_scene& s, const double from[3], const double to[3], double aspect, double vfov, double aperture, double focus, double t0, double t1) {
    double up[3] = {0., 1., 0.};
    rt::camera_look_at(from, to, up, aspect, vfov, aperture, focus, t0, t1, &s.camera);
}
uint32_t push_hdr_texture(rth_scene& s, uint32_t image) { rt_texture t{}; t.type = RT_TEX_HDR; t.image = image; s.textures.push_back(t); return (uint32_t)s.textures.size() - 1; }      // Texture::new_hdr
uint32_t push_image_texture(rth_scene& s, uint32_t image) { rt_texture t{}; t.type = RT_TEX_IMAGE; t.image = image; s.textures.push_back(t); return (uint32_t)s.textures.size() - 1; }  // Texture::new_texture
uint32_t push_infinite_light(rth_scene& s, uint32_t text_id, uint32_t xf = RT_NONE) {     // Light::make_infinite_light(to_world, n_samples, text_id) light.rs:608
    rt_light l{}; l.type = RT_LIGHT_INFINITE; l.prim_index = RT_NONE; l.text_id = text_id; l.transform = xf; s.lights.push_back(l); return (uint32_t)s.lights.size() - 1;
}

// ---- 4x4 helpers (row-major)
void mat_identity(double m[16]) { for (int i = 0; i < 16; i++) m[i] = (i % 5 == 0) ? 1. : 0.; }
void mat_mul(const double a[16], const double b[16], double o[16]) {
    double r[16];
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) { double acc = 0.; for (int k = 0; k < 4; k++) acc += a[4 * i + k] * b[4 * k + j]; r[4 * i + j] = acc; }
    memcpy(o, r, sizeof(r));
}
void mat_translate(double x, double y, double z, double m[16]) { mat_identity(m); m[3] = x; m[7] = y; m[11] = z; }
void mat_scale(double s, double m[16]) { mat_identity(m); m[0] = m[5] = m[10] = s; }
// Rotation3::from_euler_angles(roll, pitch, yaw) = Rz(yaw) * Ry(pitch) * Rx(roll) (nalgebra)
void mat_euler(double roll, double pitch, double yaw, double m[16]) {
    double sr = std::sin(roll), cr = std::cos(roll), sp = std::sin(pitch), cp = std::cos(pitch), sy = std::sin(yaw), cy = std::cos(yaw);
    mat_identity(m);
    m[0] = cy * cp; m[1] = cy * sp * sr - sy * cr; m[2] = cy * sp * cr + sy * sr;
    m[4] = sy * cp; m[5] = sy * sp * sr + cy * cr; m[6] = sy * sp * cr - cy * sr;
    m[8] = -sp;     m[9] = cp * sr;                m[10] = cp * cr;
}
void xf_point(const double* m, const double p[3], double o[3]) {
    double x = m[0] * p[0] + m[1] * p[1] + m[2] * p[2] + m[3], y = m[4] * p[0] + m[5] * p[1] + m[6] * p[2] + m[7];
    double z = m[8] * p[0] + m[9] * p[1] + m[10] * p[2] + m[11], w = m[12] * p[0] + m[13] * p[1] + m[14] * p[2] + m[15];
    o[0] = x / w; o[1] = y / w; o[2] = z / w;
}
void xf_vector(const double* m, const double v[3], double o[3]) {
    o[0] = m[0] * v[0] + m[1] * v[1] + m[2] * v[2]; o[1] = m[4] * v[0] + m[5] * v[1] + m[6] * v[2]; o[2] = m[8] * v[0] + m[9] * v[1] + m[10] * v[2];
}

// ---- Philox4x32-10 for the generators
void philox(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1, n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3; k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
struct Gen {   // sequential uniform stream
    uint64_t key; uint64_t ctr = 0; uint32_t buf[4]; int have = 0;
    explicit Gen(uint64_t k) : key(k) {}
    uint32_t u32() { if (!have) { buf[0] = (uint32_t)ctr; buf[1] = (uint32_t)(ctr >> 32); buf[2] = 0; buf[3] = 0; ctr++; philox(buf, (uint32_t)key, (uint32_t)(key >> 32)); have = 4; } return buf[--have]; }
    double uniform() { return (double)(u32() >> 8) * (1.0 / 16777216.0); }            // 24-bit, f32-exact
    double range(double a, double b) { return a + (b - a) * uniform(); }
};

// ---- stand-in meshes --------------------------------------------------------------------------
// Closed genus-0 "blob": a lat-long sphere (n_u around, n_v rings incl. poles) with multi-octave
// sinusoidal displacement, scaled into a bounding box.  Triangles: 2*n_u*(n_v-1).  Smooth vertex
// normals (area-weighted), no uvs — like the Stanford dragon OBJ the reference loads.
void append_blob(rt::HostMesh& m, uint32_t n_u, uint32_t n_v, uint64_t seed, const double lo[3], const double hi[3]) {
    Gen g(seed);
    const int OCT = 6;
    double amp[OCT], fu[OCT], fv[OCT], pu[OCT], pv[OCT];
    double a = 0.22;
    for (int k = 0; k < OCT; k++) { amp[k] = a; a *= 0.55; fu[k] = (double)(2 + (int)(g.uniform() * 3.) + 3 * k * k); fv[k] = (double)(1 + (int)(g.uniform() * 3.) + 2 * k * k); pu[k] = g.range(0., 6.28318); pv[k] = g.range(0., 6.28318); }
    uint32_t base = (uint32_t)(m.p.size() / 3);
    auto radius = [&](double phi, double theta) {
        double r = 1.0;
        for (int k = 0; k < OCT; k++) r += amp[k] * std::sin(fu[k] * phi + pu[k]) * std::sin(fv[k] * theta + pv[k]) * std::sin(theta);
        return r;
    };
    std::vector<double> pts;
    auto put = [&](double phi, double theta) {
        double r = radius(phi, theta), st = std::sin(theta);
        pts.push_back(r * st * std::cos(phi)); pts.push_back(r * std::cos(theta)); pts.push_back(r * st * std::sin(phi));
    };
    put(0., 0.);                                                       // north pole
    for (uint32_t j = 1; j < n_v; j++) for (uint32_t i = 0; i < n_u; i++) put(2. * M_PI * i / n_u, M_PI * j / n_v);
    put(0., M_PI);                                                     // south pole
    // fit into [lo,hi]
    double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
    for (size_t i = 0; i < pts.size(); i += 3) for (int k = 0; k < 3; k++) { mn[k] = std::min(mn[k], pts[i + k]); mx[k] = std::max(mx[k], pts[i + k]); }
    for (size_t i = 0; i < pts.size(); i += 3) for (int k = 0; k < 3; k++) {
        double v = lo[k] + (pts[i + k] - mn[k]) / (mx[k] - mn[k]) * (hi[k] - lo[k]);
        pts[i + k] = (double)(float)v;                                 // OBJ values are f32 (tobj) widened to f64
    }
    uint32_t nvert = (uint32_t)(pts.size() / 3), south = nvert - 1;
    auto ring = [&](uint32_t j, uint32_t i) { return 1 + (j - 1) * n_u + (i % n_u); };
    std::vector<uint32_t> idx;
    for (uint32_t i = 0; i < n_u; i++) { idx.push_back(0); idx.push_back(ring(1, i + 1)); idx.push_back(ring(1, i)); }
    for (uint32_t j = 1; j + 1 < n_v; j++) for (uint32_t i = 0; i < n_u; i++) {
        uint32_t a0 = ring(j, i), a1 = ring(j, i + 1), b0 = ring(j + 1, i), b1 = ring(j + 1, i + 1);
        idx.push_back(a0); idx.push_back(a1); idx.push_back(b0);
        idx.push_back(a1); idx.push_back(b1); idx.push_back(b0);
    }
    for (uint32_t i = 0; i < n_u; i++) { idx.push_back(south); idx.push_back(ring(n_v - 1, i)); idx.push_back(ring(n_v - 1, i + 1)); }
    std::vector<double> nrm(pts.size(), 0.);
    for (size_t t = 0; t < idx.size(); t += 3) {
        const double* p0 = &pts[3 * idx[t]]; const double* p1 = &pts[3 * idx[t + 1]]; const double* p2 = &pts[3 * idx[t + 2]];
        double e1[3] = {p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2]}, e2[3] = {p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2]};
        double c[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
        for (int v = 0; v < 3; v++) for (int k = 0; k < 3; k++) nrm[3 * idx[t + v] + k] += c[k];
    }
    for (size_t i = 0; i < nrm.size(); i += 3) {
        double l = std::sqrt(nrm[i] * nrm[i] + nrm[i + 1] * nrm[i + 1] + nrm[i + 2] * nrm[i + 2]);
        if (l > 0.) for (int k = 0; k < 3; k++) nrm[i + k] = (double)(float)(nrm[i + k] / l); else { nrm[i] = 0.; nrm[i + 1] = 1.; nrm[i + 2] = 0.; }
    }
    m.p.insert(m.p.end(), pts.begin(), pts.end());
    m.n.insert(m.n.end(), nrm.begin(), nrm.end());
    for (uint32_t v : idx) m.ind.push_back(base + v);
}

// Stand-in for data/dragon/dragon.obj (missing blob): exactly 871,414 triangles =
// 2*660*659 (body) + 2*13*59 (a small second component), in the box the reference's camera frames.
void dragon_standin(rt::HostMesh& m, uint64_t seed, double tx) {
    double lo[3] = {-3.8 + tx, -2.83, -1.7}, hi[3] = {3.8 + tx, 2.53, 1.7};
    append_blob(m, 660, 660, seed, lo, hi);
    double lo2[3] = {2.6 + tx, 2.0, -0.4}, hi2[3] = {3.4 + tx, 2.9, 0.4};
    append_blob(m, 13, 60, seed + 1, lo2, hi2);
}
// Stand-in for data/statue.obj inside the Cornell box (scenes.rs:289-292 places it around (374,*,130))
void statue_standin(rt::HostMesh& m, uint64_t seed, uint64_t n_tris) {
    uint32_t n = n_tris ? (uint32_t)std::max(8.0, std::sqrt((double)n_tris / 2.)) : 500;
    double lo[3] = {190., 0., 140.}, hi[3] = {365., 420., 300.};
    append_blob(m, n, n + 1, seed, lo, hi);
}

// ---- parser::parse_obj (parser.rs:8-87) over a tobj-2.0.2-style loader -------------------------
// Two passes.  (1) text -> numbers, PARALLEL: the file is cut into one chunk per host thread at line boundaries; every chunk parses its `v` / `vn` / `vt`
// lines into its own arrays and keeps its faces as raw (v, vt, vn) index tokens together with how many v / vt / vn lines of the chunk precede the face
// (relative indices count back from there).  (2) single-index re-indexing, parallel as well: tobj numbers the (v, vt, vn) triples in order of first
// appearance, which is the rank of each triple's smallest corner number — a concurrent hash table finds those minima, a prefix sum ranks them.  Only models[0] is used (parser.rs:21): reading stops at the first `o` / `g`
// line that follows a face.  Values are f32 (tobj) widened to f64; strtof and Rust's str::parse::<f32> both round correctly, so they agree.
struct ObjCorner { int iv, it, in; bool ht, hn; };
struct ObjChunk {
    std::vector<float> v, vn, vt;
    std::vector<ObjCorner> corners;
    struct Face { size_t off; uint32_t first, n, nv, nt, nn; };      // byte offset of the line; corners [first, first + n); v / vt / vn lines of this chunk seen before it
    std::vector<Face> faces;
    std::vector<size_t> og;                                          // byte offsets of `o` / `g` lines
};
static void obj_parse_chunk(const char* buf, size_t begin, size_t end, ObjChunk& c) {
    size_t pos = begin;
    while (pos < end) {
        size_t eol = pos; while (eol < end && buf[eol] != '\n') eol++;
        const char* p = buf + pos; const char* le = buf + eol;
        while (p < le && (*p == ' ' || *p == '\t')) p++;
        auto num = [&](const char*& q) -> float {            // next float of the line, 0 when there is none (sscanf leaves its default)
            while (q < le && (*q == ' ' || *q == '\t')) q++;
            if (q >= le) return 0.f;
            char* e = nullptr; float x = strtof(q, &e);
            if (e == q) { q = le; return 0.f; }
            q = e; return x;
        };
        if (le - p >= 2 && p[0] == 'v' && (p[1] == ' ' || p[1] == '\t')) { const char* q = p + 2; float a = num(q), b = num(q), d = num(q); c.v.push_back(a); c.v.push_back(b); c.v.push_back(d); }
        else if (le - p >= 2 && p[0] == 'v' && p[1] == 'n') { const char* q = p + 2; float a = num(q), b = num(q), d = num(q); c.vn.push_back(a); c.vn.push_back(b); c.vn.push_back(d); }
        else if (le - p >= 2 && p[0] == 'v' && p[1] == 't') { const char* q = p + 2; float a = num(q), b = num(q); c.vt.push_back(a); c.vt.push_back(b); }
        else if (le - p >= 2 && p[0] == 'f' && (p[1] == ' ' || p[1] == '\t')) {
            ObjChunk::Face f; f.off = pos; f.first = (uint32_t)c.corners.size(); f.nv = (uint32_t)(c.v.size() / 3); f.nt = (uint32_t)(c.vt.size() / 2); f.nn = (uint32_t)(c.vn.size() / 3);
            const char* q = p + 1;
            for (;;) {
                while (q < le && (*q == ' ' || *q == '\t')) q++;
                if (q >= le || *q == '\r' || *q == '#' || *q == 0) break;
                ObjCorner k{0, 0, 0, false, false};
                char* e = nullptr;
                k.iv = (int)strtol(q, &e, 10); if (e == q) break; q = e;
                if (q < le && *q == '/') { q++; if (q < le && *q != '/') { k.it = (int)strtol(q, &e, 10); q = e; k.ht = true; } if (q < le && *q == '/') { q++; k.in = (int)strtol(q, &e, 10); q = e; k.hn = true; } }
                c.corners.push_back(k);
            }
            f.n = (uint32_t)c.corners.size() - f.first;
            c.faces.push_back(f);
        } else if (le - p >= 1 && (p[0] == 'o' || p[0] == 'g') && (le - p == 1 || p[1] == ' ' || p[1] == '\t' || p[1] == '\r')) c.og.push_back(pos);
        pos = eol + 1;
    }
}
bool load_obj(const char* path, const double* xf, rt::HostMesh& out, std::string& err) {
    FILE* f = fopen(path, "rb");
    if (!f) { err = std::string("cannot open ") + path; return false; }
    fseek(f, 0, SEEK_END); const long fsz = ftell(f); fseek(f, 0, SEEK_SET);
    std::vector<char> buf((size_t)std::max(fsz, 0L) + 1);
    const size_t got = fsz > 0 ? fread(buf.data(), 1, (size_t)fsz, f) : 0;
    fclose(f);
    buf[got] = '\n';                                           // strtof / strtol always meet a terminator
    const size_t size = got;
    // pass 1: chunks at line boundaries
    size_t n_chunks = std::max<size_t>(1, std::min<size_t>(rt::host_threads(), size / (256 * 1024) + 1));
    std::vector<size_t> cut(n_chunks + 1, size);
    cut[0] = 0;
    for (size_t k = 1; k < n_chunks; k++) { size_t q = std::max(cut[k - 1], size * k / n_chunks); while (q < size && buf[q] != '\n') q++; cut[k] = std::min(size, q + 1); }
    std::vector<ObjChunk> chunks(n_chunks);
    {
        std::vector<std::thread> pool;
        for (size_t k = 0; k < n_chunks; k++) pool.emplace_back([&, k]() { obj_parse_chunk(buf.data(), cut[k], cut[k + 1], chunks[k]); });
        for (auto& t : pool) t.join();
    }
    // where models[0] ends: the first o / g line after the first face
    size_t first_face = SIZE_MAX, stop = SIZE_MAX;
    for (auto& c : chunks) if (!c.faces.empty()) { first_face = c.faces[0].off; break; }
    if (first_face != SIZE_MAX) for (auto& c : chunks) { for (size_t o : c.og) if (o > first_face) { stop = o; break; } if (stop != SIZE_MAX) break; }
    // concatenated attribute arrays + per-chunk bases
    std::vector<size_t> bv(n_chunks + 1, 0), bt(n_chunks + 1, 0), bn(n_chunks + 1, 0);
    size_t n_corners = 0;
    for (size_t k = 0; k < n_chunks; k++) { bv[k + 1] = bv[k] + chunks[k].v.size() / 3; bt[k + 1] = bt[k] + chunks[k].vt.size() / 2; bn[k + 1] = bn[k] + chunks[k].vn.size() / 3; n_corners += chunks[k].corners.size(); }
    std::vector<float> v(3 * bv[n_chunks]), vt(2 * bt[n_chunks]), vn(3 * bn[n_chunks]);
    for (size_t k = 0; k < n_chunks; k++) {
        if (!chunks[k].v.empty()) memcpy(&v[3 * bv[k]], chunks[k].v.data(), chunks[k].v.size() * 4);
        if (!chunks[k].vt.empty()) memcpy(&vt[2 * bt[k]], chunks[k].vt.data(), chunks[k].vt.size() * 4);
        if (!chunks[k].vn.empty()) memcpy(&vn[3 * bn[k]], chunks[k].vn.data(), chunks[k].vn.size() * 4);
    }
    // pass 2: (v, vt, vn) triples -> single index in order of first appearance (tobj's numbering), without a sequential walk:
    //   a. every corner of models[0] gets its absolute key — relative indices resolved against the v / vt / vn counts in front of its face (per chunk)
    //   b. concurrent open-addressing table on the triple; a slot ends up holding the SMALLEST global corner number that carries its key
    //      (compare-and-swap into an empty slot, atomic minimum afterwards: whatever sits in a slot always has the slot's key)
    //   c. a corner is a first appearance iff it is its slot's minimum; the exclusive prefix sum of those flags in corner order is the numbering
    //   d. attributes of the first appearances and the fan triangulation are written straight to their final places
    std::vector<size_t> nf(n_chunks, 0), cb(n_chunks + 1, 0), tb(n_chunks + 1, 0);      // faces in front of `stop`, corner / triangle bases per chunk
    for (size_t k = 0; k < n_chunks; k++) {
        const ObjChunk& c = chunks[k];
        size_t f = 0, tris = 0;
        while (f < c.faces.size() && c.faces[f].off < stop) { tris += c.faces[f].n >= 3 ? c.faces[f].n - 2 : 0; f++; }
        nf[k] = f;
        cb[k + 1] = cb[k] + (f ? (size_t)c.faces[f - 1].first + c.faces[f - 1].n : 0);
        tb[k + 1] = tb[k] + tris;
    }
    const size_t NC = cb[n_chunks], NT = tb[n_chunks];
    if (NT == 0) { err = std::string("no faces in ") + path; return false; }
    if (NC >= 0x7fffffffull) { err = "too many face corners"; return false; }
    rt::HVec<int> kv, kt, kn; rt::HVec<uint8_t> cflag; rt::HVec<uint32_t> slot_of, cid;
    kv.resize(NC); kt.resize(NC); kn.resize(NC); cflag.resize(NC); slot_of.resize(NC); cid.resize(NC);
    std::atomic<int> bad{0};
    rt::par_for(n_chunks, 1, [&](size_t k0, size_t k1) {
        for (size_t k = k0; k < k1; k++) {
            const ObjChunk& c = chunks[k];
            for (size_t f = 0; f < nf[k]; f++) {
                const ObjChunk::Face& fc = c.faces[f];
                const int nv = (int)(bv[k] + fc.nv), nt = (int)(bt[k] + fc.nt), nn = (int)(bn[k] + fc.nn);
                for (uint32_t j = fc.first; j < fc.first + fc.n; j++) {
                    const ObjCorner& q = c.corners[j];
                    const int iv = q.iv < 0 ? nv + q.iv : q.iv - 1, it = q.ht ? (q.it < 0 ? nt + q.it : q.it - 1) : -1, in = q.hn ? (q.in < 0 ? nn + q.in : q.in - 1) : -1;
                    if (iv < 0 || iv >= nv) { bad.store(1); return; }
                    const size_t g = cb[k] + j;
                    kv[g] = iv; kt[g] = it; kn[g] = in;
                    cflag[g] = (uint8_t)((it >= 0 && it < nt ? 1 : 0) | (in >= 0 && in < nn ? 2 : 0));
                }
            }
        }
    });
    if (bad.load()) { err = "face index out of range"; return false; }
    size_t cap = 16; while (cap < 2 * NC + 16) cap <<= 1;
    const uint32_t EMPTY = 0xffffffffu;
    std::unique_ptr<std::atomic<uint32_t>[]> table(new std::atomic<uint32_t>[cap]);
    rt::par_for(cap, (size_t)1 << 18, [&](size_t b0, size_t e0) { for (size_t i = b0; i < e0; i++) table[i].store(EMPTY, std::memory_order_relaxed); });
    rt::par_for(NC, (size_t)1 << 15, [&](size_t b0, size_t e0) {
        for (size_t g = b0; g < e0; g++) {
            const int iv = kv[g], it = kt[g], in = kn[g];
            const uint64_t h = ((uint64_t)(uint32_t)iv * 0x9E3779B97F4A7C15ull) ^ ((uint64_t)(uint32_t)it * 0xC2B2AE3D27D4EB4Full) ^ ((uint64_t)(uint32_t)in * 0x165667B19E3779F9ull);
            size_t slot = (size_t)(h ^ (h >> 29)) & (cap - 1);
            for (;;) {
                uint32_t cur = table[slot].load(std::memory_order_relaxed);
                if (cur == EMPTY && table[slot].compare_exchange_strong(cur, (uint32_t)g)) break;
                if (kv[cur] == iv && kt[cur] == it && kn[cur] == in) {            // the slot's key is ours: keep the smaller corner number
                    while ((uint32_t)g < cur && !table[slot].compare_exchange_weak(cur, (uint32_t)g)) {}
                    break;
                }
                slot = (slot + 1) & (cap - 1);
            }
            slot_of[g] = (uint32_t)slot;
        }
    });
    // first appearances, numbered in corner order
    const size_t NB = std::max<size_t>(1, std::min<size_t>(rt::host_threads() * 4, NC / 4096 + 1)), blk = (NC + NB - 1) / NB;
    std::vector<size_t> first_base(NB + 1, 0), tex_ok(NB, 0), nor_ok(NB, 0);
    rt::par_for(NB, 1, [&](size_t q0, size_t q1) {
        for (size_t q = q0; q < q1; q++) {
            size_t n = 0, t = 0, m = 0;
            for (size_t g = q * blk, e = std::min(NC, g + blk); g < e; g++) if (table[slot_of[g]].load(std::memory_order_relaxed) == (uint32_t)g) { n++; t += cflag[g] & 1; m += (cflag[g] >> 1) & 1; }
            first_base[q + 1] = n; tex_ok[q] = t; nor_ok[q] = m;
        }
    });
    size_t n_tex = 0, n_nor = 0;
    for (size_t q = 0; q < NB; q++) { first_base[q + 1] += first_base[q]; n_tex += tex_ok[q]; n_nor += nor_ok[q]; }
    const size_t np = first_base[NB];
    const bool with_uv = n_tex == np, with_n = n_nor == np;        // parser.rs keeps normals / uvs only when every vertex has one
    out.p.resize(3 * np);
    if (with_n) out.n.resize(3 * np);
    if (with_uv) out.uv.resize(2 * np);
    rt::par_for(NB, 1, [&](size_t q0, size_t q1) {
        for (size_t q = q0; q < q1; q++) {
            size_t id = first_base[q];
            for (size_t g = q * blk, e = std::min(NC, g + blk); g < e; g++) {
                if (table[slot_of[g]].load(std::memory_order_relaxed) != (uint32_t)g) continue;
                cid[g] = (uint32_t)id;
                const float* pv = &v[3 * (size_t)kv[g]];
                double pt[3] = {(double)pv[0], (double)pv[1], (double)pv[2]}, o[3];
                xf_point(xf, pt, o); out.p[3 * id] = o[0]; out.p[3 * id + 1] = o[1]; out.p[3 * id + 2] = o[2];
                if (with_n) { const float* pn = &vn[3 * (size_t)kn[g]]; double vv[3] = {(double)pn[0], (double)pn[1], (double)pn[2]}; xf_vector(xf, vv, o); out.n[3 * id] = o[0]; out.n[3 * id + 1] = o[1]; out.n[3 * id + 2] = o[2]; }
                if (with_uv) { out.uv[2 * id] = (double)vt[2 * (size_t)kt[g]]; out.uv[2 * id + 1] = (double)vt[2 * (size_t)kt[g] + 1]; }
                id++;
            }
        }
    });
    out.ind.resize(3 * NT);
    rt::par_for(n_chunks, 1, [&](size_t k0, size_t k1) {
        for (size_t k = k0; k < k1; k++) {
            const ObjChunk& c = chunks[k];
            size_t t = tb[k];
            for (size_t f = 0; f < nf[k]; f++) {
                const ObjChunk::Face& fc = c.faces[f];
                auto id_of = [&](uint32_t j) { return cid[table[slot_of[cb[k] + j]].load(std::memory_order_relaxed)]; };
                if (fc.n < 3) continue;
                const uint32_t i0 = id_of(fc.first);
                uint32_t prev = id_of(fc.first + 1);
                for (uint32_t j = fc.first + 2; j < fc.first + fc.n; j++) {                 // fan triangulation
                    const uint32_t cur = id_of(j);
                    out.ind[3 * t] = i0; out.ind[3 * t + 1] = prev; out.ind[3 * t + 2] = cur; t++; prev = cur;
                }
            }
        }
    });
    return true;
}
bool file_exists(const std::string& p) { FILE* f = fopen(p.c_str(), "rb"); if (f) { fclose(f); return true; } return false; }

}  // namespace

namespace {

double aspect_of(const rth_params& p) { return (double)p.width / (double)p.height; }   // GLOBAL_STATE.get_aspect_ratio()

// dragon mesh `which` (0: at origin, 1: translated by (5,0,0) — scenes.rs:551-552), real OBJ when available
uint32_t add_dragon(rth_scene& s, const rth_params& p, double tx) {
    rt::HostMesh m;
    std::string path = p.data_dir ? std::string(p.data_dir) + "/dragon/dragon.obj" : std::string();
    bool loaded = false;
    if (!path.empty() && file_exists(path)) {
        double xf[16]; mat_scale(10., xf); xf[3] = tx;      // Similarity3::new(translation, 0, 10.).to_homogeneous()
        std::string err; loaded = load_obj(path.c_str(), xf, m, err);
    }
    if (!loaded) { dragon_standin(m, p.seed ? p.seed : 7, tx); s.standin = true; }
    s.meshes.push_back(std::move(m));
    return (uint32_t)s.meshes.size() - 1;
}

// scenes.rs:474-546
void build_sphere_roughness(rth_scene& s, const rth_params& p) {
    double from[3] = {-8.5, 5., 0.}, to[3] = {0., -0.15, -0.08};
    camera(s, from, to, aspect_of(p), 70., 0.0, 10., 0., 0.);
    uint32_t t0 = push_solid(s, 0.4 * 2., 0.15 * 2., 0.15 * 2.);        // light_gray
    uint32_t t1 = push_solid(s, 0.15 * 2., 0.15 * 2., 0.4 * 2.);        // dark_gray
    push_checkered(s, t0, t1, 0.1);                                      // 2
    push_solid(s, 15., 15., 15.);                                        // 3
    push_matte(s, 2, 0.);                                                // mat 0
    const double space = 2.8;
    for (int i = 1; i < 9; i++) {
        uint32_t len = (uint32_t)s.textures.size();
        push_solid(s, ((double)(i - 1)) / 90. + SMALL, 0., 0.);          // roughness
        push_solid(s, 0., 0., 0.);                                       // eta = black
        push_solid(s, 1., 1., 1.);                                       // k = white
        push_metal(s, len + 1, len + 2, RT_NONE, RT_NONE, len, true);
        push_prim(s, make_sphere(0., 1., -space * 4.5 + space * (double)i, 1., (uint32_t)i));
    }
    push_prim(s, make_rect(RT_PRIM_XZ_RECT, -10000., -10000., 10000., 10000., -0.01, 0));
    rt_primitive light = make_rect(RT_PRIM_XZ_RECT, -10., -10., 10., 10., 50., (uint32_t)s.materials.size() - 1);
    light.light_index = 0;
    uint64_t li = push_prim(s, flip(light));
    push_diffuse_light(s, li, 10., 10., 10., false);
    s.png = "metal_spheres.png";
}

// the commented make_world (scenes.rs:15-86), regenerated with a fixed seed (its util::rand is unseeded and its
// sky no longer exists) plus sphere_roughness's rect light: Lambertian + mirror + glass spheres
void build_make_world(rth_scene& s, const rth_params& p) {
    double from[3] = {13., 2., 3.}, to[3] = {0., 0., 0.};
    camera(s, from, to, aspect_of(p), 20., 0.16 * 0., 10., 0., 0.);      // aperture 0: the lens path draws but does not move rays
    Gen g(p.seed ? p.seed : 1);
    push_solid(s, 201. / 256., 201. / 256., 201. / 256.);
    push_solid(s, 51. / 256., 75.5 / 256., 25.5 / 256.);
    push_checkered(s, 0, 1, 4.);
    push_matte(s, 2, 0.);
    push_prim(s, make_sphere(0., -10000., 0., 10000., 0));
    for (int a = -11; a < 11; a++) for (int b = -11; b < 11; b++) {
        double cx = a + 0.9 * g.uniform(), cz = b + 0.9 * g.uniform();
        double dx = cx - 4., dz = cz;
        if (std::sqrt(dx * dx + dz * dz) <= 0.9) continue;
        double mat = g.uniform();
        uint32_t m;
        if (mat < 0.6) { uint32_t t = push_solid(s, 255. * g.uniform() * g.range(0.4, 0.8) / 256., 255. * g.uniform() * g.range(0.4, 0.8) / 256., 255. * g.uniform() * g.range(0.4, 0.8) / 256.); m = push_matte(s, t, 0.); }
        else if (mat < 0.86) { uint32_t t = push_solid(s, g.range(30., 200.) / 256., g.range(30., 200.) / 256., g.range(30., 200.) / 256.); m = push_mirror(s, t); }
        else { uint32_t t = push_solid(s, g.range(0.5, 1.), g.range(0.5, 1.), g.range(0.5, 1.)); m = push_glass(s, t, t, 0., 0., g.range(1.2, 1.8), true); }
        push_prim(s, make_sphere(cx, 0.2, cz, 0.2, m));
    }
    { uint32_t t = push_solid(s, 0.5, 0.5, 0.5); push_prim(s, make_sphere(-4., 1., 0., 1., push_matte(s, t, 45.))); }   // Perlin texture is out of scope: grey Oren–Nayar
    { uint32_t t = push_solid(s, 1., 1., 1.); push_prim(s, make_sphere(0., 1., 0., 1., push_glass(s, t, t, 0., 0., 1.3, true))); }
    { uint32_t e = push_solid(s, 223.5 / 256., 85. / 256., 32.5 / 256.), k = push_solid(s, 0., 0., 0.), r = push_solid(s, 0.02, 0., 0.);
      push_prim(s, make_sphere(4., 1., 0., 1., push_metal(s, e, k, r, r, r, true))); }
    uint32_t lt = push_solid(s, 15., 15., 15.);
    rt_primitive light = make_rect(RT_PRIM_XZ_RECT, -10., -10., 10., 10., 50., push_light_mat(s, lt));
    light.light_index = 0;
    uint64_t li = push_prim(s, flip(light));
    push_diffuse_light(s, li, 10., 10., 10., false);
    s.png = "sphere_world.png";
}

void cornell_walls(rth_scene& s, uint32_t green, uint32_t red, uint32_t white_m, uint32_t light_mat, bool flip_light) {
    push_prim(s, flip(make_rect(RT_PRIM_YZ_RECT, 0., 0., 555., 555., 555., green)));
    push_prim(s, make_rect(RT_PRIM_YZ_RECT, 0., 0., 555., 555., 0., red));
    rt_primitive light = make_rect(RT_PRIM_XZ_RECT, 213., 227., 343., 332., 554.9, light_mat);
    light.light_index = 0;
    push_prim(s, flip_light ? flip(light) : light);
    push_prim(s, make_rect(RT_PRIM_XZ_RECT, 0., 0., 555., 555., 0., white_m));
    push_prim(s, flip(make_rect(RT_PRIM_XZ_RECT, 0., 0., 555., 555., 555., white_m)));
    push_prim(s, flip(make_rect(RT_PRIM_XY_RECT, 0., 0., 555., 555., 555., white_m)));
}
// Cube::get_sides (hittable.rs:787-846): z0, z1, y0, y1, x0, x1
void push_cube(rth_scene& s, const double mn[3], const double mx[3], uint32_t mat, uint32_t xf) {
    push_prim(s, flip(make_rect(RT_PRIM_XY_RECT, mn[0], mn[1], mx[0], mx[1], mn[2], mat, xf)));
    push_prim(s, make_rect(RT_PRIM_XY_RECT, mn[0], mn[1], mx[0], mx[1], mx[2], mat, xf));
    push_prim(s, flip(make_rect(RT_PRIM_XZ_RECT, mn[0], mn[2], mx[0], mx[2], mn[1], mat, xf)));
    push_prim(s, make_rect(RT_PRIM_XZ_RECT, mn[0], mn[2], mx[0], mx[2], mx[1], mat, xf));
    push_prim(s, flip(make_rect(RT_PRIM_YZ_RECT, mn[1], mn[2], mx[1], mx[2], mn[0], mat, xf)));
    push_prim(s, make_rect(RT_PRIM_YZ_RECT, mn[1], mn[2], mx[1], mx[2], mx[0], mat, xf));
}
// scenes.rs:89-197
void build_cornell_box(rth_scene& s, const rth_params& p) {
    double from[3] = {278., 278., -800.}, to[3] = {278., 278., 0.};
    camera(s, from, to, aspect_of(p), 40., 0., 10., 0., 1.);
    push_solid(s, 0.65, 0.05, 0.05); push_solid(s, 0.73, 0.73, 0.73); push_solid(s, 0.12, 0.45, 0.15); push_solid(s, 28.0, 28.0, 28.0);
    push_matte(s, 0, 0.); push_matte(s, 1, 0.); push_matte(s, 2, 0.); push_light_mat(s, 3);
    cornell_walls(s, 2, 0, 1, 3, true);
    push_diffuse_light(s, 2, 15., 15., 15., false);
    double t1[16], t2[16], r1[16], r2[16], first[16], second[16];
    mat_translate(265., 0., 295., t1); mat_translate(130., 0., 65., t2);
    mat_euler(0., 15. * PI / 180., 0., r1); mat_euler(0., -18. * PI / 180., 0., r2);
    mat_mul(t1, r1, first); mat_mul(t2, r2, second);
    double z[3] = {0., 0., 0.}, c1[3] = {165., 165., 165.}, c2[3] = {165., 330., 165.};
    uint32_t xs = push_xform(s, second); push_cube(s, z, c1, 1, xs);
    uint32_t xf = push_xform(s, first); push_cube(s, z, c2, 1, xf);
    s.png = "cornell_box.png";
}
// scenes.rs:200-307
void build_cornell_box_statue(rth_scene& s, const rth_params& p) {
    double from[3] = {278., 278., -800.}, to[3] = {278., 278., 0.};
    camera(s, from, to, aspect_of(p), 40., 0., 10., 0., 1.);
    push_solid(s, 0.65, 0.05, 0.05); push_solid(s, 0.73, 0.73, 0.73); push_solid(s, 0.12, 0.45, 0.15);
    push_solid(s, 1., 1., 1.); push_solid(s, 0.3, 0.3, 0.3); push_solid(s, 0.01, 0., 0.);
    double sc = 0.2 / 0.3765;
    push_solid(s, 0.1514 * sc, 0.0139 * sc, 0.3765 * sc); push_solid(s, 1.13, 1.13, 1.13);
    push_matte(s, 1, 0.); push_matte(s, 0, 0.); push_matte(s, 2, 0.);
    if (p.variant == 1) push_metal(s, 5, 3, 5, 5, 5, true);               // scenes.rs:244-246 as committed
    else push_matte(s, 1, 0.);                                            // the commented matte alternative (scenes.rs:243), BASELINE config 2 "Lambertian"
    cornell_walls(s, 2, 1, 0, 0, false);
    push_diffuse_light(s, 2, 0.97 * 25., 0.92 * 25., 0.23 * 25., true);
    rt::HostMesh m; bool loaded = false;
    std::string path = p.data_dir ? std::string(p.data_dir) + "/statue.obj" : std::string();
    if (!path.empty() && file_exists(path)) {
        double tr[16], r1[16], scl[16], tmp[16], xf[16];
        mat_translate(374., 435., 130., tr); mat_euler(0., 0., PI, r1); mat_scale(0.86, scl);
        mat_mul(tr, r1, tmp); mat_mul(tmp, scl, xf);
        std::string err; loaded = load_obj(path.c_str(), xf, m, err);
    }
    if (!loaded) { statue_standin(m, p.seed ? p.seed : 11, p.n_tris); s.standin = true; }
    s.meshes.push_back(std::move(m));
    push_mesh_triangles(s, 0, 3);
    s.png = "cornell_statue.png";
}
// scenes.rs:310-375 (variant 0) and the commented metal_dragon (scenes.rs:424-471, variant 1) with the
// two_dragons metal parameters (scenes.rs:582-606) and plastic_dragon's light
void build_plastic_dragon(rth_scene& s, const rth_params& p) {
    uint32_t mesh = add_dragon(s, p, 0.);
    double radians = 5. * (PI / 180.), r = std::sqrt(82.26);
    double from[3] = {r * std::sin(radians + PI / 4.4), 4., r * std::cos(radians + PI / 4.4)}, to[3] = {0., -0.15, -0.08};
    camera(s, from, to, aspect_of(p), 70., 0.0, 10., 0., 0.);
    uint32_t t0 = push_solid(s, 0.4 * 2., 0.15 * 2., 0.15 * 2.), t1 = push_solid(s, 0.15 * 2., 0.15 * 2., 0.4 * 2.);
    push_checkered(s, t0, t1, p.variant == 1 ? 0.1 : 10000.);                                     // 2
    uint32_t dragon_mat;
    if (p.variant == 1) {
        uint32_t eta = push_solid(s, 0.05, 0.5, 0.75), k = push_solid(s, 0., 0., 0.), rough = push_solid(s, 0.1, 0., 0.);
        push_matte(s, 2, 0.);
        dragon_mat = push_metal(s, eta, k, rough, rough, rough, true);
    } else {
        double sc = 0.56 / 0.3765;
        push_solid(s, 0.1514 * sc, 0.0139 * sc, 0.3765 * sc);                                     // 3 purple
        push_solid(s, 1., 1., 1.);                                                                // 4 spec
        push_matte(s, 2, 0.);
        dragon_mat = push_plastic(s, 3, 4, 0.001, true);
    }
    push_prim(s, make_rect(RT_PRIM_XZ_RECT, -10000., -10000., 10000., 10000., -2.83, 0));
    push_mesh_triangles(s, mesh, dragon_mat);
    rt_primitive light = make_rect(RT_PRIM_XZ_RECT, -5., -5., 5., 5., 15., 0);
    light.light_index = 0;
    uint64_t li = push_prim(s, flip(light));
    push_diffuse_light(s, li, 4., 4., 4., false);
    s.png = p.variant == 1 ? "metal_dragon.png" : "plastic_dragon.png";
}
// scenes.rs:549-624; variant 1 enables the commented block :608-611 (glass dragon = mesh 0, material 2)
void build_two_dragons(rth_scene& s, const rth_params& p) {
    uint32_t m0 = add_dragon(s, p, 0.), m1 = add_dragon(s, p, 5.);
    double from[3] = {-8.5, 5., 0.}, to[3] = {0., -0.15, -0.08};
    camera(s, from, to, aspect_of(p), 60., 0.0, 10., 0., 0.);
    uint32_t t0 = push_solid(s, 0.4 * 2., 0.15 * 2., 0.15 * 2.), t1 = push_solid(s, 0.15 * 2., 0.15 * 2., 0.4 * 2.);
    push_checkered(s, t0, t1, 0.1);                       // 2
    push_solid(s, 15., 15., 15.);                         // 3
    push_matte(s, 2, 0.);                                 // mat 0
    push_prim(s, make_rect(RT_PRIM_XZ_RECT, -10000., -10000., 10000., 10000., -2.83, 0));
    uint32_t lm = push_light_mat(s, 3);                   // mat 1
    rt_primitive light = make_rect(RT_PRIM_XZ_RECT, -10., -10., 10., 10., 50., lm);
    light.light_index = 0;
    push_prim(s, light);                                  // prim 1, not flipped
    push_diffuse_light(s, 1, 12., 12., 12., false);
    push_solid(s, 1., 1., 1.); push_solid(s, 1., 1., 1.); // 4, 5
    push_solid(s, 0.05, 0.5, 0.75); push_solid(s, 0., 0., 0.); push_solid(s, 0.1, 0., 0.);   // 6 eta, 7 k, 8 rough
    push_glass(s, 4, 5, 0.0, 0.0, 1.5, true);             // mat 2
    push_metal(s, 6, 7, 8, 8, 8, true);                   // mat 3
    if (p.variant == 1) push_mesh_triangles(s, m0, 2);
    push_mesh_triangles(s, m1, 3);
    s.png = "two_dragons.png";
}
// ---- Radiance .hdr (RGBE) reader: what Texture::new_hdr gets from image 0.23's HdrDecoder::read_image_hdr
// (material.rs:626-642): header lines up to a blank line, a "-Y H +X W" resolution line, then scanlines that are either
// new-style RLE (2 2 hi lo, four component planes of runs / literals) or flat RGBE quadruples; each pixel becomes
// Rgb<f32> = c * 2^(e-136), or 0 when e == 0 (`Rgbe8Pixel::to_hdr`).
bool load_hdr(const char* path, rth_scene::Image& img, std::string& err) {
    FILE* f = fopen(path, "rb");
    if (!f) { err = std::string("cannot open ") + path; return false; }
    char line[512]; bool magic = false; int W = 0, H = 0;
    while (fgets(line, sizeof(line), f)) {
        if (!magic) { if (strncmp(line, "#?", 2) != 0) { err = "not a Radiance HDR file"; fclose(f); return false; } magic = true; continue; }
        if (line[0] == '\n' || (line[0] == '\r' && line[1] == '\n')) break;
    }
    if (!fgets(line, sizeof(line), f) || sscanf(line, "-Y %d +X %d", &H, &W) != 2 || W <= 0 || H <= 0) { err = "unsupported HDR orientation"; fclose(f); return false; }
    std::vector<uint8_t> rgbe((size_t)W * H * 4), scan((size_t)W * 4);
    for (int y = 0; y < H; y++) {
        uint8_t* row = &rgbe[(size_t)y * W * 4];
        uint8_t h4[4];
        if (fread(h4, 1, 4, f) != 4) { err = "truncated HDR"; fclose(f); return false; }
        if (W >= 8 && W < 32768 && h4[0] == 2 && h4[1] == 2 && ((h4[2] << 8) | h4[3]) == W) {
            for (int c = 0; c < 4; c++) {
                int x = 0;
                while (x < W) {
                    int n = fgetc(f); if (n == EOF) { err = "truncated HDR"; fclose(f); return false; }
                    if (n > 128) { n -= 128; int v = fgetc(f); if (v == EOF || x + n > W) { err = "bad HDR run"; fclose(f); return false; } for (int k = 0; k < n; k++) scan[(size_t)c * W + x++] = (uint8_t)v; }
                    else { if (n == 0 || x + n > W || fread(&scan[(size_t)c * W + x], 1, n, f) != (size_t)n) { err = "bad HDR literal"; fclose(f); return false; } x += n; }
                }
            }
            for (int x = 0; x < W; x++) for (int c = 0; c < 4; c++) row[4 * x + c] = scan[(size_t)c * W + x];
        } else {
            memcpy(row, h4, 4);
            if (W > 1 && fread(row + 4, 1, (size_t)(W - 1) * 4, f) != (size_t)(W - 1) * 4) { err = "truncated HDR"; fclose(f); return false; }
        }
    }
    fclose(f);
    img.w = (uint32_t)W; img.h = (uint32_t)H; img.hdr = true; img.f.resize((size_t)W * H * 3);
    for (size_t i = 0; i < (size_t)W * H; i++) {
        uint8_t e = rgbe[4 * i + 3];
        float sc = e == 0 ? 0.0f : std::exp2((float)e - 136.0f);
        for (int c = 0; c < 3; c++) img.f[3 * i + c] = e == 0 ? 0.0f : sc * (float)rgbe[4 * i + c];
    }
    return true;
}
// deterministic stand-in for data/material/textures/envmap.hdr when the file is not there (the GPU box has no reference
// checkout): a sky gradient with a warm horizon band, a small very bright sun and a dim ground, W x H = 256 x 128,
// snapped to RGBE8-representable f32 values like a decoded file would be.
void envmap_standin(rth_scene::Image& img, uint32_t W = 256, uint32_t H = 128) {
    img.w = W; img.h = H; img.hdr = true; img.f.resize((size_t)W * H * 3);
    for (uint32_t y = 0; y < H; y++) for (uint32_t x = 0; x < W; x++) {
        double v = (y + 0.5) / H, u = (x + 0.5) / W;
        double up = std::cos(v * 3.14159265358979);                     // +1 zenith .. -1 nadir
        double c[3];
        if (up >= 0.) { double t = std::pow(1. - up, 3.); c[0] = 0.25 + 0.9 * t; c[1] = 0.45 + 0.6 * t; c[2] = 0.9 + 0.1 * t; }
        else { double t = std::pow(1. + up, 6.); c[0] = 0.08 + 0.5 * t; c[1] = 0.07 + 0.4 * t; c[2] = 0.06 + 0.3 * t; }
        double du = (u - 0.30) * 2., dv = (v - 0.28) * 1.;                // sun
        double d2 = du * du + dv * dv;
        double sun = 60. * std::exp(-d2 / (2. * 0.02 * 0.02)) + 2. * std::exp(-d2 / (2. * 0.08 * 0.08));
        c[0] += sun; c[1] += sun * 0.92; c[2] += sun * 0.8;
        double mx = std::fmax(c[0], std::fmax(c[1], c[2]));
        int e; std::frexp(mx, &e);
        for (int k = 0; k < 3; k++) { double q = std::floor(std::ldexp(c[k], 8 - e)); img.f[3 * ((size_t)y * W + x) + k] = (float)std::ldexp(q, e - 8); }
    }
}
uint32_t add_envmap(rth_scene& s, const rth_params& p, const char* rel) {
    rth_scene::Image img; bool loaded = false;
    if (p.data_dir) { std::string path = std::string(p.data_dir) + "/" + rel, err; if (file_exists(path)) loaded = load_hdr(path.c_str(), img, err); }
    if (!loaded) { envmap_standin(img); s.standin = true; }
    s.images.push_back(std::move(img));
    return (uint32_t)s.images.size() - 1;
}
// an OBJ of the reference's data/ directory under `xf`, or a blob stand-in in [lo,hi]
uint32_t add_obj_or_blob(rth_scene& s, const rth_params& p, const char* rel, const double xf[16], uint32_t nu, uint32_t nv, uint64_t seed, const double lo[3], const double hi[3]) {
    rt::HostMesh m; bool loaded = false;
    if (p.data_dir) { std::string path = std::string(p.data_dir) + "/" + rel, err; if (file_exists(path)) loaded = load_obj(path.c_str(), xf, m, err); }
    if (!loaded) { append_blob(m, nu, nv, seed, lo, hi); s.standin = true; }
    s.meshes.push_back(std::move(m));
    return (uint32_t)s.meshes.size() - 1;
}
void camera_up(rth_scene& s, const double from[3], const double to[3], const double up[3], double aspect, double vfov) {
    rt::camera_look_at(from, to, up, aspect, vfov, 0.0, 10., 0., 0., &s.camera);     // Camera::new → t0 = t1 = 0
}
// scenes.rs:627-741 `material_hdr(mat_num)`: the material-test ball under an HDR environment (Light::Infinite), variant =
// mat_num: 0 smooth_plastic (:824), 1 rosegold_metal (:839), 2 mirror (:876), 3 glass (:857); 4 = disney_mat (:867-904), which the
// reference defines next to the other four but never wires into the match (its `_ =>` arm is glass)
void build_material_hdr(rth_scene& s, const rth_params& p) {
    const double from[3] = {3.04068, 3.17153, 3.20454}, dir[3] = {-0.583445, -0.538765, -0.60772}, up[3] = {-0.373123, 0.842456, -0.388647};
    const double to[3] = {from[0] + dir[0], from[1] + dir[1], from[2] + dir[2]};
    camera_up(s, from, to, up, aspect_of(p), 20.);
    uint32_t env = push_hdr_texture(s, add_envmap(s, p, "material/textures/envmap.hdr"));     // texts[0]
    push_infinite_light(s, env);
    auto sim = [](double sc, double tx, double ty, double tz, double m[16]) { mat_identity(m); m[0] = m[5] = m[10] = sc; m[3] = tx; m[7] = ty; m[11] = tz; };
    double m1[16], m2[16], m0[16];
    sim(0.482906, 0.0571719, 0.213656, 0.0682078, m1); sim(0.482906, 0.156382, 0.777229, 0.161698, m2); sim(0.482906, 0.110507, 0.494301, 0.126194, m0);
    const double rect_m[16] = {-1.88298, 1.9602, 2.50299e-007, -0.708772, -2.37623e-007, 1.18811e-007, -2.71809, 0., -1.9602, -1.88298, 8.90586e-008, -0.732108, 0., 0., 0., 1.};
    // stand-in boxes: a base, a cap and the ball between them, where the transformed reference meshes sit
    const double lo1[3] = {-0.38, 0.005, -0.37}, hi1[3] = {0.50, 0.42, 0.51}, lo2[3] = {-0.12, 0.60, -0.11}, hi2[3] = {0.43, 0.95, 0.44}, lo0[3] = {-0.30, 0.10, -0.28}, hi0[3] = {0.52, 0.90, 0.53};
    uint32_t mesh1 = add_obj_or_blob(s, p, "material/models/Mesh001.obj", m1, 96, 95, 11, lo1, hi1);
    uint32_t mesh2 = add_obj_or_blob(s, p, "material/models/Mesh002.obj", m2, 64, 63, 12, lo2, hi2);
    uint32_t mesh0 = add_obj_or_blob(s, p, "material/models/Mesh000.obj", m0, 84, 83, 13, lo0, hi0);
    (void)mesh2;
    switch (p.variant) {
        case 0: { uint32_t a = push_solid(s, 0.1608, 0.0014767, 0.4), b = push_solid(s, 1., 1., 1.); push_plastic(s, a, b, 0.002, false); break; }
        case 1: { uint32_t a = push_solid(s, 1. - 0.718, 1. - 0.431, 1. - 0.475), b = push_solid(s, 1., 1., 1.), c = push_solid(s, 0.002, 0., 0.); push_metal(s, a, b, c, c, c, true); break; }
        case 2: { uint32_t a = push_solid(s, 1., 1., 1.); push_mirror(s, a); break; }
        case 4: {   // disney_mat (scenes.rs:867-904): 14 solid textures in declaration order, thin = false
            const uint32_t c = (uint32_t)s.textures.size();
            push_solid(s, 0.1608, 0.0014767, 0.4); push_solid(s, 0., 0., 0.); push_solid(s, 1.5, 0., 0.); push_solid(s, 0.5, 0., 0.); push_solid(s, 0., 0., 0.);
            push_solid(s, 0., 0., 0.); push_solid(s, 0., 0., 0.); push_solid(s, 0.5, 0., 0.); push_solid(s, 0., 0., 0.); push_solid(s, 1., 0., 0.);
            push_solid(s, 0., 0., 0.); push_solid(s, 0., 0., 0.); push_solid(s, 0., 0., 0.); push_solid(s, 1., 0., 0.);
            rt_disney d{}; uint32_t* ids = &d.color; for (uint32_t k = 0; k < 14; k++) ids[k] = c + k;      // color .. clearcoat_gloss, spec_trans, scatter_distance, flatness, diff_trans
            d.thin = 0;
            s.disneys.push_back(d);
            rt_material m{}; m.type = RT_MAT_DISNEY; for (int k = 0; k < 5; k++) m.tex[k] = RT_NONE; m.tex[0] = (uint32_t)s.disneys.size() - 1;
            s.materials.push_back(m);
            break;
        }
        default: { push_solid(s, 1. - 0.718, 1. - 0.431, 1. - 0.475); uint32_t b = push_solid(s, 1., 1., 1.); push_solid(s, 0.002, 0., 0.); push_glass(s, b, b, 0.01, 0.01, 1.5, true); break; }
    }
    uint32_t len = (uint32_t)s.textures.size();
    push_solid(s, 0.2, 0.2, 0.2); push_solid(s, 0.325, 0.31, 0.325); push_solid(s, 0.725, 0.71, 0.68);
    push_checkered(s, len + 1, len + 2, 10.);
    push_matte(s, len, 0.);          // material 1
    push_matte(s, len + 3, 0.);      // material 2
    push_mesh_triangles(s, 0, 0);    // generate_triangles(meshes, 0, 0)
    push_mesh_triangles(s, 1, 0);    // generate_triangles(meshes, 1, 0)
    push_mesh_triangles(s, 2, 1);    // generate_triangles(meshes, 2, 1)
    (void)mesh1; (void)mesh0;
    push_prim(s, make_rect(RT_PRIM_XY_RECT, -1., -1., 1., 1., 0., 2, push_xform(s, rect_m)));
    s.png = "material.png";
}

// scenes.rs:744-811 `teapot_hdr()`: a nearly smooth white plastic teapot (lid = Mesh000, body = Mesh001; stand-in blobs where the
// OBJ files are missing) on a checkered floor — an XY rect under the Tungsten floor transform — lit only by the HDR environment.
void build_teapot_hdr(rth_scene& s, const rth_params& p) {
    const double from[3] = {23.895, 11.2207, 0.0400773}, dir[3] = {-0.939631, -0.342149, -0.00519335}, up[3] = {-0.342144, 0.939646, -0.00189103};
    const double to[3] = {from[0] + dir[0], from[1] + dir[1], from[2] + dir[2]};
    camera_up(s, from, to, up, aspect_of(p), 35. / 2.);
    uint32_t env = push_hdr_texture(s, add_envmap(s, p, "teapot/textures/envmap.hdr"));     // texts[0]
    push_infinite_light(s, env);
    double ident[16]; mat_identity(ident);
    const double floor_m[16] = {-39.9766, 39.9766, -1.74743e-006, 0., 4.94249e-006, 2.47125e-006, -56.5355, 0., -39.9766, -39.9766, -5.2423e-006, 0., 0., 0., 0., 1.};
    const double lo_lid[3] = {-1.35, 2.36, -1.35}, hi_lid[3] = {1.35, 3.15, 1.35}, lo_body[3] = {-3.0, 0.0, -2.0}, hi_body[3] = {3.43, 2.45, 2.0};
    add_obj_or_blob(s, p, "teapot/models/Mesh000.obj", ident, 64, 63, 21, lo_lid, hi_lid);
    add_obj_or_blob(s, p, "teapot/models/Mesh001.obj", ident, 128, 127, 22, lo_body, hi_body);
    uint32_t len = (uint32_t)s.textures.size();
    push_solid(s, 0.9, 0.9, 0.9); push_solid(s, 1., 1., 1.);
    push_plastic(s, len, len + 1, 0.00001, true);      // material 0
    len = (uint32_t)s.textures.size();
    push_solid(s, 0.2, 0.2, 0.2); push_solid(s, 0.325, 0.31, 0.325); push_solid(s, 0.725, 0.71, 0.68);
    push_checkered(s, len + 1, len + 2, 10.);
    push_matte(s, len, 0.);          // material 1
    push_matte(s, len + 3, 0.);      // material 2
    push_mesh_triangles(s, 0, 0);
    push_mesh_triangles(s, 1, 0);
    push_prim(s, make_rect(RT_PRIM_XY_RECT, -1., -1., 1., 1., 0., 2, push_xform(s, floor_m)));
    s.png = "teapot_hdr.png";
}

// triangle soup; centre ~ U[-1,1]^3, edges ~ U[-s,s]^3, s = 2 N^(-1/3), f32-exact values
void build_soup(rth_scene& s, const rth_params& p) {
    uint64_t n = p.n_tris ? p.n_tris : 1000000;
    double from[3] = {0., 0., -4.}, to[3] = {0., 0., 0.};
    camera(s, from, to, aspect_of(p), 40., 0.0, 10., 0., 0.);
    int bucket = 0; while ((1ull << (bucket + 1)) <= n) bucket++;
    uint64_t key = 0x5EED0000ull + (uint64_t)bucket + (p.seed << 32);
    double sz = 2. * std::pow((double)n, -1. / 3.);
    rt::HostMesh m; m.p.resize(9 * n); m.ind.resize(3 * n);
    rt::par_for(n, 1 << 16, [&](size_t tb, size_t te) {       // counter-based generator: triangle i depends on i alone
    for (uint64_t i = tb; i < te; i++) {
        uint32_t c0[4] = {(uint32_t)i, (uint32_t)(i >> 32), 0, 0}, c1[4] = {(uint32_t)i, (uint32_t)(i >> 32), 1, 0}, c2[4] = {(uint32_t)i, (uint32_t)(i >> 32), 2, 0};
        philox(c0, (uint32_t)key, (uint32_t)(key >> 32)); philox(c1, (uint32_t)key, (uint32_t)(key >> 32)); philox(c2, (uint32_t)key, (uint32_t)(key >> 32));
        auto u = [](uint32_t x) { return (double)(x >> 8) * (1.0 / 16777216.0); };
        double c[3] = {2. * u(c0[0]) - 1., 2. * u(c0[1]) - 1., 2. * u(c0[2]) - 1.};
        double e1[3] = {sz * (2. * u(c1[0]) - 1.), sz * (2. * u(c1[1]) - 1.), sz * (2. * u(c1[2]) - 1.)};
        double e2[3] = {sz * (2. * u(c2[0]) - 1.), sz * (2. * u(c2[1]) - 1.), sz * (2. * u(c2[2]) - 1.)};
        for (int k = 0; k < 3; k++) {
            m.p[9 * i + k] = (double)(float)c[k]; m.p[9 * i + 3 + k] = (double)(float)(c[k] + e1[k]); m.p[9 * i + 6 + k] = (double)(float)(c[k] + e2[k]);
        }
        m.ind[3 * i] = (uint32_t)(3 * i); m.ind[3 * i + 1] = (uint32_t)(3 * i + 1); m.ind[3 * i + 2] = (uint32_t)(3 * i + 2);
    }
    });
    s.meshes.push_back(std::move(m));
    uint32_t t = push_solid(s, 0.7, 0.7, 0.7);
    push_mesh_triangles(s, 0, push_matte(s, t, 0.));
    s.png = "soup.png";
}

}  // namespace

extern "C" {

rt_status rth_scene_create_empty(rth_scene** out) { if (!out) return RT_ERR_INVALID; *out = new rth_scene(); return RT_OK; }
rt_status rth_scene_build(const char* name, const rth_params* params, rth_scene** out) {
    if (!name || !params || !out) { host_error("rth_scene_build: null argument"); return RT_ERR_INVALID; }
    if (params->width == 0 || params->height == 0) { host_error("rth_scene_build: zero image size"); return RT_ERR_INVALID; }
    if (params->n_tris > 170000000ull) { host_error("rth_scene_build: n_tris too large"); return RT_ERR_INVALID; }
    rth_scene* s = new rth_scene();
    std::string n(name);
    if (n == "sphere_roughness") build_sphere_roughness(*s, *params);
    else if (n == "make_world") build_make_world(*s, *params);
    else if (n == "cornell_box") build_cornell_box(*s, *params);
    else if (n == "cornell_box_statue") build_cornell_box_statue(*s, *params);
    else if (n == "plastic_dragon") build_plastic_dragon(*s, *params);
    else if (n == "two_dragons") build_two_dragons(*s, *params);
    else if (n == "soup") build_soup(*s, *params);
    else if (n == "material_hdr") build_material_hdr(*s, *params);
    else if (n == "teapot_hdr") build_teapot_hdr(*s, *params);
    else { delete s; host_error("rth_scene_build: unknown scene '" + n + "'"); return RT_ERR_INVALID; }
    *out = s;
    return RT_OK;
}
void rth_scene_destroy(rth_scene* s) { delete s; }
rt_status rth_scene_camera(const rth_scene* s, rt_camera* out) { if (!s || !out) return RT_ERR_INVALID; *out = s->camera; return RT_OK; }
const char* rth_scene_png_name(const rth_scene* s) { return s ? s->png.c_str() : ""; }
int32_t rth_scene_uses_standin(const rth_scene* s) { return s && s->standin ? 1 : 0; }
rt_status rth_scene_counts(const rth_scene* s, uint64_t c[6]) {
    if (!s || !c) return RT_ERR_INVALID;
    c[0] = s->meshes.size(); c[1] = s->n_prims(); c[2] = s->xforms.size() / 16; c[3] = s->textures.size(); c[4] = s->materials.size(); c[5] = s->lights.size();
    return RT_OK;
}
rt_status rth_scene_mesh_sizes(const rth_scene* s, uint32_t mesh, uint64_t sz[4]) {
    if (!s || !sz || mesh >= s->meshes.size()) return RT_ERR_INVALID;
    const rt::HostMesh& m = s->meshes[mesh]; sz[0] = m.p.size() / 3; sz[1] = m.n.size() / 3; sz[2] = m.uv.size() / 2; sz[3] = m.ind.size();
    return RT_OK;
}
rt_status rth_scene_mesh_copy(const rth_scene* s, uint32_t mesh, double* p, double* n, double* uv, uint32_t* ind) {
    if (!s || mesh >= s->meshes.size()) return RT_ERR_INVALID;
    const rt::HostMesh& m = s->meshes[mesh];
    if (p) memcpy(p, m.p.data(), m.p.size() * 8); if (n) memcpy(n, m.n.data(), m.n.size() * 8);
    if (uv) memcpy(uv, m.uv.data(), m.uv.size() * 8); if (ind) memcpy(ind, m.ind.data(), m.ind.size() * 4);
    return RT_OK;
}
rt_status rth_scene_copy_primitives(const rth_scene* s, rt_primitive* out) {
    if (!s || !out) return RT_ERR_INVALID;
    uint64_t k = 0;
    for (auto& it : s->items) {
        if (!it.run) { out[k++] = it.prim; continue; }
        uint64_t nt = s->meshes[it.mesh].ind.size() / 3;
        for (uint64_t t = 0; t < nt; t++) { rt_primitive p = prim_base(RT_PRIM_TRIANGLE, it.mat); p.mesh_index = it.mesh; p.ind = (uint32_t)(3 * t); out[k++] = p; }
    }
    return RT_OK;
}
rt_status rth_scene_copy_transforms(const rth_scene* s, double* out) { if (!s || !out) return RT_ERR_INVALID; memcpy(out, s->xforms.data(), s->xforms.size() * 8); return RT_OK; }
rt_status rth_scene_copy_textures(const rth_scene* s, rt_texture* out) { if (!s || !out) return RT_ERR_INVALID; memcpy(out, s->textures.data(), s->textures.size() * sizeof(rt_texture)); return RT_OK; }
rt_status rth_scene_copy_materials(const rth_scene* s, rt_material* out) { if (!s || !out) return RT_ERR_INVALID; memcpy(out, s->materials.data(), s->materials.size() * sizeof(rt_material)); return RT_OK; }
rt_status rth_scene_copy_lights(const rth_scene* s, rt_light* out) { if (!s || !out) return RT_ERR_INVALID; memcpy(out, s->lights.data(), s->lights.size() * sizeof(rt_light)); return RT_OK; }

rt_status rth_scene_num_disney(const rth_scene* s, uint64_t* n) { if (!s || !n) return RT_ERR_INVALID; *n = s->disneys.size(); return RT_OK; }
rt_status rth_scene_copy_disney(const rth_scene* s, rt_disney* out) { if (!s || !out) return RT_ERR_INVALID; memcpy(out, s->disneys.data(), s->disneys.size() * sizeof(rt_disney)); return RT_OK; }
rt_status rth_scene_num_images(const rth_scene* s, uint64_t* n) { if (!s || !n) return RT_ERR_INVALID; *n = s->images.size(); return RT_OK; }
rt_status rth_scene_image_info(const rth_scene* s, uint32_t i, uint32_t info[3]) {
    if (!s || !info || i >= s->images.size()) return RT_ERR_INVALID;
    info[0] = s->images[i].w; info[1] = s->images[i].h; info[2] = s->images[i].hdr ? 1u : 0u; return RT_OK;
}
rt_status rth_scene_image_copy(const rth_scene* s, uint32_t i, void* out) {
    if (!s || !out || i >= s->images.size()) return RT_ERR_INVALID;
    const auto& im = s->images[i];
    if (im.hdr) memcpy(out, im.f.data(), im.f.size() * sizeof(float)); else memcpy(out, im.b.data(), im.b.size());
    return RT_OK;
}
rt_status rth_load_hdr(const char* path, uint32_t* w, uint32_t* h, float* rgb_out, uint64_t capacity_floats) {
    if (!path || !w || !h) { host_error("rth_load_hdr: null argument"); return RT_ERR_INVALID; }
    rth_scene::Image img; std::string err;
    if (!load_hdr(path, img, err)) { host_error("rth_load_hdr: " + err); return RT_ERR_IO; }
    *w = img.w; *h = img.h;
    if (rgb_out) { if (capacity_floats < img.f.size()) { host_error("rth_load_hdr: buffer too small"); return RT_ERR_INVALID; } memcpy(rgb_out, img.f.data(), img.f.size() * sizeof(float)); }
    return RT_OK;
}

// The core library's scene setters, bound at the first rth_scene_upload: librthost.so has NO link-time dependency on librtcore.so (the CPU
// reference arm of bench.py builds its scene tables with this library and must not map the CUDA one).  A caller that holds an rt_scene* has
// librtcore.so in the process already; it is looked up next to this library (RTLD_NOLOAD first).
namespace {
struct CoreApi {
    rt_status (*add_image_hdr)(rt_scene*, uint32_t, uint32_t, const float*, uint32_t*) = nullptr;
    rt_status (*add_image_rgb8)(rt_scene*, uint32_t, uint32_t, const uint8_t*, uint32_t*) = nullptr;
    rt_status (*add_mesh)(rt_scene*, uint64_t, const double*, uint64_t, const double*, uint64_t, const double*, uint64_t, const uint32_t*, uint32_t*) = nullptr;
    rt_status (*add_transform)(rt_scene*, const double*, uint32_t*) = nullptr;
    rt_status (*set_textures)(rt_scene*, uint64_t, const rt_texture*) = nullptr;
    rt_status (*set_materials)(rt_scene*, uint64_t, const rt_material*) = nullptr;
    rt_status (*set_lights)(rt_scene*, uint64_t, const rt_light*) = nullptr;
    rt_status (*add_mesh_triangles)(rt_scene*, uint32_t, uint32_t, uint32_t*) = nullptr;
    rt_status (*add_primitives)(rt_scene*, uint64_t, const rt_primitive*, uint32_t*) = nullptr;
    rt_status (*add_disney)(rt_scene*, const rt_disney*, uint32_t*) = nullptr;
    bool ok = false;
};
const CoreApi& core_api() {
    static CoreApi api;
    static std::once_flag once;
    std::call_once(once, []() {
        std::string path = "librtcore.so";
        if (const char* e = getenv("RT_LIBRTCORE")) path = e;
        else { Dl_info info; if (dladdr((const void*)&core_api, &info) && info.dli_fname) { std::string me = info.dli_fname; size_t sl = me.rfind('/'); if (sl != std::string::npos) path = me.substr(0, sl + 1) + "librtcore.so"; } }
        void* h = dlopen(path.c_str(), RTLD_NOW | RTLD_NOLOAD);
        if (!h) h = dlopen(path.c_str(), RTLD_NOW | RTLD_LOCAL);
        if (!h) { host_error(std::string("rth_scene_upload: cannot open ") + path + ": " + dlerror()); return; }
        bool all = true;
        auto bind = [&](auto& fn, const char* name) { fn = reinterpret_cast<std::remove_reference_t<decltype(fn)>>(dlsym(h, name)); if (!fn) { all = false; host_error(std::string("rth_scene_upload: missing symbol ") + name); } };
        bind(api.add_image_hdr, "rt_scene_add_image_hdr"); bind(api.add_image_rgb8, "rt_scene_add_image_rgb8"); bind(api.add_mesh, "rt_scene_add_mesh");
        bind(api.add_transform, "rt_scene_add_transform"); bind(api.set_textures, "rt_scene_set_textures"); bind(api.set_materials, "rt_scene_set_materials");
        bind(api.set_lights, "rt_scene_set_lights"); bind(api.add_mesh_triangles, "rt_scene_add_mesh_triangles"); bind(api.add_primitives, "rt_scene_add_primitives"); bind(api.add_disney, "rt_scene_add_disney");
        api.ok = all;
    });
    return api;
}
}  // namespace

rt_status rth_scene_upload(const rth_scene* s, rt_scene* dst) {
    if (!s || !dst) { host_error("rth_scene_upload: null argument"); return RT_ERR_INVALID; }
    const CoreApi& core = core_api();
    if (!core.ok) return RT_ERR_STATE;
    rt_status st;
    for (auto& im : s->images) {
        if (im.hdr) st = core.add_image_hdr(dst, im.w, im.h, im.f.data(), nullptr); else st = core.add_image_rgb8(dst, im.w, im.h, im.b.data(), nullptr);
        if (st) return st;
    }
    for (auto& m : s->meshes) {
        uint32_t idx;
        if ((st = core.add_mesh(dst, m.p.size() / 3, m.p.data(), m.n.size() / 3, m.n.empty() ? nullptr : m.n.data(), m.uv.size() / 2,
                                m.uv.empty() ? nullptr : m.uv.data(), m.ind.size(), m.ind.data(), &idx))) return st;
    }
    for (size_t i = 0; i < s->xforms.size() / 16; i++) if ((st = core.add_transform(dst, &s->xforms[16 * i], nullptr))) return st;
    if ((st = core.set_textures(dst, s->textures.size(), s->textures.data()))) return st;
    for (auto& d : s->disneys) if ((st = core.add_disney(dst, &d, nullptr))) return st;
    if ((st = core.set_materials(dst, s->materials.size(), s->materials.data()))) return st;
    if ((st = core.set_lights(dst, s->lights.size(), s->lights.data()))) return st;
    for (auto& it : s->items) {
        if (it.run) st = core.add_mesh_triangles(dst, it.mesh, it.mat, nullptr);
        else st = core.add_primitives(dst, 1, &it.prim, nullptr);
        if (st) return st;
    }
    return RT_OK;
}
const char* rth_last_error_string(void) { return g_host_error.c_str(); }
rt_status rth_scene_add_obj(rth_scene* s, const char* path, const double xf[16], uint32_t* out) {
    if (!s || !path || !xf) { host_error("rth_scene_add_obj: null argument"); return RT_ERR_INVALID; }
    rt::HostMesh m; std::string err;
    if (!load_obj(path, xf, m, err)) { host_error("Failed to parse obj " + std::string(path) + ": " + err); return RT_ERR_IO; }   // parser.rs:83-85 panics
    s->meshes.push_back(std::move(m));
    if (out) *out = (uint32_t)s->meshes.size() - 1;
    return RT_OK;
}
// pixel-centre primary rays (lens_radius = 0): Camera::get_ray (geometry.rs:177-190) at u=(x+0.5)/w, v=(y+0.5)/h
rt_status rth_primary_rays(const rt_camera* c, uint32_t w, uint32_t h, rt_ray* out) {
    if (!c || !out) return RT_ERR_INVALID;
    for (uint32_t y = 0; y < h; y++) for (uint32_t x = 0; x < w; x++) {
        double u = ((double)x + 0.5) / (double)w, v = ((double)y + 0.5) / (double)h;
        rt_ray& r = out[(size_t)y * w + x];
        for (int k = 0; k < 3; k++) {
            double to = c->upper_left_corner[k] + c->horizontal_offset[k] * u - c->vertical_offset[k] * v;
            r.origin[k] = c->origin[k]; r.dir[k] = to - c->origin[k];
        }
    }
    return RT_OK;
}

}  // extern "C"
