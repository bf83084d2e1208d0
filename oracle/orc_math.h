// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product; nothing under
// rustraytracer_b200/ may include, link or call this.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs use it, as the checker / CPU baseline.
//
// PARITY UNPINNED: the reference (calvin-godfrey/RustRaytracer) has no tests, golden vectors
// or fixtures, and cannot be compiled here (no Rust toolchain), so this restatement is pinned
// only by hand-derived known-answer tests (tests/test_oracle_kat.py).
//
// Plain scalar f64, left-to-right evaluation, compiled with -ffp-contract=off, restating the
// arithmetic of nalgebra 0.21.1 as the reference's call sites use it (dot, cross,
// Unit::new_normalize = divide by norm(), argmax = first maximum, Projective3 transforms).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>
#include <vector>

namespace orc {

// consts.rs:3-42 — copied values are parity-critical
constexpr double GAMMA = 2.2;
constexpr double INF = 1e308;                 // consts.rs:30
constexpr double PI = 3.14159265358979;       // consts.rs:31 (15 digits, not M_PI)
constexpr double SMALL = 0.001;               // consts.rs:32
constexpr double INV_PI = 1.0 / PI;           // consts.rs:41
constexpr double INV_2PI = 1.0 / (2.0 * PI);  // consts.rs:42
constexpr uint8_t BSDF_REFLECTION = 1, BSDF_TRANSMISSION = 2, BSDF_DIFFUSE = 4, BSDF_GLOSSY = 8,
                  BSDF_SPECULAR = 16, BSDF_ALL = 31;  // consts.rs:14-20
constexpr uint8_t RADIANCE = 0, IMPORTANCE = 1;        // consts.rs:27-28

struct V3 {
    double x, y, z;
    V3() : x(0), y(0), z(0) {}
    V3(double a, double b, double c) : x(a), y(b), z(c) {}
    double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    double& at(int i) { return i == 0 ? x : (i == 1 ? y : z); }
};
struct V2 { double x, y; V2() : x(0), y(0) {} V2(double a, double b) : x(a), y(b) {} };

inline V3 operator+(const V3& a, const V3& b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(const V3& a, const V3& b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(const V3& a) { return V3(-a.x, -a.y, -a.z); }
inline V3 operator*(const V3& a, double s) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(double s, const V3& a) { return V3(s * a.x, s * a.y, s * a.z); }
inline V3 operator/(const V3& a, double s) { return V3(a.x / s, a.y / s, a.z / s); }
inline bool operator==(const V3& a, const V3& b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
inline bool operator!=(const V3& a, const V3& b) { return !(a == b); }
inline V3 cmul(const V3& a, const V3& b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 cdiv(const V3& a, const V3& b) { return V3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline double dot(const V3& a, const V3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(const V3& a, const V3& b) {
    return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
inline double norm2(const V3& a) { return dot(a, a); }
inline double norm(const V3& a) { return std::sqrt(norm2(a)); }
inline V3 normalize(const V3& a) { double n = norm(a); return V3(a.x / n, a.y / n, a.z / n); }  // Unit::new_normalize
inline V3 black() { return V3(0, 0, 0); }
inline V3 white() { return V3(1, 1, 1); }
// Rust f64::min / f64::max ignore a NaN operand, as C fmin/fmax do
inline double rmin(double a, double b) { return std::fmin(a, b); }
inline double rmax(double a, double b) { return std::fmax(a, b); }
inline double clampd(double x, double lo, double hi) { return rmin(rmax(x, lo), hi); }  // util.rs:28-30

// Row-major 4x4, `Projective3<f64>`
struct M4 {
    double m[16];
    static M4 identity() { M4 r; for (int i = 0; i < 16; i++) r.m[i] = (i % 5 == 0) ? 1.0 : 0.0; return r; }
};
// general 4x4 inverse by cofactors (the closed form nalgebra uses for 4x4: `do_inverse4`)
inline bool inverse(const M4& a, M4& out) {
    const double* m = a.m;  // treat as a flat 16 array; the cofactor formula is layout-symmetric under transpose
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
    for (int i = 0; i < 16; i++) out.m[i] = inv[i] * inv_det;
    return true;
}
// Projective3::transform_vector: upper-left 3x3 times v (w = 0), no divide
inline V3 xf_vector(const M4& a, const V3& v) {
    const double* m = a.m;
    return V3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z, m[8] * v.x + m[9] * v.y + m[10] * v.z);
}
// Projective3::transform_point: homogeneous divide
inline V3 xf_point(const M4& a, const V3& p) {
    const double* m = a.m;
    double x = m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3];
    double y = m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7];
    double z = m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11];
    double w = m[12] * p.x + m[13] * p.y + m[14] * p.z + m[15];
    return V3(x / w, y / w, z / w);
}

// ---- counter-based RNG (replaces the reference's OS-seeded thread_rng / StdRng::from_entropy,
// sampler.rs:318,334-337, util.rs:32-34).  Philox4x32-10 (Salmon et al., SC'11):
// key = (seed_lo, seed_hi), counter = (pixel, sample, dim >> 1, 0x52544358); each block yields two
// uniforms (words 0,1 for even dim, 2,3 for odd dim), each a 53-bit fraction.
inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
struct Rng {
    uint64_t seed; uint32_t pixel, sample, dim;
    Rng(uint64_t s, uint32_t px, uint32_t smp) : seed(s), pixel(px), sample(smp), dim(0) {}
    // dimension d: Philox block d >> 1, words (0,1) for even d, (2,3) for odd d
    double next() {
        uint32_t d = dim++;
        uint32_t c[4] = {pixel, sample, d >> 1, 0x52544358u};
        philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
        uint32_t a = (d & 1u) ? c[2] : c[0], b = (d & 1u) ? c[3] : c[1];
        return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) / 9007199254740992.0;
    }
};

}  // namespace orc
