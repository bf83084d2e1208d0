// film.cuh — the film's resolve step, shared by the one-GPU resolve (render.cu:k_resolve) and the fused N-GPU reduce + resolve (multi.cu).
#pragma once
#include <cstdint>

namespace rt {

// util::point_to_color (util.rs:441-471) with aces_comp (util.rs:400-408): running sums (r, g, b, n) -> three bytes
__device__ __forceinline__ void resolve_pixel(double r, double g, double b, double n, uint8_t* __restrict__ rgb) {
    const double scale = 1. / n;
    const double c[3] = {r, g, b};
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const double x = c[k] * scale * 0.6;
        const double a = 2.51, bb = 0.03, cc = 2.43, dd = 0.59, e = 0.14;
        const double v = fmin(fmax((x * (a * x + bb)) / (x * (cc * x + dd) + e), 0.), 1.);
        const double q = round(pow(v, 1. / 2.2) * 256.);
        rgb[k] = !(q > 0.) ? 0 : (q > 255. ? 255 : (uint8_t)q);          // Rust `as u8` saturates, NaN -> 0
    }
}

}  // namespace rt
