// Camera::new_motion_blur (geometry.rs:133-175): shared by librtcore.so (rt_camera_look_at) and the host scene layer librthost.so,
// which must not depend on the CUDA library (the scene functions of scenes.rs build their camera with it).
#pragma once
#include <cmath>
#include "../../include/rtcore.h"
namespace rt {
inline void camera_look_at(const double from[3], const double to[3], const double up[3], double aspect, double vfov,
                           double aperture, double focus, double t0, double t1, rt_camera* out) {
    const double PI_REF = 3.14159265358979;     // consts.rs:31
    auto normalize = [](double* v) { double n = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); v[0] /= n; v[1] /= n; v[2] /= n; };
    auto cross = [](const double* a, const double* b, double* c) { c[0] = a[1] * b[2] - a[2] * b[1]; c[1] = a[2] * b[0] - a[0] * b[2]; c[2] = a[0] * b[1] - a[1] * b[0]; };
    double w[3] = {to[0] - from[0], to[1] - from[1], to[2] - from[2]}; normalize(w);
    double u[3]; cross(up, w, u); normalize(u); for (int k = 0; k < 3; k++) u[k] = -u[k];
    double v[3]; cross(w, u, v); normalize(v); for (int k = 0; k < 3; k++) v[k] = -v[k];
    double theta = vfov * PI_REF / 180., h = std::tan(theta / 2.);
    double vh = 2. * h, vw = vh * aspect;
    for (int k = 0; k < 3; k++) {
        out->origin[k] = from[k];
        out->horizontal_offset[k] = u[k] * (vw * focus);
        out->vertical_offset[k] = v[k] * (vh * focus);
        out->upper_left_corner[k] = from[k] - out->horizontal_offset[k] * 0.5 + out->vertical_offset[k] * 0.5 + w[k] * focus;
        out->u[k] = u[k]; out->v[k] = v[k]; out->w[k] = w[k];
    }
    out->lens_radius = aperture / 2.; out->t0 = t0; out->t1 = t1;
}
}  // namespace rt
