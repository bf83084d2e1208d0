// The production-mode shade kernels: k_shade<float, MAT, CLS, ENVL> (wavefront.cuh), in their own translation unit.
// Compiled with -prec-div=false -prec-sqrt=false -ftz=true (Makefile): those flags touch f32 arithmetic only — the
// radiometric lobes, BSDF frame and MIS weights of RT_SHADE_F32 mode, whose contract is an image within the stated relative
// MSE — and leave every f64 expression (hit points, normals, light sample points, ray origins and directions) rounding like
// the reference.  ncu r01: IEEE f32 sqrt alone was 8.7 % of this kernel's executed instructions, f32 division more.
#define RT_PLAIN_NORMALIZE 1        // geom.cuh:div_by_norm — the zero-numerator shortcut pays in the f64 kernels only
#include "wavefront.cuh"

namespace rt {
void launch_shade_f32(const ShadeLaunch& L, bool envl) {
    if (envl) launch_shade_bins<float, true>(L); else launch_shade_bins<float, false>(L);
}
}  // namespace rt
