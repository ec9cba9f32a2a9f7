// gelu_fast.cuh — exact-erf-form GELU with a 2-MUFU erf (shared by the tensor-core kernels).
#pragma once
#include "tc.cuh"

namespace b200mar {

// GELU(x) = 0.5 x (1 + erf(x / sqrt 2)) with erf from Abramowitz & Stegun 7.1.26
// (|error| <= 1.5e-7, two MUFU ops) instead of the ~25-instruction erff.
__device__ __forceinline__ float gelu_fast(float x, float half_scale) {
  const float z = fabsf(x) * 0.70710678118654752440f;
  const float t = __fdividef(1.0f, fmaf(0.3275911f, z, 1.0f));
  float p = fmaf(1.061405429f, t, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  const float e = tc::ex2_approx(-z * z * 1.4426950408889634f);
  const float erf_abs = fmaf(-p * t, e, 1.0f);       // erf(|x|/sqrt2)
  const float hx = half_scale * x;
  return fmaf(hx, copysignf(erf_abs, x), hx);
}

}  // namespace b200mar
