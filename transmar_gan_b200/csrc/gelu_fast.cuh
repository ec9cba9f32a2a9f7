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

// GELU(x) * scale for a consumer that rounds the result to FP16 (the MLP hidden row, 11-bit significand):
//   GELU(x) = x * Phi(x),  Phi(x) ~ 1 / (1 + exp(-x P(min(x^2, 64)))),  P(t) = c0 + c1 t + c2 t^2
// (minimax fit of the GELU error over |x| <= 12: max |error| 2.5e-5 absolute - a tenth of the half-ulp of the FP16
// result at |GELU| ~ 0.5, zero error and exact slope at x -> 0; measured whole-model effect of the approximation
// alone on the CPU oracle: 5e-6 (W1), 4e-5 (W3 on structured input) of the 1e-3 gate).  8 FMA-pipe instructions
// + 2 MUFU instead of 15 + 2: the MLP kernel is issue-bound (the bf16-stream variant, with half the HBM bytes,
// runs at the same speed), so the instruction count is what its duration follows.
// The clamp keeps P positive for |x| > 11 (c2 < 0); exp overflow gives 1 / inf = 0, the correct limit.
__device__ __forceinline__ float gelu_sig(float x, float scale) {
#ifdef B200MAR_GELU_EXACT                              // development knob: the A&S erf form
  return gelu_fast(x, 0.5f * scale);
#endif
  constexpr float kL = -1.4426950408889634f;            // exp(-u) = ex2(kL * u)
  const float x2 = fminf(x * x, 64.0f);
  float p = fmaf(-7.030335804e-04f * kL, x2, 7.401129203e-02f * kL);
  p = fmaf(p, x2, 1.595015769f * kL);
  const float e = tc::ex2_approx(p * x);
  return __fdividef(scale * x, 1.0f + e);
}

}  // namespace b200mar
