// bwd_common.cuh — helpers shared by the tensor-core backward kernels (bwd_tc.cu, attn_bwd_tc.cu).
#pragma once
#include "common.cuh"
#include "tc.cuh"

namespace b200mar {

// ------------------------------------------------------------------------------------------------
// small helpers
__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xFFFF0000u); }

// Column sums over the 32 rows held by the lanes of a warp (one 64-vector per lane): after five
// halving exchanges lane l holds the sums of columns 2l and 2l+1, which are added to s0 / s1.
// `f(i)` yields element i of this lane's vector.
template <int H>
__device__ __forceinline__ void colsum_step(float* v, bool up, int mask) {
#pragma unroll
  for (int i = 0; i < H; ++i) {
    const float send = up ? v[i] : v[i + H];
    const float keep = up ? v[i + H] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, mask);
  }
}
template <class F>
__device__ __forceinline__ void warp_colsum64(F f, int lane, float& s0, float& s1) {
  float v[32];
  {
    const bool up = lane & 16;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const float lo = f(i), hi = f(i + 32);
      v[i] = (up ? hi : lo) + __shfl_xor_sync(0xffffffffu, up ? lo : hi, 16);
    }
  }
  colsum_step<16>(v, lane & 8, 8);
  colsum_step<8>(v, lane & 4, 4);
  colsum_step<4>(v, lane & 2, 2);
  colsum_step<2>(v, lane & 1, 1);
  s0 += v[0];
  s1 += v[1];
}

// GELU(x) = x Phi(x) and its derivative Phi(x) + x phi(x); erf from Abramowitz & Stegun 7.1.26
// (|error| <= 1.5e-7), the Gaussian shares the exponential of the erf.
__device__ __forceinline__ void gelu_and_grad(float x, float& g, float& gp) {
  const float z = fabsf(x) * 0.70710678118654752440f;
  const float t = __fdividef(1.0f, fmaf(0.3275911f, z, 1.0f));
  float p = fmaf(1.061405429f, t, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  const float e = tc::ex2_approx(-z * z * 1.4426950408889634f);     // exp(-x^2/2)
  const float erf_abs = fmaf(-p * t, e, 1.0f);
  const float cdf = fmaf(0.5f, copysignf(erf_abs, x), 0.5f);
  g = x * cdf;
  gp = fmaf(x * 0.39894228040143267794f, e, cdf);
}


// launcher of wgrad_tc_kernel (bwd_tc.cu).  hs > 0: the columns of g16 are per-head packed
// [q|k|v] blocks (column n = h*hs + part*ps + d) and are scattered to dw rows part*A + h*hd + d.
int launch_wgrad(const uint16_t* in16, const uint16_t* g16, float* dw, float* dbias, long long n_tok,
                 int cin, int cout, int ld, int cin_valid, int hs, int ps, int hd, int A, cudaStream_t st);

}  // namespace b200mar
