// ln64.cu — LayerNorm over the 64 channels of a token-major fp32 matrix [n_tok][64], forward and backward,
// for the stage glue of the TRAINING path: SCDPBottleneck.norm (components/encoder_decoder.py:199-200) and
// NGswin.norm (ngswin.py:204).  (Inference folds these into bottleneck_tc / final_dot9; the LayerNorms of the
// blocks live inside the attention / MLP kernels.)
//
// HBM-bound elementwise work: 16 lanes per row, one float4 per lane, so a warp instruction touches two rows
// = 512 contiguous bytes; row statistics by four xor-shuffles inside the half-warp; two-pass variance.
// Backward recomputes mean / rstd from x and ACCUMULATES dweight / dbias with atomics (the convention of
// every backward entry point of this library), one atomic per channel and CTA.
#include "common.cuh"

namespace b200mar {

__device__ __forceinline__ float half_warp_sum(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(256) ln64_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                       const float* __restrict__ b, float* __restrict__ y,
                                                       long long n_tok) {
  const int sub = threadIdx.x & 15;
  const float4 wv = reinterpret_cast<const float4*>(w)[sub], bv = reinterpret_cast<const float4*>(b)[sub];
  // a warp owns the row pair (base, base + 1): the loop bound is warp-uniform (full-mask shuffles inside)
  const int half = (threadIdx.x >> 4) & 1;
  const long long base0 = (((long long)blockIdx.x * 256 + threadIdx.x) >> 5) * 2, stride = (long long)gridDim.x * 16;
  for (long long base = base0; base < n_tok; base += stride) {
    const long long r = base + half;
    const bool valid = r < n_tok;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid) v = reinterpret_cast<const float4*>(x + r * 64)[sub];
    const float mean = half_warp_sum((v.x + v.y) + (v.z + v.w)) * (1.0f / 64.f);
    v.x -= mean; v.y -= mean; v.z -= mean; v.w -= mean;
    const float var = half_warp_sum(fmaf(v.x, v.x, v.y * v.y) + fmaf(v.z, v.z, v.w * v.w)) * (1.0f / 64.f);
    const float rstd = 1.0f / sqrtf(var + kLnEps);
    if (valid) reinterpret_cast<float4*>(y + r * 64)[sub] =
        make_float4(fmaf(v.x * rstd, wv.x, bv.x), fmaf(v.y * rstd, wv.y, bv.y), fmaf(v.z * rstd, wv.z, bv.z),
                    fmaf(v.w * rstd, wv.w, bv.w));
  }
}

// dx = rstd * (g - mean(g) - xhat * mean(g * xhat)),  g = dy * w;  dw += sum_rows dy * xhat;  db += sum_rows dy
__global__ void __launch_bounds__(256) ln64_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                       const float* __restrict__ w, float* __restrict__ dx,
                                                       float* __restrict__ dw, float* __restrict__ db,
                                                       long long n_tok) {
  __shared__ float4 sDw[256], sDb[256];
  const int sub = threadIdx.x & 15;
  const float4 wv = reinterpret_cast<const float4*>(w)[sub];
  float4 aw = make_float4(0.f, 0.f, 0.f, 0.f), ab = aw;
  const int half = (threadIdx.x >> 4) & 1;
  const long long base0 = (((long long)blockIdx.x * 256 + threadIdx.x) >> 5) * 2, stride = (long long)gridDim.x * 16;
  for (long long base = base0; base < n_tok; base += stride) {
    const long long r = base + half;
    const bool valid = r < n_tok;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f), d = v;              // an invalid row contributes zeros
    if (valid) {
      v = reinterpret_cast<const float4*>(x + r * 64)[sub];
      d = reinterpret_cast<const float4*>(dy + r * 64)[sub];
    }
    const float mean = half_warp_sum((v.x + v.y) + (v.z + v.w)) * (1.0f / 64.f);
    v.x -= mean; v.y -= mean; v.z -= mean; v.w -= mean;
    const float var = half_warp_sum(fmaf(v.x, v.x, v.y * v.y) + fmaf(v.z, v.z, v.w * v.w)) * (1.0f / 64.f);
    const float rstd = 1.0f / sqrtf(var + kLnEps);
    v.x *= rstd; v.y *= rstd; v.z *= rstd; v.w *= rstd;                       // xhat
    const float4 g = make_float4(d.x * wv.x, d.y * wv.y, d.z * wv.z, d.w * wv.w);
    const float mg = half_warp_sum((g.x + g.y) + (g.z + g.w)) * (1.0f / 64.f);
    const float mgx = half_warp_sum(fmaf(g.x, v.x, g.y * v.y) + fmaf(g.z, v.z, g.w * v.w)) * (1.0f / 64.f);
    if (valid) reinterpret_cast<float4*>(dx + r * 64)[sub] =
        make_float4(rstd * (g.x - mg - v.x * mgx), rstd * (g.y - mg - v.y * mgx), rstd * (g.z - mg - v.z * mgx),
                    rstd * (g.w - mg - v.w * mgx));
    aw.x = fmaf(d.x, v.x, aw.x); aw.y = fmaf(d.y, v.y, aw.y); aw.z = fmaf(d.z, v.z, aw.z); aw.w = fmaf(d.w, v.w, aw.w);
    ab.x += d.x; ab.y += d.y; ab.z += d.z; ab.w += d.w;
  }
  sDw[threadIdx.x] = aw; sDb[threadIdx.x] = ab;
  __syncthreads();
  if (threadIdx.x < 16) {                       // thread t < 16 owns channels 4t..4t+3 of the CTA
    for (int i = threadIdx.x + 16; i < 256; i += 16) {
      const float4 p = sDw[i], q = sDb[i];
      aw.x += p.x; aw.y += p.y; aw.z += p.z; aw.w += p.w;
      ab.x += q.x; ab.y += q.y; ab.z += q.z; ab.w += q.w;
    }
    float* pw = dw + 4 * threadIdx.x; float* pb = db + 4 * threadIdx.x;
    atomicAdd(pw, aw.x); atomicAdd(pw + 1, aw.y); atomicAdd(pw + 2, aw.z); atomicAdd(pw + 3, aw.w);
    atomicAdd(pb, ab.x); atomicAdd(pb + 1, ab.y); atomicAdd(pb + 2, ab.z); atomicAdd(pb + 3, ab.w);
  }
}

static int ln64_grid(long long n_tok) {
  long long g = (n_tok * 16 + 255) / 256;
  const long long cap = 8ll * sm_count();
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace b200mar

extern "C" int b200mar_ln64_fwd(const float* x, const float* weight, const float* bias, float* y,
                                long long n_tok, void* stream) {
  using namespace b200mar;
  if (n_tok <= 0) return -1;
  ln64_fwd_kernel<<<ln64_grid(n_tok), 256, 0, (cudaStream_t)stream>>>(x, weight, bias, y, n_tok);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_ln64_bwd(const float* x, const float* dy, const float* weight, float* dx,
                                float* dweight, float* dbias, long long n_tok, void* stream) {
  using namespace b200mar;
  if (n_tok <= 0) return -1;
  ln64_bwd_kernel<<<ln64_grid(n_tok), 256, 0, (cudaStream_t)stream>>>(x, dy, weight, dx, dweight, dbias, n_tok);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------- stream casts (precision mode "bf16")
// fp32 <-> bf16 of a token-major stream tensor at stage entry / exit: 8 elements per thread (one 32-byte read or
// write per thread on the fp32 side, 16 bytes on the bf16 side), grid-stride.  n must be a multiple of 8.
namespace b200mar {
__global__ void __launch_bounds__(256) cast_f32_bf16_kernel(const float4* __restrict__ src, uint4* __restrict__ dst,
                                                            long long n8) {
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n8; i += (long long)gridDim.x * 256) {
    const float4 a = __ldg(src + 2 * i), b = __ldg(src + 2 * i + 1);
    uint32_t r0, r1, r2, r3;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r0) : "f"(a.y), "f"(a.x));
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r1) : "f"(a.w), "f"(a.z));
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r2) : "f"(b.y), "f"(b.x));
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r3) : "f"(b.w), "f"(b.z));
    dst[i] = make_uint4(r0, r1, r2, r3);
  }
}
__global__ void __launch_bounds__(256) cast_bf16_f32_kernel(const uint4* __restrict__ src, float4* __restrict__ dst,
                                                            long long n8) {
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n8; i += (long long)gridDim.x * 256) {
    const uint4 v = __ldg(src + i);
    dst[2 * i] = make_float4(__uint_as_float(v.x << 16), __uint_as_float(v.x & 0xFFFF0000u),
                             __uint_as_float(v.y << 16), __uint_as_float(v.y & 0xFFFF0000u));
    dst[2 * i + 1] = make_float4(__uint_as_float(v.z << 16), __uint_as_float(v.z & 0xFFFF0000u),
                                 __uint_as_float(v.w << 16), __uint_as_float(v.w & 0xFFFF0000u));
  }
}
}  // namespace b200mar

extern "C" int b200mar_cast_stream(const void* src, void* dst, long long n, int to_bf16, void* stream) {
  using namespace b200mar;
  if (n <= 0 || n % 8) return -1;
  const long long n8 = n / 8;
  long long grid = (n8 + 255) / 256;
  if (grid > 16ll * sm_count()) grid = 16ll * sm_count();
  if (to_bf16)
    cast_f32_bf16_kernel<<<(int)grid, 256, 0, (cudaStream_t)stream>>>((const float4*)src, (uint4*)dst, n8);
  else
    cast_bf16_f32_kernel<<<(int)grid, 256, 0, (cudaStream_t)stream>>>((const uint4*)src, (float4*)dst, n8);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
