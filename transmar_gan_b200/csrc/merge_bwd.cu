// merge_bwd.cu — PatchMerging backward (components/encoder_decoder.py:51-74), fp32.
// y = LN256(cat2x2(x)) W^T.  The normalised activations are recomputed; the weight gradient is
// accumulated as G1[n][k] = sum dy[m][n] xhat[m][k] plus colsum(dy) and the LayerNorm affine is
// applied once at flush time: dW[n][k] = g[k] G1[n][k] + b[k] sum_m dy[m][n].
#include "common.cuh"
#include "merge_common.cuh"

namespace b200mar {

__global__ void __launch_bounds__(kThreads, 1) patch_merge_bwd_kernel(MergeParams p) {
  using L = MergeSmem;
  extern __shared__ __align__(16) float smem[];
  float* sW = smem + L::oW;
  float* sLnW = smem + L::oLnW;
  float* sLnB = smem + L::oLnB;
  float* sX = smem + L::oX;        // xhat (pre-affine normalised)
  float* sDy = smem + L::oDy;
  float* sRstd = smem + L::oStat;
  float* sColsum = sRstd + kTok;
  float* sDxn = smem + L::oDxn;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  load_matrix(sW, kLd256, p.red_w, kDim, 256, kDim);
  for (int i = tid; i < 256; i += kThreads) { sLnW[i] = p.ln_w[i]; sLnB[i] = p.ln_b[i]; }
  __syncthreads();

  float accW[4][4][4];
#pragma unroll
  for (int c = 0; c < 4; ++c) zero_acc(accW[c]);
  float colsum = 0.f;               // sum_m dy[m][tid], tid < 64
  float dg[8], db[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { dg[i] = 0.f; db[i] = 0.f; }

  const int oh = p.ph / 2, ow = p.pw / 2;
  const long long n_out = (long long)p.B * oh * ow;
  const long long ntile = (n_out + kTok - 1) / kTok;
  for (long long tile = blockIdx.x; tile < ntile; tile += gridDim.x) {
    // ---- gather x (2x2) and the dy tile
    {
      const int sub = lane >> 4, c = (lane & 15) << 2;
#pragma unroll 4
      for (int it = 0; it < 16; ++it) {
        const int pair = (warp * 16 + it) * 2 + sub;
        const int m = pair >> 2, q = pair & 3;
        const long long o = tile * kTok + m;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (o < n_out) v = *reinterpret_cast<const float4*>(p.x + merge_src(p, o, q) * kDim + c);
        *reinterpret_cast<float4*>(sX + m * kLd256 + q * kDim + c) = v;
      }
      for (int idx = tid; idx < kTok * kDim / 4; idx += kThreads) {
        const int m = idx >> 4, cc = (idx & 15) << 2;
        const long long o = tile * kTok + m;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (o < n_out) v = *reinterpret_cast<const float4*>(p.dy + o * kDim + cc);
        *reinterpret_cast<float4*>(sDy + m * kLd64 + cc) = v;
      }
    }
    __syncthreads();
    // ---- normalise (no affine): sX <- xhat
#pragma unroll 1
    for (int rr = 0; rr < 8; ++rr) {
      const int m = warp * 8 + rr;
      float* row = sX + m * kLd256;
      float v[8], s = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) { v[i] = row[lane + 32 * i]; s += v[i]; }
      const float mean = warp_sum(s) * (1.0f / 256.f);
      float q = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) { v[i] -= mean; q = fmaf(v[i], v[i], q); }
      const float rstd = 1.0f / sqrtf(warp_sum(q) * (1.0f / 256.f) + kLnEps);
#pragma unroll
      for (int i = 0; i < 8; ++i) row[lane + 32 * i] = v[i] * rstd;
      if (lane == 0) sRstd[m] = rstd;
    }
    __syncthreads();
    // ---- G1 += dy^T xhat ; colsum(dy) ; dxn = dy W -> sDxn
#pragma unroll
    for (int chunk = 0; chunk < 4; ++chunk)
      gemm64<true, true>(accW[chunk], sDy, kLd64, sX + chunk * 64, kLd256, kTok);
    if (tid < kDim) {
      float s = 0.f;
#pragma unroll 8
      for (int m = 0; m < kTok; ++m) s += sDy[m * kLd64 + tid];
      colsum += s;
    }
#pragma unroll 1
    for (int chunk = 0; chunk < 4; ++chunk) {
      float acc[4][4]; zero_acc(acc);
      gemm64<false, true>(acc, sDy, kLd64, sW + chunk * 64, kLd256, kDim);
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<float4*>(sDxn + gemm_row<false>(i) * kLd256 + chunk * 64 + gemm_col<true>(0)) =
            make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    }
    __syncthreads();
    // ---- LayerNorm backward over 256 channels (in place in sDxn)
#pragma unroll 1
    for (int rr = 0; rr < 8; ++rr) {
      const int m = warp * 8 + rr;
      float* drow = sDxn + m * kLd256;
      const float* xrow = sX + m * kLd256;
      float t[8], xh[8], s1 = 0.f, s2 = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float v = drow[lane + 32 * i];
        xh[i] = xrow[lane + 32 * i];
        dg[i] = fmaf(v, xh[i], dg[i]);
        db[i] += v;
        t[i] = v * sLnW[lane + 32 * i];
        s1 += t[i]; s2 = fmaf(t[i], xh[i], s2);
      }
      const float m1 = warp_sum(s1) * (1.0f / 256.f), m2 = warp_sum(s2) * (1.0f / 256.f);
      const float rstd = sRstd[m];
#pragma unroll
      for (int i = 0; i < 8; ++i) drow[lane + 32 * i] = rstd * (t[i] - m1 - xh[i] * m2);
    }
    __syncthreads();
    // ---- scatter dx back to the 2x2 source tokens
    {
      const int sub = lane >> 4, c = (lane & 15) << 2;
#pragma unroll 4
      for (int it = 0; it < 16; ++it) {
        const int pair = (warp * 16 + it) * 2 + sub;
        const int m = pair >> 2, q = pair & 3;
        const long long o = tile * kTok + m;
        if (o < n_out)
          *reinterpret_cast<float4*>(p.dx + merge_src(p, o, q) * kDim + c) =
              *reinterpret_cast<const float4*>(sDxn + m * kLd256 + q * kDim + c);
      }
    }
    __syncthreads();
  }
  // ---- flush
  if (tid < kDim) sColsum[tid] = colsum;
  __syncthreads();
  // a thread's four columns are consecutive: one 16-byte vector reduction per (chunk, row) when the gradient
  // buffer is 16-byte aligned (dp.FlatGradBucket views are), 4x fewer L2 operations than scalar atomics
  const bool vec = (reinterpret_cast<uintptr_t>(p.d_red_w) & 15) == 0;
#pragma unroll
  for (int chunk = 0; chunk < 4; ++chunk)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int n = gemm_row<true>(i), k0 = chunk * 64 + gemm_col<true>(0);
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = sLnW[k0 + j] * accW[chunk][i][j] + sLnB[k0 + j] * sColsum[n];
      if (vec) {
        atomicAdd(reinterpret_cast<float4*>(p.d_red_w + n * 256 + k0), make_float4(v[0], v[1], v[2], v[3]));
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) atomicAdd(p.d_red_w + n * 256 + k0 + j, v[j]);
      }
    }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    atomicAdd(p.d_ln_w + lane + 32 * i, dg[i]);
    atomicAdd(p.d_ln_b + lane + 32 * i, db[i]);
  }
}

}  // namespace b200mar

extern "C" int b200mar_patch_merge_bwd(const float* x, const float* dy, const float* ln_w,
                                       const float* ln_b, const float* red_w, float* dx,
                                       float* d_ln_w, float* d_ln_b, float* d_red_w, int B, int ph,
                                       int pw, void* stream) {
  using namespace b200mar;
  if (ph % 2 || pw % 2) return -1;
  MergeParams p{};
  p.x = x; p.dy = dy; p.ln_w = ln_w; p.ln_b = ln_b; p.red_w = red_w; p.dx = dx;
  p.d_ln_w = d_ln_w; p.d_ln_b = d_ln_b; p.d_red_w = d_red_w; p.B = B; p.ph = ph; p.pw = pw;
  const size_t smem = MergeSmem::kFloatsBwd * sizeof(float);
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(patch_merge_bwd_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured.done();
  }
  const long long ntile = ((long long)B * (ph / 2) * (pw / 2) + kTok - 1) / kTok;
  const int grid = (int)(ntile < sm_count() ? ntile : sm_count());
  patch_merge_bwd_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
