// mlp_bwd.cu — backward of the fused MLP(GELU) -> LayerNorm(norm2) -> residual block
// (fp32 SIMT path).  Recomputes the hidden activation from `a`; parameter gradients are
// accumulated in registers across the persistent tile loop and flushed once per CTA with
// atomics.  Reference forward: components/transformer_blocks.py:119.
#include "common.cuh"
#include "mlp_common.cuh"

namespace b200mar {

__global__ void __launch_bounds__(kThreads, 1) mlp_bwd_kernel(MlpParams p) {
  using L = MlpSmem;
  extern __shared__ __align__(16) float smem[];
  float* sW1 = smem + L::oW1;
  float* sW2 = smem + L::oW2;
  float* sB1 = smem + L::oB1;
  float* sB2 = smem + L::oB2;
  float* sLnW = smem + L::oLnW;
  float* sA = smem + L::oA;
  float* sH = smem + L::oH;
  float* sY = smem + L::oY;      // m, then dm
  float* sPre = smem + L::oPre;  // pre-activation, then dpre
  float* sDy = smem + L::oDy;    // dout tile
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  load_mlp_weights(smem, p);
  __syncthreads();

  float accW1[2][4][4], accW2[2][4][4];
#pragma unroll
  for (int c = 0; c < 2; ++c) { zero_acc(accW1[c]); zero_acc(accW2[c]); }
  float db1 = 0.f;                                  // column tid of fc1 bias (tid < 128)
  float db2[2] = {0.f, 0.f}, dg[2] = {0.f, 0.f}, dbe[2] = {0.f, 0.f};  // columns lane, lane+32

  const long long ntile = p.n_tok / kTok;
  for (long long tile = blockIdx.x; tile < ntile; tile += gridDim.x) {
    const size_t gbase = (size_t)tile * kTok * kDim;
    for (int idx = tid; idx < kTok * kDim / 4; idx += kThreads) {
      const int m = idx >> 4, c = (idx & 15) << 2;
      *reinterpret_cast<float4*>(sA + m * kLd64 + c) =
          *reinterpret_cast<const float4*>(p.a + gbase + m * kDim + c);
      *reinterpret_cast<float4*>(sDy + m * kLd64 + c) =
          *reinterpret_cast<const float4*>(p.dout + gbase + m * kDim + c);
    }
    __syncthreads();
    // ---- recompute pre = a W1^T + b1, h = gelu(pre)
#pragma unroll 1
    for (int chunk = 0; chunk < 2; ++chunk) {
      float acc[4][4]; zero_acc(acc);
      gemm64<false, false>(acc, sA, kLd64, sW1 + chunk * 64 * kLd64, kLd64, kDim);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = chunk * 64 + gemm_col<false>(j);
        const float b = sB1[n];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int m = gemm_row<false>(i);
          const float pre = acc[i][j] + b;
          sPre[m * kLd128 + n] = pre;
          sH[m * kLd128 + n] = gelu_erf(pre);
        }
      }
    }
    __syncthreads();
    // ---- recompute m = h W2^T + b2
    {
      float acc[4][4]; zero_acc(acc);
      gemm64<false, false>(acc, sH, kLd128, sW2, kLd128, kHid);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = gemm_col<false>(j);
        const float b = sB2[n];
#pragma unroll
        for (int i = 0; i < 4; ++i) sY[gemm_row<false>(i) * kLd64 + n] = acc[i][j] + b;
      }
    }
    __syncthreads();
    // ---- LayerNorm backward (rows of this warp): sY <- dm
#pragma unroll 1
    for (int rr = 0; rr < 8; ++rr) {
      const int m = warp * 8 + rr;
      const float y0 = sY[m * kLd64 + lane], y1 = sY[m * kLd64 + lane + 32];
      const float mean = warp_sum(y0 + y1) * (1.0f / kDim);
      const float c0 = y0 - mean, c1 = y1 - mean;
      const float rstd = 1.0f / sqrtf(warp_sum(c0 * c0 + c1 * c1) * (1.0f / kDim) + kLnEps);
      const float n0 = c0 * rstd, n1 = c1 * rstd;
      const float g0 = sDy[m * kLd64 + lane], g1 = sDy[m * kLd64 + lane + 32];
      dg[0] = fmaf(g0, n0, dg[0]); dg[1] = fmaf(g1, n1, dg[1]);
      dbe[0] += g0; dbe[1] += g1;
      const float t0 = g0 * sLnW[lane], t1 = g1 * sLnW[lane + 32];
      const float m1 = warp_sum(t0 + t1) * (1.0f / kDim);
      const float m2 = warp_sum(t0 * n0 + t1 * n1) * (1.0f / kDim);
      const float d0 = rstd * (t0 - m1 - n0 * m2), d1 = rstd * (t1 - m1 - n1 * m2);
      sY[m * kLd64 + lane] = d0; sY[m * kLd64 + lane + 32] = d1;
      db2[0] += d0; db2[1] += d1;
    }
    __syncthreads();
    // ---- dW2[n][k] += sum_m dm[m][n] h[m][k]
#pragma unroll
    for (int chunk = 0; chunk < 2; ++chunk)
      gemm64<true, true>(accW2[chunk], sY, kLd64, sH + chunk * 64, kLd128, kTok);
    __syncthreads();   // all reads of sH (as h) are done before it is... (sH is not overwritten; sPre is)
    // ---- dpre = (dm W2) * gelu'(pre)   (in place over sPre)
#pragma unroll 1
    for (int chunk = 0; chunk < 2; ++chunk) {
      float acc[4][4]; zero_acc(acc);
      gemm64<false, true>(acc, sY, kLd64, sW2 + chunk * 64, kLd128, kDim);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int m = gemm_row<false>(i);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int k = chunk * 64 + gemm_col<true>(j);
          sPre[m * kLd128 + k] = acc[i][j] * gelu_erf_grad(sPre[m * kLd128 + k]);
        }
      }
    }
    __syncthreads();
    // ---- db1, dW1[k][c] += sum_m dpre[m][k] a[m][c]
    if (tid < kHid) {
      float s = 0.f;
#pragma unroll 8
      for (int m = 0; m < kTok; ++m) s += sPre[m * kLd128 + tid];
      db1 += s;
    }
#pragma unroll
    for (int chunk = 0; chunk < 2; ++chunk)
      gemm64<true, true>(accW1[chunk], sPre + chunk * 64, kLd128, sA, kLd64, kTok);
    // ---- da = dout + dpre W1
    {
      float acc[4][4]; zero_acc(acc);
      gemm64<false, true>(acc, sPre, kLd128, sW1, kLd64, kHid);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int m = gemm_row<false>(i);
        const int c = gemm_col<true>(0);
        const float4 g = *reinterpret_cast<const float4*>(sDy + m * kLd64 + c);
        *reinterpret_cast<float4*>(p.da + gbase + m * kDim + c) =
            make_float4(g.x + acc[i][0], g.y + acc[i][1], g.z + acc[i][2], g.w + acc[i][3]);
      }
    }
    __syncthreads();
  }

  // ---- flush parameter gradients
#pragma unroll
  for (int chunk = 0; chunk < 2; ++chunk)
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        // dW2[n][k]: rows n = 4ty+i, cols k = chunk*64 + 4tx+j
        atomicAdd(p.d_fc2_w + gemm_row<true>(i) * kHid + chunk * 64 + gemm_col<true>(j),
                  accW2[chunk][i][j]);
        // dW1[k][c]: rows k = chunk*64 + 4ty+i, cols c = 4tx+j
        atomicAdd(p.d_fc1_w + (chunk * 64 + gemm_row<true>(i)) * kDim + gemm_col<true>(j),
                  accW1[chunk][i][j]);
      }
  if (tid < kHid) atomicAdd(p.d_fc1_b + tid, db1);
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    atomicAdd(p.d_fc2_b + lane + 32 * e, db2[e]);
    atomicAdd(p.d_ln_w + lane + 32 * e, dg[e]);
    atomicAdd(p.d_ln_b + lane + 32 * e, dbe[e]);
  }
}

}  // namespace b200mar

extern "C" int b200mar_mlp_bwd(const float* a, const float* dout, const float* fc1_w,
                               const float* fc1_b, const float* fc2_w, const float* fc2_b,
                               const float* ln_w, const float* ln_b, float* da, float* d_fc1_w,
                               float* d_fc1_b, float* d_fc2_w, float* d_fc2_b, float* d_ln_w,
                               float* d_ln_b, long long n_tok, void* stream) {
  using namespace b200mar;
  if (n_tok <= 0 || n_tok % kTok) return -1;
  MlpParams p{};
  p.a = a; p.dout = dout; p.fc1_w = fc1_w; p.fc1_b = fc1_b; p.fc2_w = fc2_w; p.fc2_b = fc2_b;
  p.ln_w = ln_w; p.ln_b = ln_b; p.da = da; p.d_fc1_w = d_fc1_w; p.d_fc1_b = d_fc1_b;
  p.d_fc2_w = d_fc2_w; p.d_fc2_b = d_fc2_b; p.d_ln_w = d_ln_w; p.d_ln_b = d_ln_b; p.n_tok = n_tok;
  const size_t smem = MlpSmem::kFloatsBwd * sizeof(float);
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(mlp_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    configured.done();
  }
  const long long ntile = n_tok / kTok;
  const int grid = (int)(ntile < sm_count() ? ntile : sm_count());
  mlp_bwd_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
