// mlp_fwd.cu — fused MLP(GELU) -> LayerNorm(norm2) -> residual (+ within-stage residual sum),
// forward (fp32 SIMT path).  Persistent CTAs stream 64-token tiles; fc1/fc2 stay in smem and
// the 128-wide hidden activation never leaves the SM.
//
// Reference: components/transformer_blocks.py:119 (x + norm2(ffn(x))), timm Mlp
// (fc1 -> GELU(erf) -> fc2), components/encoder_decoder.py:138-141 (next input = out + z).
#include "common.cuh"
#include "mlp_common.cuh"

namespace b200mar {

__global__ void __launch_bounds__(kThreads, 1) mlp_fwd_kernel(MlpParams p) {
  using L = MlpSmem;
  extern __shared__ __align__(16) float smem[];
  float* sW1 = smem + L::oW1;
  float* sW2 = smem + L::oW2;
  float* sB1 = smem + L::oB1;
  float* sB2 = smem + L::oB2;
  float* sLnW = smem + L::oLnW;
  float* sLnB = smem + L::oLnB;
  float* sA = smem + L::oA;
  float* sH = smem + L::oH;
  float* sY = smem + L::oY;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  load_mlp_weights(smem, p);
  __syncthreads();

  const long long ntile = p.n_tok / kTok;
  for (long long tile = blockIdx.x; tile < ntile; tile += gridDim.x) {
    const float* a_g = p.a + (size_t)tile * kTok * kDim;
    // ---- load the 64x64 token tile (contiguous 16 KB)
    for (int idx = tid; idx < kTok * kDim / 4; idx += kThreads) {
      const int m = idx >> 4, c = (idx & 15) << 2;
      *reinterpret_cast<float4*>(sA + m * kLd64 + c) =
          *reinterpret_cast<const float4*>(a_g + m * kDim + c);
    }
    __syncthreads();
    // ---- fc1 + GELU -> sH [64][128]
#pragma unroll 1
    for (int chunk = 0; chunk < 2; ++chunk) {
      float acc[4][4]; zero_acc(acc);
      gemm64<false, false>(acc, sA, kLd64, sW1 + chunk * 64 * kLd64, kLd64, kDim);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = chunk * 64 + gemm_col<false>(j);
        const float b = sB1[n];
#pragma unroll
        for (int i = 0; i < 4; ++i) sH[gemm_row<false>(i) * kLd128 + n] = gelu_erf(acc[i][j] + b);
      }
    }
    __syncthreads();
    // ---- fc2 -> sY [64][64]
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
    // ---- LayerNorm(norm2) + residual (+ z for the next block's input)
#pragma unroll 1
    for (int rr = 0; rr < 8; ++rr) {
      const int m = warp * 8 + rr;
      const float y0 = sY[m * kLd64 + lane], y1 = sY[m * kLd64 + lane + 32];
      const float mean = warp_sum(y0 + y1) * (1.0f / kDim);
      const float d0 = y0 - mean, d1 = y1 - mean;
      const float var = warp_sum(d0 * d0 + d1 * d1) * (1.0f / kDim);
      const float rstd = 1.0f / sqrtf(var + kLnEps);
      const size_t base = ((size_t)tile * kTok + m) * kDim;
      const float o0 = sA[m * kLd64 + lane] + (d0 * rstd * sLnW[lane] + sLnB[lane]);
      const float o1 = sA[m * kLd64 + lane + 32] + (d1 * rstd * sLnW[lane + 32] + sLnB[lane + 32]);
      if (p.out) { p.out[base + lane] = o0; p.out[base + lane + 32] = o1; }
      if (p.zsum) {
        p.zsum[base + lane] = o0 + p.z[base + lane];
        p.zsum[base + lane + 32] = o1 + p.z[base + lane + 32];
      }
    }
    __syncthreads();
  }
}

}  // namespace b200mar

extern "C" int b200mar_mlp_fwd(const float* a, const float* z, const float* fc1_w,
                               const float* fc1_b, const float* fc2_w, const float* fc2_b,
                               const float* ln_w, const float* ln_b, float* out, float* zsum,
                               long long n_tok, void* stream) {
  using namespace b200mar;
  if (n_tok <= 0 || n_tok % kTok) return -1;
  if (zsum && !z) return -1;
  MlpParams p{};
  p.a = a; p.z = z; p.fc1_w = fc1_w; p.fc1_b = fc1_b; p.fc2_w = fc2_w; p.fc2_b = fc2_b;
  p.ln_w = ln_w; p.ln_b = ln_b; p.out = out; p.zsum = zsum; p.n_tok = n_tok;
  const size_t smem = MlpSmem::kFloatsFwd * sizeof(float);
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(mlp_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    configured.done();
  }
  const long long ntile = n_tok / kTok;
  const int grid = (int)(ntile < sm_count() ? ntile : sm_count());
  mlp_fwd_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
