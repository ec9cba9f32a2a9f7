// mlp_tc.cu — fused MLP(GELU) -> LayerNorm(norm2) -> residual (+ within-stage sum), forward,
// on the 5th-gen tensor cores: tcgen05.mma kind::tf32 with fp32 accumulators in TMEM.
//
// One persistent CTA per SM = two independent 128-thread workers that share the resident
// weights (fc1 32 KB + fc2 32 KB, chunk-major K-major operands) and each stream 128-token
// tiles: fc1 (two N=64 halves) -> TMEM -> GELU epilogue (thread = token row) -> smem hidden
// half -> fc2 partial (K=64, accumulating in TMEM) -> LayerNorm in registers (a whole row per
// thread, no shuffles) -> residual -> coalesced store.  While one worker runs its CUDA-core
// epilogue the other has the tensor pipe / load path, so the SM overlaps without warp
// specialisation.  Reference: components/transformer_blocks.py:119; encoder_decoder.py:138-141.
#include "common.cuh"
#include "mlp_common.cuh"
#include "tc.cuh"

namespace b200mar {

struct MlpTcSmem {                                   // float offsets
  static constexpr int oW1 = 0;                      // [128 rows][64 k] chunk-major, R=128: 8192
  static constexpr int oW2 = oW1 + 128 * 64;         // [64 rows][128 k] chunk-major, R=64 : 8192
  static constexpr int oB1 = oW2 + 64 * 128;
  static constexpr int oB2 = oB1 + 128;
  static constexpr int oLnW = oB2 + 64;
  static constexpr int oLnB = oLnW + 64;
  static constexpr int oBar = oLnB + 64;             // 2 workers x 2 mbarriers (uint64) + tmem slot
  static constexpr int oWork = oBar + 16;
  static constexpr int kWorkFloats = 128 * 64 + 128 * 64;   // A tile + hidden half, chunk-major R=128
  static constexpr int kFloats = oWork + 2 * kWorkFloats;
};

B200MAR_DEFINE_TC_ERROR_GETTER(tc_error_mlp)

__global__ void __launch_bounds__(256, 1) mlp_tc_fwd_kernel(MlpParams p) {
  using L = MlpTcSmem;
  extern __shared__ __align__(128) float smem[];
  float* sW1 = smem + L::oW1;
  float* sW2 = smem + L::oW2;
  float* sB1 = smem + L::oB1;
  float* sB2 = smem + L::oB2;
  float* sLnW = smem + L::oLnW;
  float* sLnB = smem + L::oLnB;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L::oBar + 8);
  const int tid = threadIdx.x, wk = tid >> 7, t = tid & 127, lane = tid & 31, wq = (tid >> 5) & 3;
  float* sA = smem + L::oWork + wk * L::kWorkFloats;
  float* sH = sA + 128 * 64;
  uint64_t* bar1 = bars + wk * 2;
  uint64_t* bar2 = bar1 + 1;

  // ---- stage weights as chunk-major K-major operands: W[n][k] -> (k/4)*(R*4) + n*4 + k%4
  for (int idx = tid; idx < 128 * 16; idx += 256) {           // fc1 [128][64]: 16 chunks per row
    const int n = idx & 127, kc = idx >> 7;
    *reinterpret_cast<float4*>(sW1 + kc * 512 + n * 4) =
        *reinterpret_cast<const float4*>(p.fc1_w + n * 64 + kc * 4);
  }
  for (int idx = tid; idx < 64 * 32; idx += 256) {            // fc2 [64][128]: 32 chunks per row
    const int n = idx & 63, kc = idx >> 6;
    *reinterpret_cast<float4*>(sW2 + kc * 256 + n * 4) =
        *reinterpret_cast<const float4*>(p.fc2_w + n * 128 + kc * 4);
  }
  for (int i = tid; i < 128; i += 256) sB1[i] = p.fc1_b[i];
  if (tid < 64) { sB2[tid] = p.fc2_b[tid]; sLnW[tid] = p.ln_w[tid]; sLnB[tid] = p.ln_b[tid]; }
  if (tid == 0) {
    for (int i = 0; i < 4; ++i) tc::mbar_init(bars + i, 1);
    tc::fence_barrier_init();
  }
  if (tid < 32) tc::tmem_alloc<512>(tmem_slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot + wk * 256;            // worker's columns: D1 [0,128) D2 [128,192)
  const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;        // this warp's TMEM lane quadrant
  const uint32_t aS = tc::smem_u32(sA), hS = tc::smem_u32(sH);
  const uint32_t w1S = tc::smem_u32(sW1), w2S = tc::smem_u32(sW2);
  constexpr uint32_t kIdesc = tc::idesc_tf32(128, 64);
  uint32_t ph1 = 0, ph2 = 0;

  const long long ntile = (p.n_tok + 127) / 128;
  for (long long tile = (long long)blockIdx.x * 2 + wk; tile < ntile; tile += (long long)gridDim.x * 2) {
    const long long tok0 = tile * 128;
    const int valid = (int)((p.n_tok - tok0) < 128 ? (p.n_tok - tok0) : 128);
    // ---- load the 128x64 token tile, coalesced 64-byte segments -> conflict-free 16-byte slots
    {
      const int q = lane >> 3, m8 = lane & 7;
#pragma unroll
      for (int it = 0; it < 16; ++it) {
        const int m = wq * 32 + (it >> 2) * 8 + m8, kc = (it & 3) * 4 + q;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (m < valid) v = *reinterpret_cast<const float4*>(p.a + (size_t)(tok0 + m) * 64 + kc * 4);
        *reinterpret_cast<float4*>(sA + kc * 512 + m * 4) = v;
      }
    }
    tc::fence_async_smem();
    tc::worker_sync(1 + wk);
    // ---- fc1: D1[:, 0:64] and D1[:, 64:128]  (A = tile, B = fc1 rows 0..63 / 64..127)
    if (t == 0) {
      tc::tc_fence_after();
      tc::mma_tile_tf32(tmem_base, aS, 128, w1S, 128, 64, kIdesc, false);
      tc::mma_tile_tf32(tmem_base + 64, aS, 128, w1S + 64 * 16, 128, 64, kIdesc, false);
      tc::mma_commit(bar1);
    }
    tc::mbar_wait(bar1, ph1, &g_tc_error); ph1 ^= 1;
    tc::tc_fence_after();
    const int row = wq * 32 + lane;                    // token row owned by this thread = TMEM lane
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
      // ---- GELU epilogue of hidden columns [64*half, 64*half+64) -> sH (chunk-major, K = 64)
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float v[16];
        tc::tmem_ld16(tmem_base + lane_addr + half * 64 + g * 16, v);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int n = half * 64 + g * 16 + c * 4;
          *reinterpret_cast<float4*>(sH + (g * 4 + c) * 512 + row * 4) =
              make_float4(gelu_erf(v[c * 4 + 0] + sB1[n + 0]), gelu_erf(v[c * 4 + 1] + sB1[n + 1]),
                          gelu_erf(v[c * 4 + 2] + sB1[n + 2]), gelu_erf(v[c * 4 + 3] + sB1[n + 3]));
        }
      }
      tc::fence_async_smem();
      tc::tc_fence_before();
      tc::worker_sync(1 + wk);
      // ---- fc2 partial: D2 (+)= H_half [128x64] * fc2[:, 64*half : 64*half+64]^T
      if (t == 0) {
        tc::tc_fence_after();
        tc::mma_tile_tf32(tmem_base + 128, hS, 128, w2S + half * 16 * 1024, 64, 64, kIdesc, half == 1);
        tc::mma_commit(bar2);
      }
      tc::mbar_wait(bar2, ph2, &g_tc_error); ph2 ^= 1;
      tc::tc_fence_after();
    }
    // ---- LayerNorm(norm2) of the row held in registers + residual, result written over sA
    {
      float y[64];
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float v[16];
        tc::tmem_ld16(tmem_base + lane_addr + 128 + g * 16, v);
#pragma unroll
        for (int i = 0; i < 16; ++i) y[g * 16 + i] = v[i] + sB2[g * 16 + i];
      }
      float mean = 0.f;
#pragma unroll
      for (int i = 0; i < 64; ++i) mean += y[i];
      mean *= (1.0f / 64.f);
      float var = 0.f;
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] -= mean; var = fmaf(y[i], y[i], var); }
      const float rstd = 1.0f / sqrtf(var * (1.0f / 64.f) + kLnEps);
#pragma unroll
      for (int kc = 0; kc < 16; ++kc) {
        float4 a4 = *reinterpret_cast<const float4*>(sA + kc * 512 + row * 4);
        a4.x += y[kc * 4 + 0] * rstd * sLnW[kc * 4 + 0] + sLnB[kc * 4 + 0];
        a4.y += y[kc * 4 + 1] * rstd * sLnW[kc * 4 + 1] + sLnB[kc * 4 + 1];
        a4.z += y[kc * 4 + 2] * rstd * sLnW[kc * 4 + 2] + sLnB[kc * 4 + 2];
        a4.w += y[kc * 4 + 3] * rstd * sLnW[kc * 4 + 3] + sLnB[kc * 4 + 3];
        *reinterpret_cast<float4*>(sA + kc * 512 + row * 4) = a4;
      }
    }
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    // ---- coalesced copy-out (same mapping as the load): out, and zsum = out + z
    {
      const int q = lane >> 3, m8 = lane & 7;
#pragma unroll
      for (int it = 0; it < 16; ++it) {
        const int m = wq * 32 + (it >> 2) * 8 + m8, kc = (it & 3) * 4 + q;
        if (m < valid) {
          const float4 o = *reinterpret_cast<const float4*>(sA + kc * 512 + m * 4);
          const size_t gi = (size_t)(tok0 + m) * 64 + kc * 4;
          if (p.out) *reinterpret_cast<float4*>(p.out + gi) = o;
          if (p.zsum) {
            const float4 zv = *reinterpret_cast<const float4*>(p.z + gi);
            *reinterpret_cast<float4*>(p.zsum + gi) = make_float4(o.x + zv.x, o.y + zv.y, o.z + zv.z, o.w + zv.w);
          }
        }
      }
    }
    tc::worker_sync(1 + wk);
  }
  tc::tc_fence_before();
  __syncthreads();
  if (tid < 32) tc::tmem_dealloc<512>(*tmem_slot);
}

}  // namespace b200mar

extern "C" int b200mar_mlp_tc_fwd(const float* a, const float* z, const float* fc1_w,
                                  const float* fc1_b, const float* fc2_w, const float* fc2_b,
                                  const float* ln_w, const float* ln_b, float* out, float* zsum,
                                  long long n_tok, void* stream) {
  using namespace b200mar;
  if (n_tok <= 0 || n_tok % kTok) return -1;
  if (zsum && !z) return -1;
  MlpParams p{};
  p.a = a; p.z = z; p.fc1_w = fc1_w; p.fc1_b = fc1_b; p.fc2_w = fc2_w; p.fc2_b = fc2_b;
  p.ln_w = ln_w; p.ln_b = ln_b; p.out = out; p.zsum = zsum; p.n_tok = n_tok;
  const size_t smem = MlpTcSmem::kFloats * sizeof(float);
  static bool configured = false;
  if (!configured) {
    B200MAR_CHECK(cudaFuncSetAttribute(mlp_tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    configured = true;
  }
  const long long ntile = (n_tok + 127) / 128;
  long long grid = (ntile + 1) / 2;
  if (grid > sm_count()) grid = sm_count();
  mlp_tc_fwd_kernel<<<(int)grid, 256, smem, (cudaStream_t)stream>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

