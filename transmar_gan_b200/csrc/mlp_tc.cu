// mlp_tc.cu — fused MLP(GELU) -> LayerNorm(norm2) -> residual (+ within-stage sum), forward, on
// the 5th-gen tensor cores: tcgen05.mma kind::f16 with FP16 operands (11-bit significand, as TF32)
// and fp32 accumulators in TMEM.  Range control: every token row of `a` is scaled by a power of two
// to max|a| in [1,2) and each weight matrix by one power of two; the exact inverses are applied to
// the fp32 accumulators.  GELU (exact-erf form), LayerNorm and the residual adds are fp32.
//
// One persistent CTA per SM = FOUR independent 128-thread workers sharing the resident weights
// (fc1 16 KB + fc2 16 KB).  A worker streams 128-token tiles; thread t owns token row t = TMEM lane:
//   coalesced load of the tile -> fp16 A operand in smem
//   D1 = A W1^T            4 MMAs 128x128x16 -> TMEM columns [0,128)
//   GELU epilogue in two 64-column halves; the packed-fp16 hidden row is written back to TMEM over
//   the consumed accumulator columns ([0,64) = 128 hidden values) - the hidden never touches smem
//   D2 = H W2^T            8 MMAs 128x64x16, A operand = H from TMEM -> TMEM columns [64,128)
//   LayerNorm in registers (a whole row per thread) -> fp32 staging tile (over the dead A operand)
//   coalesced copy-out: out = a + LN(.), zsum = out + z.
// Reference: components/transformer_blocks.py:119; components/encoder_decoder.py:138-141.
#include "common.cuh"
#include "mlp_common.cuh"
#include "tc.cuh"
#include <cuda_fp16.h>

namespace b200mar {

B200MAR_DEFINE_TC_ERROR_GETTER(tc_error_mlp)

constexpr int kMlpWorkers = 4;

struct MlpTcSmem {                                   // BYTE offsets
  static constexpr int oW1 = 0;                      // fp16 [128 rows][64 k], 8-element k-chunks, R = 128
  static constexpr int oW2 = oW1 + 128 * 64 * 2;     // fp16 [64 rows][128 k], R = 64
  static constexpr int oB1 = oW2 + 64 * 128 * 2;     // fp32 [128]
  static constexpr int oB2 = oB1 + 512;
  static constexpr int oLnW = oB2 + 256;
  static constexpr int oLnB = oLnW + 256;
  static constexpr int oMisc = oLnB + 256;           // [0]: 1/scale(W1), [1]: 1/scale(W2), [2],[3]: scratch
  static constexpr int oBar = oMisc + 32;            // 4 workers x 2 mbarriers + tmem slot
  static constexpr int oWork = ((oBar + 128 + 127) / 128) * 128;
  static constexpr int wA = 0;                       // fp16 [128][64] A operand; later the fp32 staging tile
  static constexpr int wRow = 128 * 64 * 4;          // fp32 [128] inverse row scales
  static constexpr int kWorkBytes = wRow + 512;
  static constexpr int kBytes = oWork + kMlpWorkers * kWorkBytes;
  static_assert(kBytes <= 232448, "mlp_tc shared memory exceeds 227 KB");
};

// GELU(x) = 0.5 x (1 + erf(x / sqrt 2)) with erf from Abramowitz & Stegun 7.1.26
// (|error| <= 1.5e-7, two MUFU ops) instead of the ~25-instruction erff.
__device__ __forceinline__ float gelu_fast(float x) {
  const float z = fabsf(x) * 0.70710678118654752440f;
  const float t = __fdividef(1.0f, fmaf(0.3275911f, z, 1.0f));
  float p = fmaf(1.061405429f, t, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  const float e = tc::ex2_approx(-z * z * 1.4426950408889634f);
  const float erf_abs = fmaf(-p * t, e, 1.0f);       // erf(|x|/sqrt2)
  const float hx = 0.5f * x;
  return fmaf(hx, copysignf(erf_abs, x), hx);
}

__global__ void __launch_bounds__(128 * kMlpWorkers, 1) mlp_tc_fwd_kernel(MlpParams p) {
  using L = MlpTcSmem;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __half* sW1 = reinterpret_cast<__half*>(smem_raw + L::oW1);
  __half* sW2 = reinterpret_cast<__half*>(smem_raw + L::oW2);
  float* sB1 = reinterpret_cast<float*>(smem_raw + L::oB1);
  float* sB2 = reinterpret_cast<float*>(smem_raw + L::oB2);
  float* sLnW = reinterpret_cast<float*>(smem_raw + L::oLnW);
  float* sLnB = reinterpret_cast<float*>(smem_raw + L::oLnB);
  float* sMisc = reinterpret_cast<float*>(smem_raw + L::oMisc);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem_raw + L::oBar + kMlpWorkers * 16);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);      // warp-uniform ids
  const int wk = warp >> 2, wq = warp & 3;
  unsigned char* sWk = smem_raw + L::oWork + wk * L::kWorkBytes;
  unsigned char* sA = sWk + L::wA;
  float* sStage = reinterpret_cast<float*>(sWk);                // fp32 [128][64] chunk-major
  float* sRowInv = reinterpret_cast<float*>(sWk + L::wRow);
  uint64_t* bar1 = bars + wk * 2;
  uint64_t* bar2 = bar1 + 1;

  // ---- stage the weights as fp16 K-major operands, each scaled by one power of two
  uint32_t* sMax = reinterpret_cast<uint32_t*>(sMisc + 2);
  if (tid < 2) sMax[tid] = 0u;
  __syncthreads();
  {
    float m1 = 0.f, m2 = 0.f;
    for (int i = tid; i < 128 * 64; i += 128 * kMlpWorkers) {
      m1 = fmaxf(m1, fabsf(p.fc1_w[i]));
      m2 = fmaxf(m2, fabsf(p.fc2_w[i]));
    }
    atomicMax(sMax, __float_as_uint(m1));
    atomicMax(sMax + 1, __float_as_uint(m2));
  }
  __syncthreads();
  uint32_t e1 = (sMax[0] >> 23) & 0xFFu, e2 = (sMax[1] >> 23) & 0xFFu;
  e1 = e1 < 1u ? 127u : (e1 > 253u ? 253u : e1);
  e2 = e2 < 1u ? 127u : (e2 > 253u ? 253u : e2);
  const float sc1 = __uint_as_float((254u - e1) << 23), sc2 = __uint_as_float((254u - e2) << 23);
  __syncthreads();
  if (tid == 0) { sMisc[0] = __uint_as_float(e1 << 23); sMisc[1] = __uint_as_float(e2 << 23); }
  for (int idx = tid; idx < 128 * 8; idx += 128 * kMlpWorkers) {        // fc1 [128][64]
    const int n = idx & 127, kc = idx >> 7;
    const float4* src = reinterpret_cast<const float4*>(p.fc1_w + n * 64 + kc * 8);
    const float4 a = src[0], b = src[1];
    *reinterpret_cast<uint4*>(sW1 + kc * 1024 + n * 8) =
        make_uint4(tc::pack_f16(a.x * sc1, a.y * sc1), tc::pack_f16(a.z * sc1, a.w * sc1),
                   tc::pack_f16(b.x * sc1, b.y * sc1), tc::pack_f16(b.z * sc1, b.w * sc1));
  }
  for (int idx = tid; idx < 64 * 16; idx += 128 * kMlpWorkers) {        // fc2 [64][128]
    const int n = idx & 63, kc = idx >> 6;
    const float4* src = reinterpret_cast<const float4*>(p.fc2_w + n * 128 + kc * 8);
    const float4 a = src[0], b = src[1];
    *reinterpret_cast<uint4*>(sW2 + kc * 512 + n * 8) =
        make_uint4(tc::pack_f16(a.x * sc2, a.y * sc2), tc::pack_f16(a.z * sc2, a.w * sc2),
                   tc::pack_f16(b.x * sc2, b.y * sc2), tc::pack_f16(b.z * sc2, b.w * sc2));
  }
  if (tid < 128) sB1[tid] = p.fc1_b[tid];
  if (tid < 64) { sB2[tid] = p.fc2_b[tid]; sLnW[tid] = p.ln_w[tid]; sLnB[tid] = p.ln_b[tid]; }
  if (tid == 0) {
    for (int i = 0; i < 2 * kMlpWorkers; ++i) tc::mbar_init(bars + i, 1);
    tc::fence_barrier_init();
  }
  if (warp == 0) tc::tmem_alloc<512>(tmem_slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();

  const uint32_t tbase = *tmem_slot + wk * 128;                 // D1 / H at [0,128), D2 at [64,128)
  const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
  constexpr uint32_t kI128 = tc::idesc_f16(128, 128), kI64 = tc::idesc_f16(128, 64);
  const uint64_t dA = tc::smem_desc(tc::smem_u32(sA), 2048, 128),
                 dW1 = tc::smem_desc(tc::smem_u32(sW1), 2048, 128),
                 dW2 = tc::smem_desc(tc::smem_u32(sW2), 1024, 128);
  uint32_t ph1 = 0, ph2 = 0;
  const int row = wq * 32 + lane;                               // token row of this thread = TMEM lane
  const float w1inv = sMisc[0], w2inv = sMisc[1];

  const long long ntile = (p.n_tok + 127) / 128;
  for (long long tile = (long long)blockIdx.x * kMlpWorkers + wk; tile < ntile;
       tile += (long long)gridDim.x * kMlpWorkers) {
    const long long tok0 = tile * 128;
    const int valid = (int)((p.n_tok - tok0) < 128 ? (p.n_tok - tok0) : 128);
    // ---- load the 128x64 tile (coalesced 64-byte segments), row-scale, convert to fp16
    {
      const int q = lane >> 3, m8 = lane & 7;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float4 xv[8];
        {
          float4 zv[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int it = half * 8 + i;
            const int m = wq * 32 + (it >> 2) * 8 + m8, kc = (it & 3) * 4 + q;
            xv[i] = zv[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (m < valid) {
              const size_t gi = (size_t)(tok0 + m) * 64 + kc * 4;
              xv[i] = __ldg(reinterpret_cast<const float4*>(p.a + gi));
              if (p.a_is_delta) zv[i] = __ldg(reinterpret_cast<const float4*>(p.z + gi));
            }
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            xv[i].x += zv[i].x; xv[i].y += zv[i].y; xv[i].z += zv[i].z; xv[i].w += zv[i].w;
          }
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          float mxv = 0.f;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 v = xv[j * 4 + i];
            mxv = fmaxf(mxv, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
          }
          mxv = fmaxf(mxv, __shfl_xor_sync(0xffffffffu, mxv, 8));
          mxv = fmaxf(mxv, __shfl_xor_sync(0xffffffffu, mxv, 16));
          uint32_t e = (__float_as_uint(mxv) >> 23) & 0xFFu;
          e = e < 1u ? 127u : (e > 253u ? 253u : e);
          const float sc = __uint_as_float((254u - e) << 23);
          const int m = wq * 32 + (half * 2 + j) * 8 + m8;
          if (q == 0) sRowInv[m] = __uint_as_float(e << 23);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int kc = i * 4 + q;
            const float4 v = xv[j * 4 + i];
            *reinterpret_cast<uint2*>(sA + (kc >> 1) * 2048 + m * 16 + (kc & 1) * 8) =
                make_uint2(tc::pack_f16(v.x * sc, v.y * sc), tc::pack_f16(v.z * sc, v.w * sc));
          }
        }
      }
    }
    tc::fence_async_smem();
    tc::worker_sync(1 + wk);
    // ---- fc1: D1 [128 x 128]
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) tc::mma_f16(tbase, dA + ks * 256, dW1 + ks * 256, kI128, ks > 0);
        tc::mma_commit(bar1);
      }
      __syncwarp();
    }
    const float ainv = sRowInv[row] * w1inv;
    tc::mbar_wait(bar1, ph1, &g_tc_error); ph1 ^= 1;
    tc::tc_fence_after();
    // ---- GELU epilogue, two halves of 64 hidden columns; packed fp16 hidden -> TMEM [0,64)
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
      float v[64];
      tc::tmem_ld64(tbase + lane_addr + half * 64, v);
      float hp[32];
      const float* b1 = sB1 + half * 64;
#pragma unroll
      for (int c = 0; c < 64; c += 2) {
        const float g0 = gelu_fast(fmaf(v[c], ainv, b1[c]));
        const float g1 = gelu_fast(fmaf(v[c + 1], ainv, b1[c + 1]));
        hp[c >> 1] = __uint_as_float(tc::pack_f16(g0, g1));
      }
      tc::tmem_st32(tbase + lane_addr + half * 32, hp);
    }
    tc::tmem_wait_st();
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    // ---- fc2: D2 [128 x 64] = H [128 x 128] W2^T, A operand from TMEM
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) tc::mma_f16_ts(tbase + 64, tbase + ks * 8, dW2 + ks * 128, kI64, ks > 0);
        tc::mma_commit(bar2);
      }
      __syncwarp();
    }
    tc::mbar_wait(bar2, ph2, &g_tc_error); ph2 ^= 1;
    tc::tc_fence_after();
    // ---- LayerNorm(norm2) of the row held in registers -> fp32 staging tile (over the A operand)
    {
      float y[64];
      tc::tmem_ld64(tbase + lane_addr + 64, y);
      float mean = 0.f;
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] = fmaf(y[i], w2inv, sB2[i]); mean += y[i]; }
      mean *= (1.0f / 64.f);
      float var = 0.f;
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] -= mean; var = fmaf(y[i], y[i], var); }
      const float rstd = 1.0f / sqrtf(var * (1.0f / 64.f) + kLnEps);
#pragma unroll
      for (int kc = 0; kc < 16; ++kc)
        *reinterpret_cast<float4*>(sStage + kc * 512 + row * 4) =
            make_float4(y[kc * 4 + 0] * rstd * sLnW[kc * 4 + 0] + sLnB[kc * 4 + 0],
                        y[kc * 4 + 1] * rstd * sLnW[kc * 4 + 1] + sLnB[kc * 4 + 1],
                        y[kc * 4 + 2] * rstd * sLnW[kc * 4 + 2] + sLnB[kc * 4 + 2],
                        y[kc * 4 + 3] * rstd * sLnW[kc * 4 + 3] + sLnB[kc * 4 + 3]);
    }
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    // ---- coalesced copy-out (same mapping as the load): out = a + LN(.), zsum = out + z
    {
      const int q = lane >> 3, m8 = lane & 7;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float4 av[8], zv[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int it = half * 8 + i;
          const int m = wq * 32 + (it >> 2) * 8 + m8, kc = (it & 3) * 4 + q;
          av[i] = zv[i] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (m < valid) {
            const size_t gi = (size_t)(tok0 + m) * 64 + kc * 4;
            av[i] = __ldg(reinterpret_cast<const float4*>(p.a + gi));
            if (p.zsum || p.a_is_delta) zv[i] = __ldg(reinterpret_cast<const float4*>(p.z + gi));
          }
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int it = half * 8 + i;
          const int m = wq * 32 + (it >> 2) * 8 + m8, kc = (it & 3) * 4 + q;
          if (m < valid) {
            const float4 y = *reinterpret_cast<const float4*>(sStage + kc * 512 + m * 4);
            const size_t gi = (size_t)(tok0 + m) * 64 + kc * 4;
            if (p.a_is_delta) {
              av[i].x += zv[i].x; av[i].y += zv[i].y; av[i].z += zv[i].z; av[i].w += zv[i].w;
              if (p.a_out) *reinterpret_cast<float4*>(p.a_out + gi) = av[i];
            }
            const float4 o = make_float4(av[i].x + y.x, av[i].y + y.y, av[i].z + y.z, av[i].w + y.w);
            if (p.out) *reinterpret_cast<float4*>(p.out + gi) = o;
            if (p.zsum)
              *reinterpret_cast<float4*>(p.zsum + gi) =
                  make_float4(o.x + zv[i].x, o.y + zv[i].y, o.z + zv[i].z, o.w + zv[i].w);
          }
        }
      }
    }
    tc::worker_sync(1 + wk);
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(*tmem_slot);
}

}  // namespace b200mar

extern "C" int b200mar_mlp_tc_fwd(const float* a, const float* z, const float* fc1_w,
                                  const float* fc1_b, const float* fc2_w, const float* fc2_b,
                                  const float* ln_w, const float* ln_b, float* out, float* zsum,
                                  float* a_out, int a_is_delta, long long n_tok, void* stream) {
  using namespace b200mar;
  if (n_tok <= 0 || n_tok % kTok) return -1;
  if ((zsum || a_is_delta) && !z) return -1;
  if (a_out && !a_is_delta) return -1;
  MlpParams p{};
  p.a = a; p.z = z; p.fc1_w = fc1_w; p.fc1_b = fc1_b; p.fc2_w = fc2_w; p.fc2_b = fc2_b;
  p.ln_w = ln_w; p.ln_b = ln_b; p.out = out; p.zsum = zsum; p.n_tok = n_tok;
  p.a_out = a_out; p.a_is_delta = a_is_delta;
  const size_t smem = MlpTcSmem::kBytes;
  static bool configured = false;
  if (!configured) {
    B200MAR_CHECK(cudaFuncSetAttribute(mlp_tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    configured = true;
  }
  const long long ntile = (n_tok + 127) / 128;
  long long grid = (ntile + kMlpWorkers - 1) / kMlpWorkers;
  if (grid > sm_count()) grid = sm_count();
  mlp_tc_fwd_kernel<<<(int)grid, 128 * kMlpWorkers, smem, (cudaStream_t)stream>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
