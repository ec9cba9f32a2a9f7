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
#include "tma.cuh"
#include "gelu_fast.cuh"
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

// Power-of-two scale for a hidden row so that it cannot overflow fp16: |pre_j| <= amax * maxL1 + maxb
// (amax = the power-of-two upper bound of the row's max|a|, maxL1 = max_j sum_k |W1[j][k]|,
// maxb = max_j |b1[j]|) and |GELU(x)| <= |x|.  Returns 0.5 * scale (the GELU's half factor folded
// in); *inv receives 1 / scale.  scale = 1 whenever the bound is below 2^15, i.e. always in practice.
__device__ __forceinline__ float hidden_half_scale(float amax, float maxL1, float maxb, float* inv) {
  const float bound = fmaf(amax, maxL1, maxb);
  uint32_t eb = (__float_as_uint(bound) >> 23) & 0xFFu;            // bound < 2^(eb-126)
  eb = eb < 141u ? 141u : (eb > 254u ? 254u : eb);
  *inv = __uint_as_float((eb - 14u) << 23);
  return __uint_as_float((267u - eb) << 23);                       // 0.5 * 2^(141-eb)
}

// Stage fc1/fc2 as fp16 K-major operands, each matrix scaled by one power of two (the inverse
// scales go to sMisc[0], sMisc[1]); biases and LN affine as fp32.  Called by all `nthreads` threads;
// contains __syncthreads().
template <int NT>
__device__ __forceinline__ void stage_mlp_weights_f16(const MlpParams& p, __half* sW1, __half* sW2,
                                                      float* sB1, float* sB2, float* sLnW,
                                                      float* sLnB, float* sMisc,
                                                      __half* sW1lo = nullptr, __half* sW2lo = nullptr) {
  // One pass over global memory: every thread fetches its share of the two matrices (4096 float4 in all) with
  // back-to-back 16-byte loads and KEEPS them in registers while the matrix maxima / row L1 norms are reduced, then
  // converts and stores them.  (The first version read the matrices three times - a scalar loop for the maxima, a
  // 64-load strided loop per row for the L1 norms, the staging loops - each a chain of global-latency rounds:
  // 20 K cycles = 11 us of prologue per launch, a quarter of the kernel at the 16 x 128^2 training size.)
  static_assert(NT % 32 == 0, "whole warps");
  constexpr int NPER = (4096 + NT - 1) / NT;
  const int tid = threadIdx.x;
  uint32_t* sMax = reinterpret_cast<uint32_t*>(sMisc + 2);     // [2],[3]: scratch; [4]: maxL1; [5]: maxb
  if (tid < 4) sMax[tid] = 0u;
  __syncthreads();
  float4 w[NPER];
#pragma unroll
  for (int j = 0; j < NPER; ++j) {
    const int f = tid + j * NT;                                // [0, 2048): fc1 [128][64]; [2048, 4096): fc2 [64][128]
    w[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (f < 4096)
      w[j] = __ldg(reinterpret_cast<const float4*>(f < 2048 ? p.fc1_w : p.fc2_w) + (f & 2047));
  }
  {
    float m1 = 0.f, m2 = 0.f;
#pragma unroll
    for (int j = 0; j < NPER; ++j) {
      const int f = tid + j * NT;                              // (warp-uniform: 2048 and 4096 are multiples of 32)
      const float ax = fabsf(w[j].x), ay = fabsf(w[j].y), az = fabsf(w[j].z), aw = fabsf(w[j].w);
      const float am = fmaxf(fmaxf(ax, ay), fmaxf(az, aw));
      if (f < 2048) {
        m1 = fmaxf(m1, am);
        float l1 = (ax + ay) + (az + aw);                      // L1 norm of an fc1 row: its 16 float4 sit in 16 lanes
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) l1 += __shfl_xor_sync(0xffffffffu, l1, o);
        if ((f & 15) == 0) atomicMax(sMax + 2, __float_as_uint(l1));
      } else if (f < 4096) {
        m2 = fmaxf(m2, am);
      }
    }
    atomicMax(sMax, __float_as_uint(m1));                     // non-negative floats order like unsigned ints
    atomicMax(sMax + 1, __float_as_uint(m2));
  }
  if (tid < 128) atomicMax(sMax + 3, __float_as_uint(fabsf(p.fc1_b[tid])));
  __syncthreads();
  uint32_t e1 = (sMax[0] >> 23) & 0xFFu, e2 = (sMax[1] >> 23) & 0xFFu;
  e1 = e1 < 1u ? 127u : (e1 > 253u ? 253u : e1);
  e2 = e2 < 1u ? 127u : (e2 > 253u ? 253u : e2);
  const float sc1 = __uint_as_float((254u - e1) << 23), sc2 = __uint_as_float((254u - e2) << 23);
  if (tid == 0) { sMisc[0] = __uint_as_float(e1 << 23); sMisc[1] = __uint_as_float(e2 << 23); }
#pragma unroll
  for (int j = 0; j < NPER; ++j) {
    const int f = tid + j * NT;
    if (f < 4096) {
      const bool one = f < 2048;
      const int g = f & 2047;
      // fc1: row n = g / 16, k0 = 4 (g % 16), chunk-major [k/8][128 rows][8]; fc2: n = g / 32, k0 = 4 (g % 32), [k/8][64][8]
      const int n = one ? g >> 4 : g >> 5, k0 = one ? (g & 15) << 2 : (g & 31) << 2;
      const int off = (k0 >> 3) * (one ? 1024 : 512) + n * 8 + (k0 & 7);
      const float sc = one ? sc1 : sc2;
      const float a = w[j].x * sc, b = w[j].y * sc, c = w[j].z * sc, d = w[j].w * sc;
      const uint2 hv = make_uint2(tc::pack_f16(a, b), tc::pack_f16(c, d));
      *reinterpret_cast<uint2*>((one ? sW1 : sW2) + off) = hv;
      if (sW1lo)                                               // compensated level: lo = fp16(w - hi)
        *reinterpret_cast<uint2*>((one ? sW1lo : sW2lo) + off) = make_uint2(tc::lo_f16(a, b, hv.x), tc::lo_f16(c, d, hv.y));
    }
  }
  if (tid < 128) sB1[tid] = p.fc1_b[tid];
  if (tid < 64) { sB2[tid] = p.fc2_b[tid]; sLnW[tid] = p.ln_w[tid]; sLnB[tid] = p.ln_b[tid]; }
}

// Optional phase profile (compile with -DB200MAR_MLP_PROF): see attn_tc.cu.
#ifdef B200MAR_MLP_PROF
__device__ unsigned long long g_mlp_prof[32];
#define PROF_DECL __shared__ unsigned long long s_prof[32]; if (threadIdx.x < 32) s_prof[threadIdx.x] = 0; __syncthreads(); \
  long long prof_t = clock64(); const bool prof_on = blockIdx.x == 0 && threadIdx.x == 32;
#define PROF(i) do { if (prof_on) { const long long t_ = clock64(); s_prof[i] += (unsigned long long)(t_ - prof_t); prof_t = clock64(); } } while (0)
#define PROF_FLUSH do { if (prof_on) for (int i_ = 0; i_ < 32; ++i_) g_mlp_prof[i_] += s_prof[i_]; } while (0)
#else
#define PROF_DECL
#define PROF(i)
#define PROF_FLUSH
#endif

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

  stage_mlp_weights_f16<128 * kMlpWorkers>(p, sW1, sW2, sB1, sB2, sLnW, sLnB, sMisc);
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
  PROF_DECL
  for (long long tile = (long long)blockIdx.x * kMlpWorkers + wk; tile < ntile;
       tile += (long long)gridDim.x * kMlpWorkers) {
    PROF(0);
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
    PROF(1);                                          // load + convert
    tc::fence_async_smem();
    tc::worker_sync(1 + wk);
    PROF(2);
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
    float hinv;
    const float hhalf = hidden_half_scale(2.0f * sRowInv[row], sMisc[4], sMisc[5], &hinv);
    tc::mbar_wait(bar1, ph1, &g_tc_error); ph1 ^= 1;
    tc::tc_fence_after();
    PROF(3);                                          // fc1 issue + wait
    // ---- GELU epilogue, two halves of 64 hidden columns; packed fp16 hidden -> TMEM [0,64)
    const float hfull = hhalf + hhalf;                // gelu_sig takes the full hidden scale
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
      float v[64];
      tc::tmem_ld64(tbase + lane_addr + half * 64, v);
      float hp[32];
      const float* b1 = sB1 + half * 64;
#pragma unroll
      for (int c = 0; c < 64; c += 2) {
        const float g0 = gelu_sig(fmaf(v[c], ainv, b1[c]), hfull);
        const float g1 = gelu_sig(fmaf(v[c + 1], ainv, b1[c + 1]), hfull);
        hp[c >> 1] = __uint_as_float(tc::pack_f16(g0, g1));
      }
      tc::tmem_st32(tbase + lane_addr + half * 32, hp);
    }
    tc::tmem_wait_st();
    PROF(4);                                          // GELU
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    PROF(5);
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
    PROF(6);                                          // fc2 issue + wait
    // ---- LayerNorm(norm2) of the row held in registers -> fp32 staging tile (over the A operand)
    {
      float y[64];
      tc::tmem_ld64(tbase + lane_addr + 64, y);
      float mean4[4] = {0.f, 0.f, 0.f, 0.f};          // four partial sums: short dependency chains
      const float yinv = w2inv * hinv;
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] = fmaf(y[i], yinv, sB2[i]); mean4[i & 3] += y[i]; }
      const float mean = ((mean4[0] + mean4[1]) + (mean4[2] + mean4[3])) * (1.0f / 64.f);
      float var4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] -= mean; var4[i & 3] = fmaf(y[i], y[i], var4[i & 3]); }
      const float var = (var4[0] + var4[1]) + (var4[2] + var4[3]);
      const float rstd = 1.0f / sqrtf(var * (1.0f / 64.f) + kLnEps);
#pragma unroll
      for (int kc = 0; kc < 16; ++kc)
        *reinterpret_cast<float4*>(sStage + kc * 512 + row * 4) =
            make_float4(y[kc * 4 + 0] * rstd * sLnW[kc * 4 + 0] + sLnB[kc * 4 + 0],
                        y[kc * 4 + 1] * rstd * sLnW[kc * 4 + 1] + sLnB[kc * 4 + 1],
                        y[kc * 4 + 2] * rstd * sLnW[kc * 4 + 2] + sLnB[kc * 4 + 2],
                        y[kc * 4 + 3] * rstd * sLnW[kc * 4 + 3] + sLnB[kc * 4 + 3]);
    }
    PROF(7);                                          // LN
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    PROF(8);
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
    PROF(9);                                          // copy-out
    tc::worker_sync(1 + wk);
    PROF(10);
  }
  PROF_FLUSH;
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(*tmem_slot);
}

#ifdef B200MAR_MLP_PROF
}  // namespace b200mar
extern "C" int b200mar_mlp_prof(unsigned long long* out32, int reset) {
  unsigned long long z[32] = {0};
  if (cudaMemcpyFromSymbol(out32, b200mar::g_mlp_prof, sizeof(z)) != cudaSuccess) return -1;
  if (reset && cudaMemcpyToSymbol(b200mar::g_mlp_prof, z, sizeof(z)) != cudaSuccess) return -1;
  return 0;
}
namespace b200mar {
#endif


// ================================================================================================
// mlp_tma_fwd_kernel — the same arithmetic with the residual stream moved by the TMA engine, for the
// in-place form of the within-stage sum (z += out):
//   * the fp32 `a` tile arrives by two bulk tensor loads (128 B-swizzled boxes, tma.cuh) into one of
//     TWO 32 KB buffers per worker, the next tile's load being issued while this tile computes;
//   * thread t reads its row from the swizzled tile (conflict-free), scales/packs it to fp16 and puts it
//     in TMEM as the A operand of fc1 (no shared-memory A tile);
//   * out = a + LN2(.) is written back IN PLACE over the a tile by the same thread, then one bulk
//     tensor store (out) and/or one bulk tensor REDUCE-ADD into z (z += out: the next block's input)
//     drain it - z itself is never read by the SMs.
// Three workers per CTA, TMEM 160 columns each: D1 [0,128) -> H [0,64) / D2 [64,128), A [128,160).
constexpr int kTmaWorkers = 3;

// COMP (precision level "tf32c" of the MLP half): fc1 and fc2 on fp16 hi + lo operand PAIRS, three MMA passes
// Ah Wh + Al Wh + Ah Wl into the same fp32 accumulator (~22 significant bits instead of 11), the A&S-erf GELU
// (1.5e-7) instead of the sigmoid-form fit.  The lo weight copies cost 32 KB and the A / H operand pairs 192 TMEM
// columns per worker: D1 [0,128) -> H hi | lo per 64-column half ([0,32) | [32,64), [64,96) | [96,128)),
// A hi | lo [128,160) | [160,192) -> D2 [128,192): TWO workers.
template <bool COMP>
struct MlpTmaSmemT {                                 // BYTE offsets (weights as MlpTcSmem)
  static constexpr int kWorkers = COMP ? 2 : kTmaWorkers;
  static constexpr int oBar = MlpTcSmem::oBar;       // 3 workers x 4 mbarriers + tmem slot
  static constexpr int oW1lo = ((oBar + 128 + 127) / 128) * 128;     // COMP: fp16 lo parts, layouts of oW1 / oW2
  static constexpr int oW2lo = oW1lo + 128 * 64 * 2;
  static constexpr int oWork = (((COMP ? oW2lo + 64 * 128 * 2 : oBar + 128) + 1023) / 1024) * 1024;
  static constexpr int kWorkBytes = 2 * 32768;       // two swizzled fp32 [128][64] tiles
  static constexpr int kBytes = oWork + kWorkers * kWorkBytes;
  static_assert(kBytes <= 232448, "mlp_tma shared memory exceeds 227 KB");
};
using MlpTmaSmem = MlpTmaSmemT<false>;

// S16 = true (precision mode "bf16"): a, z and out are BF16 in HBM; a tile is ONE 16 KB box of [128 rows][64 bf16
// = 128 B] (128 B swizzle); rows are widened to fp32 on read, out = a + LN2(.) is formed in fp32 and rounded to
// bf16 once, and z += out is a bf16 bulk reduce-add.
template <bool S16, bool COMP = false>
__global__ void __launch_bounds__(128 * MlpTmaSmemT<COMP>::kWorkers, 1)
mlp_tma_fwd_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapZ,
                   const __grid_constant__ CUtensorMap mapOut, MlpParams p) {
  using L = MlpTcSmem;
  using MlpTmaSmem = MlpTmaSmemT<COMP>;
  constexpr int kTmaWorkers = MlpTmaSmem::kWorkers;
  constexpr uint32_t kTStride = COMP ? 192 : 160, kTA = 128, kTD2 = COMP ? 128 : 64;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __half* sW1 = reinterpret_cast<__half*>(smem_raw + L::oW1);
  __half* sW2 = reinterpret_cast<__half*>(smem_raw + L::oW2);
  float* sB1 = reinterpret_cast<float*>(smem_raw + L::oB1);
  float* sB2 = reinterpret_cast<float*>(smem_raw + L::oB2);
  float* sLnW = reinterpret_cast<float*>(smem_raw + L::oLnW);
  float* sLnB = reinterpret_cast<float*>(smem_raw + L::oLnB);
  float* sMisc = reinterpret_cast<float*>(smem_raw + L::oMisc);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + MlpTmaSmem::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem_raw + MlpTmaSmem::oBar + kTmaWorkers * 32);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int wk = warp >> 2, wq = warp & 3;
  unsigned char* sWk = smem_raw + MlpTmaSmem::oWork + wk * MlpTmaSmem::kWorkBytes;
  uint64_t* full = bars + wk * 4;                    // [0],[1]: tile buffers; [2]: fc1; [3]: fc2
  uint64_t* bar1 = full + 2;
  uint64_t* bar2 = full + 3;

  stage_mlp_weights_f16<128 * kTmaWorkers>(p, sW1, sW2, sB1, sB2, sLnW, sLnB, sMisc,
                        COMP ? reinterpret_cast<__half*>(smem_raw + MlpTmaSmem::oW1lo) : nullptr,
                        COMP ? reinterpret_cast<__half*>(smem_raw + MlpTmaSmem::oW2lo) : nullptr);
  if (tid == 0) {
    for (int i = 0; i < 4 * kTmaWorkers; ++i) tc::mbar_init(bars + i, 1);
    tc::fence_barrier_init();
    tma::prefetch_desc(&mapA); tma::prefetch_desc(&mapZ); tma::prefetch_desc(&mapOut);
  }
  if (warp == 0) tc::tmem_alloc<512>(tmem_slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();

  const uint32_t tbase = *tmem_slot + wk * kTStride;
  const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
  constexpr uint32_t kI128 = tc::idesc_f16(128, 128), kI64 = tc::idesc_f16(128, 64);
  const uint64_t dW1 = tc::smem_desc(tc::smem_u32(sW1), 2048, 128),
                 dW2 = tc::smem_desc(tc::smem_u32(sW2), 1024, 128),
                 dW1lo = tc::smem_desc(tc::smem_u32(smem_raw + MlpTmaSmem::oW1lo), 2048, 128),
                 dW2lo = tc::smem_desc(tc::smem_u32(smem_raw + MlpTmaSmem::oW2lo), 1024, 128);
  const int row = wq * 32 + lane;
  const float w1inv = sMisc[0], w2inv = sMisc[1];
  const bool leader = wq == 0 && lane == 0;          // bulk groups are per thread: one fixed thread
  uint32_t ph1 = 0, ph2 = 0;

  const long long ntile = (p.n_tok + 127) / 128;
  const long long stride = (long long)gridDim.x * kTmaWorkers;
  long long tile = (long long)blockIdx.x * kTmaWorkers + wk;
  constexpr uint32_t kTileBytes = S16 ? 16384 : 32768;
  // byte offset of 16-byte chunk i of row r in a bf16 tile (8 chunks of 8 bf16 per 128 B row)
  auto off16 = [](int r, int i) -> uint32_t { return (uint32_t)(r * 128 + ((i ^ (r & 7)) << 4)); };
  if (leader && tile < ntile) {
    tma::mbar_expect_tx(full, kTileBytes);
    tma::load_2d(&mapA, full, sWk, 0, (int)(tile * 128));
    if constexpr (!S16) tma::load_2d(&mapA, full, sWk + 16384, 32, (int)(tile * 128));
  }
  PROF_DECL
  for (uint32_t it = 0; tile < ntile; tile += stride, ++it) {
    const int b = it & 1;
    unsigned char* buf = sWk + b * 32768;
    PROF(0);
    tc::mbar_wait(full + b, (it >> 1) & 1, &g_tc_error);
    PROF(1);                                          // wait for the a tile
    // ---- own row -> power-of-two scale -> packed fp16 A operand in TMEM [128,160)
    float ainv, hhalf, hinv;
    {
      float x[64];
      if constexpr (S16) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint4 v = *reinterpret_cast<const uint4*>(buf + off16(row, i));
          x[8 * i + 0] = __uint_as_float(v.x << 16); x[8 * i + 1] = __uint_as_float(v.x & 0xFFFF0000u);
          x[8 * i + 2] = __uint_as_float(v.y << 16); x[8 * i + 3] = __uint_as_float(v.y & 0xFFFF0000u);
          x[8 * i + 4] = __uint_as_float(v.z << 16); x[8 * i + 5] = __uint_as_float(v.z & 0xFFFF0000u);
          x[8 * i + 6] = __uint_as_float(v.w << 16); x[8 * i + 7] = __uint_as_float(v.w & 0xFFFF0000u);
        }
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float4 v = *reinterpret_cast<const float4*>(buf + tma::tile_off(row, i));
          x[4 * i] = v.x; x[4 * i + 1] = v.y; x[4 * i + 2] = v.z; x[4 * i + 3] = v.w;
        }
      }
      float m0 = 0.f, m1 = 0.f, m2 = 0.f, m3 = 0.f;
#pragma unroll
      for (int i = 0; i < 64; i += 4) {
        m0 = fmaxf(m0, fabsf(x[i])); m1 = fmaxf(m1, fabsf(x[i + 1]));
        m2 = fmaxf(m2, fabsf(x[i + 2])); m3 = fmaxf(m3, fabsf(x[i + 3]));
      }
      uint32_t e = (__float_as_uint(fmaxf(fmaxf(m0, m1), fmaxf(m2, m3))) >> 23) & 0xFFu;
      e = e < 1u ? 127u : (e > 253u ? 253u : e);
      const float sc = __uint_as_float((254u - e) << 23);
      ainv = __uint_as_float(e << 23) * w1inv;
      hhalf = hidden_half_scale(2.0f * __uint_as_float(e << 23), sMisc[4], sMisc[5], &hinv);
      float hp[32];
#pragma unroll
      for (int i = 0; i < 32; ++i)
        hp[i] = __uint_as_float(tc::pack_f16(x[2 * i] * sc, x[2 * i + 1] * sc));
      tc::tmem_st32(tbase + kTA + lane_addr, hp);
      if constexpr (COMP) {
        float lp[32];
#pragma unroll
        for (int i = 0; i < 32; ++i)
          lp[i] = __uint_as_float(tc::lo_f16(x[2 * i] * sc, x[2 * i + 1] * sc, __float_as_uint(hp[i])));
        tc::tmem_st32(tbase + kTA + 32 + lane_addr, lp);
      }
    }
    tc::tmem_wait_st();
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    PROF(2);                                          // row convert
    // ---- fc1: D1 [128 x 128] = A (TMEM) W1^T; then prefetch the next tile into the other buffer
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          tc::mma_f16_ts(tbase, tbase + kTA + ks * 8, dW1 + ks * 256, kI128, ks > 0);
          if constexpr (COMP) {
            tc::mma_f16_ts(tbase, tbase + kTA + 32 + ks * 8, dW1 + ks * 256, kI128, 1u);
            tc::mma_f16_ts(tbase, tbase + kTA + ks * 8, dW1lo + ks * 256, kI128, 1u);
          }
        }
        tc::mma_commit(bar1);
      }
      __syncwarp();
      if (lane == 0 && tile + stride < ntile) {
        tma::wait_read_all();                         // the stores that drained the other buffer
        unsigned char* nb = sWk + (b ^ 1) * 32768;
        tma::mbar_expect_tx(full + (b ^ 1), kTileBytes);
        tma::load_2d(&mapA, full + (b ^ 1), nb, 0, (int)((tile + stride) * 128));
        if constexpr (!S16) tma::load_2d(&mapA, full + (b ^ 1), nb + 16384, 32, (int)((tile + stride) * 128));
      }
      __syncwarp();
    }
    tc::mbar_wait(bar1, ph1, &g_tc_error); ph1 ^= 1;
    tc::tc_fence_after();
    PROF(3);
    // ---- GELU epilogue, two halves of 64 hidden columns; packed fp16 hidden -> TMEM [0,64)
    const float hfull = hhalf + hhalf;                // gelu_sig takes the full hidden scale
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
      float v[64];
      tc::tmem_ld64(tbase + lane_addr + half * 64, v);
      float hp[32];
      const float* b1 = sB1 + half * 64;
      if constexpr (COMP) {                           // exact-erf GELU, hidden row as an fp16 hi | lo pair
        float lp[32];
#pragma unroll
        for (int c = 0; c < 64; c += 2) {
          const float g0 = gelu_fast(fmaf(v[c], ainv, b1[c]), hhalf);
          const float g1 = gelu_fast(fmaf(v[c + 1], ainv, b1[c + 1]), hhalf);
          const uint32_t hv = tc::pack_f16(g0, g1);
          hp[c >> 1] = __uint_as_float(hv);
          lp[c >> 1] = __uint_as_float(tc::lo_f16(g0, g1, hv));
        }
        tc::tmem_st32(tbase + lane_addr + half * 64, hp);
        tc::tmem_st32(tbase + lane_addr + half * 64 + 32, lp);
      } else {
#pragma unroll
        for (int c = 0; c < 64; c += 2) {
          const float g0 = gelu_sig(fmaf(v[c], ainv, b1[c]), hfull);
          const float g1 = gelu_sig(fmaf(v[c + 1], ainv, b1[c + 1]), hfull);
          hp[c >> 1] = __uint_as_float(tc::pack_f16(g0, g1));
        }
        tc::tmem_st32(tbase + lane_addr + half * 32, hp);
      }
    }
    tc::tmem_wait_st();
    PROF(4);
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    PROF(5);
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          if constexpr (COMP) {
            const uint32_t ah = tbase + (ks < 4 ? 0 : 64) + (ks & 3) * 8;
            tc::mma_f16_ts(tbase + kTD2, ah, dW2 + ks * 128, kI64, ks > 0);
            tc::mma_f16_ts(tbase + kTD2, ah + 32, dW2 + ks * 128, kI64, 1u);
            tc::mma_f16_ts(tbase + kTD2, ah, dW2lo + ks * 128, kI64, 1u);
          } else {
            tc::mma_f16_ts(tbase + 64, tbase + ks * 8, dW2 + ks * 128, kI64, ks > 0);
          }
        }
        tc::mma_commit(bar2);
      }
      __syncwarp();
    }
    tc::mbar_wait(bar2, ph2, &g_tc_error); ph2 ^= 1;
    tc::tc_fence_after();
    PROF(6);
    // ---- LayerNorm(norm2) in registers; out = a + LN(.) in place over the a tile
    {
      float y[64];
      tc::tmem_ld64(tbase + lane_addr + kTD2, y);
      float mean4[4] = {0.f, 0.f, 0.f, 0.f};          // four partial sums: short dependency chains
      const float yinv = w2inv * hinv;
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] = fmaf(y[i], yinv, sB2[i]); mean4[i & 3] += y[i]; }
      const float mean = ((mean4[0] + mean4[1]) + (mean4[2] + mean4[3])) * (1.0f / 64.f);
      float var4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] -= mean; var4[i & 3] = fmaf(y[i], y[i], var4[i & 3]); }
      const float var = (var4[0] + var4[1]) + (var4[2] + var4[3]);
      const float rstd = 1.0f / sqrtf(var * (1.0f / 64.f) + kLnEps);
      if constexpr (S16) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          uint4* q = reinterpret_cast<uint4*>(buf + off16(row, i));
          const uint4 v = *q;
          const uint32_t w[4] = {v.x, v.y, v.z, v.w};
          uint32_t r[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int c = 8 * i + 2 * j;
            const float lo = __uint_as_float(w[j] << 16) + (y[c] * rstd * sLnW[c] + sLnB[c]);
            const float hi = __uint_as_float(w[j] & 0xFFFF0000u) + (y[c + 1] * rstd * sLnW[c + 1] + sLnB[c + 1]);
            r[j] = tc::pack_bf16(lo, hi);
          }
          *q = make_uint4(r[0], r[1], r[2], r[3]);
        }
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float4* q = reinterpret_cast<float4*>(buf + tma::tile_off(row, i));
          const float4 av = *q;
          *q = make_float4(av.x + (y[4 * i + 0] * rstd * sLnW[4 * i + 0] + sLnB[4 * i + 0]),
                           av.y + (y[4 * i + 1] * rstd * sLnW[4 * i + 1] + sLnB[4 * i + 1]),
                           av.z + (y[4 * i + 2] * rstd * sLnW[4 * i + 2] + sLnB[4 * i + 2]),
                           av.w + (y[4 * i + 3] * rstd * sLnW[4 * i + 3] + sLnB[4 * i + 3]));
        }
      }
    }
    PROF(7);
    tc::fence_async_smem();                           // generic writes -> visible to the TMA engine
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    if (leader) {
      const int r0 = (int)(tile * 128);
      if (p.out) { tma::store_2d(&mapOut, buf, 0, r0); if constexpr (!S16) tma::store_2d(&mapOut, buf + 16384, 32, r0); }
      if (p.zsum) { tma::reduce_add_2d(&mapZ, buf, 0, r0); if constexpr (!S16) tma::reduce_add_2d(&mapZ, buf + 16384, 32, r0); }
      tma::commit_group();
    }
    PROF(8);
  }
  if (leader) tma::wait_all();
  PROF_FLUSH;
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
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(mlp_tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    configured.done();
  }
  const long long ntile = (n_tok + 127) / 128;
  long long grid = (ntile + kMlpWorkers - 1) / kMlpWorkers;
  if (grid > sm_count()) grid = sm_count();
  mlp_tc_fwd_kernel<<<(int)grid, 128 * kMlpWorkers, smem, (cudaStream_t)stream>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_mlp_tma_fwd(const float* a, float* z_inout, const float* fc1_w,
                                   const float* fc1_b, const float* fc2_w, const float* fc2_b,
                                   const float* ln_w, const float* ln_b, float* out,
                                   long long n_tok, void* stream) {
  using namespace b200mar;
  if (n_tok <= 0 || n_tok % kTok || n_tok * 1 > 0x7fffffffLL) return -1;
  if (!out && !z_inout) return -1;
  MlpParams p{};
  p.a = a; p.z = z_inout; p.fc1_w = fc1_w; p.fc1_b = fc1_b; p.fc2_w = fc2_w; p.fc2_b = fc2_b;
  p.ln_w = ln_w; p.ln_b = ln_b; p.out = out; p.zsum = z_inout; p.n_tok = n_tok;
  CUtensorMap mA, mZ, mO;
  if (int rc = tma::encode_rows64(&mA, a, n_tok)) return rc;
  if (int rc = tma::encode_rows64(&mZ, z_inout ? z_inout : a, n_tok)) return rc;
  if (int rc = tma::encode_rows64(&mO, out ? out : a, n_tok)) return rc;
  const size_t smem = MlpTmaSmem::kBytes;
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(mlp_tma_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    configured.done();
  }
  const long long ntile = (n_tok + 127) / 128;
  long long grid = (ntile + kTmaWorkers - 1) / kTmaWorkers;
  if (grid > sm_count()) grid = sm_count();
  mlp_tma_fwd_kernel<false><<<(int)grid, 128 * kTmaWorkers, smem, (cudaStream_t)stream>>>(mA, mZ, mO, p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

// Precision level "tf32c" of the MLP half: same contract as b200mar_mlp_tma_fwd, fc1 / fc2 on compensated fp16
// hi + lo operand pairs and the 1.5e-7 erf GELU (two workers per SM).
extern "C" int b200mar_mlp_tma_fwd_comp(const float* a, float* z_inout, const float* fc1_w,
                                        const float* fc1_b, const float* fc2_w, const float* fc2_b,
                                        const float* ln_w, const float* ln_b, float* out,
                                        long long n_tok, void* stream) {
  using namespace b200mar;
  if (n_tok <= 0 || n_tok % kTok || n_tok * 1 > 0x7fffffffLL) return -1;
  if (!out && !z_inout) return -1;
  MlpParams p{};
  p.a = a; p.z = z_inout; p.fc1_w = fc1_w; p.fc1_b = fc1_b; p.fc2_w = fc2_w; p.fc2_b = fc2_b;
  p.ln_w = ln_w; p.ln_b = ln_b; p.out = out; p.zsum = z_inout; p.n_tok = n_tok;
  CUtensorMap mA, mZ, mO;
  if (int rc = tma::encode_rows64(&mA, a, n_tok)) return rc;
  if (int rc = tma::encode_rows64(&mZ, z_inout ? z_inout : a, n_tok)) return rc;
  if (int rc = tma::encode_rows64(&mO, out ? out : a, n_tok)) return rc;
  using S = MlpTmaSmemT<true>;
  const size_t smem = S::kBytes;
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(mlp_tma_fwd_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    configured.done();
  }
  const long long ntile = (n_tok + 127) / 128;
  long long grid = (ntile + S::kWorkers - 1) / S::kWorkers;
  if (grid > sm_count()) grid = sm_count();
  mlp_tma_fwd_kernel<false, true><<<(int)grid, 128 * S::kWorkers, smem, (cudaStream_t)stream>>>(mA, mZ, mO, p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

// Precision mode "bf16": a16, z16 (in/out), out16 are BF16 [n_tok][64]; parameters fp32.
extern "C" int b200mar_mlp_tma_fwd_s16(const void* a16, void* z16_inout, const float* fc1_w,
                                       const float* fc1_b, const float* fc2_w, const float* fc2_b,
                                       const float* ln_w, const float* ln_b, void* out16,
                                       long long n_tok, void* stream) {
  using namespace b200mar;
  if (n_tok <= 0 || n_tok % kTok || n_tok * 1 > 0x7fffffffLL) return -1;
  if (!out16 && !z16_inout) return -1;
  MlpParams p{};
  p.a = (const float*)a16; p.z = (const float*)z16_inout; p.fc1_w = fc1_w; p.fc1_b = fc1_b; p.fc2_w = fc2_w;
  p.fc2_b = fc2_b; p.ln_w = ln_w; p.ln_b = ln_b; p.out = (float*)out16; p.zsum = (float*)z16_inout; p.n_tok = n_tok;
  CUtensorMap mA, mZ, mO;
  if (int rc = tma::encode_rows64_16(&mA, a16, n_tok)) return rc;
  if (int rc = tma::encode_rows64_16(&mZ, z16_inout ? z16_inout : a16, n_tok)) return rc;
  if (int rc = tma::encode_rows64_16(&mO, out16 ? out16 : a16, n_tok)) return rc;
  const size_t smem = MlpTmaSmem::kBytes;
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(mlp_tma_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    configured.done();
  }
  const long long ntile = (n_tok + 127) / 128;
  long long grid = (ntile + kTmaWorkers - 1) / kTmaWorkers;
  if (grid > sm_count()) grid = sm_count();
  mlp_tma_fwd_kernel<true><<<(int)grid, 128 * kTmaWorkers, smem, (cudaStream_t)stream>>>(mA, mZ, mO, p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
