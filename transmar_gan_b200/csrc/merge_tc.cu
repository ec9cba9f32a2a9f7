// merge_tc.cu — PatchMerging forward (components/encoder_decoder.py:51-74) on TMA + tcgen05:
//   y = Linear_{256->64, no bias}( LayerNorm_256( cat_q x[2r + (q>>1), 2c + (q&1)] ) )
//
// Tile = 16 x 8 output tokens (128 = the MMA M).  For each of the four 2x2 parity classes q the TMA
// engine fetches the 128 source tokens with a stride-2 tensor-map box straight into the canonical
// 128 B-swizzled K-major layout, i.e. the box IS the A operand of a kind::tf32 MMA
// (K = 64 channels of quadrant q).  LayerNorm is folded around the GEMM so that no thread ever needs
// the whole 256-channel row at once:
//   row thread t, quadrant q:  m_q = mean of its 64 values, x <- x - m_q (in place, tf32-rounded),
//                              ssq_q = sum (x - m_q)^2
//   D  += (x_q - m_q) (gamma_q * W_q)^T                        (tensor cores, fp32 accumulate in TMEM)
//   mean = avg m_q ; var = (sum ssq_q + 64 sum (m_q - mean)^2) / 256
//   y_n = rstd * (D_n + sum_q (m_q - mean) G[q][n]) + C[n],   G[q][n] = sum_{k in q} gamma_k W[n][k],
//                                                              C[n]    = sum_k beta_k W[n][k]
// which is exactly LN followed by the Linear (per-quadrant centring keeps the operands well
// conditioned).  Two 128-thread workers per persistent CTA, each with two 32 KB quadrant buffers
// (load of step s+1 overlaps compute of step s); the result row is written over the last buffer and
// leaves by a TMA store.
#include "common.cuh"
#include "merge_common.cuh"
#include "tc.cuh"
#include "tma.cuh"

namespace b200mar {

B200MAR_DEFINE_TC_ERROR_GETTER(tc_error_merge)

constexpr int kMergeWorkers = 2;

struct MergeTcSmem {                                 // BYTE offsets
  static constexpr int oW = 0;                       // tf32 [64 n][256 k] K-major, 4-element k-chunks, R = 64
  static constexpr int oG = oW + 64 * 256 * 4;       // fp32 [4][64]
  static constexpr int oC = oG + 4 * 64 * 4;         // fp32 [64]
  static constexpr int oBar = oC + 256;              // 2 workers x 4 mbarriers + tmem slot
  static constexpr int oWork = ((oBar + 128 + 1023) / 1024) * 1024;
  static constexpr int kWorkBytes = 2 * 32768;       // two swizzled [2 halves][128 rows][128 B] buffers
  static constexpr int kBytes = oWork + kMergeWorkers * kWorkBytes;
  static_assert(kBytes <= 232448, "merge_tc shared memory exceeds 227 KB");
};

using tc::round_tf32;

__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
  // K-major, 128 B swizzle: 8-row groups of 1024 B (SBO); LBO unused (1); layout type 2
  return tc::smem_desc(saddr, 16, 1024) | ((uint64_t)2 << 61);
}

__global__ void __launch_bounds__(128 * kMergeWorkers, 1)
patch_merge_tc_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapY,
                      MergeParams p) {
  using L = MergeTcSmem;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  float* sWt = reinterpret_cast<float*>(smem_raw + L::oW);
  float* sG = reinterpret_cast<float*>(smem_raw + L::oG);
  float* sC = reinterpret_cast<float*>(smem_raw + L::oC);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem_raw + L::oBar + kMergeWorkers * 32);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int wk = warp >> 2, wq = warp & 3;
  unsigned char* sWk = smem_raw + L::oWork + wk * L::kWorkBytes;
  uint64_t* full = bars + wk * 4;                    // [0],[1]: quadrant buffers landed
  uint64_t* barM = full + 2;                         // [0],[1]: MMAs that read the buffer are complete

  // ---- stage W' = gamma (.) W as tf32 K-major; G, C from the rounded values
  for (int idx = tid; idx < 64 * 64; idx += 128 * kMergeWorkers) {
    const int n = idx & 63, kc = idx >> 6;           // 4-element k-chunk kc
    const float4 w = *reinterpret_cast<const float4*>(p.red_w + n * 256 + kc * 4);
    const float4 g = *reinterpret_cast<const float4*>(p.ln_w + kc * 4);
    *reinterpret_cast<float4*>(sWt + kc * 256 + n * 4) =
        make_float4(round_tf32(w.x * g.x), round_tf32(w.y * g.y), round_tf32(w.z * g.z),
                    round_tf32(w.w * g.w));
  }
  __syncthreads();
  for (int idx = tid; idx < 5 * 64; idx += 128 * kMergeWorkers) {
    const int n = idx & 63, q = idx >> 6;
    float a = 0.f;
    if (q < 4) {
      for (int k = q * 64; k < q * 64 + 64; ++k) a += sWt[(k >> 2) * 256 + n * 4 + (k & 3)];
      sG[q * 64 + n] = a;
    } else {
      for (int k = 0; k < 256; ++k) a = fmaf(p.ln_b[k], p.red_w[n * 256 + k], a);
      sC[n] = a;
    }
  }
  if (tid == 0) {
    for (int i = 0; i < 4 * kMergeWorkers; ++i) tc::mbar_init(bars + i, 1);
    tc::fence_barrier_init();
    tma::prefetch_desc(&mapX); tma::prefetch_desc(&mapY);
  }
  if (warp == 0) tc::tmem_alloc<128>(tmem_slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();

  const uint32_t tbase = *tmem_slot + wk * 64;
  const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
  constexpr uint32_t kT64 = tc::idesc_tf32(128, 64);
  const uint64_t dW = tc::smem_desc(tc::smem_u32(sWt), 1024, 128);
  const int row = wq * 32 + lane;
  const bool leader = wq == 0 && lane == 0;
  const uint32_t roff = (uint32_t)(row * 128);
  const int sw = row & 7;

  const int oh = p.ph / 2, ow = p.pw / 2;
  const int tw = (ow + 15) / 16, th = (oh + 7) / 8;
  const int ntile = p.B * th * tw;
  const int stride = gridDim.x * kMergeWorkers;
  auto issue_load = [&](int tile_, int q, int b) {   // leader only
    const int tx = tile_ % tw, ty = (tile_ / tw) % th, bb = tile_ / (tw * th);
    unsigned char* dst = sWk + b * 32768;
    tma::mbar_expect_tx(full + b, 32768);
    tma::load_4d(&mapX, full + b, dst, 0, tx * 32 + (q & 1), ty * 16 + (q >> 1), bb);
    tma::load_4d(&mapX, full + b, dst + 16384, 32, tx * 32 + (q & 1), ty * 16 + (q >> 1), bb);
  };
  int tile = blockIdx.x * kMergeWorkers + wk;
  if (leader && tile < ntile) issue_load(tile, 0, 0);
  uint32_t step = 0;
  for (; tile < ntile; tile += stride) {
    float mq[4], ssq = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q, ++step) {
      const int b = step & 1;
      unsigned char* buf = sWk + b * 32768;
      tc::mbar_wait(full + b, (step >> 1) & 1, &g_tc_error);
      // ---- own row of quadrant q: centre in place
      {
        float x[64];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float4 v = *reinterpret_cast<const float4*>(buf + (i >> 3) * 16384 + roff + (((i & 7) ^ sw) << 4));
          x[4 * i] = v.x; x[4 * i + 1] = v.y; x[4 * i + 2] = v.z; x[4 * i + 3] = v.w;
        }
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
        for (int i = 0; i < 64; i += 4) { s0 += x[i]; s1 += x[i + 1]; s2 += x[i + 2]; s3 += x[i + 3]; }
        const float m = ((s0 + s1) + (s2 + s3)) * (1.0f / 64.f);
        mq[q] = m;
        float v0 = 0.f, v1 = 0.f;
#pragma unroll
        for (int i = 0; i < 64; i += 2) {
          x[i] -= m; x[i + 1] -= m;
          v0 = fmaf(x[i], x[i], v0); v1 = fmaf(x[i + 1], x[i + 1], v1);
        }
        ssq += v0 + v1;
#pragma unroll
        for (int i = 0; i < 16; ++i)
          *reinterpret_cast<float4*>(buf + (i >> 3) * 16384 + roff + (((i & 7) ^ sw) << 4)) =
              make_float4(round_tf32(x[4 * i]), round_tf32(x[4 * i + 1]), round_tf32(x[4 * i + 2]),
                          round_tf32(x[4 * i + 3]));
      }
      tc::fence_async_smem();
      tc::tc_fence_before();
      tc::worker_sync(1 + wk);
      if (wq == 0) {
        tc::tc_fence_after();
        if (tc::elect_one()) {
          const uint64_t dA0 = smem_desc_sw128(tc::smem_u32(buf));
          const uint64_t dA1 = smem_desc_sw128(tc::smem_u32(buf + 16384));
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            tc::mma_tf32(tbase, (ks < 4 ? dA0 : dA1) + (uint64_t)((ks & 3) * 2),
                         dW + (uint64_t)((q * 8 + ks) * 128), kT64, (q | ks) != 0);
          tc::mma_commit(barM + b);
        }
        __syncwarp();
        if (lane == 0) {
          // ---- next step's quadrant into the other buffer (its MMAs / its store must be done)
          const int ntile_ = q < 3 ? tile : tile + stride;
          if (ntile_ < ntile) {
            if (step > 0) tc::mbar_wait(barM + (b ^ 1), ((step - 1) >> 1) & 1, &g_tc_error);
            tma::wait_read_all();
            issue_load(ntile_, (q + 1) & 3, b ^ 1);
          }
        }
        __syncwarp();
      }
    }
    // ---- all four quadrants accumulated: LayerNorm fold + store
    const uint32_t last = step - 1;
    unsigned char* obuf = sWk + (last & 1) * 32768;
    tc::mbar_wait(barM + (last & 1), (last >> 1) & 1, &g_tc_error);
    tc::tc_fence_after();
    {
      float y[64];
      tc::tmem_ld64(tbase + lane_addr, y);
      const float mean = 0.25f * ((mq[0] + mq[1]) + (mq[2] + mq[3]));
      const float d0 = mq[0] - mean, d1 = mq[1] - mean, d2 = mq[2] - mean, d3 = mq[3] - mean;
      const float var = (ssq + 64.f * ((d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3))) * (1.0f / 256.f);
      const float rstd = 1.0f / sqrtf(var + kLnEps);
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int n = 4 * i + j;
          const float corr = fmaf(d0, sG[n], fmaf(d1, sG[64 + n], fmaf(d2, sG[128 + n], d3 * sG[192 + n])));
          o[j] = fmaf(rstd, y[n] + corr, sC[n]);
        }
        *reinterpret_cast<float4*>(obuf + (i >> 3) * 16384 + roff + (((i & 7) ^ sw) << 4)) =
            make_float4(o[0], o[1], o[2], o[3]);
      }
    }
    tc::fence_async_smem();
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    if (leader) {
      const int tx = tile % tw, ty = (tile / tw) % th, bb = tile / (tw * th);
      tma::store_4d(&mapY, obuf, 0, tx * 16, ty * 8, bb);
      tma::store_4d(&mapY, obuf + 16384, 32, tx * 16, ty * 8, bb);
      tma::commit_group();
    }
  }
  if (leader) tma::wait_all();
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<128>(*tmem_slot);
}

}  // namespace b200mar

extern "C" int b200mar_patch_merge_tc_fwd(const float* x, const float* ln_w, const float* ln_b,
                                          const float* red_w, float* y, int B, int ph, int pw,
                                          void* stream) {
  using namespace b200mar;
  if (B <= 0 || ph % 2 || pw % 2 || ph < 2 || pw < 2) return -1;
  MergeParams p{};
  p.x = x; p.ln_w = ln_w; p.ln_b = ln_b; p.red_w = red_w; p.y = y; p.B = B; p.ph = ph; p.pw = pw;
  CUtensorMap mX, mY;
  if (int rc = tma::encode_grid(&mX, x, B, ph, pw, 16, 8, 2)) return rc;
  if (int rc = tma::encode_grid(&mY, y, B, ph / 2, pw / 2, 16, 8, 1)) return rc;
  const size_t smem = MergeTcSmem::kBytes;
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(patch_merge_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    configured.done();
  }
  const int ntile = B * ((ph / 2 + 7) / 8) * ((pw / 2 + 15) / 16);
  int grid = (ntile + kMergeWorkers - 1) / kMergeWorkers;
  if (grid > sm_count()) grid = sm_count();
  patch_merge_tc_kernel<<<grid, 128 * kMergeWorkers, smem, (cudaStream_t)stream>>>(mX, mY, p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
