// bottleneck_tc.cu — SCDP bottleneck tail (components/encoder_decoder.py:194-200, ngswin.py:189) on
// TMA + tcgen05:
//   out = LN( Linear_{84->64}( GELU( depthwise3x3(cat) + b_dw ) ) + b_pw ) + e1
// Tile = 16 x 8 tokens (128 = the MMA M); thread t owns token (y, x) = (t / 16, t % 16).
//   * the 18 x 10 x 84 halo of the concat tensor arrives as ONE bulk tensor load; the conv's zero
//     padding is the tensor map's out-of-bounds fill;
//   * each thread runs the depthwise 3x3 + GELU of its token (756 FMAs) out of the halo (the 336-byte
//     token stride makes whole-token float4 reads conflict-free), scales its 84 results by a power
//     of two and stores them as the packed-fp16 A operand in TMEM (K padded to 96);
//   * pointwise 84 -> 64: 6 MMAs 128x64x16 (A from TMEM, fp16 weights in smem), fp32 in TMEM;
//   * bias + LayerNorm in registers, + e1 (its tile came by TMA, 128 B-swizzled), in place, TMA store.
// Two 128-thread workers per persistent CTA.
#include "common.cuh"
#include "tc.cuh"
#include "tma.cuh"
#include "gelu_fast.cuh"
#include <cuda_fp16.h>

namespace b200mar {

B200MAR_DEFINE_TC_ERROR_GETTER(tc_error_bottleneck)

constexpr int kBnWorkers = 2;
constexpr int kHaloW = 18, kHaloH = 10, kCat = 84;

struct BnParams {
  const float* dw_w; const float* dw_b; const float* pw_w; const float* pw_b;
  const float* ln_w; const float* ln_b;
  int B, H, W;
};

struct BnSmem {                                       // BYTE offsets
  static constexpr int oW = 0;                        // fp16 [64 n][96 k], 8-element k-chunks, R = 64
  static constexpr int oDw = oW + 64 * 96 * 2;        // fp32 [9 taps][84]
  static constexpr int oDb = oDw + 9 * kCat * 4;      // fp32 [84]
  static constexpr int oPb = oDb + kCat * 4;          // fp32 [64]
  static constexpr int oLnW = oPb + 256;
  static constexpr int oLnB = oLnW + 256;
  static constexpr int oMisc = oLnB + 256;            // [0]: 1/scale(W)
  static constexpr int oBar = oMisc + 32;             // 2 workers x 4 mbarriers + tmem slot
  static constexpr int oWork = ((oBar + 128 + 1023) / 1024) * 1024;
  static constexpr int wE1 = 0;                       // swizzled fp32 [2 halves][128 rows][128 B]
  static constexpr int wHalo = 32768;                 // fp32 [180 positions][84]
  static constexpr int kWorkBytes = ((wHalo + kHaloW * kHaloH * kCat * 4 + 1023) / 1024) * 1024;
  static constexpr int kBytes = oWork + kBnWorkers * kWorkBytes;
  static_assert(kBytes <= 232448, "bottleneck_tc shared memory exceeds 227 KB");
};

__global__ void __launch_bounds__(128 * kBnWorkers, 1)
bottleneck_tc_kernel(const __grid_constant__ CUtensorMap mapCat, const __grid_constant__ CUtensorMap mapE1,
                     const __grid_constant__ CUtensorMap mapOut, BnParams p) {
  using L = BnSmem;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __half* sW = reinterpret_cast<__half*>(smem_raw + L::oW);
  float* sDw = reinterpret_cast<float*>(smem_raw + L::oDw);
  float* sDb = reinterpret_cast<float*>(smem_raw + L::oDb);
  float* sPb = reinterpret_cast<float*>(smem_raw + L::oPb);
  float* sLnW = reinterpret_cast<float*>(smem_raw + L::oLnW);
  float* sLnB = reinterpret_cast<float*>(smem_raw + L::oLnB);
  float* sMisc = reinterpret_cast<float*>(smem_raw + L::oMisc);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem_raw + L::oBar + kBnWorkers * 32);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int wk = warp >> 2, wq = warp & 3;
  unsigned char* sWk = smem_raw + L::oWork + wk * L::kWorkBytes;
  unsigned char* sE1 = sWk + L::wE1;
  const float* sHalo = reinterpret_cast<const float*>(sWk + L::wHalo);
  uint64_t* barIn = bars + wk * 4;                    // halo + e1 tile landed
  uint64_t* barM = barIn + 1;

  // ---- weights: pointwise as fp16 K-major (K 84 -> 96, zero padded), scaled by one power of two
  uint32_t* sMax = reinterpret_cast<uint32_t*>(sMisc + 2);
  if (tid == 0) *sMax = 0u;
  __syncthreads();
  {
    float m = 0.f;
    for (int i = tid; i < 64 * kCat; i += 128 * kBnWorkers) m = fmaxf(m, fabsf(p.pw_w[i]));
    atomicMax(sMax, __float_as_uint(m));
  }
  __syncthreads();
  uint32_t we = (*sMax >> 23) & 0xFFu;
  we = we < 1u ? 127u : (we > 253u ? 253u : we);
  const float wsc = __uint_as_float((254u - we) << 23);
  if (tid == 0) sMisc[0] = __uint_as_float(we << 23);
  for (int idx = tid; idx < 64 * 12; idx += 128 * kBnWorkers) {
    const int n = idx & 63, kc = idx >> 6;
    uint32_t pk[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = kc * 8 + 2 * j;
      const float a = k < kCat ? p.pw_w[n * kCat + k] * wsc : 0.f;
      const float b = k + 1 < kCat ? p.pw_w[n * kCat + k + 1] * wsc : 0.f;
      pk[j] = tc::pack_f16(a, b);
    }
    *reinterpret_cast<uint4*>(sW + kc * 512 + n * 8) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
  }
  for (int idx = tid; idx < 9 * kCat; idx += 128 * kBnWorkers) sDw[idx] = p.dw_w[(idx % kCat) * 9 + idx / kCat];
  for (int idx = tid; idx < kCat; idx += 128 * kBnWorkers) sDb[idx] = p.dw_b[idx];
  if (tid < 64) { sPb[tid] = p.pw_b[tid]; sLnW[tid] = p.ln_w[tid]; sLnB[tid] = p.ln_b[tid]; }
  if (tid == 0) {
    for (int i = 0; i < 4 * kBnWorkers; ++i) tc::mbar_init(bars + i, 1);
    tc::fence_barrier_init();
    tma::prefetch_desc(&mapCat); tma::prefetch_desc(&mapE1); tma::prefetch_desc(&mapOut);
  }
  if (warp == 0) tc::tmem_alloc<256>(tmem_slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();

  const uint32_t tbase = *tmem_slot + wk * 128;       // D at [0,64), A (packed fp16, K = 96) at [64,112)
  const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
  constexpr uint32_t kI64 = tc::idesc_f16(128, 64);
  const uint64_t dW = tc::smem_desc(tc::smem_u32(sW), 1024, 128);
  const int row = wq * 32 + lane;
  const int ty_ = row >> 4, tx_ = row & 15;
  const bool leader = wq == 0 && lane == 0;
  const uint32_t roff = (uint32_t)(row * 128);
  const int sw = row & 7;
  const float winv = sMisc[0];
  uint32_t phIn = 0, phM = 0;

  const int tw = (p.W + 15) / 16, th = (p.H + 7) / 8;
  const int ntile = p.B * th * tw;
  const int stride = gridDim.x * kBnWorkers;
  auto issue_loads = [&](int tile_) {
    const int tx = tile_ % tw, ty = (tile_ / tw) % th, bb = tile_ / (tw * th);
    tma::mbar_expect_tx(barIn, kHaloW * kHaloH * kCat * 4 + 32768);
    tma::load_4d(&mapCat, barIn, sWk + L::wHalo, 0, tx * 16 - 1, ty * 8 - 1, bb);
    tma::load_4d(&mapE1, barIn, sE1, 0, tx * 16, ty * 8, bb);
    tma::load_4d(&mapE1, barIn, sE1 + 16384, 32, tx * 16, ty * 8, bb);
  };
  int tile = blockIdx.x * kBnWorkers + wk;
  if (leader && tile < ntile) issue_loads(tile);
  for (; tile < ntile; tile += stride) {
    tc::mbar_wait(barIn, phIn, &g_tc_error); phIn ^= 1;
    // ---- depthwise 3x3 + GELU of this thread's token, all 84 channels
    float ainv;
    {
      float g[96];
      const float* hp = sHalo + (ty_ * kHaloW + tx_) * kCat;
#pragma unroll
      for (int c4 = 0; c4 < 21; ++c4) {
        float4 acc = *reinterpret_cast<const float4*>(sDb + c4 * 4);
#pragma unroll
        for (int dy = 0; dy < 3; ++dy)
#pragma unroll
          for (int dx = 0; dx < 3; ++dx) {
            const float4 h = *reinterpret_cast<const float4*>(hp + (dy * kHaloW + dx) * kCat + c4 * 4);
            const float4 w = *reinterpret_cast<const float4*>(sDw + (dy * 3 + dx) * kCat + c4 * 4);
            acc.x = fmaf(h.x, w.x, acc.x); acc.y = fmaf(h.y, w.y, acc.y);
            acc.z = fmaf(h.z, w.z, acc.z); acc.w = fmaf(h.w, w.w, acc.w);
          }
        g[c4 * 4 + 0] = gelu_fast(acc.x, 0.5f); g[c4 * 4 + 1] = gelu_fast(acc.y, 0.5f);
        g[c4 * 4 + 2] = gelu_fast(acc.z, 0.5f); g[c4 * 4 + 3] = gelu_fast(acc.w, 0.5f);
      }
#pragma unroll
      for (int i = kCat; i < 96; ++i) g[i] = 0.f;
      float m0 = 0.f, m1 = 0.f;
#pragma unroll
      for (int i = 0; i < kCat; i += 2) { m0 = fmaxf(m0, fabsf(g[i])); m1 = fmaxf(m1, fabsf(g[i + 1])); }
      uint32_t e = (__float_as_uint(fmaxf(m0, m1)) >> 23) & 0xFFu;
      e = e < 1u ? 127u : (e > 253u ? 253u : e);
      const float sc = __uint_as_float((254u - e) << 23);
      ainv = __uint_as_float(e << 23) * winv;
      float hp32[32], hp16[16];
#pragma unroll
      for (int i = 0; i < 32; ++i) hp32[i] = __uint_as_float(tc::pack_f16(g[2 * i] * sc, g[2 * i + 1] * sc));
#pragma unroll
      for (int i = 0; i < 16; ++i)
        hp16[i] = __uint_as_float(tc::pack_f16(g[64 + 2 * i] * sc, g[64 + 2 * i + 1] * sc));
      tc::tmem_st32(tbase + 64 + lane_addr, hp32);
      tc::tmem_st16(tbase + 96 + lane_addr, hp16);
    }
    tc::tmem_wait_st();
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 6; ++ks)
          tc::mma_f16_ts(tbase, tbase + 64 + ks * 8, dW + (uint64_t)(ks * 128), kI64, ks > 0);
        tc::mma_commit(barM);
      }
      __syncwarp();
    }
    tc::mbar_wait(barM, phM, &g_tc_error); phM ^= 1;
    tc::tc_fence_after();
    // ---- bias + LayerNorm + e1, in place over the e1 tile
    {
      float y[64];
      tc::tmem_ld64(tbase + lane_addr, y);
      float mean4[4] = {0.f, 0.f, 0.f, 0.f};      // four partial sums: short dependency chains
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] = fmaf(y[i], ainv, sPb[i]); mean4[i & 3] += y[i]; }
      const float mean = ((mean4[0] + mean4[1]) + (mean4[2] + mean4[3])) * (1.0f / 64.f);
      float var4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] -= mean; var4[i & 3] = fmaf(y[i], y[i], var4[i & 3]); }
      const float var = (var4[0] + var4[1]) + (var4[2] + var4[3]);
      const float rstd = 1.0f / sqrtf(var * (1.0f / 64.f) + kLnEps);
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        float4* q = reinterpret_cast<float4*>(sE1 + (i >> 3) * 16384 + roff + (((i & 7) ^ sw) << 4));
        const float4 v = *q;
        *q = make_float4(v.x + (y[4 * i + 0] * rstd * sLnW[4 * i + 0] + sLnB[4 * i + 0]),
                         v.y + (y[4 * i + 1] * rstd * sLnW[4 * i + 1] + sLnB[4 * i + 1]),
                         v.z + (y[4 * i + 2] * rstd * sLnW[4 * i + 2] + sLnB[4 * i + 2]),
                         v.w + (y[4 * i + 3] * rstd * sLnW[4 * i + 3] + sLnB[4 * i + 3]));
      }
    }
    tc::fence_async_smem();
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    if (leader) {
      const int tx = tile % tw, ty = (tile / tw) % th, bb = tile / (tw * th);
      tma::store_4d(&mapOut, sE1, 0, tx * 16, ty * 8, bb);
      tma::store_4d(&mapOut, sE1 + 16384, 32, tx * 16, ty * 8, bb);
      tma::commit_group();
      if (tile + stride < ntile) {
        tma::wait_read_all();
        issue_loads(tile + stride);
      }
    }
  }
  if (leader) tma::wait_all();
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<256>(*tmem_slot);
}

}  // namespace b200mar

extern "C" int b200mar_bottleneck_tail_tc_fwd(const float* cat, const float* dw_w, const float* dw_b,
                                              const float* pw_w, const float* pw_b, const float* ln_w,
                                              const float* ln_b, const float* e1, float* out, int B,
                                              int H, int W, void* stream) {
  using namespace b200mar;
  if (B <= 0 || H <= 0 || W <= 0) return -1;
  BnParams p{dw_w, dw_b, pw_w, pw_b, ln_w, ln_b, B, H, W};
  CUtensorMap mC, mE, mO;
  if (int rc = tma::encode_halo(&mC, cat, B, H, W, kCat, kHaloW, kHaloH)) return rc;
  if (int rc = tma::encode_grid(&mE, e1, B, H, W, 16, 8, 1)) return rc;
  if (int rc = tma::encode_grid(&mO, out, B, H, W, 16, 8, 1)) return rc;
  const size_t smem = BnSmem::kBytes;
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(bottleneck_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    configured.done();
  }
  const int ntile = B * ((H + 7) / 8) * ((W + 15) / 16);
  int grid = (ntile + kBnWorkers - 1) / kBnWorkers;
  if (grid > sm_count()) grid = sm_count();
  bottleneck_tc_kernel<<<grid, 128 * kBnWorkers, smem, (cudaStream_t)stream>>>(mC, mE, mO, p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
