// cascade_tc.cu — InterPool (2x2 max) of the cascade + concat + across_cascade_proj Linear
// (components/pooling.py:13-17, ngswin.py:175-178, components/encoder_decoder.py:117,136) on
// TMA + tcgen05:
//   out[t] = W [64, K] . cat( max2x2(srcA)[t], max2x2(srcB)[t] (optional), direct[t] ) + bias
// and, optionally, poolA[t] = max2x2(srcA)[t] (the pooled shallow features the next stage and the
// bottleneck reuse).  Tile = 16 x 8 output tokens (128 = the MMA M); thread t owns output token t.
// A pooled operand is built from four stride-2 tensor-map boxes (one per 2x2 parity class): the
// thread keeps the running maximum of its row in registers and writes it over the fourth box, which
// then IS the 128 B-swizzled K-major A operand of a kind::tf32 MMA (rounded to tf32; poolA receives the
// exact maxima straight from the registers); the `direct` tile is rounded in place and consumed by
// the MMA where the TMA engine put it.
// Two 128-thread workers per persistent CTA, two 32 KB buffers each (load of step s+1 overlaps step s).
#include "common.cuh"
#include "tc.cuh"
#include "tma.cuh"

namespace b200mar {

B200MAR_DEFINE_TC_ERROR_GETTER(tc_error_cascade)

constexpr int kCasWorkers = 2;

struct CasParams {
  const float* w; const float* bias; float* pool;
  int B, h, wd, K, has_b;
};

struct CasSmem {                                      // BYTE offsets
  static constexpr int oW = 0;                        // tf32 [64 n][192 k] K-major, 4-element k-chunks, R = 64
  static constexpr int oBias = oW + 64 * 192 * 4;
  static constexpr int oBar = oBias + 256;            // 2 workers x 4 mbarriers + tmem slot
  static constexpr int oWork = ((oBar + 128 + 1023) / 1024) * 1024;
  static constexpr int kWorkBytes = 2 * 32768;
  static constexpr int kBytes = oWork + kCasWorkers * kWorkBytes;
  static_assert(kBytes <= 232448, "cascade_tc shared memory exceeds 227 KB");
};

__device__ __forceinline__ uint64_t cas_desc_sw128(uint32_t saddr) {
  return tc::smem_desc(saddr, 16, 1024) | ((uint64_t)2 << 61);
}

__global__ void __launch_bounds__(128 * kCasWorkers, 1)
cascade_tc_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                  const __grid_constant__ CUtensorMap mapD, const __grid_constant__ CUtensorMap mapOut,
                  CasParams p) {
  using L = CasSmem;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  float* sWt = reinterpret_cast<float*>(smem_raw + L::oW);
  float* sBias = reinterpret_cast<float*>(smem_raw + L::oBias);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem_raw + L::oBar + kCasWorkers * 32);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int wk = warp >> 2, wq = warp & 3;
  unsigned char* sWk = smem_raw + L::oWork + wk * L::kWorkBytes;
  uint64_t* full = bars + wk * 4;                     // [0],[1]
  uint64_t* barM = full + 2;                          // [0],[1]: the MMA group that read buffer b is complete

  for (int idx = tid; idx < 64 * (p.K / 4); idx += 128 * kCasWorkers) {
    const int n = idx & 63, kc = idx >> 6;
    *reinterpret_cast<float4*>(sWt + kc * 256 + n * 4) =
        tc::round_tf32(*reinterpret_cast<const float4*>(p.w + (size_t)n * p.K + kc * 4));
  }
  if (tid < 64) sBias[tid] = p.bias[tid];
  if (tid == 0) {
    for (int i = 0; i < 4 * kCasWorkers; ++i) tc::mbar_init(bars + i, 1);
    tc::fence_barrier_init();
    tma::prefetch_desc(&mapA); tma::prefetch_desc(&mapD); tma::prefetch_desc(&mapOut);
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
  auto off = [&](int i) -> uint32_t { return (uint32_t)((i >> 3) * 16384) + roff + (uint32_t)(((i & 7) ^ sw) << 4); };

  const int tw = (p.wd + 15) / 16, th = (p.h + 7) / 8;
  const int ntile = p.B * th * tw;
  const int stride = gridDim.x * kCasWorkers;
  const int nsrc = p.has_b ? 2 : 1;
  const int nstep = 4 * nsrc + 1;                     // quadrant steps of each pooled source, then direct
  // step s of tile_: s < 4*nsrc -> pooled source s/4, parity class s%4; else the direct tile
  auto issue_load = [&](int tile_, int s, int b) {    // leader only
    const int tx = tile_ % tw, ty = (tile_ / tw) % th, bb = tile_ / (tw * th);
    unsigned char* dst = sWk + b * 32768;
    tma::mbar_expect_tx(full + b, 32768);
    if (s < 4 * nsrc) {
      const CUtensorMap* m = s < 4 ? &mapA : &mapB;
      const int q = s & 3;
      tma::load_4d(m, full + b, dst, 0, tx * 32 + (q & 1), ty * 16 + (q >> 1), bb);
      tma::load_4d(m, full + b, dst + 16384, 32, tx * 32 + (q & 1), ty * 16 + (q >> 1), bb);
    } else {
      tma::load_4d(&mapD, full + b, dst, 0, tx * 16, ty * 8, bb);
      tma::load_4d(&mapD, full + b, dst + 16384, 32, tx * 16, ty * 8, bb);
    }
  };
  int tile = blockIdx.x * kCasWorkers + wk;
  if (leader && tile < ntile) issue_load(tile, 0, 0);
  uint32_t gs = 0;                                    // global step counter: buffer = gs & 1
  uint32_t cnt0 = 0, cnt1 = 0;                        // MMA groups committed on barM[0] / barM[1] (uniform)
  for (; tile < ntile; tile += stride) {
    const int tx = tile % tw, ty = (tile / tw) % th, bb = tile / (tw * th);
    float m[64];
    for (int s = 0; s < nstep; ++s, ++gs) {
      const int b = gs & 1;
      unsigned char* buf = sWk + b * 32768;
      tc::mbar_wait(full + b, (gs >> 1) & 1, &g_tc_error);
      const bool pooled = s < 4 * nsrc;
      const int q = s & 3;
      bool operand = !pooled;                         // does this step end with an MMA on `buf`?
      if (pooled) {
        if (q == 0) {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float4 v = *reinterpret_cast<const float4*>(buf + off(i));
            m[4 * i] = v.x; m[4 * i + 1] = v.y; m[4 * i + 2] = v.z; m[4 * i + 3] = v.w;
          }
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float4 v = *reinterpret_cast<const float4*>(buf + off(i));
            m[4 * i] = fmaxf(m[4 * i], v.x); m[4 * i + 1] = fmaxf(m[4 * i + 1], v.y);
            m[4 * i + 2] = fmaxf(m[4 * i + 2], v.z); m[4 * i + 3] = fmaxf(m[4 * i + 3], v.w);
          }
        }
        if (q == 3) {
          // the operand is rounded to tf32 (the MMA would truncate); poolA gets the exact maxima
          float* pool_row = nullptr;
          if (s < 4 && p.pool) {
            const int ox = tx * 16 + (row & 15), oy = ty * 8 + (row >> 4);
            if (ox < p.wd && oy < p.h) pool_row = p.pool + (((size_t)bb * p.h + oy) * p.wd + ox) * 64;
          }
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float4 v = make_float4(m[4 * i], m[4 * i + 1], m[4 * i + 2], m[4 * i + 3]);
            *reinterpret_cast<float4*>(buf + off(i)) = tc::round_tf32(v);
            if (pool_row) *reinterpret_cast<float4*>(pool_row + 4 * i) = v;
          }
          tc::fence_async_smem();
          operand = true;
        }
      } else {
        // direct tile: round to tf32 in place (own row)
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float4* q4 = reinterpret_cast<float4*>(buf + off(i));
          *q4 = tc::round_tf32(*q4);
        }
        tc::fence_async_smem();
      }
      tc::tc_fence_before();
      tc::worker_sync(1 + wk);
      if (wq == 0) {
        if (operand) {
          tc::tc_fence_after();
          if (tc::elect_one()) {
            const int slab = pooled ? (s >> 2) : nsrc;
            const uint64_t dA0 = cas_desc_sw128(tc::smem_u32(buf));
            const uint64_t dA1 = cas_desc_sw128(tc::smem_u32(buf + 16384));
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)
              tc::mma_tf32(tbase, (ks < 4 ? dA0 : dA1) + (uint64_t)((ks & 3) * 2),
                           dW + (uint64_t)((slab * 8 + ks) * 128), kT64, (slab | ks) != 0);
            tc::mma_commit(barM + b);
          }
          __syncwarp();
        }
      }
      if (operand) { if (b) ++cnt1; else ++cnt0; }
      if (leader) {
        // ---- next step's load into the other buffer: the MMA group / store that read it must be done
        const bool last = s == nstep - 1;
        const int ntile_ = last ? tile + stride : tile;
        if (ntile_ < ntile) {
          const uint32_t co = b ? cnt0 : cnt1;
          if (co > 0) tc::mbar_wait(barM + (b ^ 1), (co - 1) & 1, &g_tc_error);
          tma::wait_read_all();
          issue_load(ntile_, last ? 0 : s + 1, b ^ 1);
        }
      }
    }
    // ---- all K slabs accumulated: + bias, stage over the direct tile's buffer, TMA store
    {
      const int bl = (gs - 1) & 1;                     // the direct step's buffer: last MMA group of the tile
      tc::mbar_wait(barM + bl, ((bl ? cnt1 : cnt0) - 1) & 1, &g_tc_error);
    }
    tc::tc_fence_after();
    unsigned char* obuf = sWk + ((gs - 1) & 1) * 32768;
    {
      float y[64];
      tc::tmem_ld64(tbase + lane_addr, y);
#pragma unroll
      for (int i = 0; i < 16; ++i)
        *reinterpret_cast<float4*>(obuf + off(i)) =
            make_float4(y[4 * i] + sBias[4 * i], y[4 * i + 1] + sBias[4 * i + 1],
                        y[4 * i + 2] + sBias[4 * i + 2], y[4 * i + 3] + sBias[4 * i + 3]);
    }
    tc::fence_async_smem();
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    if (leader) {
      tma::store_4d(&mapOut, obuf, 0, tx * 16, ty * 8, bb);
      tma::store_4d(&mapOut, obuf + 16384, 32, tx * 16, ty * 8, bb);
      tma::commit_group();
    }
  }
  if (leader) tma::wait_all();
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<128>(*tmem_slot);
}

}  // namespace b200mar

extern "C" int b200mar_cascade_proj_tc_fwd(const float* srcA, const float* srcB, const float* direct,
                                           const float* w, const float* bias, float* out,
                                           float* poolA, int B, int h, int wd, void* stream) {
  using namespace b200mar;
  if (B <= 0 || h <= 0 || wd <= 0 || !srcA || !direct) return -1;
  CasParams p{w, bias, poolA, B, h, wd, srcB ? 192 : 128, srcB ? 1 : 0};
  CUtensorMap mA, mB, mD, mO;
  if (int rc = tma::encode_grid(&mA, srcA, B, 2 * h, 2 * wd, 16, 8, 2)) return rc;
  if (int rc = tma::encode_grid(&mB, srcB ? srcB : srcA, B, 2 * h, 2 * wd, 16, 8, 2)) return rc;
  if (int rc = tma::encode_grid(&mD, direct, B, h, wd, 16, 8, 1)) return rc;
  if (int rc = tma::encode_grid(&mO, out, B, h, wd, 16, 8, 1)) return rc;
  const size_t smem = CasSmem::kBytes;
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(cascade_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    configured.done();
  }
  const int ntile = B * ((h + 7) / 8) * ((wd + 15) / 16);
  int grid = (ntile + kCasWorkers - 1) / kCasWorkers;
  if (grid > sm_count()) grid = sm_count();
  cascade_tc_kernel<<<grid, 128 * kCasWorkers, smem, (cudaStream_t)stream>>>(mA, mB, mD, mO, p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
