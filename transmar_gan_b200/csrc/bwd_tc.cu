// bwd_tc.cu — tensor-core (tcgen05, BF16 operands, fp32 accumulate in TMEM) backward kernels shared by
// the NSTB blocks:
//
//   wgrad_tc_kernel     dW[out][in] += sum_t G[t][out] * IN[t][in]  (+ dbias[out] += sum_t G[t][out])
//                       The tokens are the K dimension of the MMA: a [128 tokens][C] chunk-major tile is
//                       read as an MN-major operand (= its transpose), so neither operand is ever
//                       transposed in memory; the partial dW of a persistent CTA stays in TMEM across all
//                       of its tiles and is flushed once with coalesced atomics.  The bias gradient comes
//                       out of the same MMAs through a block of ones appended to the IN tile (lane 64).
//
//   mlp_bwd_tc_kernel   backward of out = a + LN2(fc2(GELU(fc1(a)))) (components/transformer_blocks.py:119)
//                       with one thread per token row (TMEM lane = row): fc1/fc2 are recomputed, LayerNorm
//                       backward runs on the row held in registers, dg = dy W2 and da = dout + dh W1 are
//                       MMAs whose A operand is written to TMEM by the row threads and whose B operand is
//                       the SAME resident weight tile read MN-major.  The per-token operands of the weight
//                       gradients (a, g, dh, dy as bf16) are exported for wgrad_tc_kernel; the column sums
//                       (d_fc2_b, d_ln_w, d_ln_b) are reduced across the warp with a halving exchange.
//
// BF16 (fp32 range, 8-bit significand) is used for every backward operand: gradients span many orders of
// magnitude, so the forward kernels' fp16 + power-of-two scaling is not carried over.
#include <cstdlib>
#include "common.cuh"
#include "tc.cuh"
#include <cuda_bf16.h>
#include "bwd_common.cuh"
#include "tma.cuh"

namespace b200mar {

B200MAR_DEFINE_TC_ERROR_GETTER(tc_error_bwd)

// ================================================================================================
// wgrad_tc_kernel
struct WgradParams {
  const uint16_t* in;   // bf16 [n_tok][cin]
  const uint16_t* g;    // bf16 [n_tok][cout]
  float* dw;            // fp32 [cout][cin], accumulated with atomics
  float* dbias;         // fp32 [cout] or null (cin == 64 only)
  long long n_tok;
  int cin, cout;
  int ld, cin_valid;    // row stride of dw, input channels actually written
  int hs, ps, hd, A;    // hs > 0: per-head packed [q|k|v] columns (see launch_wgrad)
  int nstage;           // depth of the TMA ring (2 or 3)
};

constexpr int kWgStages = 3;      // ring depth (2 when three stages of the largest shape do not fit)

// One CTA per SM: the operand tiles arrive by TMA (3-D tensor maps that deliver a row-major bf16 matrix
// chunk-major, tma.cuh) through a 3-stage ring; one elected thread is producer and MMA issuer; the partial
// dW lives in TMEM until the CTA's last tile and is flushed with coalesced atomics.
__global__ void __launch_bounds__(128) wgrad_tc_kernel(const __grid_constant__ CUtensorMap mapIn,
                                                       const __grid_constant__ CUtensorMap mapG, WgradParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t full[kWgStages], empty[kWgStages], done;
  __shared__ uint32_t slot;
  __shared__ __align__(16) int sRowMap[256];     // accumulator column -> row of dw / dbias (-1: padding)
  const int tid = threadIdx.x, lane = tid & 31;
  for (int c = tid; c < p.cout; c += 128) {      // (per-head packed [q|k|v] columns: divisions done once, here)
    int orow = c;
    if (p.hs > 0) {
      const int h = c / p.hs, r = c - h * p.hs, part = r / p.ps, d = r - part * p.ps;
      orow = (part < 3 && d < p.hd) ? part * p.A + h * p.hd + d : -1;
    }
    sRowMap[c] = orow;
  }
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  // stage = [128][cin] IN tile | [128][8] ones (lane 64 of the accumulator = column sums of G) | [128][cout] G tile
  const uint32_t in_bytes = p.cin * 256, g_bytes = p.cout * 256, stage_bytes = in_bytes + 2048 + g_bytes;
  const int nstage = p.nstage;
  for (int s = 0; s < nstage; ++s)
    *reinterpret_cast<uint4*>(smem_raw + s * stage_bytes + in_bytes + tid * 16) =
        make_uint4(0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u);
  if (tid == 0) {
    for (int s = 0; s < kWgStages; ++s) { tc::mbar_init(full + s, 1); tc::mbar_init(empty + s, 1); }
    tc::mbar_init(&done, 1);
    tc::fence_barrier_init();
    tma::prefetch_desc(&mapIn); tma::prefetch_desc(&mapG);
  }
  if (warp == 0) tc::tmem_alloc<256>(&slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tbase = slot;
  const long long ntile = (p.n_tok + 127) / 128;
  const long long my = (ntile - blockIdx.x + gridDim.x - 1) / gridDim.x;     // tiles of this CTA
  if (warp == 0) {
    if (tc::elect_one()) {
      const uint32_t idesc = tc::idesc_bf16(128, p.cout, true, true);
      auto load = [&](long long i) {
        const int s = (int)(i % nstage);
        unsigned char* st = smem_raw + s * stage_bytes;
        const int r0 = (int)((blockIdx.x + i * gridDim.x) * 128);
        tma::mbar_expect_tx(full + s, in_bytes + g_bytes);
        tma::load_3d(&mapIn, full + s, st, 0, r0, 0);
        tma::load_3d(&mapG, full + s, st + in_bytes + 2048, 0, r0, 0);
      };
      for (long long i = 0; i < nstage && i < my; ++i) load(i);
      for (long long i = 0; i < my; ++i) {
        const int s = (int)(i % nstage);
        const uint32_t par = (uint32_t)((i / nstage) & 1);
        tc::mbar_wait(full + s, par, &g_tc_error);
        tc::tc_fence_after();
        unsigned char* st = smem_raw + s * stage_bytes;
        const uint64_t dA = tc::desc_mnmajor16(tc::smem_u32(st), 128),
                       dB = tc::desc_mnmajor16(tc::smem_u32(st + in_bytes + 2048), 128);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) tc::mma_f16(tbase, dA + ks * 16, dB + ks * 16, idesc, (i > 0 || ks > 0) ? 1u : 0u);
        tc::mma_commit(empty + s);
        if (i + nstage < my) {                       // refill this stage once its MMAs have read it
          tc::mbar_wait(empty + s, par, &g_tc_error);
          load(i + nstage);
        }
      }
      tc::mma_commit(&done);
    }
    __syncwarp();
  }
  tc::mbar_wait(&done, 0, &g_tc_error);
  tc::tc_fence_after();
  // ---- flush: lane = input channel, columns = output channels
  if (my > 0) {
    const int ci = warp * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
    // every CTA starts at a different 16-column group: the CTAs reach the flush together, and in the
    // same order all of them would queue on the same L2 addresses
    // (branch-free inside a group: the destination of element i is base + orow_i * stride for weight and bias
    // lanes alike, the 16 row indices of a group are read with four 16-byte loads, and the reductions are
    // predicated - a "load index, branch, compute address, reduce" chain per element cost 39 K cycles = 20 us
    // per launch, two thirds of a small launch)
    const int ngroup = p.cout >> 4;
    const bool isw = ci < p.cin_valid, isb = ci == 64 && p.cin == 64 && p.dbias;
    float* const dst = isw ? p.dw + ci : p.dbias;
    const int stride = isw ? p.ld : 1;
    int c0 = ((int)blockIdx.x % ngroup) << 4;
    for (int it = 0; it < ngroup; ++it) {
      float v[16];
      tc::tmem_ld16(tbase + lane_addr + c0, v);
      if (isw || isb) {
        int orow[16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int4 r = *reinterpret_cast<const int4*>(sRowMap + c0 + 4 * i);
          orow[4 * i] = r.x; orow[4 * i + 1] = r.y; orow[4 * i + 2] = r.z; orow[4 * i + 3] = r.w;
        }
#pragma unroll
        for (int i = 0; i < 16; ++i)                   // predicated reduction: no branch per element
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ge.s32 p, %2, 0;\n\t@p red.global.add.f32 [%0], %1;\n\t}"
                       :: "l"(dst + (long long)orow[i] * (long long)stride), "f"(v[i]), "r"(orow[i]) : "memory");
      }
      c0 += 16;
      if (c0 >= p.cout) c0 = 0;
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<256>(slot);
}

// ================================================================================================
// mlp_bwd_tc_kernel
struct MlpBwdTcParams {
  const float* a;       // [n_tok][64]  block-tail input (post-attention stream)
  const float* dout;    // [n_tok][64]  gradient of out = a + LN2(mlp(a))
  const float* fc1_w; const float* fc1_b; const float* fc2_w; const float* fc2_b; const float* ln_w;
  float* da;            // [n_tok][64]
  uint16_t* a16;        // bf16 exports for the weight gradients
  uint16_t* g16;        // [n_tok][128] GELU output
  uint16_t* dh16;       // [n_tok][128] gradient of the fc1 pre-activation
  uint16_t* dy16;       // [n_tok][64]  gradient of the fc2 output
  float* d_fc2_b; float* d_ln_w; float* d_ln_b;
  long long n_tok;
};

constexpr int kBwdWorkers = 2;

struct MlpBwdSmem {                                  // BYTE offsets
  static constexpr int oW1 = 0;                      // bf16 [128 hid rows][64 in ch], chunk-major R = 128
  static constexpr int oW2 = oW1 + 128 * 64 * 2;     // bf16 [64 out rows][128 hid ch], R = 64
  static constexpr int oB1 = oW2 + 64 * 128 * 2;     // fp32 [128]
  static constexpr int oB2 = oB1 + 512;              // fp32 [64]
  static constexpr int oLnW = oB2 + 256;             // fp32 [64]
  static constexpr int oBar = oLnW + 256;            // kBwdWorkers x 4 mbarriers + tmem slot
  static constexpr int kBytes = oBar + kBwdWorkers * 32 + 16;
};

__global__ void __launch_bounds__(128 * kBwdWorkers, 1) mlp_bwd_tc_kernel(MlpBwdTcParams p) {
  using L = MlpBwdSmem;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char* sW1 = smem_raw + L::oW1;
  unsigned char* sW2 = smem_raw + L::oW2;
  float* sB1 = reinterpret_cast<float*>(smem_raw + L::oB1);
  float* sB2 = reinterpret_cast<float*>(smem_raw + L::oB2);
  float* sLnW = reinterpret_cast<float*>(smem_raw + L::oLnW);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem_raw + L::oBar + kBwdWorkers * 32);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int wk = warp >> 2, wq = warp & 3;
  uint64_t* bar = bars + wk * 4;                     // [0] fc1  [1] fc2  [2] dg  [3] da

  // ---- stage the weights as bf16 chunk-major tiles
  for (int idx = tid; idx < 128 * 8; idx += 128 * kBwdWorkers) {          // fc1 [128][64]
    const int n = idx & 127, kc = idx >> 7;
    const float4* src = reinterpret_cast<const float4*>(p.fc1_w + n * 64 + kc * 8);
    const float4 x = src[0], y = src[1];
    *reinterpret_cast<uint4*>(sW1 + kc * 2048 + n * 16) =
        make_uint4(tc::pack_bf16(x.x, x.y), tc::pack_bf16(x.z, x.w), tc::pack_bf16(y.x, y.y), tc::pack_bf16(y.z, y.w));
  }
  for (int idx = tid; idx < 64 * 16; idx += 128 * kBwdWorkers) {          // fc2 [64][128]
    const int n = idx & 63, kc = idx >> 6;
    const float4* src = reinterpret_cast<const float4*>(p.fc2_w + n * 128 + kc * 8);
    const float4 x = src[0], y = src[1];
    *reinterpret_cast<uint4*>(sW2 + kc * 1024 + n * 16) =
        make_uint4(tc::pack_bf16(x.x, x.y), tc::pack_bf16(x.z, x.w), tc::pack_bf16(y.x, y.y), tc::pack_bf16(y.z, y.w));
  }
  if (tid < 128) sB1[tid] = p.fc1_b[tid];
  if (tid < 64) { sB2[tid] = p.fc2_b[tid]; sLnW[tid] = p.ln_w[tid]; }
  if (tid == 0) {
    for (int i = 0; i < 4 * kBwdWorkers; ++i) tc::mbar_init(bars + i, 1);
    tc::fence_barrier_init();
  }
  if (warp == 0) tc::tmem_alloc<512>(tmem_slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();

  // TMEM columns of a worker (256):
  //   [0,128)   D1 = fc1 pre-activation -> { g0 [0,32) | gp0 [32,64) | g1 [64,96) | gp1 [96,128) } packed bf16
  //             (g: GELU output, A operand of fc2; gp: GELU derivative, later overwritten by dh, A operand of da)
  //             -> dg quarters 0,1 over g0, g1
  //   [128,192) D2 = fc2 output -> dg quarters 2,3 -> da accumulator
  //   [192,224) a packed (A operand of fc1) -> dy packed (A operand of dg)
  const uint32_t tbase = *tmem_slot + wk * 256;
  const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
  const uint32_t tl = tbase + lane_addr;
  constexpr uint32_t cD2 = 128, cA = 192;
  constexpr uint32_t kI128 = tc::idesc_bf16(128, 128), kI64 = tc::idesc_bf16(128, 64),
                     kI32mn = tc::idesc_bf16(128, 32, false, true), kI64mn = tc::idesc_bf16(128, 64, false, true);
  const uint64_t dW1k = tc::desc_kmajor16(tc::smem_u32(sW1), 128), dW2k = tc::desc_kmajor16(tc::smem_u32(sW2), 64),
                 dW1mn = tc::desc_mnmajor16(tc::smem_u32(sW1), 128), dW2mn = tc::desc_mnmajor16(tc::smem_u32(sW2), 64);
  const int row = wq * 32 + lane;
  uint32_t ph = 0;
  float db2_0 = 0.f, db2_1 = 0.f, dg_0 = 0.f, dg_1 = 0.f, dbe_0 = 0.f, dbe_1 = 0.f;   // columns 2*lane, 2*lane+1

  const long long ntile = (p.n_tok + 127) / 128;
  for (long long tile = (long long)blockIdx.x * kBwdWorkers + wk; tile < ntile;
       tile += (long long)gridDim.x * kBwdWorkers) {
    const long long tok = tile * 128 + row;
    const bool ok = tok < p.n_tok;
    // ---- P0: a row -> packed bf16 A operand in TMEM + export
    {
      uint32_t ap[32];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ok) v = __ldg(reinterpret_cast<const float4*>(p.a + (size_t)tok * 64) + i);
        ap[2 * i] = tc::pack_bf16(v.x, v.y);
        ap[2 * i + 1] = tc::pack_bf16(v.z, v.w);
      }
      tc::tmem_st32(tl + cA, reinterpret_cast<const float(&)[32]>(ap));
      if (ok) {
        uint4* dst = reinterpret_cast<uint4*>(p.a16 + (size_t)tok * 64);
#pragma unroll
        for (int i = 0; i < 8; ++i) dst[i] = make_uint4(ap[4 * i], ap[4 * i + 1], ap[4 * i + 2], ap[4 * i + 3]);
      }
    }
    tc::tmem_wait_st();
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
          tc::mma_f16_ts(tbase, tbase + cA + ks * 8, dW1k + ks * 256, kI128, ks > 0);
        tc::mma_commit(bar + 0);
      }
      __syncwarp();
    }
    tc::mbar_wait(bar + 0, ph, &g_tc_error);
    tc::tc_fence_after();
    // ---- P1: GELU and its derivative, two halves of 64 hidden channels
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
      float v[64];
      tc::tmem_ld64(tl + half * 64, v);
      uint32_t gq[32], gpq[32];
      const float* b1 = sB1 + half * 64;
#pragma unroll
      for (int c = 0; c < 64; c += 2) {
        float g0, g1, d0, d1;
        gelu_and_grad(v[c] + b1[c], g0, d0);
        gelu_and_grad(v[c + 1] + b1[c + 1], g1, d1);
        gq[c >> 1] = tc::pack_bf16(g0, g1);
        gpq[c >> 1] = tc::pack_bf16(d0, d1);
      }
      tc::tmem_st32(tl + half * 64, reinterpret_cast<const float(&)[32]>(gq));
      tc::tmem_st32(tl + half * 64 + 32, reinterpret_cast<const float(&)[32]>(gpq));
      if (ok) {
        uint4* dst = reinterpret_cast<uint4*>(p.g16 + (size_t)tok * 128 + half * 64);
#pragma unroll
        for (int i = 0; i < 8; ++i) dst[i] = make_uint4(gq[4 * i], gq[4 * i + 1], gq[4 * i + 2], gq[4 * i + 3]);
      }
    }
    tc::tmem_wait_st();
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          tc::mma_f16_ts(tbase + cD2, tbase + (ks < 4 ? ks * 8 : 64 + (ks - 4) * 8), dW2k + ks * 128, kI64, ks > 0);
        tc::mma_commit(bar + 1);
      }
      __syncwarp();
    }
    tc::mbar_wait(bar + 1, ph, &g_tc_error);
    tc::tc_fence_after();
    // ---- P2: LayerNorm backward on the row -> dy (packed bf16 A operand + export), column sums
    {
      float y[64];
      tc::tmem_ld64(tl + cD2, y);
      float m4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] += sB2[i]; m4[i & 3] += y[i]; }
      const float mean = ((m4[0] + m4[1]) + (m4[2] + m4[3])) * (1.0f / 64.f);
      float v4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] -= mean; v4[i & 3] = fmaf(y[i], y[i], v4[i & 3]); }
      const float rstd = 1.0f / sqrtf(((v4[0] + v4[1]) + (v4[2] + v4[3])) * (1.0f / 64.f) + kLnEps);
      float d[64];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ok) v = __ldg(reinterpret_cast<const float4*>(p.dout + (size_t)tok * 64) + i);
        d[4 * i] = v.x; d[4 * i + 1] = v.y; d[4 * i + 2] = v.z; d[4 * i + 3] = v.w;
      }
      float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 64; ++i) {
        y[i] *= rstd;                                  // normalised row
        const float t = d[i] * sLnW[i];
        s1[i & 3] += t;
        s2[i & 3] = fmaf(t, y[i], s2[i & 3]);
      }
      const float m1 = ((s1[0] + s1[1]) + (s1[2] + s1[3])) * (1.0f / 64.f);
      const float m2 = ((s2[0] + s2[1]) + (s2[2] + s2[3])) * (1.0f / 64.f);
      warp_colsum64([&](int i) { return d[i] * y[i]; }, lane, dg_0, dg_1);      // d_ln_w
      warp_colsum64([&](int i) { return d[i]; }, lane, dbe_0, dbe_1);           // d_ln_b
#pragma unroll
      for (int i = 0; i < 64; ++i) y[i] = rstd * (fmaf(d[i], sLnW[i], -m1) - y[i] * m2);   // dy
      warp_colsum64([&](int i) { return y[i]; }, lane, db2_0, db2_1);           // d_fc2_b
      uint32_t dq[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) dq[i] = tc::pack_bf16(y[2 * i], y[2 * i + 1]);
      tc::tmem_st32(tl + cA, reinterpret_cast<const float(&)[32]>(dq));
      if (ok) {
        uint4* dst = reinterpret_cast<uint4*>(p.dy16 + (size_t)tok * 64);
#pragma unroll
        for (int i = 0; i < 8; ++i) dst[i] = make_uint4(dq[4 * i], dq[4 * i + 1], dq[4 * i + 2], dq[4 * i + 3]);
      }
    }
    tc::tmem_wait_st();
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    // ---- dg = dy W2 in four quarters of 32 hidden channels (B = the fc2 tile read MN-major)
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t dcol = q == 0 ? 0u : (q == 1 ? 64u : (q == 2 ? 128u : 160u));
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)
            tc::mma_f16_ts(tbase + dcol, tbase + cA + ks * 8, dW2mn + q * 256 + ks * 16, kI32mn, ks > 0);
        }
        tc::mma_commit(bar + 2);
      }
      __syncwarp();
    }
    tc::mbar_wait(bar + 2, ph, &g_tc_error);
    tc::tc_fence_after();
    // ---- P3: dh = dg * GELU'(pre) -> packed bf16 over gp (A operand of da) + export
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t dcol = q == 0 ? 0u : (q == 1 ? 64u : (q == 2 ? 128u : 160u));
      const uint32_t gcol = q == 0 ? 32u : (q == 1 ? 48u : (q == 2 ? 96u : 112u));
      float dg[32], gp[16];
      tc::tmem_ld32(tl + dcol, dg);
      tc::tmem_ld16(tl + gcol, gp);
      uint32_t hq[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const uint32_t u = __float_as_uint(gp[i]);
        hq[i] = tc::pack_bf16(dg[2 * i] * bf16_lo(u), dg[2 * i + 1] * bf16_hi(u));
      }
      tc::tmem_st16(tl + gcol, reinterpret_cast<const float(&)[16]>(hq));
      if (ok) {
        uint4* dst = reinterpret_cast<uint4*>(p.dh16 + (size_t)tok * 128 + q * 32);
#pragma unroll
        for (int i = 0; i < 4; ++i) dst[i] = make_uint4(hq[4 * i], hq[4 * i + 1], hq[4 * i + 2], hq[4 * i + 3]);
      }
    }
    tc::tmem_wait_st();
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    // ---- da = dout + dh W1 (B = the fc1 tile read MN-major)
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          tc::mma_f16_ts(tbase + cD2, tbase + (ks < 4 ? 32 + ks * 8 : 96 + (ks - 4) * 8), dW1mn + ks * 16, kI64mn, ks > 0);
        tc::mma_commit(bar + 3);
      }
      __syncwarp();
    }
    tc::mbar_wait(bar + 3, ph, &g_tc_error);
    tc::tc_fence_after();
    {
      float v[64];
      tc::tmem_ld64(tl + cD2, v);
      if (ok) {
        const float4* src = reinterpret_cast<const float4*>(p.dout + (size_t)tok * 64);
        float4* dst = reinterpret_cast<float4*>(p.da + (size_t)tok * 64);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float4 g = __ldg(src + i);
          dst[i] = make_float4(g.x + v[4 * i], g.y + v[4 * i + 1], g.z + v[4 * i + 2], g.w + v[4 * i + 3]);
        }
      }
    }
    ph ^= 1;
  }
  // ---- flush the column sums (lane l of every warp holds columns 2l, 2l+1)
  atomicAdd(p.d_fc2_b + 2 * lane, db2_0); atomicAdd(p.d_fc2_b + 2 * lane + 1, db2_1);
  atomicAdd(p.d_ln_w + 2 * lane, dg_0);   atomicAdd(p.d_ln_w + 2 * lane + 1, dg_1);
  atomicAdd(p.d_ln_b + 2 * lane, dbe_0);  atomicAdd(p.d_ln_b + 2 * lane + 1, dbe_1);
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(*tmem_slot);
}

// ================================================================================================
// mlp_bwd_tma_kernel - the same arithmetic with every global access of the tile moved by the TMA engine.
// The plain kernel above reads / writes each token row from its own thread (32 different 128-byte lines per warp
// instruction, 144 such instructions per thread and tile): ~3 K LSU cycles per warp and tile, i.e. ~85 us of its
// 130 us at 262 k tokens.  Here the a and dout tiles arrive as 128 B-swizzled boxes (prefetched one tile ahead), the
// four bf16 exports and da are staged in shared memory in the same swizzled row layout (a row thread touches only
// its own 128-byte rows, conflict-free) and leave as bulk tensor stores.  Three 32 KB buffers per worker, each
// reused twice per tile; a buffer written in phase k is stored after the worker barrier that ends phase k and is
// known to be free again after the barrier that ends phase k + 1 (the leader waits for the bulk read before it):
//   bufA: a tile -> a16 (over box 0, own rows) -> next a tile (loaded after the barrier of phase 1)
//   bufD: dout tile (row kept in registers) -> dy16 (over box 0) -> da tile -> next dout tile
//   bufG: g16 -> dh16
struct MlpBwdTmaSmem {
  static constexpr int oWork = ((MlpBwdSmem::oBar + kBwdWorkers * 64 + 16 + 1023) / 1024) * 1024;   // 8 mbarrier slots per worker
  static constexpr int kWorkBytes = 3 * 32768;
  static constexpr int kBytes = oWork + kBwdWorkers * kWorkBytes;
  static_assert(kBytes <= 232448, "mlp_bwd_tma shared memory exceeds 227 KB");
};

__global__ void __launch_bounds__(128 * kBwdWorkers, 1)
mlp_bwd_tma_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapDout,
                   const __grid_constant__ CUtensorMap mapDa, const __grid_constant__ CUtensorMap mapA16,
                   const __grid_constant__ CUtensorMap mapG16, const __grid_constant__ CUtensorMap mapDh16,
                   const __grid_constant__ CUtensorMap mapDy16, MlpBwdTcParams p) {
  using L = MlpBwdSmem;
  using LT = MlpBwdTmaSmem;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* sW1 = smem_raw + L::oW1;
  unsigned char* sW2 = smem_raw + L::oW2;
  float* sB1 = reinterpret_cast<float*>(smem_raw + L::oB1);
  float* sB2 = reinterpret_cast<float*>(smem_raw + L::oB2);
  float* sLnW = reinterpret_cast<float*>(smem_raw + L::oLnW);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem_raw + L::oBar + kBwdWorkers * 64);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int wk = warp >> 2, wq = warp & 3;
  uint64_t* bar = bars + wk * 8;                     // [0] fc1  [1] fc2  [2] dg  [3] da  [4] a tile  [5] dout tile
  uint64_t* barA = bar + 4;
  uint64_t* barD = bar + 5;
  unsigned char* bufA = smem_raw + LT::oWork + wk * LT::kWorkBytes;   // a tile (fp32) -> a16 staging (box-0 region)
  unsigned char* bufD = bufA + 32768;                                 // dout tile -> dy16 staging -> da tile
  unsigned char* bufG = bufA + 65536;                                 // g16 staging -> dh16 staging
  const bool leader = wq == 0 && lane == 0;          // bulk groups are per thread: one fixed thread
  // byte offset of 16-byte chunk i (8 bf16) of row r in a bf16 [128 rows][128 B] tile (128 B swizzle)
  auto off16 = [](int r, int i) -> uint32_t { return (uint32_t)(r * 128 + ((i ^ (r & 7)) << 4)); };

  // ---- stage the weights as bf16 chunk-major tiles
  for (int idx = tid; idx < 128 * 8; idx += 128 * kBwdWorkers) {          // fc1 [128][64]
    const int n = idx & 127, kc = idx >> 7;
    const float4* src = reinterpret_cast<const float4*>(p.fc1_w + n * 64 + kc * 8);
    const float4 x = src[0], y = src[1];
    *reinterpret_cast<uint4*>(sW1 + kc * 2048 + n * 16) =
        make_uint4(tc::pack_bf16(x.x, x.y), tc::pack_bf16(x.z, x.w), tc::pack_bf16(y.x, y.y), tc::pack_bf16(y.z, y.w));
  }
  for (int idx = tid; idx < 64 * 16; idx += 128 * kBwdWorkers) {          // fc2 [64][128]
    const int n = idx & 63, kc = idx >> 6;
    const float4* src = reinterpret_cast<const float4*>(p.fc2_w + n * 128 + kc * 8);
    const float4 x = src[0], y = src[1];
    *reinterpret_cast<uint4*>(sW2 + kc * 1024 + n * 16) =
        make_uint4(tc::pack_bf16(x.x, x.y), tc::pack_bf16(x.z, x.w), tc::pack_bf16(y.x, y.y), tc::pack_bf16(y.z, y.w));
  }
  if (tid < 128) sB1[tid] = p.fc1_b[tid];
  if (tid < 64) { sB2[tid] = p.fc2_b[tid]; sLnW[tid] = p.ln_w[tid]; }
  if (tid == 0) {
    for (int i = 0; i < 8 * kBwdWorkers; ++i) tc::mbar_init(bars + i, 1);
    tc::fence_barrier_init();
    tma::prefetch_desc(&mapA); tma::prefetch_desc(&mapDout); tma::prefetch_desc(&mapDa); tma::prefetch_desc(&mapA16);
    tma::prefetch_desc(&mapG16); tma::prefetch_desc(&mapDh16); tma::prefetch_desc(&mapDy16);
  }
  if (warp == 0) tc::tmem_alloc<512>(tmem_slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();

  // TMEM columns of a worker (256):
  //   [0,128)   D1 = fc1 pre-activation -> { g0 [0,32) | gp0 [32,64) | g1 [64,96) | gp1 [96,128) } packed bf16
  //             (g: GELU output, A operand of fc2; gp: GELU derivative, later overwritten by dh, A operand of da)
  //             -> dg quarters 0,1 over g0, g1
  //   [128,192) D2 = fc2 output -> dg quarters 2,3 -> da accumulator
  //   [192,224) a packed (A operand of fc1) -> dy packed (A operand of dg)
  const uint32_t tbase = *tmem_slot + wk * 256;
  const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
  const uint32_t tl = tbase + lane_addr;
  constexpr uint32_t cD2 = 128, cA = 192;
  constexpr uint32_t kI128 = tc::idesc_bf16(128, 128), kI64 = tc::idesc_bf16(128, 64),
                     kI32mn = tc::idesc_bf16(128, 32, false, true), kI64mn = tc::idesc_bf16(128, 64, false, true);
  const uint64_t dW1k = tc::desc_kmajor16(tc::smem_u32(sW1), 128), dW2k = tc::desc_kmajor16(tc::smem_u32(sW2), 64),
                 dW1mn = tc::desc_mnmajor16(tc::smem_u32(sW1), 128), dW2mn = tc::desc_mnmajor16(tc::smem_u32(sW2), 64);
  const int row = wq * 32 + lane;
  uint32_t ph = 0;
  float db2_0 = 0.f, db2_1 = 0.f, dg_0 = 0.f, dg_1 = 0.f, dbe_0 = 0.f, dbe_1 = 0.f;   // columns 2*lane, 2*lane+1

  const long long ntile = (p.n_tok + 127) / 128;
  const long long tstride = (long long)gridDim.x * kBwdWorkers;
  auto load_tile = [&](const CUtensorMap* m, uint64_t* b, unsigned char* dst, long long t) {
    tma::mbar_expect_tx(b, 32768);
    tma::load_2d(m, b, dst, 0, (int)(t * 128));
    tma::load_2d(m, b, dst + 16384, 32, (int)(t * 128));
  };
  if (leader && (long long)blockIdx.x * kBwdWorkers + wk < ntile) {
    load_tile(&mapA, barA, bufA, (long long)blockIdx.x * kBwdWorkers + wk);
    load_tile(&mapDout, barD, bufD, (long long)blockIdx.x * kBwdWorkers + wk);
  }
  for (long long tile = (long long)blockIdx.x * kBwdWorkers + wk; tile < ntile; tile += tstride) {
    const int r0 = (int)(tile * 128);
    // ---- P0: a row -> packed bf16 A operand in TMEM + export
    {
      uint32_t ap[32];
      tc::mbar_wait(barA, ph, &g_tc_error);             // the a tile (rows past n_tok are zero-filled by the TMA)
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float4 v = *reinterpret_cast<const float4*>(bufA + tma::tile_off(row, i));
        ap[2 * i] = tc::pack_bf16(v.x, v.y);
        ap[2 * i + 1] = tc::pack_bf16(v.z, v.w);
      }
      tc::tmem_st32(tl + cA, reinterpret_cast<const float(&)[32]>(ap));
      // a16 export: the bf16 row over this thread's OWN fp32 row bytes of box 0 (already in registers)
#pragma unroll
      for (int i = 0; i < 8; ++i)
        *reinterpret_cast<uint4*>(bufA + off16(row, i)) = make_uint4(ap[4 * i], ap[4 * i + 1], ap[4 * i + 2], ap[4 * i + 3]);
    }
    tc::tmem_wait_st();
    tc::fence_async_smem();
    tc::tc_fence_before();
    if (leader) tma::wait_read_all();                  // every store of the previous tile has left bufG / bufD
    tc::worker_sync(1 + wk);
    if (leader) { tma::store_2d(&mapA16, bufA, 0, r0); tma::commit_group(); }
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
          tc::mma_f16_ts(tbase, tbase + cA + ks * 8, dW1k + ks * 256, kI128, ks > 0);
        tc::mma_commit(bar + 0);
      }
      __syncwarp();
    }
    tc::mbar_wait(bar + 0, ph, &g_tc_error);
    tc::tc_fence_after();
    // ---- P1: GELU and its derivative, two halves of 64 hidden channels
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
      float v[64];
      tc::tmem_ld64(tl + half * 64, v);
      uint32_t gq[32], gpq[32];
      const float* b1 = sB1 + half * 64;
#pragma unroll
      for (int c = 0; c < 64; c += 2) {
        float g0, g1, d0, d1;
        gelu_and_grad(v[c] + b1[c], g0, d0);
        gelu_and_grad(v[c + 1] + b1[c + 1], g1, d1);
        gq[c >> 1] = tc::pack_bf16(g0, g1);
        gpq[c >> 1] = tc::pack_bf16(d0, d1);
      }
      tc::tmem_st32(tl + half * 64, reinterpret_cast<const float(&)[32]>(gq));
      tc::tmem_st32(tl + half * 64 + 32, reinterpret_cast<const float(&)[32]>(gpq));
#pragma unroll
      for (int i = 0; i < 8; ++i)
        *reinterpret_cast<uint4*>(bufG + half * 16384 + off16(row, i)) =
            make_uint4(gq[4 * i], gq[4 * i + 1], gq[4 * i + 2], gq[4 * i + 3]);
    }
    tc::tmem_wait_st();
    tc::fence_async_smem();
    tc::tc_fence_before();
    if (leader) tma::wait_read_all();                  // the a16 store has left bufA
    tc::worker_sync(1 + wk);
    if (leader) {
      tma::store_2d(&mapG16, bufG, 0, r0); tma::store_2d(&mapG16, bufG + 16384, 64, r0); tma::commit_group();
      if (tile + tstride < ntile) load_tile(&mapA, barA, bufA, tile + tstride);
    }
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          tc::mma_f16_ts(tbase + cD2, tbase + (ks < 4 ? ks * 8 : 64 + (ks - 4) * 8), dW2k + ks * 128, kI64, ks > 0);
        tc::mma_commit(bar + 1);
      }
      __syncwarp();
    }
    tc::mbar_wait(bar + 1, ph, &g_tc_error);
    tc::tc_fence_after();
    // ---- P2: LayerNorm backward on the row -> dy (packed bf16 A operand + export), column sums
    float d[64];                                       // the dout row: read once, reused for da at the tile's end
    {
      float y[64];
      tc::tmem_ld64(tl + cD2, y);
      float m4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] += sB2[i]; m4[i & 3] += y[i]; }
      const float mean = ((m4[0] + m4[1]) + (m4[2] + m4[3])) * (1.0f / 64.f);
      float v4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] -= mean; v4[i & 3] = fmaf(y[i], y[i], v4[i & 3]); }
      const float rstd = 1.0f / sqrtf(((v4[0] + v4[1]) + (v4[2] + v4[3])) * (1.0f / 64.f) + kLnEps);
      tc::mbar_wait(barD, ph, &g_tc_error);             // the dout tile
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float4 v = *reinterpret_cast<const float4*>(bufD + tma::tile_off(row, i));
        d[4 * i] = v.x; d[4 * i + 1] = v.y; d[4 * i + 2] = v.z; d[4 * i + 3] = v.w;
      }
      float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 64; ++i) {
        y[i] *= rstd;                                  // normalised row
        const float t = d[i] * sLnW[i];
        s1[i & 3] += t;
        s2[i & 3] = fmaf(t, y[i], s2[i & 3]);
      }
      const float m1 = ((s1[0] + s1[1]) + (s1[2] + s1[3])) * (1.0f / 64.f);
      const float m2 = ((s2[0] + s2[1]) + (s2[2] + s2[3])) * (1.0f / 64.f);
      warp_colsum64([&](int i) { return d[i] * y[i]; }, lane, dg_0, dg_1);      // d_ln_w
      warp_colsum64([&](int i) { return d[i]; }, lane, dbe_0, dbe_1);           // d_ln_b
#pragma unroll
      for (int i = 0; i < 64; ++i) y[i] = rstd * (fmaf(d[i], sLnW[i], -m1) - y[i] * m2);   // dy
      warp_colsum64([&](int i) { return y[i]; }, lane, db2_0, db2_1);           // d_fc2_b
      uint32_t dq[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) dq[i] = tc::pack_bf16(y[2 * i], y[2 * i + 1]);
      tc::tmem_st32(tl + cA, reinterpret_cast<const float(&)[32]>(dq));
      // dy16 export over this thread's own dout row bytes of box 0 (the row stays in registers)
#pragma unroll
      for (int i = 0; i < 8; ++i)
        *reinterpret_cast<uint4*>(bufD + off16(row, i)) = make_uint4(dq[4 * i], dq[4 * i + 1], dq[4 * i + 2], dq[4 * i + 3]);
    }
    tc::tmem_wait_st();
    tc::fence_async_smem();
    tc::tc_fence_before();
    if (leader) tma::wait_read_all();                  // the g16 store has left bufG
    tc::worker_sync(1 + wk);
    if (leader) { tma::store_2d(&mapDy16, bufD, 0, r0); tma::commit_group(); }
    // ---- dg = dy W2 in four quarters of 32 hidden channels (B = the fc2 tile read MN-major)
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t dcol = q == 0 ? 0u : (q == 1 ? 64u : (q == 2 ? 128u : 160u));
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)
            tc::mma_f16_ts(tbase + dcol, tbase + cA + ks * 8, dW2mn + q * 256 + ks * 16, kI32mn, ks > 0);
        }
        tc::mma_commit(bar + 2);
      }
      __syncwarp();
    }
    tc::mbar_wait(bar + 2, ph, &g_tc_error);
    tc::tc_fence_after();
    // ---- P3: dh = dg * GELU'(pre) -> packed bf16 over gp (A operand of da) + export
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t dcol = q == 0 ? 0u : (q == 1 ? 64u : (q == 2 ? 128u : 160u));
      const uint32_t gcol = q == 0 ? 32u : (q == 1 ? 48u : (q == 2 ? 96u : 112u));
      float dg[32], gp[16];
      tc::tmem_ld32(tl + dcol, dg);
      tc::tmem_ld16(tl + gcol, gp);
      uint32_t hq[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const uint32_t u = __float_as_uint(gp[i]);
        hq[i] = tc::pack_bf16(dg[2 * i] * bf16_lo(u), dg[2 * i + 1] * bf16_hi(u));
      }
      tc::tmem_st16(tl + gcol, reinterpret_cast<const float(&)[16]>(hq));
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<uint4*>(bufG + (q >> 1) * 16384 + off16(row, (q & 1) * 4 + i)) =
            make_uint4(hq[4 * i], hq[4 * i + 1], hq[4 * i + 2], hq[4 * i + 3]);
    }
    tc::tmem_wait_st();
    tc::fence_async_smem();
    tc::tc_fence_before();
    if (leader) tma::wait_read_all();                  // the dy16 store has left bufD
    tc::worker_sync(1 + wk);
    if (leader) { tma::store_2d(&mapDh16, bufG, 0, r0); tma::store_2d(&mapDh16, bufG + 16384, 64, r0); tma::commit_group(); }
    // ---- da = dout + dh W1 (B = the fc1 tile read MN-major)
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          tc::mma_f16_ts(tbase + cD2, tbase + (ks < 4 ? 32 + ks * 8 : 96 + (ks - 4) * 8), dW1mn + ks * 16, kI64mn, ks > 0);
        tc::mma_commit(bar + 3);
      }
      __syncwarp();
    }
    tc::mbar_wait(bar + 3, ph, &g_tc_error);
    tc::tc_fence_after();
    {
      float v[64];
      tc::tmem_ld64(tl + cD2, v);
#pragma unroll
      for (int i = 0; i < 16; ++i)
        *reinterpret_cast<float4*>(bufD + tma::tile_off(row, i)) =
            make_float4(d[4 * i] + v[4 * i], d[4 * i + 1] + v[4 * i + 1], d[4 * i + 2] + v[4 * i + 2], d[4 * i + 3] + v[4 * i + 3]);
    }
    tc::fence_async_smem();
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    if (leader) {
      tma::store_2d(&mapDa, bufD, 0, r0); tma::store_2d(&mapDa, bufD + 16384, 32, r0); tma::commit_group();
      if (tile + tstride < ntile) {
        tma::wait_read_all();                          // the da store has left bufD
        load_tile(&mapDout, barD, bufD, tile + tstride);
      }
    }
    ph ^= 1;
  }
  if (leader) tma::wait_all();
  // ---- flush the column sums (lane l of every warp holds columns 2l, 2l+1)
  atomicAdd(p.d_fc2_b + 2 * lane, db2_0); atomicAdd(p.d_fc2_b + 2 * lane + 1, db2_1);
  atomicAdd(p.d_ln_w + 2 * lane, dg_0);   atomicAdd(p.d_ln_w + 2 * lane + 1, dg_1);
  atomicAdd(p.d_ln_b + 2 * lane, dbe_0);  atomicAdd(p.d_ln_b + 2 * lane + 1, dbe_1);
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(*tmem_slot);
}

int launch_wgrad(const uint16_t* in16, const uint16_t* g16, float* dw, float* dbias, long long n_tok,
                 int cin, int cout, int ld, int cin_valid, int hs, int ps, int hd, int A, cudaStream_t st) {
  if (n_tok <= 0 || (cin != 64 && cin != 128) || cout % 32 || cout < 64 || cout > 256) return -1;
  if (dbias && cin != 64) return -1;
  if (cin_valid <= 0 || cin_valid > cin || ld < cin_valid) return -1;
  const size_t stage = (size_t)cin * 256 + 2048 + (size_t)cout * 256;
  const int nstage = kWgStages * stage <= 224 * 1024 ? kWgStages : 2;
  WgradParams p{in16, g16, dw, dbias, n_tok, cin, cout, ld, cin_valid, hs, ps, hd, A, nstage};
  const size_t smem = nstage * stage;
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    configured.done();
  }
  if (smem > 224 * 1024) return -1;
  CUtensorMap mIn, mG;
  if (int rc = tma::encode_rows16_chunkmajor(&mIn, in16, n_tok, cin)) return rc;
  if (int rc = tma::encode_rows16_chunkmajor(&mG, g16, n_tok, cout)) return rc;
  const long long ntile = (n_tok + 127) / 128;
  long long grid = (ntile + 3) / 4;                  // >= 4 tiles per CTA: the flush is ~cout*cin atomics per CTA
  if (grid > sm_count()) grid = sm_count();
  wgrad_tc_kernel<<<(int)grid, 128, smem, st>>>(mIn, mG, p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace b200mar

extern "C" int b200mar_wgrad_tc(const void* in16, const void* g16, float* dw, float* dbias, long long n_tok,
                                int cin, int cout, void* stream) {
  using namespace b200mar;
  return launch_wgrad((const uint16_t*)in16, (const uint16_t*)g16, dw, dbias, n_tok, cin, cout, cin, cin, 0, 0, 0, 0,
                      (cudaStream_t)stream);
}

extern "C" int b200mar_mlp_bwd_tc(const float* a, const float* dout, const float* fc1_w, const float* fc1_b,
                                  const float* fc2_w, const float* fc2_b, const float* ln_w, const float* ln_b,
                                  float* da, float* d_fc1_w, float* d_fc1_b, float* d_fc2_w, float* d_fc2_b,
                                  float* d_ln_w, float* d_ln_b, void* scratch, long long n_tok, void* stream) {
  using namespace b200mar;
  (void)ln_b;
  if (n_tok <= 0 || n_tok % kTok || !scratch) return -1;
  uint16_t* s = (uint16_t*)scratch;                  // n_tok * 384 bf16
  MlpBwdTcParams p{};
  p.a = a; p.dout = dout; p.fc1_w = fc1_w; p.fc1_b = fc1_b; p.fc2_w = fc2_w; p.fc2_b = fc2_b; p.ln_w = ln_w;
  p.da = da; p.a16 = s; p.g16 = s + n_tok * 64; p.dh16 = s + n_tok * 192; p.dy16 = s + n_tok * 320;
  p.d_fc2_b = d_fc2_b; p.d_ln_w = d_ln_w; p.d_ln_b = d_ln_b; p.n_tok = n_tok;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = MlpBwdSmem::kBytes;
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(mlp_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured.done();
  }
  const long long ntile = (n_tok + 127) / 128;
  long long grid = (ntile + kBwdWorkers - 1) / kBwdWorkers;
  if (grid > sm_count()) grid = sm_count();
  static const bool plain = getenv("B200MAR_MLP_BWD_PLAIN") != nullptr;   // development knob: the plain-load kernel
  CUtensorMap mA, mDo, mDa, mA16, mG16, mDh16, mDy16;
  if (!plain && n_tok * 1 <= 0x7fffffffLL &&
      !tma::encode_rows64(&mA, a, n_tok) && !tma::encode_rows64(&mDo, dout, n_tok) && !tma::encode_rows64(&mDa, da, n_tok) &&
      !tma::encode_rows64_16(&mA16, p.a16, n_tok) && !tma::encode_rows16_box64(&mG16, p.g16, n_tok, 128) &&
      !tma::encode_rows16_box64(&mDh16, p.dh16, n_tok, 128) && !tma::encode_rows64_16(&mDy16, p.dy16, n_tok)) {
    static DeviceOnce configured_tma;
    if (configured_tma.first()) {
      B200MAR_CHECK(cudaFuncSetAttribute(mlp_bwd_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)MlpBwdTmaSmem::kBytes));
      configured_tma.done();
    }
    mlp_bwd_tma_kernel<<<(int)grid, 128 * kBwdWorkers, MlpBwdTmaSmem::kBytes, st>>>(mA, mDo, mDa, mA16, mG16, mDh16, mDy16, p);
  } else {
    mlp_bwd_tc_kernel<<<(int)grid, 128 * kBwdWorkers, smem, st>>>(p);
  }
  B200MAR_CHECK(cudaGetLastError());
  if (int rc = launch_wgrad(p.a16, p.dh16, d_fc1_w, d_fc1_b, n_tok, 64, 128, 64, 64, 0, 0, 0, 0, st)) return rc;   // dW1 [128][64], db1
  if (int rc = launch_wgrad(p.g16, p.dy16, d_fc2_w, nullptr, n_tok, 128, 64, 128, 128, 0, 0, 0, 0, st)) return rc;  // dW2 [64][128]
  return 0;
}
