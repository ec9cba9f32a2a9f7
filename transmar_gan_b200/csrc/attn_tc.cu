// attn_tc.cu — fused N-gram window partition + scaled-cosine window attention + LN1 + residual,
// forward, on the 5th-gen tensor cores (tcgen05.mma kind::tf32, fp32 accumulators in TMEM).
//
// One persistent CTA per SM = two independent 128-thread workers sharing the resident weights
// (qkv 48 KB + proj 16 KB as chunk-major K-major operands, bias table, LN affine).  A worker
// streams tiles of TWO (shifted) windows = 128 tokens; thread t owns token row t = TMEM lane t.
// Per tile:
//   gather X = z + ctx (cyclic shift folded into the address math)            -> smem A operand
//   for each head h (QKV of head h+1 is issued while head h is in its epilogue):
//     [q_h|k_h|v_h] = X * Wqkv[h]^T        1 MMA 128x48x64 -> TMEM (double-buffered); the weight
//                                           tile is re-packed per head as [q16|k16|v16] rows
//     row-wise: +bias, F.normalize(q), F.normalize(k), q *= exp(min(logit_scale,ln100))*log2e
//               -> smem operands Q_h, K_h (K-major) and V_h^T (K-major, padded chunk stride)
//     S = Q_h K_h^T                         1 MMA 128x128x16 (both windows; the off-diagonal
//                                           64x64 blocks are never read)
//     row-wise: + relative-position bias (smem gather) + analytic shift mask, max, ex2, sum;
//               unnormalised P written back to TMEM over S (tcgen05.st)
//     O_h = P [V_h(w0) | V_h(w1)]           1 MMA 128x32x64, A operand = P from TMEM; each row
//                                           keeps the 16 columns of its own window
//   Y = O Wp^T (1 MMA 128x64x64), +bias, LayerNorm (a whole row per thread), + z, scatter.
// While one worker runs a CUDA-core epilogue the other owns the tensor pipe / load path.
//
// Reference semantics: components/window_utils.py:127-147, components/attention.py:70-101,
// components/transformer_blocks.py:63-112,117 (same contract as attn_fwd.cu, the fp32 SIMT kernel).
#include "common.cuh"
#include "attn_common.cuh"
#include "tc.cuh"

namespace b200mar {

B200MAR_DEFINE_TC_ERROR_GETTER(tc_error_attn)

constexpr float kLog2e = 1.4426950408889634f;
constexpr int kVLd = 132;                      // floats between successive 16-byte key-chunks of V^T
                                               // (32 rows x 4 + 4: == 4 mod 32, conflict-free scalar stores)

struct AttnTcSmem {                            // float offsets
  static constexpr int kWRows = 288;           // 6 heads x [q16 | k16 | v16] (4 heads: 192 used)
  static constexpr int oWqkv = 0;              // [288 rows][64 k] chunk-major (R = 288), per-head packed
  static constexpr int oWp = oWqkv + kWRows * 64; // [64 rows][64 k]  chunk-major (R = 64)
  static constexpr int oBqkv = oWp + 64 * 64;
  static constexpr int oBp = oBqkv + kWRows;
  static constexpr int oLnW = oBp + 64;
  static constexpr int oLnB = oLnW + 64;
  static constexpr int oScale = oLnB + 64;
  static constexpr int oBias = oScale + 8;     // [6][228], pre-multiplied by log2(e)
  static constexpr int oBar = oBias + 6 * 228; // 2 workers x 4 mbarriers (uint64) + tmem slot
  static constexpr int oWork = oBar + 32;
  // per worker
  static constexpr int wX = 0;                 // [128][64] chunk-major (R = 128): X, then O, then LN out
  static constexpr int wQ = wX + 128 * 64;     // [128][16] chunk-major
  static constexpr int wK = wQ + 128 * 16;     // [128][16] chunk-major
  static constexpr int wV = wK + 128 * 16;     // [V^T(w0) 16 d | V^T(w1) 16 d][64 keys], chunk stride kVLd
  static constexpr int wTok = wV + 16 * kVLd;
  static constexpr int kWorkFloats = wTok + 3 * 128;
  static constexpr int kFloats = oWork + 2 * kWorkFloats;
  static_assert(oWork % 32 == 0 && kWorkFloats % 32 == 0, "128-byte aligned worker blocks");
  static_assert(kFloats * 4 <= 232448, "attn_tc shared memory exceeds 227 KB");
};

// TMEM columns of one worker (256 of the CTA's 512)
constexpr uint32_t kTQ = 0;      // 2 x 48: [q_h | k_h | v_h] accumulators, double-buffered over heads
constexpr uint32_t kTS = 96;     // 128: S = Q K^T (both windows); P overwrites columns [0,64); Y at the end
constexpr uint32_t kTO = 224;    // 32: O_h, columns [0,16) from window 0's V, [16,32) from window 1's

template <int NH>
__global__ void __launch_bounds__(256, 1) attn_tc_fwd_kernel(AttnParams p) {
  using L = AttnTcSmem;
  constexpr int HD = kDim / NH, A = HD * NH;
  extern __shared__ __align__(128) float smem[];
  float* sWqkv = smem + L::oWqkv;
  float* sWp = smem + L::oWp;
  float* sBqkv = smem + L::oBqkv;
  float* sBp = smem + L::oBp;
  float* sLnW = smem + L::oLnW;
  float* sLnB = smem + L::oLnB;
  float* sScale = smem + L::oScale;
  float* sBias = smem + L::oBias;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L::oBar + 16);
  const int tid = threadIdx.x, lane = tid & 31;
  // warp-uniform ids (broadcast from lane 0 so that everything derived from them - smem / TMEM
  // addresses, MMA descriptors - stays in uniform registers)
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int wk = warp >> 2, wq = warp & 3;
  float* sW = smem + L::oWork + wk * L::kWorkFloats;
  float* sX = sW + L::wX;
  float* sQ = sW + L::wQ;
  float* sK = sW + L::wK;
  float* sV = sW + L::wV;
  int* sTok = reinterpret_cast<int*>(sW + L::wTok);
  int* sCwin = sTok + 128;
  int* sReg = sCwin + 128;
  uint64_t* barQ = bars + wk * 4;       // [2]
  uint64_t* barS = barQ + 2;
  uint64_t* barO = barQ + 3;

  // ---- stage the block's weights as K-major chunk-major operands (once per CTA)
  for (int idx = tid; idx < L::kWRows * 16; idx += 256) {
    const int n = idx % L::kWRows, kc = idx / L::kWRows;      // packed row n = h*48 + part*16 + d
    const int h = n / 48, part = (n % 48) >> 4, d = n & 15;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (h < NH && d < HD)
      v = *reinterpret_cast<const float4*>(p.qkv_w + (part * A + h * HD + d) * 64 + kc * 4);
    *reinterpret_cast<float4*>(sWqkv + kc * (L::kWRows * 4) + n * 4) = v;
  }
  for (int idx = tid; idx < 64 * 16; idx += 256) {
    const int n = idx & 63, kc = idx >> 6;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (kc * 4 < A) v = *reinterpret_cast<const float4*>(p.proj_w + n * A + kc * 4);
    *reinterpret_cast<float4*>(sWp + kc * 256 + n * 4) = v;
  }
  for (int n = tid; n < L::kWRows; n += 256) {                 // bias in the same packed order
    const int h = n / 48, part = (n % 48) >> 4, d = n & 15;
    sBqkv[n] = (h < NH && d < HD) ? p.qkv_b[part * A + h * HD + d] : 0.f;
  }
  if (tid < 64) { sBp[tid] = p.proj_b[tid]; sLnW[tid] = p.ln_w[tid]; sLnB[tid] = p.ln_b[tid]; }
  if (tid < NH) sScale[tid] = expf(fminf(p.logit_scale[tid], kLogitMax)) * kLog2e;
  for (int i = tid; i < NH * 225; i += 256) {
    const int h = i / 225, e = i - h * 225;
    sBias[h * 228 + e] = p.bias_table[e * NH + h] * kLog2e;
  }
  for (int i = tid; i < 2 * L::kWorkFloats; i += 256) smem[L::oWork + i] = 0.f;
  if (tid == 0) {
    for (int i = 0; i < 8; ++i) tc::mbar_init(bars + i, 1);
    tc::fence_barrier_init();
  }
  if (tid < 32) tc::tmem_alloc<512>(tmem_slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();

  const uint32_t tbase = *tmem_slot + wk * 256;
  const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
  const uint32_t xS = tc::smem_u32(sX), qS = tc::smem_u32(sQ), kS = tc::smem_u32(sK), vS = tc::smem_u32(sV);
  const uint32_t wqkvS = tc::smem_u32(sWqkv), wpS = tc::smem_u32(sWp);
  constexpr uint32_t kI32 = tc::idesc_tf32(128, 32), kI48 = tc::idesc_tf32(128, 48),
                     kI64 = tc::idesc_tf32(128, 64), kI128 = tc::idesc_tf32(128, 128);
  uint32_t phQ = 0, phS = 0, phO = 0;

  const Geom g = make_geom(p.B, p.ph, p.pw, p.shift);
  const int nwin = g.B * g.wh * g.ww;
  const int ntile = (nwin + 1) >> 1;
  const int row = wq * 32 + lane;                 // token row of this thread == its TMEM lane
  const int wsel = row >> 6, lrow = row & 63;     // window of the pair, token inside the window
  const int bias_base = ((lrow >> 3) * 15 + (lrow & 7)) + 112;   // rel_index(r, c) = bias_base - (15 ci + cj)

  // operand descriptors are built once; a K step of 8 advances the start-address field (>>4) by
  // two k-chunks
  const uint64_t dX = tc::smem_desc(xS, 2048, 128), dWqkv = tc::smem_desc(wqkvS, L::kWRows * 16, 128),
                 dQ = tc::smem_desc(qS, 2048, 128), dK = tc::smem_desc(kS, 2048, 128),
                 dV = tc::smem_desc(vS, kVLd * 4, 128), dWp = tc::smem_desc(wpS, 1024, 128);
  // [q_h | k_h | v_h] = X * Wqkv[rows of head h]^T -> TMEM buffer `buf` (issued by one thread)
  auto issue_qkv = [&](int h, int buf) {
    const uint64_t db = dWqkv + (uint64_t)(h * 48);
#pragma unroll
    for (int ks = 0; ks < 8; ++ks)
      tc::mma_tf32(tbase + kTQ + buf * 48, dX + ks * 256, db + ks * (2 * L::kWRows), kI48, ks > 0);
  };

  for (int tile = blockIdx.x * 2 + wk; tile < ntile; tile += gridDim.x * 2) {
    // ---- slot table of the two windows
    {
      const int win = tile * 2 + wsel;
      if (win < nwin) {
        const Slot s = slot_of(g, win, lrow);
        sTok[row] = (int)s.tok; sCwin[row] = s.cwin; sReg[row] = s.region;
      } else {
        sTok[row] = -1; sCwin[row] = 0; sReg[row] = 0;
      }
    }
    tc::worker_sync(1 + wk);
    // ---- gather X = z + ctx: coalesced 64-byte segments -> conflict-free 16-byte slots
    {
      const int q = lane >> 3, m8 = lane & 7;
      float4 zv[16], cv[16];
#pragma unroll
      for (int it = 0; it < 16; ++it) {                    // all 32 loads in flight before any use
        const int m = wq * 32 + (it >> 2) * 8 + m8, kc = (it & 3) * 4 + q;
        const int tok = sTok[m];
        zv[it] = cv[it] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tok >= 0) {
          zv[it] = __ldg(reinterpret_cast<const float4*>(p.z + (size_t)tok * 64 + kc * 4));
          cv[it] = __ldg(reinterpret_cast<const float4*>(p.ctx + (size_t)sCwin[m] * 64 + kc * 4));
        }
      }
#pragma unroll
      for (int it = 0; it < 16; ++it) {
        const int m = wq * 32 + (it >> 2) * 8 + m8, kc = (it & 3) * 4 + q;
        *reinterpret_cast<float4*>(sX + kc * 512 + m * 4) =
            make_float4(zv[it].x + cv[it].x, zv[it].y + cv[it].y, zv[it].z + cv[it].z, zv[it].w + cv[it].w);
      }
    }
    // ---- shift mask of this row as a 64-bit key set (analytic 9-region rule)
    uint32_t mlo = 0, mhi = 0;
    if (p.shift > 0) {
      const int rg = sReg[row];
      const int* reg = sReg + wsel * 64;
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        mlo |= (uint32_t)(reg[c] != rg) << c;
        mhi |= (uint32_t)(reg[c + 32] != rg) << c;
      }
    }
    tc::fence_async_smem();
    tc::worker_sync(1 + wk);
    if (wq == 0) {                                   // warp-uniform: one elected lane issues
      tc::tc_fence_after();
      if (tc::elect_one()) { issue_qkv(0, 0); tc::mma_commit(barQ); }
      __syncwarp();
    }

    // per-head attention outputs of this row; indexed by the (rolled) head loop, so the array
    // lives in local memory: 10-16 stores per head, one 64-float read-back per tile
    float o[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) o[i] = 0.f;
    float inv_prev = 0.f;

#pragma unroll 1
    for (int h = 0; h < NH; ++h) {
      const int buf = h & 1;
      // ---- wait for [q|k|v]_h, start head h+1's projection behind it
      tc::mbar_wait(barQ + buf, (phQ >> buf) & 1u, &g_tc_error);
      phQ ^= 1u << buf;
      tc::tc_fence_after();
      float qkv[48];
      tc::tmem_ld48(tbase + lane_addr + kTQ + buf * 48, qkv);
      float qs = 0.f, ks = 0.f;
      {
        const float* b = sBqkv + h * 48;               // packed [q16 | k16 | v16]; pad rows are exact zeros
#pragma unroll
        for (int d = 0; d < HD; ++d) {
          qkv[d] += b[d];
          qkv[16 + d] += b[16 + d];
          qkv[32 + d] += b[32 + d];
          qs = fmaf(qkv[d], qkv[d], qs);
          ks = fmaf(qkv[16 + d], qkv[16 + d], ks);
        }
      }
      const float qmul = sScale[h] / fmaxf(sqrtf(qs), kNormEps);
      const float kmul = 1.0f / fmaxf(sqrtf(ks), kNormEps);
#pragma unroll
      for (int d = 0; d < HD; ++d) { qkv[d] *= qmul; qkv[16 + d] *= kmul; }
      // ---- previous head's O (its PV MMAs read P from TMEM and V^T from smem: wait before reuse)
      if (h > 0) {
        tc::mbar_wait(barO, phO, &g_tc_error); phO ^= 1;
        tc::tc_fence_after();
        float ov[16];
        tc::tmem_ld16(tbase + lane_addr + kTO + wsel * 16, ov);
#pragma unroll
        for (int d = 0; d < HD; ++d) o[(h - 1) * HD + d] = ov[d] * inv_prev;
      }
      // ---- operands of head h: Q_h, K_h (rows = tokens), V_h^T (rows = d, keys contiguous)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        *reinterpret_cast<float4*>(sQ + c * 512 + row * 4) =
            make_float4(qkv[c * 4], qkv[c * 4 + 1], qkv[c * 4 + 2], qkv[c * 4 + 3]);
        *reinterpret_cast<float4*>(sK + c * 512 + row * 4) =
            make_float4(qkv[16 + c * 4], qkv[16 + c * 4 + 1], qkv[16 + c * 4 + 2], qkv[16 + c * 4 + 3]);
      }
      {
        float* vt = sV + (lrow >> 2) * kVLd + wsel * 64 + (lrow & 3);
#pragma unroll
        for (int d = 0; d < HD; ++d) vt[d * 4] = qkv[32 + d];
      }
      tc::fence_async_smem();
      tc::tc_fence_before();
      tc::worker_sync(1 + wk);
      if (wq == 0) {
        tc::tc_fence_after();
        if (tc::elect_one()) {
          tc::mma_tf32(tbase + kTS, dQ, dK, kI128, 0u);
          tc::mma_tf32(tbase + kTS, dQ + 256, dK + 256, kI128, 1u);
          tc::mma_commit(barS);
          if (h + 1 < NH) {                            // next head's projection, in the shadow of S_h
            issue_qkv(h + 1, buf ^ 1);
            tc::mma_commit(barQ + (buf ^ 1));
          }
        }
        __syncwarp();
      }
      tc::mbar_wait(barS, phS, &g_tc_error); phS ^= 1;
      tc::tc_fence_after();
      // ---- softmax of this row over its own window's 64 keys (log2 domain)
      {
        float s[64];
        tc::tmem_ld64(tbase + lane_addr + kTS + wsel * 64, s);
        const float* bias = sBias + h * 228 + bias_base;
#pragma unroll
        for (int c = 0; c < 64; ++c) s[c] += bias[-((c >> 3) * 15 + (c & 7))];
        if (mlo | mhi) {
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            if ((mlo >> c) & 1u) s[c] += -100.0f * kLog2e;
            if ((mhi >> c) & 1u) s[c + 32] += -100.0f * kLog2e;
          }
        }
        float mx = s[0];
#pragma unroll
        for (int c = 1; c < 64; ++c) mx = fmaxf(mx, s[c]);
        float sum4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < 64; ++c) { s[c] = tc::ex2_approx(s[c] - mx); sum4[c & 3] += s[c]; }
        inv_prev = 1.0f / ((sum4[0] + sum4[1]) + (sum4[2] + sum4[3]));
        tc::tmem_st64(tbase + lane_addr + kTS, s);       // unnormalised P over S columns [0,64)
        tc::tmem_wait_st();
      }
      tc::tc_fence_before();
      tc::worker_sync(1 + wk);
      if (wq == 0) {
        tc::tc_fence_after();
        if (tc::elect_one()) {
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            tc::mma_tf32_ts(tbase + kTO, tbase + kTS + ks * 8, dV + ks * (2 * kVLd * 4 / 16), kI32, ks > 0);
          tc::mma_commit(barO);
        }
        __syncwarp();
      }
    }
    // ---- last head's O, then the output projection
    {
      tc::mbar_wait(barO, phO, &g_tc_error); phO ^= 1;
      tc::tc_fence_after();
      float ov[16];
      tc::tmem_ld16(tbase + lane_addr + kTO + wsel * 16, ov);
#pragma unroll
      for (int d = 0; d < HD; ++d) o[(NH - 1) * HD + d] = ov[d] * inv_prev;
    }
#pragma unroll
    for (int kc = 0; kc < 16; ++kc)
      *reinterpret_cast<float4*>(sX + kc * 512 + row * 4) =
          make_float4(o[kc * 4], o[kc * 4 + 1], o[kc * 4 + 2], o[kc * 4 + 3]);
    tc::fence_async_smem();
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) tc::mma_tf32(tbase + kTS, dX + ks * 256, dWp + ks * 128, kI64, ks > 0);
        tc::mma_commit(barS);
      }
      __syncwarp();
    }
    tc::mbar_wait(barS, phS, &g_tc_error); phS ^= 1;
    tc::tc_fence_after();
    // ---- + proj bias, LayerNorm(norm1) of the row held in registers -> sX (own row)
    {
      float y[64];
      tc::tmem_ld64(tbase + lane_addr + kTS, y);
      float mean = 0.f;
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] += sBp[i]; mean += y[i]; }
      mean *= (1.0f / 64.f);
      float var = 0.f;
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] -= mean; var = fmaf(y[i], y[i], var); }
      const float rstd = 1.0f / sqrtf(var * (1.0f / 64.f) + kLnEps);
#pragma unroll
      for (int kc = 0; kc < 16; ++kc)
        *reinterpret_cast<float4*>(sX + kc * 512 + row * 4) =
            make_float4(y[kc * 4 + 0] * rstd * sLnW[kc * 4 + 0] + sLnB[kc * 4 + 0],
                        y[kc * 4 + 1] * rstd * sLnW[kc * 4 + 1] + sLnB[kc * 4 + 1],
                        y[kc * 4 + 2] * rstd * sLnW[kc * 4 + 2] + sLnB[kc * 4 + 2],
                        y[kc * 4 + 3] * rstd * sLnW[kc * 4 + 3] + sLnB[kc * 4 + 3]);
    }
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    // ---- residual + scatter back to the source token positions (same mapping as the gather)
    {
      const int q = lane >> 3, m8 = lane & 7;
      float4 zv[16];
#pragma unroll
      for (int it = 0; it < 16; ++it) {
        const int m = wq * 32 + (it >> 2) * 8 + m8, kc = (it & 3) * 4 + q;
        const int tok = sTok[m];
        zv[it] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tok >= 0) zv[it] = __ldg(reinterpret_cast<const float4*>(p.z + (size_t)tok * 64 + kc * 4));
      }
#pragma unroll
      for (int it = 0; it < 16; ++it) {
        const int m = wq * 32 + (it >> 2) * 8 + m8, kc = (it & 3) * 4 + q;
        const int tok = sTok[m];
        if (tok >= 0) {
          const float4 y = *reinterpret_cast<const float4*>(sX + kc * 512 + m * 4);
          *reinterpret_cast<float4*>(p.a + (size_t)tok * 64 + kc * 4) =
              make_float4(zv[it].x + y.x, zv[it].y + y.y, zv[it].z + y.z, zv[it].w + y.w);
        }
      }
    }
    tc::worker_sync(1 + wk);
  }
  tc::tc_fence_before();
  __syncthreads();
  if (tid < 32) tc::tmem_dealloc<512>(*tmem_slot);
}

template <int NH>
static int launch_attn_tc_fwd(const AttnParams& p, cudaStream_t st) {
  const size_t smem = AttnTcSmem::kFloats * sizeof(float);
  static bool configured = false;
  if (!configured) {
    B200MAR_CHECK(cudaFuncSetAttribute(attn_tc_fwd_kernel<NH>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = true;
  }
  const int nwin = p.B * (p.ph / kWin) * (p.pw / kWin);
  const int ntile = (nwin + 1) / 2;
  int grid = (ntile + 1) / 2;
  if (grid > sm_count()) grid = sm_count();
  attn_tc_fwd_kernel<NH><<<grid, 256, smem, st>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace b200mar

extern "C" int b200mar_attn_tc_fwd(const float* z, const float* ctx, const float* qkv_w,
                                   const float* qkv_b, const float* proj_w, const float* proj_b,
                                   const float* logit_scale, const float* bias_table,
                                   const float* ln_w, const float* ln_b, float* a, int B, int ph,
                                   int pw, int nh, int shift, void* stream) {
  using namespace b200mar;
  if (ph % kWin || pw % kWin || ph < kWin || pw < kWin || (shift != 0 && shift != kShift)) return -1;
  if ((long long)B * ph * pw >= (1ll << 31)) return -1;
  AttnParams p{};
  p.z = z; p.ctx = ctx; p.qkv_w = qkv_w; p.qkv_b = qkv_b; p.proj_w = proj_w; p.proj_b = proj_b;
  p.logit_scale = logit_scale; p.bias_table = bias_table; p.ln_w = ln_w; p.ln_b = ln_b;
  p.a = a; p.B = B; p.ph = ph; p.pw = pw; p.shift = shift;
  cudaStream_t st = (cudaStream_t)stream;
  if (nh == 6) return launch_attn_tc_fwd<6>(p, st);
  if (nh == 4) return launch_attn_tc_fwd<4>(p, st);
  return -1;
}
