// attn_bwd_tc.cu — backward of the fused N-gram window partition + scaled-cosine window attention +
// LN1 + residual on the tcgen05 tensor cores.  Forward being differentiated (components/attention.py:70-101,
// components/transformer_blocks.py:117, components/window_utils.py:127-147):
//   X = gather(z) + ctx ; [q|k|v] = X Wqkv^T + b ; qn = q/max(|q|,eps), kn likewise
//   S = s_h qn kn^T + bias_h + mask ; P = softmax(S) ; O_h = P v ; Y = O Wp^T + bp ; a = z + scatter(LN1(Y))
//
// Two row-per-thread kernels (tile = two windows = 128 token rows, thread t = row t = TMEM lane t) and
// the generic weight-gradient kernel of bwd_tc.cu:
//
//   attn_bwd_core_kernel<NH>   pass 1 recomputes the forward (same fp16 / power-of-two-scaled operands as
//       attn_tc.cu for X Wqkv^T and Q'K'^T, so P matches the forward) to get O and Y; LayerNorm backward on
//       the row gives dY, dO = dY Wp.  Pass 2 walks the heads again: S and dP = dO_h V^T come from one
//       MMA group, dS = P (dP - rowdot) stays in registers, and the three "transposed" products
//       dV = P^T dO_h, dkn = dS^T qn, dqn = dS kn use the row-major P / dS / qn / kn / dO tiles the row
//       threads write with plain 16-byte stores, read MN-major by the MMA (tokens as the K dimension).
//       F.normalize backward, the relative-position-bias scatter (shared-memory reductions), the
//       logit-scale gradient (qn . dqn) and the LN affine column sums are fp32.  Exports, in window
//       order, the bf16 operands of the weight gradients: X, O, dY and the per-head packed d[q|k|v].
//   attn_bwd_dx_kernel         dX = d[q|k|v] Wqkv (A operand from TMEM, B = the packed weight tile read
//       MN-major), dz = da + scatter(dX), dctx += per-(window, quadrant) sums of dX.
//   wgrad_tc (x2)              d_qkv_w / d_qkv_b from (X, d[q|k|v]), d_proj_w / d_proj_b from (O, dY).
#include "common.cuh"
#include "attn_common.cuh"
#include "tc.cuh"
#include "bwd_common.cuh"
#include <cuda_fp16.h>
#include <cuda_bf16.h>

namespace b200mar {

B200MAR_DEFINE_TC_ERROR_GETTER(tc_error_attn_bwd)

// Optional phase profile (compile with -DB200MAR_ABWD_PROF): cycles of one thread (block 0, thread 32)
// between successive PROF(i) marks, summed per phase i; read with b200mar_abwd_prof().
#ifdef B200MAR_ABWD_PROF
__device__ unsigned long long g_abwd_prof[32];
#define PROF_DECL __shared__ unsigned long long s_prof[32]; if (threadIdx.x < 32) s_prof[threadIdx.x] = 0; __syncthreads(); \
  long long prof_t = clock64(); const bool prof_on = blockIdx.x == 0 && threadIdx.x == 32;
#define PROF(i) do { if (prof_on) { const long long t_ = clock64(); s_prof[i] += (unsigned long long)(t_ - prof_t); prof_t = clock64(); } } while (0)
#define PROF_FLUSH do { if (prof_on) for (int i_ = 0; i_ < 32; ++i_) g_abwd_prof[i_] += s_prof[i_]; } while (0)
#else
#define PROF_DECL
#define PROF(i)
#define PROF_FLUSH
#endif

constexpr float kLog2eB = 1.4426950408889634f;
constexpr int kABWorkers = 2;
constexpr int kBiasLdB = 228;

struct AttnBwdTcParams {
  AttnParams a;          // inputs, da, d_logit_scale, d_bias_table, d_ln_w, d_ln_b
  uint16_t* x16;         // bf16 [ntile*128][64]   X = z + ctx, window order
  uint16_t* o16;         // bf16 [ntile*128][64]   attention output before the projection (cols >= A zero)
  const uint16_t* o16_saved;   // the same, saved by attn_tc_fwd_save: pass 1 is skipped when given
  uint16_t* dy16;        // bf16 [ntile*128][64]   gradient of the projection output
  uint16_t* dqkv16;      // bf16 [ntile*128][192]  per-head packed d[q|k|v]
};

template <int NH> struct BCfg {
  static constexpr int HS = NH == 6 ? 32 : 48;     // packed rows / accumulator columns per head
  static constexpr int PS = NH == 6 ? 10 : 16;     // stride of q, k, v inside a head's block
  static constexpr int HD = kDim / NH;
  static constexpr int A = HD * NH;
};

struct AttnBwdSmem {                           // BYTE offsets
  static constexpr int oWqkv = 0;              // fp16 [192 packed rows][64 ch], chunk-major R = 192
  static constexpr int oWp = oWqkv + 192 * 64 * 2;      // bf16 [64 out rows][64 in ch], R = 64
  static constexpr int oBqkv = oWp + 64 * 64 * 2;       // fp32 [192] packed order
  static constexpr int oBp = oBqkv + 192 * 4;
  static constexpr int oLnW = oBp + 256;
  static constexpr int oScale = oLnW + 256;    // fp32 [8]: scale_h * log2e ; [6]: inverse of the Wqkv scale
  static constexpr int oScaleRaw = oScale + 32;         // fp32 [8]: scale_h
  static constexpr int oDScale = oScaleRaw + 32;        // fp32 [8]: accumulators of sum dS.cos
  static constexpr int oBias = oDScale + 32;   // fp32 [6][228] bias * log2e
  static constexpr int oDBias = oBias + 6 * kBiasLdB * 4;   // fp32 [2 workers x 2 windows][6][228] accumulators
  static constexpr int oBar = oDBias + 2 * kABWorkers * 6 * kBiasLdB * 4;    // workers x 4 mbarriers + tmem slot
  static constexpr int oWork = ((oBar + kABWorkers * 32 + 16 + 127) / 128) * 128;
  static constexpr int wX = 0;                 // fp16 [128][64]  R = 128
  static constexpr int wQ = wX + 16384;        // fp16 [128][32]  R = 128   Q' = [q|0] / [0|q]
  static constexpr int wK = wQ + 8192;         // fp16 [64][32]   R = 64    K' = [k(w0)|k(w1)]
  static constexpr int wV = wK + 4096;         // bf16 [64][32]   V' = [v(w0)|v(w1)]
  static constexpr int wKb = wV + 4096;        // bf16 [64][32]   kn
  static constexpr int wQT = wKb + 4096;       // bf16 [64][32]   qn (rows = query slot)
  static constexpr int wDOT = wQT + 4096;      // bf16 [64][32]   dO_h
  static constexpr int wPT = wDOT + 4096;      // bf16 [64][128]  P   (rows = query slot, ch = (window, key))
  static constexpr int wST = wPT + 16384;      // bf16 [64][128]  dS
  static constexpr int wReg = wST + 16384;     // int [128]
  static constexpr int kWorkBytes = wReg + 512;
  static constexpr int kBytes = oWork + kABWorkers * kWorkBytes;
  static_assert(oBar % 8 == 0, "mbarrier alignment");
  static_assert(kBytes <= 232448, "attn_bwd_tc shared memory exceeds 227 KB");
};

// TMEM columns of one worker (256)
constexpr uint32_t cQKV = 0;     // 48: [q|k|v]_h accumulator; pass 2: dV
constexpr uint32_t cS = 48;      // 64: S -> P / dS packed (32); Y and dO accumulators between the passes
constexpr uint32_t cP2 = 112;    // 64: pass 1: O_h (32); pass 2: dP (64) -> dqn (32) | dkn (32)
constexpr uint32_t cOp = 176;    // 32: O packed (A of Y) -> dY packed (A of dO)
constexpr uint32_t cDOp = 208;   // 32: dO packed (whole row), read per head in pass 2
constexpr uint32_t cDOh = 240;   // 16: [dO_h|0] / [0|dO_h] packed (A of dP)

template <int NH>
__global__ void __launch_bounds__(128 * kABWorkers, 1) attn_bwd_core_kernel(AttnBwdTcParams pp) {
  using L = AttnBwdSmem;
  constexpr int HS = BCfg<NH>::HS, PS = BCfg<NH>::PS, HD = BCfg<NH>::HD, A = BCfg<NH>::A;
  const AttnParams& p = pp.a;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __half* sWqkv = reinterpret_cast<__half*>(smem_raw + L::oWqkv);
  unsigned char* sWp = smem_raw + L::oWp;
  float* sBqkv = reinterpret_cast<float*>(smem_raw + L::oBqkv);
  float* sBp = reinterpret_cast<float*>(smem_raw + L::oBp);
  float* sLnW = reinterpret_cast<float*>(smem_raw + L::oLnW);
  float* sScale = reinterpret_cast<float*>(smem_raw + L::oScale);
  float* sScaleRaw = reinterpret_cast<float*>(smem_raw + L::oScaleRaw);
  float* sDScale = reinterpret_cast<float*>(smem_raw + L::oDScale);
  float* sBias = reinterpret_cast<float*>(smem_raw + L::oBias);
  float* sDBias = reinterpret_cast<float*>(smem_raw + L::oDBias);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem_raw + L::oBar + kABWorkers * 32);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int wk = warp >> 2, wq = warp & 3;
  unsigned char* sW = smem_raw + L::oWork + wk * L::kWorkBytes;
  unsigned char* sX = sW + L::wX;
  unsigned char* sQ = sW + L::wQ;
  unsigned char* sK = sW + L::wK;
  unsigned char* sV = sW + L::wV;
  unsigned char* sKb = sW + L::wKb;
  unsigned char* sQT = sW + L::wQT;
  unsigned char* sDOT = sW + L::wDOT;
  unsigned char* sPT = sW + L::wPT;
  unsigned char* sST = sW + L::wST;
  int* sReg = reinterpret_cast<int*>(sW + L::wReg);
  uint64_t* barQ = bars + wk * 4;
  uint64_t* barS = barQ + 1;
  uint64_t* barO = barQ + 2;

  // ---- stage the weights (once per CTA); the qkv matrix exactly as attn_tc.cu does
  __shared__ uint32_t sWmax;
  if (tid == 0) sWmax = 0u;
  __syncthreads();
  {
    float m = 0.f;
    for (int i = tid; i < 3 * A * 64; i += 128 * kABWorkers) m = fmaxf(m, fabsf(p.qkv_w[i]));
    atomicMax(&sWmax, __float_as_uint(m));
  }
  __syncthreads();
  uint32_t we = (sWmax >> 23) & 0xFFu;
  we = we < 1u ? 127u : (we > 253u ? 253u : we);
  const float wsc = __uint_as_float((254u - we) << 23);
  for (int idx = tid; idx < 192 * 8; idx += 128 * kABWorkers) {
    const int n = idx % 192, kc = idx / 192;
    const int h = n / HS, part = (n % HS) / PS, d = (n % HS) % PS;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (h < NH && part < 3 && d < HD) {
      const float4* src = reinterpret_cast<const float4*>(p.qkv_w + (part * A + h * HD + d) * 64 + kc * 8);
      const float4 a = src[0], b = src[1];
      v = make_uint4(tc::pack_f16(a.x * wsc, a.y * wsc), tc::pack_f16(a.z * wsc, a.w * wsc),
                     tc::pack_f16(b.x * wsc, b.y * wsc), tc::pack_f16(b.z * wsc, b.w * wsc));
    }
    *reinterpret_cast<uint4*>(sWqkv + (size_t)kc * (192 * 8) + n * 8) = v;
  }
  for (int idx = tid; idx < 64 * 8; idx += 128 * kABWorkers) {          // proj [64 out][A in] -> bf16 [64][64]
    const int n = idx & 63, kc = idx >> 6;
    uint32_t w[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k0 = kc * 8 + 2 * j;
      w[j] = tc::pack_bf16(k0 < A ? p.proj_w[n * A + k0] : 0.f, k0 + 1 < A ? p.proj_w[n * A + k0 + 1] : 0.f);
    }
    *reinterpret_cast<uint4*>(sWp + kc * 1024 + n * 16) = make_uint4(w[0], w[1], w[2], w[3]);
  }
  for (int n = tid; n < 192; n += 128 * kABWorkers) {
    const int h = n / HS, part = (n % HS) / PS, d = (n % HS) % PS;
    sBqkv[n] = (h < NH && part < 3 && d < HD) ? p.qkv_b[part * A + h * HD + d] : 0.f;
  }
  if (tid < 64) { sBp[tid] = p.proj_b[tid]; sLnW[tid] = p.ln_w[tid]; }
  if (tid < 8) {
    const float sc = tid < NH ? expf(fminf(p.logit_scale[tid], kLogitMax)) : 0.f;
    sScaleRaw[tid] = sc;
    sScale[tid] = tid == 6 ? __uint_as_float(we << 23) : sc * kLog2eB;
    sDScale[tid] = 0.f;
  }
  for (int i = tid; i < 6 * kBiasLdB; i += 128 * kABWorkers) {
    const int h = i / kBiasLdB, e = i - h * kBiasLdB;
    sBias[i] = (h < NH && e < 225) ? p.bias_table[e * NH + h] * kLog2eB : 0.f;
  }
  for (int i = tid; i < 2 * kABWorkers * 6 * kBiasLdB; i += 128 * kABWorkers) sDBias[i] = 0.f;
  if (tid == 0) {
    for (int i = 0; i < 4 * kABWorkers; ++i) tc::mbar_init(bars + i, 1);
    tc::fence_barrier_init();
  }
  if (warp == 0) tc::tmem_alloc<512>(tmem_slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();

  const uint32_t tbase = *tmem_slot + wk * 256;
  const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
  const uint32_t tl = tbase + lane_addr;
  constexpr uint32_t kIqkv = tc::idesc_f16(128, HS), kIS = tc::idesc_f16(128, 64),
                     kB64 = tc::idesc_bf16(128, 64), kB64bmn = tc::idesc_bf16(128, 64, false, true),
                     kB32bmn = tc::idesc_bf16(128, 32, false, true), kB32mm = tc::idesc_bf16(128, 32, true, true);
  const uint64_t dX = tc::desc_kmajor16(tc::smem_u32(sX), 128),
                 dWqkv = tc::desc_kmajor16(tc::smem_u32(sWqkv), 192),
                 dQ = tc::desc_kmajor16(tc::smem_u32(sQ), 128), dK = tc::desc_kmajor16(tc::smem_u32(sK), 64),
                 dVk = tc::desc_kmajor16(tc::smem_u32(sV), 64), dVmn = tc::desc_mnmajor16(tc::smem_u32(sV), 64),
                 dKbmn = tc::desc_mnmajor16(tc::smem_u32(sKb), 64), dQTmn = tc::desc_mnmajor16(tc::smem_u32(sQT), 64),
                 dDOTmn = tc::desc_mnmajor16(tc::smem_u32(sDOT), 64), dPTmn = tc::desc_mnmajor16(tc::smem_u32(sPT), 64),
                 dSTmn = tc::desc_mnmajor16(tc::smem_u32(sST), 64),
                 dWpk = tc::desc_kmajor16(tc::smem_u32(sWp), 64), dWpmn = tc::desc_mnmajor16(tc::smem_u32(sWp), 64);
  uint32_t phQ = 0, phS = 0, phO = 0;

  const Geom g = make_geom(p.B, p.ph, p.pw, p.shift);
  const int nwin = g.B * g.wh * g.ww;
  const int ntile = (nwin + 1) >> 1;
  const int row = wq * 32 + lane;
  const int wsel = wq >> 1, lrow = row & 63;
  const int bias_base = ((lrow >> 3) * 15 + (lrow & 7)) + 112;   // rel_index(r, c) = bias_base - (15 ci + cj)
  float dgam0 = 0.f, dgam1 = 0.f, dbet0 = 0.f, dbet1 = 0.f;      // LN1 affine gradients, columns 2 lane, 2 lane + 1

  // one elected lane of warp 0 of the worker issues; `f` emits the MMAs and the commit
  auto issue = [&](auto f) {
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) f();
      __syncwarp();
    }
  };

  PROF_DECL
  for (int tile = blockIdx.x * kABWorkers + wk; tile < ntile; tile += gridDim.x * kABWorkers) {
    PROF(0);
    const int win = tile * 2 + wsel;
    const bool wvalid = win < nwin;
    const size_t erow = (size_t)tile * 128 + row;            // row of the window-order exports
    long long tok = 0;
    float xinv;
    // ---------------------------------------------------------------- X = z + ctx
    {
      float x[64];
      if (wvalid) {
        const Slot sl = slot_of(g, win, lrow);
        tok = sl.tok;
        sReg[row] = sl.region;
        const float4* cp = reinterpret_cast<const float4*>(p.ctx + (size_t)sl.cwin * 64);
        const float4* zp = reinterpret_cast<const float4*>(p.z + (size_t)tok * 64);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float4 c = __ldg(cp + i), z = __ldg(zp + i);
          x[4 * i] = c.x + z.x; x[4 * i + 1] = c.y + z.y; x[4 * i + 2] = c.z + z.z; x[4 * i + 3] = c.w + z.w;
        }
      } else {
        sReg[row] = 0;
#pragma unroll
        for (int i = 0; i < 64; ++i) x[i] = 0.f;
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
      xinv = __uint_as_float(e << 23) * sScale[6];
#pragma unroll
      for (int kc = 0; kc < 8; ++kc)
        *reinterpret_cast<uint4*>(sX + kc * 2048 + row * 16) =
            make_uint4(tc::pack_f16(x[8 * kc] * sc, x[8 * kc + 1] * sc), tc::pack_f16(x[8 * kc + 2] * sc, x[8 * kc + 3] * sc),
                       tc::pack_f16(x[8 * kc + 4] * sc, x[8 * kc + 5] * sc), tc::pack_f16(x[8 * kc + 6] * sc, x[8 * kc + 7] * sc));
      if (wvalid) {
        uint4* dst = reinterpret_cast<uint4*>(pp.x16 + erow * 64);
#pragma unroll
        for (int kc = 0; kc < 8; ++kc)
          dst[kc] = make_uint4(tc::pack_bf16(x[8 * kc], x[8 * kc + 1]), tc::pack_bf16(x[8 * kc + 2], x[8 * kc + 3]),
                               tc::pack_bf16(x[8 * kc + 4], x[8 * kc + 5]), tc::pack_bf16(x[8 * kc + 6], x[8 * kc + 7]));
      }
      // the other window's half of Q' stays zero for the whole tile
      const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
      *reinterpret_cast<uint4*>(sQ + (2 - 2 * wsel) * 2048 + row * 16) = z4;
      *reinterpret_cast<uint4*>(sQ + (3 - 2 * wsel) * 2048 + row * 16) = z4;
    }
    tc::fence_async_smem();
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
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

    // [q|k|v]_h of this row: MMA, bias, F.normalize.  qn, kn: unit vectors; iq, ik: 1/max(|.|,eps);
    // also writes the operands every pass needs: Q' (scaled by s_h log2e), K', V'.
    auto head_qkv = [&](int h, float (&qn)[16], float (&kn)[16], float (&vv)[16], float& iq, float& ik,
                        bool& qok, bool& kok) {
      issue([&] {
        const uint64_t db = dWqkv + (uint64_t)(h * HS);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) tc::mma_f16(tbase + cQKV, dX + ks * 256, db + ks * 384, kIqkv, ks > 0);
        tc::mma_commit(barQ);
      });
      tc::mbar_wait(barQ, phQ, &g_tc_error); phQ ^= 1;
      tc::tc_fence_after();
      float qkv[HS];
      if constexpr (HS == 32) tc::tmem_ld32(tl + cQKV, qkv);
      else tc::tmem_ld48(tl + cQKV, qkv);
      const float* b = sBqkv + h * HS;
      float qs = 0.f, ks2 = 0.f;
#pragma unroll
      for (int d = 0; d < 16; ++d) {
        if (d < HD) {
          qn[d] = fmaf(qkv[d], xinv, b[d]);
          kn[d] = fmaf(qkv[PS + d], xinv, b[PS + d]);
          vv[d] = fmaf(qkv[2 * PS + d], xinv, b[2 * PS + d]);
          qs = fmaf(qn[d], qn[d], qs);
          ks2 = fmaf(kn[d], kn[d], ks2);
        } else {
          qn[d] = 0.f; kn[d] = 0.f; vv[d] = 0.f;
        }
      }
      iq = rsqrtf(fmaxf(qs, kNormEps * kNormEps));
      ik = rsqrtf(fmaxf(ks2, kNormEps * kNormEps));
      qok = qs >= kNormEps * kNormEps;
      kok = ks2 >= kNormEps * kNormEps;
      uint32_t qp[8], kp[8], vp[8];
#pragma unroll
      for (int d = 0; d < 16; ++d) { qn[d] *= iq; kn[d] *= ik; }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        qp[j] = tc::pack_f16(qn[2 * j] * sScale[h], qn[2 * j + 1] * sScale[h]);
        kp[j] = tc::pack_f16(kn[2 * j], kn[2 * j + 1]);
        vp[j] = tc::pack_bf16(vv[2 * j], vv[2 * j + 1]);
      }
      *reinterpret_cast<uint4*>(sQ + (2 * wsel) * 2048 + row * 16) = make_uint4(qp[0], qp[1], qp[2], qp[3]);
      *reinterpret_cast<uint4*>(sQ + (2 * wsel + 1) * 2048 + row * 16) = make_uint4(qp[4], qp[5], qp[6], qp[7]);
      *reinterpret_cast<uint4*>(sK + (2 * wsel) * 1024 + lrow * 16) = make_uint4(kp[0], kp[1], kp[2], kp[3]);
      *reinterpret_cast<uint4*>(sK + (2 * wsel + 1) * 1024 + lrow * 16) = make_uint4(kp[4], kp[5], kp[6], kp[7]);
      *reinterpret_cast<uint4*>(sV + (2 * wsel) * 1024 + lrow * 16) = make_uint4(vp[0], vp[1], vp[2], vp[3]);
      *reinterpret_cast<uint4*>(sV + (2 * wsel + 1) * 1024 + lrow * 16) = make_uint4(vp[4], vp[5], vp[6], vp[7]);
    };
    // softmax of the row: s (MMA output, log2 domain without bias) -> normalised probabilities in place
    auto softmax_row = [&](int h, float (&s)[64]) {
      const float* bias = sBias + h * kBiasLdB + bias_base;
#pragma unroll
      for (int c = 0; c < 64; ++c) {
        s[c] += bias[-((c >> 3) * 15 + (c & 7))];
        if (((c < 32 ? mlo >> c : mhi >> (c - 32)) & 1u)) s[c] += -100.0f * kLog2eB;
      }
      float mx4[4] = {s[0], s[1], s[2], s[3]};
#pragma unroll
      for (int c = 4; c < 64; ++c) mx4[c & 3] = fmaxf(mx4[c & 3], s[c]);
      const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      float sum4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < 64; ++c) { s[c] = tc::ex2_approx(s[c] - mx); sum4[c & 3] += s[c]; }
      const float inv = 1.0f / ((sum4[0] + sum4[1]) + (sum4[2] + sum4[3]));
#pragma unroll
      for (int c = 0; c < 64; ++c) s[c] *= inv;
    };

    PROF(1);                                                 // X tile, masks
    // ================================================================ pass 1: O, Y, dY, dO
    if (pp.o16_saved) {
      // ---- O row as saved by the forward -> packed bf16 A operand of Y = O Wp^T
      uint32_t op[32];
      const uint4* src = reinterpret_cast<const uint4*>(pp.o16_saved + erow * 64);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (wvalid) v = __ldg(src + i);
        op[4 * i] = v.x; op[4 * i + 1] = v.y; op[4 * i + 2] = v.z; op[4 * i + 3] = v.w;
      }
      tc::tmem_st32(tl + cOp, reinterpret_cast<const float(&)[32]>(op));
      tc::tmem_wait_st();
    } else {
    float o[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) o[i] = 0.f;
#pragma unroll 1
    for (int h = 0; h < NH; ++h) {
      float qn[16], kn[16], vv[16], iq, ik; bool qok, kok;
      head_qkv(h, qn, kn, vv, iq, ik, qok, kok);
      tc::fence_async_smem();
      tc::tc_fence_before();
      tc::worker_sync(1 + wk);
      issue([&] {
        tc::mma_f16(tbase + cS, dQ, dK, kIS, 0u);
        tc::mma_f16(tbase + cS, dQ + 256, dK + 128, kIS, 1u);
        tc::mma_commit(barS);
      });
      tc::mbar_wait(barS, phS, &g_tc_error); phS ^= 1;
      tc::tc_fence_after();
      {
        float s[64];
        tc::tmem_ld64(tl + cS, s);
        softmax_row(h, s);
        uint32_t pk[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) pk[i] = tc::pack_bf16(s[2 * i], s[2 * i + 1]);
        tc::tmem_st32(tl + cS, reinterpret_cast<const float(&)[32]>(pk));
        tc::tmem_wait_st();
      }
      tc::tc_fence_before();
      tc::worker_sync(1 + wk);
      issue([&] {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) tc::mma_f16_ts(tbase + cP2, tbase + cS + ks * 8, dVmn + ks * 16, kB32bmn, ks > 0);
        tc::mma_commit(barO);
      });
      tc::mbar_wait(barO, phO, &g_tc_error); phO ^= 1;
      tc::tc_fence_after();
      float ov[16];
      tc::tmem_ld16(tl + cP2 + wsel * 16, ov);
#pragma unroll
      for (int d = 0; d < HD; ++d) o[h * HD + d] = ov[d];
    }
    // ---- O row -> packed bf16 A operand + export; Y = O Wp^T
    {
      uint32_t op[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) op[i] = tc::pack_bf16(o[2 * i], o[2 * i + 1]);
      tc::tmem_st32(tl + cOp, reinterpret_cast<const float(&)[32]>(op));
      if (wvalid) {
        uint4* dst = reinterpret_cast<uint4*>(pp.o16 + erow * 64);
#pragma unroll
        for (int i = 0; i < 8; ++i) dst[i] = make_uint4(op[4 * i], op[4 * i + 1], op[4 * i + 2], op[4 * i + 3]);
      }
      tc::tmem_wait_st();
    }
    }
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    issue([&] {
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) tc::mma_f16_ts(tbase + cS, tbase + cOp + ks * 8, dWpk + ks * 128, kB64, ks > 0);
      tc::mma_commit(barS);
    });
    tc::mbar_wait(barS, phS, &g_tc_error); phS ^= 1;
    tc::tc_fence_after();
    // ---- LayerNorm(norm1) backward on the row -> dY (packed bf16 A operand + export), column sums
    {
      float y[64];
      tc::tmem_ld64(tl + cS, y);
      float m4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] += sBp[i]; m4[i & 3] += y[i]; }
      const float mean = ((m4[0] + m4[1]) + (m4[2] + m4[3])) * (1.0f / 64.f);
      float v4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] -= mean; v4[i & 3] = fmaf(y[i], y[i], v4[i & 3]); }
      const float rstd = 1.0f / sqrtf(((v4[0] + v4[1]) + (v4[2] + v4[3])) * (1.0f / 64.f) + kLnEps);
      float d[64];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (wvalid) v = __ldg(reinterpret_cast<const float4*>(p.da + (size_t)tok * 64) + i);
        d[4 * i] = v.x; d[4 * i + 1] = v.y; d[4 * i + 2] = v.z; d[4 * i + 3] = v.w;
      }
      float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 64; ++i) {
        y[i] *= rstd;
        const float t = d[i] * sLnW[i];
        s1[i & 3] += t;
        s2[i & 3] = fmaf(t, y[i], s2[i & 3]);
      }
      const float m1 = ((s1[0] + s1[1]) + (s1[2] + s1[3])) * (1.0f / 64.f);
      const float m2 = ((s2[0] + s2[1]) + (s2[2] + s2[3])) * (1.0f / 64.f);
      warp_colsum64([&](int i) { return d[i] * y[i]; }, lane, dgam0, dgam1);
      warp_colsum64([&](int i) { return d[i]; }, lane, dbet0, dbet1);
      uint32_t dq[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const float a0 = rstd * (fmaf(d[2 * i], sLnW[2 * i], -m1) - y[2 * i] * m2);
        const float a1 = rstd * (fmaf(d[2 * i + 1], sLnW[2 * i + 1], -m1) - y[2 * i + 1] * m2);
        dq[i] = tc::pack_bf16(a0, a1);
      }
      tc::tmem_st32(tl + cOp, reinterpret_cast<const float(&)[32]>(dq));
      if (wvalid) {
        uint4* dst = reinterpret_cast<uint4*>(pp.dy16 + erow * 64);
#pragma unroll
        for (int i = 0; i < 8; ++i) dst[i] = make_uint4(dq[4 * i], dq[4 * i + 1], dq[4 * i + 2], dq[4 * i + 3]);
      }
      tc::tmem_wait_st();
    }
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    // ---- dO = dY Wp (B = the proj tile read MN-major) -> packed bf16, kept in TMEM for pass 2
    issue([&] {
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) tc::mma_f16_ts(tbase + cS, tbase + cOp + ks * 8, dWpmn + ks * 16, kB64bmn, ks > 0);
      tc::mma_commit(barS);
    });
    tc::mbar_wait(barS, phS, &g_tc_error); phS ^= 1;
    tc::tc_fence_after();
    {
      float v[64];
      tc::tmem_ld64(tl + cS, v);
      uint32_t dp[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) dp[i] = tc::pack_bf16(v[2 * i], v[2 * i + 1]);
      tc::tmem_st32(tl + cDOp, reinterpret_cast<const float(&)[32]>(dp));
      tc::tmem_wait_st();
    }

    PROF(2);                                                 // O / Y / LN backward / dO
    // ================================================================ pass 2: per-head gradients
#pragma unroll 1
    for (int h = 0; h < NH; ++h) {
      float qn[16], kn[16], vv[16], iq, ik; bool qok, kok;
      head_qkv(h, qn, kn, vv, iq, ik, qok, kok);
      PROF(3);                                               // [q|k|v]_h MMA + normalise + operand tiles
      {
        // dO_h of this row (bf16 pairs at packed column h*HD/2) -> [dO_h|0] / [0|dO_h] A operand and dO^T tile
        float dpk[16];
        tc::tmem_ld16(tl + cDOp + (h * HD) / 2, dpk);
        uint32_t dh[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) dh[j] = 2 * j < HD ? __float_as_uint(dpk[j]) : 0u;
        uint32_t a16[16];
#pragma unroll
        for (int j = 0; j < 8; ++j) { a16[wsel * 8 + j] = dh[j]; a16[(1 - wsel) * 8 + j] = 0u; }
        tc::tmem_st16(tl + cDOh, reinterpret_cast<const float(&)[16]>(a16));
        *reinterpret_cast<uint4*>(sDOT + (2 * wsel) * 1024 + lrow * 16) = make_uint4(dh[0], dh[1], dh[2], dh[3]);
        *reinterpret_cast<uint4*>(sDOT + (2 * wsel + 1) * 1024 + lrow * 16) = make_uint4(dh[4], dh[5], dh[6], dh[7]);
        uint32_t qb[8], kb[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          qb[j] = tc::pack_bf16(qn[2 * j], qn[2 * j + 1]);
          kb[j] = tc::pack_bf16(kn[2 * j], kn[2 * j + 1]);
        }
        *reinterpret_cast<uint4*>(sQT + (2 * wsel) * 1024 + lrow * 16) = make_uint4(qb[0], qb[1], qb[2], qb[3]);
        *reinterpret_cast<uint4*>(sQT + (2 * wsel + 1) * 1024 + lrow * 16) = make_uint4(qb[4], qb[5], qb[6], qb[7]);
        *reinterpret_cast<uint4*>(sKb + (2 * wsel) * 1024 + lrow * 16) = make_uint4(kb[0], kb[1], kb[2], kb[3]);
        *reinterpret_cast<uint4*>(sKb + (2 * wsel + 1) * 1024 + lrow * 16) = make_uint4(kb[4], kb[5], kb[6], kb[7]);
        tc::tmem_wait_st();
      }
      PROF(4);                                               // dO_h operands
      tc::fence_async_smem();
      tc::tc_fence_before();
      tc::worker_sync(1 + wk);
      PROF(5);                                               // sync
      // ---- S = Q'K'^T and dP = dO_h V^T
      issue([&] {
        tc::mma_f16(tbase + cS, dQ, dK, kIS, 0u);
        tc::mma_f16(tbase + cS, dQ + 256, dK + 128, kIS, 1u);
        tc::mma_f16_ts(tbase + cP2, tbase + cDOh, dVk, kB64, 0u);
        tc::mma_f16_ts(tbase + cP2, tbase + cDOh + 8, dVk + 128, kB64, 1u);
        tc::mma_commit(barS);
      });
      tc::mbar_wait(barS, phS, &g_tc_error); phS ^= 1;
      tc::tc_fence_after();
      PROF(6);                                               // S + dP MMAs
      {
        float s[64], dp[64];
        tc::tmem_ld64(tl + cS, s);
        softmax_row(h, s);                                  // s = P
        PROF(7);                                             // softmax
        tc::tmem_ld64(tl + cP2, dp);
        float r4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < 64; ++c) r4[c & 3] = fmaf(s[c], dp[c], r4[c & 3]);
        const float rowdot = (r4[0] + r4[1]) + (r4[2] + r4[3]);
        // P^T operand (rows = query slot, channels = (window, key)), then dS in place of dP
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(sPT + (wsel * 8 + j) * 1024 + lrow * 16) =
              make_uint4(tc::pack_bf16(s[8 * j], s[8 * j + 1]), tc::pack_bf16(s[8 * j + 2], s[8 * j + 3]),
                         tc::pack_bf16(s[8 * j + 4], s[8 * j + 5]), tc::pack_bf16(s[8 * j + 6], s[8 * j + 7]));
        PROF(8);                                             // dP load, rowdot, P^T tile
#pragma unroll
        for (int c = 0; c < 64; ++c) dp[c] = wvalid ? s[c] * (dp[c] - rowdot) : 0.f;   // dS
        PROF(9);                                             // dS + bias-table scatter
        uint32_t ds[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) ds[i] = tc::pack_bf16(dp[2 * i], dp[2 * i + 1]);
        tc::tmem_st32(tl + cS, reinterpret_cast<const float(&)[32]>(ds));
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(sST + (wsel * 8 + j) * 1024 + lrow * 16) =
              make_uint4(ds[4 * j], ds[4 * j + 1], ds[4 * j + 2], ds[4 * j + 3]);
        tc::tmem_wait_st();
      }
      PROF(10);                                              // dS operands
      tc::fence_async_smem();
      tc::tc_fence_before();
      tc::worker_sync(1 + wk);
      PROF(11);                                              // sync
      // ---- dV = P^T dO_h ; dqn = dS kn ; dkn = dS^T qn   (tokens of the window pair as K)
      issue([&] {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) tc::mma_f16(tbase + cQKV, dPTmn + ks * 16, dDOTmn + ks * 16, kB32mm, ks > 0);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) tc::mma_f16_ts(tbase + cP2, tbase + cS + ks * 8, dKbmn + ks * 16, kB32bmn, ks > 0);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) tc::mma_f16(tbase + cP2 + 32, dSTmn + ks * 16, dQTmn + ks * 16, kB32mm, ks > 0);
        tc::mma_commit(barO);
      });
      // ---- relative-position-bias gradient of this head and tile, in the shadow of the MMAs: the table
      //      entry of (query r, key c) depends on (ri - ci, rj - cj) only, so thread (window w, a, b) sums
      //      the 64 entries dS[(w, (ci+a)%8, (cj+b)%8)][(ci, cj)] of the dS tile - they belong to the four
      //      entries (a or a-8, b or b-8) - into its worker/window-private copy of the table: no atomics.
      {
        const int a8 = lrow >> 3, b8 = lrow & 7;
        const unsigned char* base = sST + wsel * 8192;
        float acc00 = 0.f, acc01 = 0.f, acc10 = 0.f, acc11 = 0.f;
#pragma unroll
        for (int ci = 0; ci < 8; ++ci) {
          const int ri = (ci + a8) & 7;
          float t0 = 0.f, t1 = 0.f;
#pragma unroll
          for (int cjs = 0; cjs < 8; ++cjs) {
            const int cj = (cjs + 2 * a8) & 7;     // skewed per a: the four a-groups of a warp hit distinct banks
            const int rj = (cj + b8) & 7;
            const float v = __uint_as_float((uint32_t)(*reinterpret_cast<const unsigned short*>(
                                base + ci * 1024 + (ri * 8 + rj) * 16 + cj * 2)) << 16);
            if (cj + b8 >= 8) t1 += v; else t0 += v;
          }
          if (ci + a8 >= 8) { acc10 += t0; acc11 += t1; } else { acc00 += t0; acc01 += t1; }
        }
        float* tab = sDBias + ((wk * 2 + wsel) * 6 + h) * kBiasLdB;
        tab[(a8 + 7) * 15 + b8 + 7] += acc00;
        if (b8 > 0) tab[(a8 + 7) * 15 + b8 - 1] += acc01;
        if (a8 > 0) tab[(a8 - 1) * 15 + b8 + 7] += acc10;
        if (a8 > 0 && b8 > 0) tab[(a8 - 1) * 15 + b8 - 1] += acc11;
      }
      PROF(15);                                              // bias-table gradient
      tc::mbar_wait(barO, phO, &g_tc_error); phO ^= 1;
      tc::tc_fence_after();
      PROF(12);                                              // dV / dqn / dkn MMAs
      {
        float dqn[16], dkn[16], dv[16];
        tc::tmem_ld16(tl + cP2 + wsel * 16, dqn);
        tc::tmem_ld16(tl + cP2 + 32 + wsel * 16, dkn);
        tc::tmem_ld16(tl + cQKV + wsel * 16, dv);
        // logit-scale gradient: sum_c dS_c cos_c = qn . (dS kn)
        float dsc = 0.f, dotq = 0.f, dotk = 0.f;
        const float sc = sScaleRaw[h];
#pragma unroll
        for (int d = 0; d < HD; ++d) {
          dsc = fmaf(qn[d], dqn[d], dsc);
          dqn[d] *= sc; dkn[d] *= sc;
          dotq = fmaf(qn[d], dqn[d], dotq);
          dotk = fmaf(kn[d], dkn[d], dotk);
        }
        if (!wvalid) dsc = 0.f;
        dsc = warp_sum(dsc);
        if (lane == 0) atomicAdd(sDScale + h, dsc);
        if (!qok) dotq = 0.f;
        if (!kok) dotk = 0.f;
        float outv[HS];
#pragma unroll
        for (int i = 0; i < HS; ++i) outv[i] = 0.f;
#pragma unroll
        for (int d = 0; d < HD; ++d) {
          outv[d] = (dqn[d] - qn[d] * dotq) * iq;           // F.normalize backward
          outv[PS + d] = (dkn[d] - kn[d] * dotk) * ik;
          outv[2 * PS + d] = dv[d];
        }
        if (wvalid) {
          uint4* dst = reinterpret_cast<uint4*>(pp.dqkv16 + erow * 192 + h * HS);
#pragma unroll
          for (int i = 0; i < HS / 8; ++i)
            dst[i] = make_uint4(tc::pack_bf16(outv[8 * i], outv[8 * i + 1]), tc::pack_bf16(outv[8 * i + 2], outv[8 * i + 3]),
                                tc::pack_bf16(outv[8 * i + 4], outv[8 * i + 5]), tc::pack_bf16(outv[8 * i + 6], outv[8 * i + 7]));
        }
      }
      PROF(13);                                              // normalize backward + export
      tc::tc_fence_before();
      tc::worker_sync(1 + wk);       // dV / dqn / dkn are read: the accumulators and tiles may be reused
      PROF(14);                                              // sync
    }
  }
  PROF_FLUSH;

  // ---- flush the CTA's parameter-gradient accumulators
  atomicAdd(p.d_ln_w + 2 * lane, dgam0); atomicAdd(p.d_ln_w + 2 * lane + 1, dgam1);
  atomicAdd(p.d_ln_b + 2 * lane, dbet0); atomicAdd(p.d_ln_b + 2 * lane + 1, dbet1);
  tc::tc_fence_before();
  __syncthreads();
  if (tid < NH && p.logit_scale[tid] <= kLogitMax) atomicAdd(p.d_logit_scale + tid, sDScale[tid] * sScaleRaw[tid]);
  for (int i = tid; i < NH * 225; i += 128 * kABWorkers) {
    const int h = i / 225, e = i - h * 225;
    float v = 0.f;
#pragma unroll
    for (int cpy = 0; cpy < 2 * kABWorkers; ++cpy) v += sDBias[(cpy * 6 + h) * kBiasLdB + e];
    atomicAdd(p.d_bias_table + e * NH + h, v);
  }
  if (warp == 0) tc::tmem_dealloc<512>(*tmem_slot);
}

// ================================================================================================
// attn_bwd_dx_kernel: dX = d[q|k|v] Wqkv ; dz = da + scatter(dX) ; dctx += window-quadrant sums of dX
constexpr int kDxWorkers = 3;
struct AttnDxSmem {
  static constexpr int oW = 0;                       // bf16 [192 packed rows][64 ch], R = 192
  static constexpr int oBar = oW + 192 * 64 * 2;
  static constexpr int oWork = ((oBar + kDxWorkers * 8 + 16 + 127) / 128) * 128;
  static constexpr int kLd = 68;                     // fp32 row stride of the dX tile
  static constexpr int wDx = 0;                      // fp32 [128][68]
  static constexpr int wCw = wDx + 128 * kLd * 4;    // int [128] source window of each row
  static constexpr int wTok = wCw + 512;             // long long [128] token of each row (-1: no token)
  static constexpr int kWorkBytes = wTok + 1024;
  static constexpr int kBytes = oWork + kDxWorkers * kWorkBytes;
};

template <int NH>
__global__ void __launch_bounds__(128 * kDxWorkers, 1) attn_bwd_dx_kernel(AttnBwdTcParams pp) {
  using L = AttnDxSmem;
  constexpr int HS = BCfg<NH>::HS, PS = BCfg<NH>::PS, HD = BCfg<NH>::HD, A = BCfg<NH>::A;
  const AttnParams& p = pp.a;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char* sWb = smem_raw + L::oW;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem_raw + L::oBar + kDxWorkers * 8);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int wk = warp >> 2, wq = warp & 3;
  unsigned char* sW = smem_raw + L::oWork + wk * L::kWorkBytes;
  float* sDx = reinterpret_cast<float*>(sW + L::wDx);
  int* sCw = reinterpret_cast<int*>(sW + L::wCw);
  long long* sTok = reinterpret_cast<long long*>(sW + L::wTok);
  uint64_t* bar = bars + wk;

  for (int idx = tid; idx < 192 * 8; idx += 128 * kDxWorkers) {
    const int n = idx % 192, kc = idx / 192;
    const int h = n / HS, part = (n % HS) / PS, d = (n % HS) % PS;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (h < NH && part < 3 && d < HD) {
      const float4* src = reinterpret_cast<const float4*>(p.qkv_w + (part * A + h * HD + d) * 64 + kc * 8);
      const float4 a = src[0], b = src[1];
      v = make_uint4(tc::pack_bf16(a.x, a.y), tc::pack_bf16(a.z, a.w), tc::pack_bf16(b.x, b.y), tc::pack_bf16(b.z, b.w));
    }
    *reinterpret_cast<uint4*>(sWb + (size_t)kc * (192 * 16) + n * 16) = v;
  }
  if (tid == 0) {
    for (int i = 0; i < kDxWorkers; ++i) tc::mbar_init(bars + i, 1);
    tc::fence_barrier_init();
  }
  if (warp == 0) tc::tmem_alloc<512>(tmem_slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();

  const uint32_t tbase = *tmem_slot + wk * 160;      // [0,96): d[q|k|v] packed ; [96,160): dX
  const uint32_t tl = tbase + ((uint32_t)(wq * 32) << 16);
  constexpr uint32_t kI = tc::idesc_bf16(128, 64, false, true);
  const uint64_t dWmn = tc::desc_mnmajor16(tc::smem_u32(sWb), 192);
  const Geom g = make_geom(p.B, p.ph, p.pw, p.shift);
  const int nwin = g.B * g.wh * g.ww;
  const int ntile = (nwin + 1) >> 1;
  const int row = wq * 32 + lane;
  const int wsel = wq >> 1, lrow = row & 63;
  uint32_t ph = 0;
  for (int tile = blockIdx.x * kDxWorkers + wk; tile < ntile; tile += gridDim.x * kDxWorkers) {
    const int win = tile * 2 + wsel;
    const bool wvalid = win < nwin;
    long long tok = 0;
    int cwin = 0;
    if (wvalid) { const Slot sl = slot_of(g, win, lrow); tok = sl.tok; cwin = sl.cwin; }
    sCw[row] = cwin;
    sTok[row] = wvalid ? tok : -1;
    {
      const uint4* src = reinterpret_cast<const uint4*>(pp.dqkv16 + ((size_t)tile * 128 + row) * 192);
#pragma unroll
      for (int part = 0; part < 3; ++part) {
        uint32_t a[32];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          uint4 v = make_uint4(0u, 0u, 0u, 0u);
          if (wvalid) v = __ldg(src + part * 8 + i);
          a[4 * i] = v.x; a[4 * i + 1] = v.y; a[4 * i + 2] = v.z; a[4 * i + 3] = v.w;
        }
        tc::tmem_st32(tl + part * 32, reinterpret_cast<const float(&)[32]>(a));
      }
      tc::tmem_wait_st();
    }
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 12; ++ks) tc::mma_f16_ts(tbase + 96, tbase + ks * 8, dWmn + ks * 16, kI, ks > 0);
        tc::mma_commit(bar);
      }
      __syncwarp();
    }
    tc::mbar_wait(bar, ph, &g_tc_error); ph ^= 1;
    tc::tc_fence_after();
    {
      float v[64];
      tc::tmem_ld64(tl + 96, v);
#pragma unroll
      for (int i = 0; i < 16; ++i)
        *reinterpret_cast<float4*>(sDx + row * L::kLd + 4 * i) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    // dz = da + dX, COOPERATIVELY: a warp instruction covers two whole token rows (2 x 256 contiguous bytes) instead
    // of one 16-byte piece of 32 different rows (32 L1 lines per instruction made the one-row-per-thread form
    // LSU-bound); the rows of this warp's quarter of the tile come from the fp32 dX tile in shared memory
    {
      const int half = lane >> 4, c4 = (lane & 15) * 4;
#pragma unroll 4
      for (int it = 0; it < 16; ++it) {
        const int r = wq * 32 + it * 2 + half;
        const long long t = sTok[r];
        if (t >= 0) {
          const float4 gv = __ldg(reinterpret_cast<const float4*>(p.da + (size_t)t * 64 + c4));
          const float4 dx = *reinterpret_cast<const float4*>(sDx + r * L::kLd + c4);
          *reinterpret_cast<float4*>(p.dz + (size_t)t * 64 + c4) = make_float4(gv.x + dx.x, gv.y + dx.y, gv.z + dx.z, gv.w + dx.w);
        }
      }
    }
    // dctx: the 16 slots of a window quadrant share one source window (shift 4) / all 64 do (shift 0)
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int item = it * 128 + row;               // (window of the pair, quadrant, channel)
      const int c = item & 63, quad = (item >> 6) & 3, w = item >> 8;
      const int i0 = (quad >> 1) * 4, j0 = (quad & 1) * 4;
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) s += sDx[(w * 64 + (i0 + i) * 8 + j0 + j) * L::kLd + c];
      if (tile * 2 + w < nwin) atomicAdd(p.dctx + (size_t)sCw[w * 64 + i0 * 8 + j0] * 64 + c, s);
    }
    tc::worker_sync(1 + wk);
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(*tmem_slot);
}

#ifdef B200MAR_ABWD_PROF
}  // namespace b200mar
extern "C" int b200mar_abwd_prof(unsigned long long* out32, int reset) {
  unsigned long long z[32] = {0};
  if (cudaMemcpyFromSymbol(out32, b200mar::g_abwd_prof, sizeof(z)) != cudaSuccess) return -1;
  if (reset && cudaMemcpyToSymbol(b200mar::g_abwd_prof, z, sizeof(z)) != cudaSuccess) return -1;
  return 0;
}
namespace b200mar {
#endif

template <int NH>
static int launch_attn_bwd_tc(const AttnBwdTcParams& pp, cudaStream_t st) {
  const AttnParams& p = pp.a;
  constexpr int HS = BCfg<NH>::HS, PS = BCfg<NH>::PS, HD = BCfg<NH>::HD, A = BCfg<NH>::A;
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(attn_bwd_core_kernel<NH>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)AttnBwdSmem::kBytes));
    B200MAR_CHECK(cudaFuncSetAttribute(attn_bwd_dx_kernel<NH>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)AttnDxSmem::kBytes));
    configured.done();
  }
  const int nwin = p.B * (p.ph / kWin) * (p.pw / kWin);
  const int ntile = (nwin + 1) / 2;
  int grid = (ntile + kABWorkers - 1) / kABWorkers;
  if (grid > sm_count()) grid = sm_count();
  attn_bwd_core_kernel<NH><<<grid, 128 * kABWorkers, AttnBwdSmem::kBytes, st>>>(pp);
  B200MAR_CHECK(cudaGetLastError());
  grid = (ntile + kDxWorkers - 1) / kDxWorkers;
  if (grid > sm_count()) grid = sm_count();
  attn_bwd_dx_kernel<NH><<<grid, 128 * kDxWorkers, AttnDxSmem::kBytes, st>>>(pp);
  B200MAR_CHECK(cudaGetLastError());
  const long long n_rows = (long long)nwin * 64;
  if (int rc = launch_wgrad(pp.x16, pp.dqkv16, p.d_qkv_w, p.d_qkv_b, n_rows, 64, 192, 64, 64, HS, PS, HD, A, st)) return rc;
  if (int rc = launch_wgrad(pp.o16_saved ? pp.o16_saved : pp.o16, pp.dy16, p.d_proj_w, p.d_proj_b, n_rows, 64, 64, A, A, 0, 0, 0, 0, st)) return rc;
  return 0;
}

}  // namespace b200mar

extern "C" int b200mar_attn_bwd_tc(const float* z, const float* ctx, const float* da,
                                   const float* qkv_w, const float* qkv_b, const float* proj_w,
                                   const float* proj_b, const float* logit_scale,
                                   const float* bias_table, const float* ln_w, const float* ln_b,
                                   float* dz, float* dctx, float* d_qkv_w, float* d_qkv_b,
                                   float* d_proj_w, float* d_proj_b, float* d_logit_scale,
                                   float* d_bias_table, float* d_ln_w, float* d_ln_b, void* scratch,
                                   const void* o16_saved, int B, int ph, int pw, int nh, int shift,
                                   void* stream) {
  using namespace b200mar;
  if (ph % kWin || pw % kWin || ph < kWin || pw < kWin || (shift != 0 && shift != kShift) || !scratch) return -1;
  if ((long long)B * ph * pw >= (1ll << 31)) return -1;
  AttnBwdTcParams pp{};
  AttnParams& p = pp.a;
  p.z = z; p.ctx = ctx; p.da = da; p.qkv_w = qkv_w; p.qkv_b = qkv_b; p.proj_w = proj_w;
  p.proj_b = proj_b; p.logit_scale = logit_scale; p.bias_table = bias_table; p.ln_w = ln_w;
  p.ln_b = ln_b; p.dz = dz; p.dctx = dctx; p.d_qkv_w = d_qkv_w; p.d_qkv_b = d_qkv_b;
  p.d_proj_w = d_proj_w; p.d_proj_b = d_proj_b; p.d_logit_scale = d_logit_scale;
  p.d_bias_table = d_bias_table; p.d_ln_w = d_ln_w; p.d_ln_b = d_ln_b;
  p.B = B; p.ph = ph; p.pw = pw; p.shift = shift;
  const long long nwin = (long long)B * (ph / kWin) * (pw / kWin);
  const long long rows = ((nwin + 1) / 2) * 128;      // scratch: rows * 384 bf16
  uint16_t* s = (uint16_t*)scratch;
  pp.o16_saved = (const uint16_t*)o16_saved;
  pp.x16 = s; pp.o16 = s + rows * 64; pp.dy16 = s + rows * 128; pp.dqkv16 = s + rows * 192;
  cudaStream_t st = (cudaStream_t)stream;
  if (nh == 6) return launch_attn_bwd_tc<6>(pp, st);
  if (nh == 4) return launch_attn_bwd_tc<4>(pp, st);
  return -1;
}
