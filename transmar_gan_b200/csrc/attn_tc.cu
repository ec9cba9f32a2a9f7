// attn_tc.cu — fused N-gram window partition + scaled-cosine window attention + LN1 + residual,
// forward, on the 5th-gen tensor cores with fp32 accumulators in TMEM.  Operand formats:
//   * kind::f16 with FP16 operands (11-bit significand, the same as TF32, at half the shared-
//     memory footprint and twice the MMA rate) where the range is under control: X (every row
//     scaled by a power of two to max|x| in [1,2)), the qkv weights (one power-of-two scale per
//     matrix), the unit-norm Q and K;
//   * kind::tf32 where magnitudes are data-dependent and LayerNorm / F.normalize downstream make
//     RELATIVE precision matter at any scale: P V and the output projection.
// The residual stream in HBM, q/k normalisation, softmax, LayerNorm and the residual add are fp32.
//
// One persistent CTA per SM = FOUR (6 heads) or THREE (4 heads) independent 128-thread workers sharing
// the resident weights (qkv re-packed per head as [q|k|v] rows, 24 KB; proj 16 KB; bias table copies;
// LN affine).  A worker streams tiles of TWO (shifted) windows = 128 tokens; thread t owns token row
// t = TMEM lane t.  Per tile:
//   z tile by TMA (4-D window boxes, 128 B swizzle; wrapped windows by direct row access);
//   X = z + ctx, row-scaled -> fp16 smem A operand
//   for each head h (QKV of head h+1 is issued in the shadow of S_h):
//     [q_h|k_h|v_h] = X * Wqkv[h]^T      4 MMAs 128xHSx16 -> TMEM (double-buffered over heads)
//     row-wise fp32: +bias, F.normalize(q), F.normalize(k), q *= exp(min(logit_scale,ln100))*log2e
//       -> Q' = [q | 0] / [0 | q] (window 0 / window 1 rows), K' = [k(w0) | k(w1)], so that ONE
//          128x64x32 MMA gives every row the logits against its own window's 64 keys;
//          V^T (rows = (window, d), keys contiguous)
//     S = Q' K'^T + Hot Bias_h^T          2 + 4 MMAs 128x64x16: the relative-position bias is added BY
//                                         THE TENSOR CORE as a one-hot(query token) x bias-matrix product;
//                                         both operands are Toeplitz and live compressed in shared memory
//                                         (descriptor strides LBO = SBO make core matrices alias, below)
//     row-wise fp32: (+ analytic shift mask on edge windows), ex2; unnormalised P written back to TMEM
//       over S (tcgen05.st).  No row maximum, no bias adds, no row sum: the softmax denominator is a
//       ones row of V^T (6 heads: head dim 10 of 16 accumulator columns), i.e. one more column of O_h
//     O_h = P [V_h(w0) | V_h(w1)]         8 MMAs 128x32x8 (tf32), A operand = P from TMEM, written
//                                         over the dead [q|k|v]_h accumulator columns
//   O row -> TMEM (over the dead qkv buffers) = A operand of Y = O Wp^T (8 MMAs 128x64x8, tf32);
//   meanwhile the z tile is fetched again (L2 hit); +bias, LayerNorm (a whole row per thread), + z in
//   place, TMA store.
// While one worker runs a CUDA-core epilogue the others own the tensor pipe / load path.
//
// Reference semantics: components/window_utils.py:127-147, components/attention.py:70-101,
// components/transformer_blocks.py:63-112,117 (same contract as attn_fwd.cu, the fp32 kernel).
#include "common.cuh"
#include "attn_common.cuh"
#include "tc.cuh"
#include "tma.cuh"
#include <cuda_fp16.h>

namespace b200mar {

B200MAR_DEFINE_TC_ERROR_GETTER(tc_error_attn)

// Optional phase profile (compile with -DB200MAR_ATTN_PROF): cycles spent by one thread (block 0,
// worker 0, lane 0 of warp 1) between successive PROF(i) marks, summed per phase i.
#ifdef B200MAR_ATTN_PROF
__device__ unsigned long long g_attn_prof[32];
#define PROF_DECL __shared__ unsigned long long s_prof[32]; if (threadIdx.x < 32) s_prof[threadIdx.x] = 0; __syncthreads(); \
  long long prof_t = clock64(); const bool prof_on = blockIdx.x == 0 && threadIdx.x == 32;
#define PROF(i) do { if (prof_on) { const long long t_ = clock64(); s_prof[i] += (unsigned long long)(t_ - prof_t); prof_t = clock64(); } } while (0)
#define PROF_FLUSH do { if (prof_on) for (int i_ = 0; i_ < 32; ++i_) g_attn_prof[i_] += s_prof[i_]; } while (0)
#else
#define PROF_DECL
#define PROF(i)
#define PROF_FLUSH
#endif

constexpr float kLog2e = 1.4426950408889634f;
// The kernel is bound by the latency of each worker's serial per-head chain (measured: halving the MUFU work or
// cutting 8 % of all instructions elsewhere changes nothing), so the row arithmetic ON that chain is kept short:
//  * TF32 operands are rounded by an integer add: cvt.rna.tf32.f32 compiles to FSETP + IADD + LOP3, and the tensor
//    core ignores the low 13 bits of a kind::tf32 operand, so adding half an ulp of the 10-bit significand IS
//    round-to-nearest, ties away from zero, for every finite value (results bit-identical);
//  * the per-head bias / normalisation arithmetic uses the packed .f32x2 instructions of sm_100;
//  * 1 / x and 1 / sqrt(x) are single MUFU operations (the IEEE forms add a Newton step, denormal pre-scaling and a
//    range-check branch).  The compensated kernel (COMP) keeps the IEEE forms.
// Together 1.470 / 1.529 -> 1.407 / 1.469 ms per launch at 32 x 416^2.  Packing the LayerNorm epilogue or the row
// conversion the same way (more spills) and moving the O_h scaling behind the S_h issue were measured and lost.
__device__ __forceinline__ float round_tf32_add(float x) { return __uint_as_float(__float_as_uint(x) + 0x1000u); }
template <bool COMP>
__device__ __forceinline__ float rcp_op(float x) {
  if constexpr (!COMP) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
  else return 1.0f / x;
}
template <bool COMP>
__device__ __forceinline__ float rsqrt_op(float x) {     // callers pass x >= 1e-24: no denormal pre-scaling needed
  if constexpr (!COMP) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
  else return rsqrtf(x);
}
__device__ __forceinline__ void mul2(float& a, float& b, float c) {
  const float2 r = __fmul2_rn(make_float2(a, b), make_float2(c, c));
  a = r.x; b = r.y;
}
__device__ __forceinline__ void fma2(float& a, float& b, float c, float e, float f) {   // (a, b) = (a, b) * c + (e, f)
  const float2 r = __ffma2_rn(make_float2(a, b), make_float2(c, c), make_float2(e, f));
  a = r.x; b = r.y;
}
// workers (tiles in flight) per CTA.  NH = 6: heads are packed as [q10|k10|v10|0 0] = 32 accumulator
// columns, so a worker needs 2 x 32 (qkv, double-buffered; O_h over the dead buffer) + 64 (S/P/Y) = 128
// TMEM columns and FOUR fit; NH = 4 (head dim 16): 2 x 48 + 64 = 160 columns, three workers.
#ifndef B200MAR_ATTN6_WORKERS
#define B200MAR_ATTN6_WORKERS 4          // development knob (A/B builds through B200MAR_NVCC_EXTRA)
#endif
template <int NH> struct Cfg {
  static constexpr int kWorkers = NH == 6 ? B200MAR_ATTN6_WORKERS : 3;
  static constexpr int HS = NH == 6 ? 32 : 48;     // packed rows / accumulator columns per head
  static constexpr int PS = NH == 6 ? 10 : 16;     // stride of q, k, v inside a head's block
  static constexpr uint32_t kTStride = 2 * HS + 64;
  static constexpr uint32_t kTS = 2 * HS;          // S; P (fp32) overwrites it; Y at the end
};
// Relative-position bias through the MMA.  Window POSITION q (thread row / key column) holds token
// (7 - q/8, q%8): the window's token rows are visited bottom-up, so that for query token (ri, rj) and key
// token (ci, cj) the K-chunk index of the one-hot operand (kc = ri) and the 8-row group of the key
// (n/8 = 7 - ci) ADD to ri - ci + 7 = the row block of the 15x15 bias table.  With LBO = SBO = 128 B the
// no-swizzle K-major descriptor places core matrix (group g, k-chunk kc) at byte (g + kc) * 128, hence
//   bias operand of a head : 15 core matrices, matrix m = ri - ci + 7, element (cj, rj) = T[15 m + rj - cj + 7]
//   one-hot operand        : 23 core matrices, identity at m = 7 (window 0 rows) and m = 15 (window 1), else 0
// instead of 64x64 and 128x64 fp16 tiles.
constexpr int kBiasTHead = 15 * 128;           // bytes of one head's compressed bias operand (fp16)
constexpr int kHotBytes = 23 * 128;
constexpr int kVChunk = 528;                   // bytes between successive 4-key chunks of V^T (fp32):
                                               // 32 rows x 16 B + 16 (== 4 words mod 32 banks)

// COMP (precision level "tf32c"): every GEMM of the block runs on operand PAIRS, three MMA passes
// Ah Bh + Al Bh + Ah Bl into the same fp32 accumulator:
//   qkv  : X = Xh + Xl, Wqkv = Wh + Wl (fp16, power-of-two row / matrix scales)            ~22 bits
//   q k^T: Q' = Qh + Ql, K' = Kh + Kl (fp16)                                                ~22 bits
//   P V  : P = Ph + Pm (BF16 pair in TMEM; always the max-subtracting softmax, so P <= 1),
//          V = Vh + Vm (BF16 pair: no range scaling needed)                                 ~16 bits
//   proj : O = Oh + Ol (fp16, power-of-two row scale, smem), Wp = Wh + Wl (fp16)            ~22 bits
// The lo copies cost 24 KB (weights) + 32 KB per worker, which leaves room for TWO workers.
template <int NH, bool COMP> struct AttnWorkers { static constexpr int value = COMP ? 2 : Cfg<NH>::kWorkers; };

template <int W, bool COMP = false>
struct AttnTcSmem {                            // BYTE offsets (W workers)
  static constexpr int kWRows = 192;           // 6 heads x [q10|k10|v10|0 0] or 4 heads x [q16|k16|v16]
  static constexpr int oWqkv = 0;              // fp16 [288 rows][64 k], 8-element k-chunks, R = 288
  static constexpr int oWp = oWqkv + kWRows * 64 * 2;   // fp32 (tf32) [64 rows][64 k], 4-element k-chunks, R = 64
  static constexpr int oBqkv = oWp + 64 * 64 * 4;       // fp32 [288], packed order
  static constexpr int oBp = oBqkv + kWRows * 4;
  static constexpr int oLnW = oBp + 256;
  static constexpr int oLnB = oLnW + 256;
  static constexpr int oScale = oLnB + 256;     // [0,6): logit scales, [6]: inverse of the qkv weight scale, [7]: scratch
  static constexpr int oBiasT = ((oScale + 32 + 127) / 128) * 128;   // fp16 [6 heads][15][8 cj][8 rj], * log2(e)
  static constexpr int oHot = oBiasT + 6 * kBiasTHead;               // fp16 one-hot operand (23 core matrices)
  static constexpr int oBar = oHot + kHotBytes;  // W workers x 8 mbarrier slots (5 used) + tmem slot
  static constexpr int oWqkvLo = ((oBar + 320 + 127) / 128) * 128;    // COMP: fp16 lo parts, same layout as oWqkv
  static constexpr int oWork = ((oWqkvLo + (COMP ? kWRows * 64 * 2 : 0) + 1023) / 1024) * 1024;   // 1 KB aligned: TMA 128 B swizzle
  // per worker; [X | Q | K | V] is also the 32 KB fp32 staging tile of the LayerNorm output
  static constexpr int wX = 0;                 // fp16 [128][64] (R = 128): X, later O
  static constexpr int wQ = wX + 128 * 64 * 2; // fp16 [128][32] (R = 128)
  static constexpr int wK = wQ + 128 * 32 * 2; // fp16 [64][32]  (R = 64)
  static constexpr int wV = wK + 64 * 32 * 2;  // fp32 [32 rows = (window, d)][64 keys], chunk stride kVChunk
  static constexpr int wTok = wV + 16 * kVChunk;
  // COMP only, outside the staging alias: lo parts of X, Q', K' (+ 4 KB) = 32 KB; after the last head the same
  // 32 KB hold the fp16 hi | lo parts of the attention output O, the A operands of the projection
  static constexpr int wXlo = ((wTok + 4 * 128 * 4 + 127) / 128) * 128;            // fp16 [128][64]
  static constexpr int wQlo = wXlo + 128 * 64 * 2;                                 // fp16 [128][32]
  static constexpr int wKlo = wQlo + 128 * 32 * 2;                                 // fp16 [64][32]
  static constexpr int wOhi = wXlo, wOlo = wQlo;                                   // fp16 [128][64] each
  static constexpr int kWorkBytes = ((wXlo + (COMP ? 2 * 128 * 64 * 2 : 0) + 1023) / 1024) * 1024;   // tok, cwin, region, row scale
  static constexpr int kBytes = oWork + W * kWorkBytes;
  static_assert(wTok >= 128 * 64 * 4, "fp32 staging tile must fit in [X|Q|K|V]");
  static_assert(oBar % 8 == 0, "mbarrier alignment");
  static_assert(kBytes <= 232448, "attn_tc shared memory exceeds 227 KB");
};

// TMEM columns of one worker: [0, 2 HS): [q_h | k_h | v_h] accumulators, double-buffered over heads
// (O_h, 32 columns, is written over buffer h&1 once it has been read); [2 HS, 2 HS + 64): S / P / Y
constexpr uint32_t kTQ = 0;

// RESID = true : a = z + LN1(attn)            (the op's standalone contract)
// RESID = false: a = LN1(attn) only; the consumer (mlp_tc with a_is_delta) adds z, which it reads
//                anyway, so this kernel never re-reads z for the residual.
// S16 = true : the residual stream (z in, a out) is BF16 in HBM (precision mode "bf16", BASELINE.json configs[1]):
//               one 8 KB TMA box per window (64 tokens x 64 channels x 2 B = 128 B rows, 128 B swizzle), rows are
//               widened to fp32 on read; the residual add runs in fp32 and a is rounded to bf16 once, on write.
template <int NH, bool RESID, bool SAVEO, bool S16 = false, bool COMP = false>
__global__ void __launch_bounds__(128 * AttnWorkers<NH, COMP>::value, 1)
attn_tc_fwd_kernel(const __grid_constant__ CUtensorMap mapZ, const __grid_constant__ CUtensorMap mapA,
                   AttnParams p) {
  constexpr int kWorkers = AttnWorkers<NH, COMP>::value, HS = Cfg<NH>::HS, PS = Cfg<NH>::PS;
  constexpr uint32_t kTStride = Cfg<NH>::kTStride, kTS = Cfg<NH>::kTS;
  using L = AttnTcSmem<kWorkers, COMP>;
  constexpr int HD = kDim / NH, A = HD * NH;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __half* sWqkv = reinterpret_cast<__half*>(smem_raw + L::oWqkv);
  float* sWp = reinterpret_cast<float*>(smem_raw + L::oWp);
  float* sBqkv = reinterpret_cast<float*>(smem_raw + L::oBqkv);
  float* sBp = reinterpret_cast<float*>(smem_raw + L::oBp);
  float* sLnW = reinterpret_cast<float*>(smem_raw + L::oLnW);
  float* sLnB = reinterpret_cast<float*>(smem_raw + L::oLnB);
  float* sScale = reinterpret_cast<float*>(smem_raw + L::oScale);
  __half* sBiasT = reinterpret_cast<__half*>(smem_raw + L::oBiasT);
  __half* sHot = reinterpret_cast<__half*>(smem_raw + L::oHot);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + L::oBar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem_raw + L::oBar + kWorkers * 64);
  const int tid = threadIdx.x, lane = tid & 31;
  // warp-uniform ids (broadcast from lane 0 so that everything derived from them - smem / TMEM
  // addresses, MMA descriptors - stays in uniform registers)
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int wk = warp >> 2, wq = warp & 3;
  unsigned char* sW = smem_raw + L::oWork + wk * L::kWorkBytes;
  unsigned char* sX = sW + L::wX;
  unsigned char* sQ = sW + L::wQ;
  unsigned char* sK = sW + L::wK;
  unsigned char* sV = sW + L::wV;
  int* sReg = reinterpret_cast<int*>(sW + L::wTok) + 256;
  uint64_t* barQ = bars + wk * 8;       // [2]
  uint64_t* barS = barQ + 2;
  uint64_t* barO = barQ + 3;
  uint64_t* barZ = barQ + 4;            // TMA loads of the z tile

  // ---- stage the block's weights as K-major operands (once per CTA).  The qkv matrix is scaled by
  //      one power of two to max|w| in [1,2) so that fp16 keeps 11 significant bits at any scale.
  // (One pass over global memory: every thread fetches its float4 share of the qkv matrix with back-to-back loads
  //  and keeps it in registers while the matrix maximum is reduced, then scales, converts and scatters it into the
  //  packed per-head order; the bias table is copied once, coalesced, into the - still idle - staging tile of worker 0
  //  and the compressed bias operands / bounded-logit flags are built from that copy.  The first version walked the
  //  matrix twice with scalar / strided loads and gathered the bias operands from global memory.)
  uint32_t* sWmax = reinterpret_cast<uint32_t*>(sScale + 7);
  float* sTab = reinterpret_cast<float*>(smem_raw + L::oWork);     // [225][NH] copy of the bias table (prologue only)
  constexpr int NT = 128 * kWorkers;
  constexpr int kQ4 = 3 * A * 16;                                  // float4 pieces of qkv_w [3 A][64]
  constexpr int NPER = (kQ4 + NT - 1) / NT;
  if (tid == 0) *sWmax = 0u;
  __syncthreads();
  float4 wq4[NPER];
#pragma unroll
  for (int j = 0; j < NPER; ++j) {
    const int f = tid + j * NT;
    wq4[j] = f < kQ4 ? __ldg(reinterpret_cast<const float4*>(p.qkv_w) + f) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (int i = tid; i < NH * 225; i += NT) sTab[i] = p.bias_table[i];
  for (int idx = tid; idx < L::kWRows * 8; idx += NT) {            // pad rows / columns of the packed operand are zero
    *reinterpret_cast<uint4*>(sWqkv + (size_t)idx * 8) = make_uint4(0u, 0u, 0u, 0u);
    if constexpr (COMP) *reinterpret_cast<uint4*>(smem_raw + L::oWqkvLo + (size_t)idx * 16) = make_uint4(0u, 0u, 0u, 0u);
  }
  {
    float m = 0.f;
#pragma unroll
    for (int j = 0; j < NPER; ++j)
      m = fmaxf(m, fmaxf(fmaxf(fabsf(wq4[j].x), fabsf(wq4[j].y)), fmaxf(fabsf(wq4[j].z), fabsf(wq4[j].w))));
    atomicMax(sWmax, __float_as_uint(m));               // non-negative floats order like unsigned ints
  }
  __syncthreads();
  uint32_t we = (*sWmax >> 23) & 0xFFu;
  we = we < 1u ? 127u : (we > 253u ? 253u : we);
  const float wsc = __uint_as_float((254u - we) << 23);   // 2^(127-we)
#pragma unroll
  for (int j = 0; j < NPER; ++j) {
    const int f = tid + j * NT;
    if (f < kQ4) {
      const int r = f >> 4, k0 = (f & 15) << 2;           // source row r = part * A + h * HD + d, columns k0 .. k0 + 3
      const int part = r / A, rem = r - part * A, h = rem / HD, d = rem - h * HD;
      const int n = h * HS + part * PS + d;               // packed row
      const size_t off = (size_t)(k0 >> 3) * (L::kWRows * 8) + n * 8 + (k0 & 7);
      const float a = wq4[j].x * wsc, b = wq4[j].y * wsc, c = wq4[j].z * wsc, e = wq4[j].w * wsc;
      const uint2 hv = make_uint2(tc::pack_f16(a, b), tc::pack_f16(c, e));
      *reinterpret_cast<uint2*>(sWqkv + off) = hv;
      if constexpr (COMP)                      // lo = fp16(w * wsc - hi): exact difference, 11 more bits
        *reinterpret_cast<uint2*>(smem_raw + L::oWqkvLo + off * 2) = make_uint2(tc::lo_f16(a, b, hv.x), tc::lo_f16(c, e, hv.y));
    }
  }
  __shared__ float sProjInv;                   // COMP: inverse of the projection weights' power-of-two scale
  if constexpr (COMP) {
    __shared__ uint32_t sPmax;
    if (tid == 0) sPmax = 0u;
    __syncthreads();
    float m = 0.f;
    for (int i = tid; i < 64 * A; i += 128 * kWorkers) m = fmaxf(m, fabsf(p.proj_w[i]));
    atomicMax(&sPmax, __float_as_uint(m));
    __syncthreads();
    uint32_t pe = (sPmax >> 23) & 0xFFu;
    pe = pe < 1u ? 127u : (pe > 253u ? 253u : pe);
    const float psc = __uint_as_float((254u - pe) << 23);
    if (tid == 0) sProjInv = __uint_as_float(pe << 23);
    // fp16 [64 rows n][64 k], 8-element k-chunks (R = 64): hi at oWp, lo at oWp + 8 KB
    for (int idx = tid; idx < 64 * 8; idx += 128 * kWorkers) {
      const int n = idx & 63, kc = idx >> 6;
      float w[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) w[j] = kc * 8 + j < A ? p.proj_w[n * A + kc * 8 + j] * psc : 0.f;
      const uint4 hv = make_uint4(tc::pack_f16(w[0], w[1]), tc::pack_f16(w[2], w[3]),
                                  tc::pack_f16(w[4], w[5]), tc::pack_f16(w[6], w[7]));
      unsigned char* dst = smem_raw + L::oWp + kc * 1024 + n * 16;
      *reinterpret_cast<uint4*>(dst) = hv;
      *reinterpret_cast<uint4*>(dst + 8192) =
          make_uint4(tc::lo_f16(w[0], w[1], hv.x), tc::lo_f16(w[2], w[3], hv.y),
                     tc::lo_f16(w[4], w[5], hv.z), tc::lo_f16(w[6], w[7], hv.w));
    }
  } else {
    for (int idx = tid; idx < 64 * 16; idx += 128 * kWorkers) {
      const int n = idx & 63, kc = idx >> 6;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (kc * 4 < A) v = *reinterpret_cast<const float4*>(p.proj_w + n * A + kc * 4);
      *reinterpret_cast<float4*>(sWp + kc * 256 + n * 4) = tc::round_tf32(v);   // the MMA would truncate
    }
  }
  for (int n = tid; n < L::kWRows; n += 128 * kWorkers) {      // bias in the same packed order
    const int h = n / HS, part = (n % HS) / PS, d = (n % HS) % PS;
    sBqkv[n] = (h < NH && part < 3 && d < HD) ? p.qkv_b[part * A + h * HD + d] : 0.f;
  }
  if (tid < 64) { sBp[tid] = p.proj_b[tid]; sLnW[tid] = p.ln_w[tid]; sLnB[tid] = p.ln_b[tid]; }
  if (tid < NH) sScale[tid] = expf(fminf(p.logit_scale[tid], kLogitMax)) * kLog2e;
  if (tid == 0) sScale[6] = __uint_as_float(we << 23);    // 2^(we-127)
  // A head whose logits are bounded a priori, |cos * scale + bias| <= scale + max|bias| <= 60 octaves, needs
  // no row maximum: p = ex2(s) can neither overflow (row sums <= 2^66) nor lose every term of a row (the
  // row's own key is never masked).  Any other head (logit scales near the ln 100 clamp, huge biases)
  // always takes the exact max-subtracting path.
  __shared__ int sFast[8];
  if (warp < NH) {                               // one warp per head, from the shared-memory copy of the table
    float mb = 0.f;
    for (int e = lane; e < 225; e += 32) mb = fmaxf(mb, fabsf(sTab[e * NH + warp]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mb = fmaxf(mb, __shfl_xor_sync(0xffffffffu, mb, o));
    if (lane == 0) sFast[warp] = (expf(fminf(p.logit_scale[warp], kLogitMax)) + mb) * kLog2e <= 60.0f ? 1 : 0;
  }
  for (int i = tid; i < NH * 15 * 64; i += 128 * kWorkers) {
    const int h = i / 960, r = i - h * 960, m = r >> 6, cj = (r >> 3) & 7, rj = r & 7;
    sBiasT[i] = __float2half_rn(sTab[(m * 15 + rj - cj + 7) * NH + h] * kLog2e);
  }
  for (int i = tid; i < 23 * 64; i += 128 * kWorkers) {
    const int m = i >> 6, j = (i >> 3) & 7, e = i & 7;
    sHot[i] = __float2half_rn(((m == 7 || m == 15) && j == e) ? 1.0f : 0.0f);
  }
  if (tid == 0) {
    for (int i = 0; i < 8 * kWorkers; ++i) tc::mbar_init(bars + i, 1);
    tc::fence_barrier_init();
    tma::prefetch_desc(&mapZ); tma::prefetch_desc(&mapA);
  }
  if (warp == 0) tc::tmem_alloc<512>(tmem_slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();

  const uint32_t tbase = *tmem_slot + wk * kTStride;
  const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
  constexpr uint32_t kIqkv = tc::idesc_f16(128, HS), kI64 = tc::idesc_f16(128, 64),
                     kT32 = tc::idesc_tf32(128, 32), kT64 = tc::idesc_tf32(128, 64);
  uint32_t phQ = 0, phS = 0, phO = 0, phZ = 0;

  const Geom g = make_geom(p.B, p.ph, p.pw, p.shift);
  const int nwin = g.B * g.wh * g.ww;
  const int ntile = (nwin + 1) >> 1;
  const int row = wq * 32 + lane;                 // token row of this thread == its TMEM lane
  const int wsel = wq >> 1, lrow = row & 63;      // window of the pair (warp-uniform), POSITION inside it
  const int ltok = ((7 - (lrow >> 3)) << 3) | (lrow & 7);   // the window token at that position (rows bottom-up)

  // operand descriptors are built once; one K = 16 step advances the start-address field (>>4)
  // by two 8-element k-chunks
  const uint64_t dX = tc::smem_desc(tc::smem_u32(sX), 2048, 128),   // also the fp32 O tile (same strides)
                 dWqkv = tc::smem_desc(tc::smem_u32(sWqkv), L::kWRows * 16, 128),
                 dQ = tc::smem_desc(tc::smem_u32(sQ), 2048, 128),
                 dK = tc::smem_desc(tc::smem_u32(sK), 1024, 128),
                 dV = tc::smem_desc(tc::smem_u32(sV), kVChunk, 128),
                 dWp = tc::smem_desc(tc::smem_u32(sWp), 1024, 128),
                 dHot = tc::smem_desc(tc::smem_u32(sHot), 128, 128),       // LBO = SBO: aliased core matrices
                 dBiasT = tc::smem_desc(tc::smem_u32(sBiasT), 128, 128);
  const uint64_t dXlo = tc::smem_desc(tc::smem_u32(sW + L::wXlo), 2048, 128),
                 dWqkvLo = tc::smem_desc(tc::smem_u32(smem_raw + L::oWqkvLo), L::kWRows * 16, 128),
                 dQlo = tc::smem_desc(tc::smem_u32(sW + L::wQlo), 2048, 128),
                 dKlo = tc::smem_desc(tc::smem_u32(sW + L::wKlo), 1024, 128),
                 // COMP: V^T as a BF16 pair, [32 rows = (window, d)][64 keys] in 8-key chunks of stride kVChunk
                 dVh = tc::smem_desc(tc::smem_u32(sV), kVChunk, 128),
                 dVm = tc::smem_desc(tc::smem_u32(sV + 8 * kVChunk), kVChunk, 128),
                 dOhi = tc::smem_desc(tc::smem_u32(sW + L::wOhi), 2048, 128),
                 dOlo = tc::smem_desc(tc::smem_u32(sW + L::wOlo), 2048, 128),
                 dWpH = tc::smem_desc(tc::smem_u32(smem_raw + L::oWp), 1024, 128),
                 dWpL = tc::smem_desc(tc::smem_u32(smem_raw + L::oWp + 8192), 1024, 128);
  constexpr uint32_t kIpv = tc::idesc_bf16(128, 32);   // (a mixed fp16 x bf16 descriptor is an illegal instruction)
  // [q_h | k_h | v_h] = X * Wqkv[rows of head h]^T -> TMEM buffer `buf` (issued by one thread)
  auto issue_qkv = [&](int h, int buf) {
    const uint64_t db = dWqkv + (uint64_t)(h * HS);
#pragma unroll
    for (int ks = 0; ks < 4; ++ks)
      tc::mma_f16(tbase + kTQ + buf * HS, dX + ks * 256, db + ks * (2 * L::kWRows), kIqkv, ks > 0);
    if constexpr (COMP) {                      // + Xl Wh + Xh Wl (the lo x lo term is below fp32 resolution)
      const uint64_t dl = dWqkvLo + (uint64_t)(h * HS);
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        tc::mma_f16(tbase + kTQ + buf * HS, dXlo + ks * 256, db + ks * (2 * L::kWRows), kIqkv, 1u);
        tc::mma_f16(tbase + kTQ + buf * HS, dX + ks * 256, dl + ks * (2 * L::kWRows), kIqkv, 1u);
      }
    }
  };

  // A window whose 8x8 token block is contiguous in the grid (all windows when shift = 0; all but the
  // last window row / column when shifted) moves by TMA; the cyclically wrapped ones are rare
  // (~4 %) and are read / written by their row threads directly.
  // (window index -> (w, h, b) by multiply-shift, common.cuh: the tile loop contains no integer division)
  auto tma_window = [&](int win, int& x0, int& y0, int& b) -> bool {
    if (win >= nwin) return false;
    int w, h;
    win_pos(g, win, w, h, b);
    x0 = w * kWin + g.shift;
    y0 = h * kWin + g.shift;
    return x0 + kWin <= g.pw && y0 + kWin <= g.ph;
  };
  // z tile of `tile_`: window w -> [2 channel halves][64 slots][128 B swizzled] at sW + w * 16 KB
  // (S16: one box of [64 slots][128 B = 64 bf16] at sW + w * 8 KB)
  constexpr int kWinBytes = S16 ? 8192 : 16384;
  auto issue_z_loads = [&](int tile_) {
    int x0[2], y0[2], b[2];
    const bool t0 = tma_window(tile_ * 2, x0[0], y0[0], b[0]);
    const bool t1 = tma_window(tile_ * 2 + 1, x0[1], y0[1], b[1]);
    if (!(t0 || t1)) return;
    tma::mbar_expect_tx(barZ, ((int)t0 + (int)t1) * kWinBytes);
#pragma unroll
    for (int w = 0; w < 2; ++w)
      if (w == 0 ? t0 : t1) {
        tma::load_4d(&mapZ, barZ, sW + w * kWinBytes, 0, x0[w], y0[w], b[w]);
        if constexpr (!S16) tma::load_4d(&mapZ, barZ, sW + w * kWinBytes + 8192, 32, x0[w], y0[w], b[w]);
      }
  };
  const bool leader = wq == 0 && lane == 0;       // bulk groups are per thread: one fixed thread
  const int tile_stride = gridDim.x * kWorkers;
  // byte offset of 16-byte chunk i of this thread's row inside the swizzled z tile
  const uint32_t zrow_off = (uint32_t)(wsel * kWinBytes + ltok * 128);
  auto zoff = [&](int i) -> uint32_t {      // fp32: i = 0..15 (4 floats each); S16: i = 0..7 (8 bf16 each)
    return zrow_off + (uint32_t)((i >> 3) * 8192 + (((i & 7) ^ (ltok & 7)) << 4));
  };
  // 8 bf16 (one 16-byte chunk) <-> 8 floats
  auto widen8 = [](const uint4& v, float* f) {
    f[0] = __uint_as_float(v.x << 16); f[1] = __uint_as_float(v.x & 0xFFFF0000u);
    f[2] = __uint_as_float(v.y << 16); f[3] = __uint_as_float(v.y & 0xFFFF0000u);
    f[4] = __uint_as_float(v.z << 16); f[5] = __uint_as_float(v.z & 0xFFFF0000u);
    f[6] = __uint_as_float(v.w << 16); f[7] = __uint_as_float(v.w & 0xFFFF0000u);
  };
  // S_h = Hot Bias_h^T + Q' K'^T (issued by one thread).  The bias part does not depend on the tile: it
  // is issued as soon as the S / P columns are free (tile start; PV_{h-1} complete), off the critical path.
  auto issue_bias = [&](int h) {
    const uint64_t db = dBiasT + (uint64_t)(h * (kBiasTHead >> 4));
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) tc::mma_f16(tbase + kTS, dHot + ks * 16, db + ks * 16, kI64, ks > 0);
  };
  auto issue_s = [&]() {
    tc::mma_f16(tbase + kTS, dQ, dK, kI64, 1u);
    tc::mma_f16(tbase + kTS, dQ + 256, dK + 128, kI64, 1u);
    if constexpr (COMP) {
      tc::mma_f16(tbase + kTS, dQlo, dK, kI64, 1u);
      tc::mma_f16(tbase + kTS, dQlo + 256, dK + 128, kI64, 1u);
      tc::mma_f16(tbase + kTS, dQ, dKlo, kI64, 1u);
      tc::mma_f16(tbase + kTS, dQ + 256, dKlo + 128, kI64, 1u);
    }
  };
  // L2 prefetch of a tile's z windows: the load at the end of the current tile then hits L2
  auto prefetch_z = [&](int tile_) {
    int x0, y0, b;
#pragma unroll
    for (int w = 0; w < 2; ++w)
      if (tma_window(tile_ * 2 + w, x0, y0, b)) {
        tma::prefetch_4d(&mapZ, 0, x0, y0, b);
        if constexpr (!S16) tma::prefetch_4d(&mapZ, 32, x0, y0, b);
      }
  };
  if (leader && (int)(blockIdx.x * kWorkers + wk) < ntile) issue_z_loads(blockIdx.x * kWorkers + wk);

  PROF_DECL
  for (int tile = blockIdx.x * kWorkers + wk; tile < ntile; tile += tile_stride) {
    PROF(0);
    if (leader && tile + tile_stride < ntile) prefetch_z(tile + tile_stride);
    const int win = tile * 2 + wsel;
    const bool wvalid = win < nwin;
    int wx0, wy0, wb;
    const bool use_tma = tma_window(win, wx0, wy0, wb);
    int pw_ = 0, ph_ = 0, pb_ = 0;                  // (w, h, b) of this thread's window
    if (wvalid) win_pos(g, win, pw_, ph_, pb_);
    bool any_tma;
    {
      int d0, d1, d2;
      any_tma = tma_window(tile * 2, d0, d1, d2) | tma_window(tile * 2 + 1, d0, d1, d2);
    }
    long long tok = 0;
    float xinv;
    {
      // ---- X = z + ctx of this thread's token row.  Cosine attention is scale-invariant in q and k
      //      (F.normalize, eps 1e-12), so a row of tiny activations (e.g. the zero-padded border)
      //      matters as much as any other: every row is scaled by a power of two to max|x| in [1,2)
      //      before the fp16 conversion and the exact inverse is applied to the fp32 accumulators.
      float x[64];
      if (wvalid) {
        const Slot sl = slot_of(g, pw_, ph_, pb_, ltok);
        tok = sl.tok;
        sReg[row] = sl.region;
        const float4* cp = reinterpret_cast<const float4*>(p.ctx + (size_t)sl.cwin * 64);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float4 v = __ldg(cp + i);
          x[4 * i] = v.x; x[4 * i + 1] = v.y; x[4 * i + 2] = v.z; x[4 * i + 3] = v.w;
        }
      } else {
        sReg[row] = 0;
#pragma unroll
        for (int i = 0; i < 64; ++i) x[i] = 0.f;
      }
      if (any_tma) { tc::mbar_wait(barZ, phZ, &g_tc_error); phZ ^= 1; }
      PROF(1);                                       // ctx + wait for the z tile
      if constexpr (S16) {
        if (use_tma || wvalid) {
          const uint4* zp = reinterpret_cast<const uint4*>(reinterpret_cast<const unsigned short*>(p.z) + (size_t)tok * 64);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint4 v = use_tma ? *reinterpret_cast<const uint4*>(sW + zoff(i)) : __ldg(zp + i);
            float f[8];
            widen8(v, f);
#pragma unroll
            for (int j = 0; j < 8; ++j) x[8 * i + j] += f[j];
          }
        }
      } else if (use_tma) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float4 v = *reinterpret_cast<const float4*>(sW + zoff(i));
          x[4 * i] += v.x; x[4 * i + 1] += v.y; x[4 * i + 2] += v.z; x[4 * i + 3] += v.w;
        }
      } else if (wvalid) {
        const float4* zp = reinterpret_cast<const float4*>(p.z + (size_t)tok * 64);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float4 v = __ldg(zp + i);
          x[4 * i] += v.x; x[4 * i + 1] += v.y; x[4 * i + 2] += v.z; x[4 * i + 3] += v.w;
        }
      }
      float m0 = 0.f, m1 = 0.f, m2 = 0.f, m3 = 0.f;
#pragma unroll
      for (int i = 0; i < 64; i += 4) {
        m0 = fmaxf(m0, fabsf(x[i])); m1 = fmaxf(m1, fabsf(x[i + 1]));
        m2 = fmaxf(m2, fabsf(x[i + 2])); m3 = fmaxf(m3, fabsf(x[i + 3]));
      }
      uint32_t e = (__float_as_uint(fmaxf(fmaxf(m0, m1), fmaxf(m2, m3))) >> 23) & 0xFFu;
      e = e < 1u ? 127u : (e > 253u ? 253u : e);              // all-zero / denormal row: scale 1
      const float sc = __uint_as_float((254u - e) << 23);     // 2^(127-e)
      xinv = __uint_as_float(e << 23) * sScale[6];            // undoes the row scale and the Wqkv scale
      tc::worker_sync(1 + wk);                       // every row is in registers: the tile may be overwritten
#pragma unroll
      for (int kc = 0; kc < 8; ++kc) {               // fp16 X operand, own row, one 16-byte slot per k-chunk
        const uint4 hv = make_uint4(tc::pack_f16(x[8 * kc] * sc, x[8 * kc + 1] * sc),
                                    tc::pack_f16(x[8 * kc + 2] * sc, x[8 * kc + 3] * sc),
                                    tc::pack_f16(x[8 * kc + 4] * sc, x[8 * kc + 5] * sc),
                                    tc::pack_f16(x[8 * kc + 6] * sc, x[8 * kc + 7] * sc));
        *reinterpret_cast<uint4*>(sX + kc * 2048 + row * 16) = hv;
        if constexpr (COMP)
          *reinterpret_cast<uint4*>(sW + L::wXlo + kc * 2048 + row * 16) =
              make_uint4(tc::lo_f16(x[8 * kc] * sc, x[8 * kc + 1] * sc, hv.x),
                         tc::lo_f16(x[8 * kc + 2] * sc, x[8 * kc + 3] * sc, hv.y),
                         tc::lo_f16(x[8 * kc + 4] * sc, x[8 * kc + 5] * sc, hv.z),
                         tc::lo_f16(x[8 * kc + 6] * sc, x[8 * kc + 7] * sc, hv.w));
      }
    }
    // ---- shift mask of this row as a 64-bit key set (analytic 9-region rule)
    uint32_t mlo = 0, mhi = 0;
    if (p.shift > 0 && wvalid && (pw_ == g.ww - 1 || ph_ == g.wh - 1)) {   // only the last window row / column
      const int rg = sReg[row];
      const int* reg = sReg + wsel * 64;
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        mlo |= (uint32_t)(reg[c] != rg) << c;
        mhi |= (uint32_t)(reg[c + 32] != rg) << c;
      }
    }
    PROF(17);                                        // row convert + mask
    tc::fence_async_smem();
    tc::worker_sync(1 + wk);
    if (wq == 0) {                                   // warp-uniform: one elected lane issues
      tc::tc_fence_after();
      if (tc::elect_one()) { issue_qkv(0, 0); tc::mma_commit(barQ); issue_bias(0); }
      __syncwarp();
    }
    PROF(2);                                         // sync + first qkv issue

    // per-head attention outputs of this row; indexed by the (rolled) head loop, so the array lives
    // in local memory: 10-16 stores per head, one 64-float read-back per tile
    float o[64];
#pragma unroll
    for (int i = A; i < 64; ++i) o[i] = 0.f;           // 6 heads: columns 60..63 are padding
    float inv_prev = 0.f;

#pragma unroll 1
    for (int h = 0; h < NH; ++h) {
      const int buf = h & 1;
      // ---- wait for [q|k|v]_h
      tc::mbar_wait(barQ + buf, (phQ >> buf) & 1u, &g_tc_error);
      phQ ^= 1u << buf;
      tc::tc_fence_after();
      PROF(3);                                         // wait [q|k|v]
      float qkv[HS];
      if constexpr (HS == 32) tc::tmem_ld32(tbase + lane_addr + kTQ + buf * HS, qkv);
      else tc::tmem_ld48(tbase + lane_addr + kTQ + buf * HS, qkv);
      float qs = 0.f, ks = 0.f;
      {
        const float* b = sBqkv + h * HS;               // packed [q | k | v] with stride PS; pad rows are exact zeros
        if constexpr (!COMP) {                         // two values per instruction
          static_assert(HD % 2 == 0 && PS % 2 == 0, "pairs");
          float q0 = 0.f, q1 = 0.f, k0 = 0.f, k1 = 0.f;
#pragma unroll
          for (int d = 0; d < 3 * PS; d += 2)          // undo the row's power-of-two input scale
            if (d % PS < HD) fma2(qkv[d], qkv[d + 1], xinv, b[d], b[d + 1]);
#pragma unroll
          for (int d = 0; d < HD; d += 2) {
            const float2 sq = __ffma2_rn(make_float2(qkv[d], qkv[d + 1]), make_float2(qkv[d], qkv[d + 1]), make_float2(q0, q1));
            q0 = sq.x; q1 = sq.y;
            const float2 sk = __ffma2_rn(make_float2(qkv[PS + d], qkv[PS + d + 1]), make_float2(qkv[PS + d], qkv[PS + d + 1]),
                                         make_float2(k0, k1));
            k0 = sk.x; k1 = sk.y;
          }
          qs = q0 + q1;
          ks = k0 + k1;
        } else {
#pragma unroll
          for (int d = 0; d < HD; ++d) {
            qkv[d] = fmaf(qkv[d], xinv, b[d]);           // undo the row's power-of-two input scale
            qkv[PS + d] = fmaf(qkv[PS + d], xinv, b[PS + d]);
            qkv[2 * PS + d] = fmaf(qkv[2 * PS + d], xinv, b[2 * PS + d]);
            qs = fmaf(qkv[d], qkv[d], qs);
            ks = fmaf(qkv[PS + d], qkv[PS + d], ks);
          }
        }
      }
      const float qmul = sScale[h] * rsqrt_op<COMP>(fmaxf(qs, kNormEps * kNormEps));   // v / max(|v|, eps)
      const float kmul = rsqrt_op<COMP>(fmaxf(ks, kNormEps * kNormEps));
      uint32_t qp[8], kp[8];
      [[maybe_unused]] uint32_t ql[8], kl[8];
      if constexpr (!COMP) {
#pragma unroll
        for (int d = 0; d < HD; d += 2) {
          mul2(qkv[d], qkv[d + 1], qmul);
          mul2(qkv[PS + d], qkv[PS + d + 1], kmul);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          qp[j] = 2 * j < HD ? tc::pack_f16(qkv[2 * j], qkv[2 * j + 1]) : 0u;
          kp[j] = 2 * j < HD ? tc::pack_f16(qkv[PS + 2 * j], qkv[PS + 2 * j + 1]) : 0u;
        }
      } else
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        qp[j] = 2 * j < HD ? tc::pack_f16(qkv[2 * j] * qmul, qkv[2 * j + 1] * qmul) : 0u;
        kp[j] = 2 * j < HD ? tc::pack_f16(qkv[PS + 2 * j] * kmul, qkv[PS + 2 * j + 1] * kmul) : 0u;
        if constexpr (COMP) {
          ql[j] = 2 * j < HD ? tc::lo_f16(qkv[2 * j] * qmul, qkv[2 * j + 1] * qmul, qp[j]) : 0u;
          kl[j] = 2 * j < HD ? tc::lo_f16(qkv[PS + 2 * j] * kmul, qkv[PS + 2 * j + 1] * kmul, kp[j]) : 0u;
        }
      }
      // ---- previous head's O: its PV MMAs read P (TMEM) and V^T (smem) and wrote O over the
      //      accumulator buffer that [q|k|v]_{h+1} will reuse: read it before anything is reused
      PROF(4);                                         // qkv ld + normalise + pack
      if (h > 0) {
        tc::mbar_wait(barO, phO, &g_tc_error); phO ^= 1;
        tc::tc_fence_after();
        float ov[16];
        tc::tmem_ld16(tbase + lane_addr + kTQ + (buf ^ 1) * HS + wsel * 16, ov);
        if (wq == 0) {                                 // PV_{h-1} is complete: the S / P columns are free
          if (tc::elect_one()) issue_bias(h);
          __syncwarp();
        }
        if constexpr (NH == 6) inv_prev = rcp_op<COMP>(ov[HD]);   // the ones row of V^T: sum of the row's (tf32) P
#pragma unroll
        for (int d = 0; d < HD; d += 2) {
          mul2(ov[d], ov[d + 1], inv_prev);
          o[(h - 1) * HD + d] = ov[d];
          o[(h - 1) * HD + d + 1] = ov[d + 1];
        }
      }
      PROF(5);                                         // wait PV(h-1) + O read
      // ---- operands of head h: Q' = [q|0] or [0|q], K' = [k(w0)|k(w1)], V^T rows (window, d)
      {
        const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
        uint4* q0 = reinterpret_cast<uint4*>(sQ + row * 16);                 // k-chunk stride 2048 B
        q0[(2 * wsel) * 128] = make_uint4(qp[0], qp[1], qp[2], qp[3]);
        q0[(2 * wsel + 1) * 128] = make_uint4(qp[4], qp[5], qp[6], qp[7]);
        if (h == 0) {                                  // the other window's half stays zero for all heads
          q0[(2 - 2 * wsel) * 128] = z4;
          q0[(3 - 2 * wsel) * 128] = z4;
        }
        uint4* k0 = reinterpret_cast<uint4*>(sK + lrow * 16);                // k-chunk stride 1024 B
        k0[(2 * wsel) * 64] = make_uint4(kp[0], kp[1], kp[2], kp[3]);
        k0[(2 * wsel + 1) * 64] = make_uint4(kp[4], kp[5], kp[6], kp[7]);
        if constexpr (COMP) {
          uint4* ql0 = reinterpret_cast<uint4*>(sW + L::wQlo + row * 16);
          ql0[(2 * wsel) * 128] = make_uint4(ql[0], ql[1], ql[2], ql[3]);
          ql0[(2 * wsel + 1) * 128] = make_uint4(ql[4], ql[5], ql[6], ql[7]);
          if (h == 0) {
            ql0[(2 - 2 * wsel) * 128] = z4;
            ql0[(3 - 2 * wsel) * 128] = z4;
          }
          uint4* kl0 = reinterpret_cast<uint4*>(sW + L::wKlo + lrow * 16);
          kl0[(2 * wsel) * 64] = make_uint4(kl[0], kl[1], kl[2], kl[3]);
          kl0[(2 * wsel + 1) * 64] = make_uint4(kl[4], kl[5], kl[6], kl[7]);
          // V^T as a BF16 pair: element ((window, d), key) at (key / 8) * kVChunk + (window * 16 + d) * 16 + (key % 8) * 2
          unsigned short* vh = reinterpret_cast<unsigned short*>(sV + (lrow >> 3) * kVChunk + wsel * 256) + (lrow & 7);
          unsigned short* vm = vh + 8 * kVChunk / 2;
#pragma unroll
          for (int d = 0; d < HD; ++d) {
            const float v = qkv[2 * PS + d];
            const uint32_t hb = tc::pack_bf16(v, 0.f) & 0xFFFFu;
            vh[d * 8] = (unsigned short)hb;
            vm[d * 8] = (unsigned short)(tc::pack_bf16(v - __uint_as_float(hb << 16), 0.f) & 0xFFFFu);
          }
          if (NH == 6 && h == 0) { vh[HD * 8] = 0x3F80u; vm[HD * 8] = 0u; }   // ones row: O column HD = row sum of P
        } else {
        float* vt = reinterpret_cast<float*>(sV + (lrow >> 2) * kVChunk + wsel * 256) + (lrow & 3);
#pragma unroll
        for (int d = 0; d < HD; ++d) vt[d * 4] = round_tf32_add(qkv[2 * PS + d]);
        if (NH == 6 && h == 0) vt[HD * 4] = 1.0f;      // ones row (both windows): O column HD = row sum of P
        }
      }
      PROF(6);                                         // operand stores
      tc::fence_async_smem();
      tc::tc_fence_before();
      tc::worker_sync(1 + wk);
      PROF(7);                                         // fence + sync
      if (wq == 0) {
        tc::tc_fence_after();
        if (tc::elect_one()) {
          issue_s();
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
      PROF(8);                                         // S issue + wait
      // ---- softmax of this row over its own window's 64 keys (log2 domain; S already holds the bias).
      //      Fast heads (sFast): p = ex2(s) with no row maximum; other heads subtract the exact row maximum.
      //      6 heads: the denominator comes out of the PV MMA (ones row of V^T); 4 heads: summed here.
      {
        float s[64];
        tc::tmem_ld64(tbase + lane_addr + kTS, s);
        if (mlo | mhi) {
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            if ((mlo >> c) & 1u) s[c] += -100.0f * kLog2e;
            if ((mhi >> c) & 1u) s[c + 32] += -100.0f * kLog2e;
          }
        }
        if (!COMP && sFast[h]) {
#pragma unroll
          for (int c = 0; c < 64; ++c) s[c] = tc::ex2_approx(s[c]);
        } else {
          float m4[4] = {s[0], s[1], s[2], s[3]};
#pragma unroll
          for (int c = 4; c < 64; ++c) m4[c & 3] = fmaxf(m4[c & 3], s[c]);
          const float mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
#pragma unroll
          for (int c = 0; c < 64; ++c) s[c] = tc::ex2_approx(s[c] - mx);
        }
        if constexpr (NH != 6) {
          float sum4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int c = 0; c < 64; ++c) sum4[c & 3] += s[c];
          inv_prev = rcp_op<COMP>((sum4[0] + sum4[1]) + (sum4[2] + sum4[3]));
        }
        if constexpr (COMP) {                            // P <= 1 as a BF16 pair: hi in columns [0,32), mid in [32,64)
          float ph[32], pl[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const uint32_t hv = tc::pack_bf16(s[2 * j], s[2 * j + 1]);
            ph[j] = __uint_as_float(hv);
            pl[j] = __uint_as_float(tc::lo_bf16(s[2 * j], s[2 * j + 1], hv));
          }
          tc::tmem_st32(tbase + lane_addr + kTS, ph);
          tc::tmem_st32(tbase + lane_addr + kTS + 32, pl);
        } else {
          tc::tmem_st64(tbase + lane_addr + kTS, s);     // unnormalised P (tf32 operand) over S
        }
        tc::tmem_wait_st();
      }
      PROF(9);                                         // softmax + P store
      tc::tc_fence_before();
      tc::worker_sync(1 + wk);
      PROF(10);                                        // sync
      if (wq == 0) {
        tc::tc_fence_after();
        if (tc::elect_one()) {
          if constexpr (COMP) {                          // Ph Vh + Pm Vh + Ph Vm, K = 16 keys per step
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
              tc::mma_f16_ts(tbase + kTQ + buf * HS, tbase + kTS + ks * 8, dVh + ks * (2 * kVChunk / 16), kIpv, ks > 0);
              tc::mma_f16_ts(tbase + kTQ + buf * HS, tbase + kTS + 32 + ks * 8, dVh + ks * (2 * kVChunk / 16), kIpv, 1u);
              tc::mma_f16_ts(tbase + kTQ + buf * HS, tbase + kTS + ks * 8, dVm + ks * (2 * kVChunk / 16), kIpv, 1u);
            }
          } else {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)
              tc::mma_tf32_ts(tbase + kTQ + buf * HS, tbase + kTS + ks * 8, dV + ks * (2 * kVChunk / 16), kT32, ks > 0);
          }
          tc::mma_commit(barO);
        }
        __syncwarp();
      }
      PROF(11);                                        // PV issue
    }
    PROF(16);                                          // PV issue (last head)
    // ---- last head's O, then the output projection
    {
      tc::mbar_wait(barO, phO, &g_tc_error); phO ^= 1;
      tc::tc_fence_after();
      float ov[16];
      tc::tmem_ld16(tbase + lane_addr + kTQ + ((NH - 1) & 1) * HS + wsel * 16, ov);
      if constexpr (NH == 6) inv_prev = rcp_op<COMP>(ov[HD]);
#pragma unroll
      for (int d = 0; d < HD; ++d) o[(NH - 1) * HD + d] = ov[d] * inv_prev;
    }
    // ---- every MMA that read [X|Q'|K'|V^T] is complete: fetch the z tile again for the residual now
    //      (an L2 hit), so that it lands behind the projection and the LayerNorm
    if (RESID && leader) issue_z_loads(tile);
    // ---- O row (tf32-rounded) -> TMEM over the dead [q|k|v] buffers: the A operand of the projection
    [[maybe_unused]] float oinv = 1.0f;
    {
      float ot[64];
      if constexpr (COMP) {                            // O as an fp16 pair in shared memory, row-scaled like X
        float m0 = 0.f, m1 = 0.f;
#pragma unroll
        for (int i = 0; i < 64; i += 2) { m0 = fmaxf(m0, fabsf(o[i])); m1 = fmaxf(m1, fabsf(o[i + 1])); }
        uint32_t e = (__float_as_uint(fmaxf(m0, m1)) >> 23) & 0xFFu;
        e = e < 1u ? 127u : (e > 253u ? 253u : e);
        const float sc = __uint_as_float((254u - e) << 23);
        oinv = __uint_as_float(e << 23) * sProjInv;
#pragma unroll
        for (int i = 0; i < 64; ++i) ot[i] = o[i];
#pragma unroll
        for (int kc = 0; kc < 8; ++kc) {
          const uint4 hv = make_uint4(tc::pack_f16(o[8 * kc] * sc, o[8 * kc + 1] * sc),
                                      tc::pack_f16(o[8 * kc + 2] * sc, o[8 * kc + 3] * sc),
                                      tc::pack_f16(o[8 * kc + 4] * sc, o[8 * kc + 5] * sc),
                                      tc::pack_f16(o[8 * kc + 6] * sc, o[8 * kc + 7] * sc));
          *reinterpret_cast<uint4*>(sW + L::wOhi + kc * 2048 + row * 16) = hv;
          *reinterpret_cast<uint4*>(sW + L::wOlo + kc * 2048 + row * 16) =
              make_uint4(tc::lo_f16(o[8 * kc] * sc, o[8 * kc + 1] * sc, hv.x),
                         tc::lo_f16(o[8 * kc + 2] * sc, o[8 * kc + 3] * sc, hv.y),
                         tc::lo_f16(o[8 * kc + 4] * sc, o[8 * kc + 5] * sc, hv.z),
                         tc::lo_f16(o[8 * kc + 6] * sc, o[8 * kc + 7] * sc, hv.w));
        }
        tc::fence_async_smem();
      } else {
#pragma unroll
        for (int i = 0; i < 64; ++i) ot[i] = round_tf32_add(o[i]);
        tc::tmem_st64(tbase + lane_addr + kTQ, ot);
      }
      if (SAVEO && wvalid) {                             // bf16 copy for the backward (window order)
        uint4* dst = reinterpret_cast<uint4*>(p.o16 + ((size_t)tile * 128 + wsel * 64 + ltok) * 64);
#pragma unroll
        for (int i = 0; i < 8; ++i)
          dst[i] = make_uint4(tc::pack_bf16(ot[8 * i], ot[8 * i + 1]), tc::pack_bf16(ot[8 * i + 2], ot[8 * i + 3]),
                              tc::pack_bf16(ot[8 * i + 4], ot[8 * i + 5]), tc::pack_bf16(ot[8 * i + 6], ot[8 * i + 7]));
      }
      tc::tmem_wait_st();
    }
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    if (wq == 0) {
      tc::tc_fence_after();
      if (tc::elect_one()) {
        if constexpr (COMP) {                          // Oh Wh + Ol Wh + Oh Wl (fp16, K = 16 per step)
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            tc::mma_f16(tbase + kTS, dOhi + ks * 256, dWpH + ks * 128, kI64, ks > 0);
            tc::mma_f16(tbase + kTS, dOlo + ks * 256, dWpH + ks * 128, kI64, 1u);
            tc::mma_f16(tbase + kTS, dOhi + ks * 256, dWpL + ks * 128, kI64, 1u);
          }
        } else {
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            tc::mma_tf32_ts(tbase + kTS, tbase + kTQ + ks * 8, dWp + ks * 128, kT64, ks > 0);
        }
        tc::mma_commit(barS);
      }
      __syncwarp();
    }
    tc::mbar_wait(barS, phS, &g_tc_error); phS ^= 1;
    tc::tc_fence_after();
    PROF(12);                                          // O tile + proj issue + wait
    // ---- + proj bias, LayerNorm(norm1) of the row held in registers; a = z + LN(.) in place in the
    //      z tile (TMA windows) or straight to global (wrapped windows)
    {
      float y[64];
      tc::tmem_ld64(tbase + lane_addr + kTS, y);
      float mean4[4] = {0.f, 0.f, 0.f, 0.f};      // four partial sums: short dependency chains
#pragma unroll
      for (int i = 0; i < 64; ++i) {
        if constexpr (COMP) y[i] = fmaf(y[i], oinv, sBp[i]); else y[i] += sBp[i];
        mean4[i & 3] += y[i];
      }
      const float mean = ((mean4[0] + mean4[1]) + (mean4[2] + mean4[3])) * (1.0f / 64.f);
      float var4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 64; ++i) { y[i] -= mean; var4[i & 3] = fmaf(y[i], y[i], var4[i & 3]); }
      const float var = (var4[0] + var4[1]) + (var4[2] + var4[3]);
      const float rstd = COMP ? 1.0f / sqrtf(var * (1.0f / 64.f) + kLnEps) : rsqrt_op<COMP>(var * (1.0f / 64.f) + kLnEps);
#pragma unroll
      for (int i = 0; i < 64; ++i) y[i] = y[i] * rstd * sLnW[i] + sLnB[i];
      PROF(13);                                        // LN epilogue
      if (RESID && any_tma) { tc::mbar_wait(barZ, phZ, &g_tc_error); phZ ^= 1; }
      PROF(14);
      if constexpr (S16) {
        if (use_tma || wvalid) {
          const uint4* zp = reinterpret_cast<const uint4*>(reinterpret_cast<const unsigned short*>(p.z) + (size_t)tok * 64);
          uint4* ap = reinterpret_cast<uint4*>(reinterpret_cast<unsigned short*>(p.a) + (size_t)tok * 64);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            uint4* q = reinterpret_cast<uint4*>(sW + zoff(i));
            float f[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (RESID) widen8(use_tma ? *q : __ldg(zp + i), f);
            const uint4 r = make_uint4(tc::pack_bf16(f[0] + y[8 * i], f[1] + y[8 * i + 1]),
                                       tc::pack_bf16(f[2] + y[8 * i + 2], f[3] + y[8 * i + 3]),
                                       tc::pack_bf16(f[4] + y[8 * i + 4], f[5] + y[8 * i + 5]),
                                       tc::pack_bf16(f[6] + y[8 * i + 6], f[7] + y[8 * i + 7]));
            if (use_tma) *q = r; else ap[i] = r;
          }
        }
      } else if (use_tma) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float4* q = reinterpret_cast<float4*>(sW + zoff(i));
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
          if (RESID) v = *q;
          *q = make_float4(v.x + y[4 * i], v.y + y[4 * i + 1], v.z + y[4 * i + 2], v.w + y[4 * i + 3]);
        }
      } else if (wvalid) {
        const float4* zp = reinterpret_cast<const float4*>(p.z + (size_t)tok * 64);
        float4* ap = reinterpret_cast<float4*>(p.a + (size_t)tok * 64);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
          if (RESID) v = __ldg(zp + i);
          ap[i] = make_float4(v.x + y[4 * i], v.y + y[4 * i + 1], v.z + y[4 * i + 2], v.w + y[4 * i + 3]);
        }
      }
    }
    tc::fence_async_smem();                            // generic writes -> visible to the TMA engine
    tc::tc_fence_before();
    tc::worker_sync(1 + wk);
    if (leader) {
      // one bulk group per window half, so that the next tile's window 0 can be fetched as soon as the
      // first store has drained its half of the buffer (the second store overlaps that load)
#pragma unroll
      for (int w = 0; w < 2; ++w) {
        int x0, y0, b;
        if (tma_window(tile * 2 + w, x0, y0, b)) {
          tma::store_4d(&mapA, sW + w * kWinBytes, 0, x0, y0, b);
          if constexpr (!S16) tma::store_4d(&mapA, sW + w * kWinBytes + 8192, 32, x0, y0, b);
        }
        tma::commit_group();                           // (an empty group completes immediately)
      }
      const int nt = tile + tile_stride;
      if (nt < ntile) {
        int x0[2], y0[2], b[2];
        const bool t0 = tma_window(nt * 2, x0[0], y0[0], b[0]);
        const bool t1 = tma_window(nt * 2 + 1, x0[1], y0[1], b[1]);
        if (t0 || t1) tma::mbar_expect_tx(barZ, ((int)t0 + (int)t1) * kWinBytes);
        tma::wait_read_1();                            // window-0 half drained
        if (t0) {
          tma::load_4d(&mapZ, barZ, sW, 0, x0[0], y0[0], b[0]);
          if constexpr (!S16) tma::load_4d(&mapZ, barZ, sW + 8192, 32, x0[0], y0[0], b[0]);
        }
        tma::wait_read_all();
        if (t1) {
          tma::load_4d(&mapZ, barZ, sW + kWinBytes, 0, x0[1], y0[1], b[1]);
          if constexpr (!S16) tma::load_4d(&mapZ, barZ, sW + kWinBytes + 8192, 32, x0[1], y0[1], b[1]);
        }
      }
    }
    PROF(15);                                          // store + next-tile load issue
  }
  if (leader) tma::wait_all();
  PROF_FLUSH;
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(*tmem_slot);
}

#ifdef B200MAR_ATTN_PROF
}  // namespace b200mar
extern "C" int b200mar_attn_prof(unsigned long long* out32, int reset) {
  unsigned long long z[32] = {0};
  if (cudaMemcpyFromSymbol(out32, b200mar::g_attn_prof, sizeof(z)) != cudaSuccess) return -1;
  if (reset && cudaMemcpyToSymbol(b200mar::g_attn_prof, z, sizeof(z)) != cudaSuccess) return -1;
  return 0;
}
namespace b200mar {
#endif

template <int NH, bool RESID, bool SAVEO = false, bool S16 = false, bool COMP = false>
static int launch_attn_tc_fwd(const AttnParams& p, cudaStream_t st) {
  constexpr int kWorkers = AttnWorkers<NH, COMP>::value;
  const size_t smem = AttnTcSmem<kWorkers, COMP>::kBytes;
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(attn_tc_fwd_kernel<NH, RESID, SAVEO, S16, COMP>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured.done();
  }
  const int nwin = p.B * (p.ph / kWin) * (p.pw / kWin);
  const int ntile = (nwin + 1) / 2;
  int grid = (ntile + kWorkers - 1) / kWorkers;
  if (grid > sm_count()) grid = sm_count();
  CUtensorMap mZ, mA;
  if constexpr (S16) {
    if (int rc = tma::encode_windows16(&mZ, p.z, p.B, p.ph, p.pw)) return rc;
    if (int rc = tma::encode_windows16(&mA, p.a, p.B, p.ph, p.pw)) return rc;
  } else {
    if (int rc = tma::encode_windows(&mZ, p.z, p.B, p.ph, p.pw)) return rc;
    if (int rc = tma::encode_windows(&mA, p.a, p.B, p.ph, p.pw)) return rc;
  }
  attn_tc_fwd_kernel<NH, RESID, SAVEO, S16, COMP><<<grid, 128 * kWorkers, smem, st>>>(mZ, mA, p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace b200mar

extern "C" int b200mar_attn_tc_fwd(const float* z, const float* ctx, const float* qkv_w,
                                   const float* qkv_b, const float* proj_w, const float* proj_b,
                                   const float* logit_scale, const float* bias_table,
                                   const float* ln_w, const float* ln_b, float* a, int B, int ph,
                                   int pw, int nh, int shift, int residual, void* stream) {
  using namespace b200mar;
  if (ph % kWin || pw % kWin || ph < kWin || pw < kWin || (shift != 0 && shift != kShift)) return -1;
  if ((long long)B * ph * pw >= (1ll << 31)) return -1;
  AttnParams p{};
  p.z = z; p.ctx = ctx; p.qkv_w = qkv_w; p.qkv_b = qkv_b; p.proj_w = proj_w; p.proj_b = proj_b;
  p.logit_scale = logit_scale; p.bias_table = bias_table; p.ln_w = ln_w; p.ln_b = ln_b;
  p.a = a; p.B = B; p.ph = ph; p.pw = pw; p.shift = shift;
  cudaStream_t st = (cudaStream_t)stream;
  if (nh == 6) return residual ? launch_attn_tc_fwd<6, true>(p, st) : launch_attn_tc_fwd<6, false>(p, st);
  if (nh == 4) return residual ? launch_attn_tc_fwd<4, true>(p, st) : launch_attn_tc_fwd<4, false>(p, st);
  return -1;
}

// Precision level "tf32c": as b200mar_attn_tc_fwd(residual = 1) / b200mar_attn_tc_fwd_save (o16 != NULL) with the qkv
// projection on compensated fp16 hi + lo operand pairs (two workers per SM; see AttnTcSmem).
extern "C" int b200mar_attn_tc_fwd_comp(const float* z, const float* ctx, const float* qkv_w,
                                        const float* qkv_b, const float* proj_w, const float* proj_b,
                                        const float* logit_scale, const float* bias_table,
                                        const float* ln_w, const float* ln_b, float* a, void* o16, int B,
                                        int ph, int pw, int nh, int shift, void* stream) {
  using namespace b200mar;
  if (ph % kWin || pw % kWin || ph < kWin || pw < kWin || (shift != 0 && shift != kShift)) return -1;
  if ((long long)B * ph * pw >= (1ll << 31)) return -1;
  AttnParams p{};
  p.z = z; p.ctx = ctx; p.qkv_w = qkv_w; p.qkv_b = qkv_b; p.proj_w = proj_w; p.proj_b = proj_b;
  p.logit_scale = logit_scale; p.bias_table = bias_table; p.ln_w = ln_w; p.ln_b = ln_b;
  p.a = a; p.o16 = (unsigned short*)o16; p.B = B; p.ph = ph; p.pw = pw; p.shift = shift;
  cudaStream_t st = (cudaStream_t)stream;
  if (nh == 6) return o16 ? launch_attn_tc_fwd<6, true, true, false, true>(p, st)
                          : launch_attn_tc_fwd<6, true, false, false, true>(p, st);
  if (nh == 4) return o16 ? launch_attn_tc_fwd<4, true, true, false, true>(p, st)
                          : launch_attn_tc_fwd<4, true, false, false, true>(p, st);
  return -1;
}

// Precision mode "bf16" (BASELINE.json configs[1]): the same kernel with the residual stream in BF16 - z16 in,
// a16 = bf16(z + LN1(attn)) out, both [B, ph*pw, 64] bf16; ctx and the parameters stay fp32.
extern "C" int b200mar_attn_tc_fwd_s16(const void* z16, const float* ctx, const float* qkv_w,
                                       const float* qkv_b, const float* proj_w, const float* proj_b,
                                       const float* logit_scale, const float* bias_table,
                                       const float* ln_w, const float* ln_b, void* a16, int B, int ph,
                                       int pw, int nh, int shift, void* stream) {
  using namespace b200mar;
  if (ph % kWin || pw % kWin || ph < kWin || pw < kWin || (shift != 0 && shift != kShift)) return -1;
  if ((long long)B * ph * pw >= (1ll << 31)) return -1;
  AttnParams p{};
  p.z = (const float*)z16; p.ctx = ctx; p.qkv_w = qkv_w; p.qkv_b = qkv_b; p.proj_w = proj_w; p.proj_b = proj_b;
  p.logit_scale = logit_scale; p.bias_table = bias_table; p.ln_w = ln_w; p.ln_b = ln_b;
  p.a = (float*)a16; p.B = B; p.ph = ph; p.pw = pw; p.shift = shift;
  cudaStream_t st = (cudaStream_t)stream;
  if (nh == 6) return launch_attn_tc_fwd<6, true, false, true>(p, st);
  if (nh == 4) return launch_attn_tc_fwd<4, true, false, true>(p, st);
  return -1;
}

// Training forward: a = z + LN1(attn) as b200mar_attn_tc_fwd(residual = 1) and, in addition, the
// pre-projection attention output O as bf16 [ceil(nwin/2)*128][64] in window order (o16), which
// b200mar_attn_bwd_tc consumes instead of recomputing it.
extern "C" int b200mar_attn_tc_fwd_save(const float* z, const float* ctx, const float* qkv_w,
                                        const float* qkv_b, const float* proj_w, const float* proj_b,
                                        const float* logit_scale, const float* bias_table,
                                        const float* ln_w, const float* ln_b, float* a, void* o16, int B,
                                        int ph, int pw, int nh, int shift, void* stream) {
  using namespace b200mar;
  if (ph % kWin || pw % kWin || ph < kWin || pw < kWin || (shift != 0 && shift != kShift) || !o16) return -1;
  if ((long long)B * ph * pw >= (1ll << 31)) return -1;
  AttnParams p{};
  p.z = z; p.ctx = ctx; p.qkv_w = qkv_w; p.qkv_b = qkv_b; p.proj_w = proj_w; p.proj_b = proj_b;
  p.logit_scale = logit_scale; p.bias_table = bias_table; p.ln_w = ln_w; p.ln_b = ln_b;
  p.a = a; p.o16 = (unsigned short*)o16; p.B = B; p.ph = ph; p.pw = pw; p.shift = shift;
  cudaStream_t st = (cudaStream_t)stream;
  if (nh == 6) return launch_attn_tc_fwd<6, true, true>(p, st);
  if (nh == 4) return launch_attn_tc_fwd<4, true, true>(p, st);
  return -1;
}
