// attn_common.cuh — parameter block and shared-memory carve-up shared by the fused
// window-attention forward and backward kernels.
#pragma once
#include "common.cuh"

namespace b200mar {

struct AttnParams {
  // inputs
  const float* z;            // [B, ph*pw, 64] block input
  const float* ctx;          // [B, wh*ww, 64] n-gram context
  const float* qkv_w;        // [3A, 64]
  const float* qkv_b;        // [3A]
  const float* proj_w;       // [64, A]
  const float* proj_b;       // [64]
  const float* logit_scale;  // [nh]
  const float* bias_table;   // [225, nh]
  const float* ln_w;         // [64]
  const float* ln_b;         // [64]
  float* a;                  // fwd out: z + LN1(attn)
  unsigned short* o16;       // attn_tc fwd (training): bf16 [ntile*128][64] pre-projection attention output,
                             // window order, saved for attn_bwd_tc (null: not written)
  // backward only
  const float* da;
  float* dz;
  float* dctx;
  float* d_qkv_w; float* d_qkv_b; float* d_proj_w; float* d_proj_b;
  float* d_logit_scale; float* d_bias_table; float* d_ln_w; float* d_ln_b;
  int B, ph, pw, shift;
};

template <int NH>
struct AttnSmem {
  static constexpr int HD = kDim / NH;       // 10 (6 heads) or 16 (4 heads)
  static constexpr int A = HD * NH;          // 60 or 64
  static constexpr int HDP = HD + 1;         // odd stride: conflict-free lane=token reads
  static constexpr int kBiasLd = 228;        // 225 padded
  static constexpr int kHeadFloats = ((NH * kTok * HDP + 3) / 4) * 4;
  static constexpr int oWqkv = 0;
  static constexpr int oWp = oWqkv + 192 * kLd64;
  static constexpr int oBqkv = oWp + kDim * kLd64;
  static constexpr int oBp = oBqkv + 192;
  static constexpr int oLnW = oBp + kDim;
  static constexpr int oLnB = oLnW + kDim;
  static constexpr int oScale = oLnB + kDim;
  static constexpr int oBias = oScale + 8;
  static constexpr int oX = oBias + NH * kBiasLd;
  static constexpr int oQ = oX + kTok * kLd64;
  static constexpr int oK = oQ + kHeadFloats;
  static constexpr int oV = oK + kHeadFloats;
  static constexpr int oO = oV + kHeadFloats;
  static constexpr int oP = oO + kTok * kLd64;
  static constexpr int oTok = oP + 8 * kTok;
  static constexpr int kFloatsFwd = oTok + 4 * kTok;
  // ---- backward carve-up (shares the weight block [0, oX) with the forward)
  static constexpr int kLdDY = 196;                       // [64][192] qkv-gradient tile, == 4 mod 32
  static constexpr int bXD = oX;                          // X = z+ctx (phases 0-1, 6) / dO (phases 4-5)
  static constexpr int bQ = bXD + kTok * kLd64;
  static constexpr int bK = bQ + kHeadFloats;
  static constexpr int bV = bK + kHeadFloats;
  static constexpr int bNq = bV + kHeadFloats;            // |q| per (head, token)
  static constexpr int bNk = bNq + NH * kTok;
  static constexpr int bBuf = bNk + NH * kTok;            // row buffer: P / Yp,dYp / dS / dX
  static constexpr int bDY = bBuf + kTok * kLd64;         // dY tile; aliases O (+0) and gathered da (+64*68)
  static constexpr int bDBias = bDY + kTok * kLdDY;       // [NH][228] bias-table gradient accumulators
  static constexpr int bTok = bDBias + NH * kBiasLd;
  static constexpr int kFloatsBwd = bTok + 4 * kTok;
  static_assert(2 * kTok * kLd64 <= kTok * kLdDY, "O and da must fit inside the dY tile");
  static_assert(kFloatsBwd * 4 <= 232448, "attn_bwd shared memory exceeds 227 KB");
};

// Stage the block's weights in shared memory (once per persistent CTA).
template <int NH>
__device__ __forceinline__ void load_attn_weights(float* smem, const AttnParams& p) {
  using L = AttnSmem<NH>;
  load_matrix(smem + L::oWqkv, kLd64, p.qkv_w, 3 * L::A, kDim, 192);
  // proj weight [64][A]: rows of A floats (A % 4 == 0)
  load_matrix(smem + L::oWp, kLd64, p.proj_w, kDim, L::A, kDim);
  for (int i = threadIdx.x; i < 192; i += blockDim.x)
    smem[L::oBqkv + i] = i < 3 * L::A ? p.qkv_b[i] : 0.f;
  for (int i = threadIdx.x; i < kDim; i += blockDim.x) {
    smem[L::oBp + i] = p.proj_b[i];
    smem[L::oLnW + i] = p.ln_w[i];
    smem[L::oLnB + i] = p.ln_b[i];
  }
  if (threadIdx.x < NH)
    smem[L::oScale + threadIdx.x] = expf(fminf(p.logit_scale[threadIdx.x], kLogitMax));
  for (int i = threadIdx.x; i < NH * 225; i += blockDim.x) {
    const int h = i / 225, e = i - h * 225;
    smem[L::oBias + h * L::kBiasLd + e] = p.bias_table[e * NH + h];
  }
}

}  // namespace b200mar
