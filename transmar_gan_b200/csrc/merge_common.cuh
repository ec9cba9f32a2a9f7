// merge_common.cuh — PatchMerging parameter block / smem layout.
#pragma once
#include "common.cuh"

namespace b200mar {

struct MergeParams {
  const float* x;       // [B, ph, pw, 64]
  const float* ln_w;    // [256]
  const float* ln_b;    // [256]
  const float* red_w;   // [64, 256]
  float* y;             // [B, ph/2, pw/2, 64]
  // backward only
  const float* dy;
  float* dx;
  float* d_ln_w; float* d_ln_b; float* d_red_w;
  int B, ph, pw;
};

struct MergeSmem {
  static constexpr int oW = 0;                          // [64][260]
  static constexpr int oLnW = oW + kDim * kLd256;
  static constexpr int oLnB = oLnW + 256;
  static constexpr int oX = oLnB + 256;                 // [64][260]
  static constexpr int kFloatsFwd = oX + kTok * kLd256;
  // backward extras
  static constexpr int oDy = kFloatsFwd;                // [64][68]
  static constexpr int oStat = oDy + kTok * kLd64;      // rstd[64], colsum(dy)[64]
  static constexpr int oDxn = oStat + 2 * kTok;         // [64][260]
  static constexpr int kFloatsBwd = oDxn + kTok * kLd256;
  static_assert(kFloatsBwd * 4 <= 232448, "patch_merge_bwd shared memory exceeds 227 KB");
};

// flat source token of output token o and quadrant q (q = 2*row_parity + col_parity):
// channel block q holds x[2r + (q>>1), 2c + (q&1)]  (encoder_decoder.py:63-67)
__device__ __forceinline__ long long merge_src(const MergeParams& p, long long o, int q) {
  const int ow = p.pw / 2, oh = p.ph / 2;
  const int c2 = (int)(o % ow);
  const int r2 = (int)((o / ow) % oh);
  const long long b = o / ((long long)ow * oh);
  return (b * p.ph + 2 * r2 + (q >> 1)) * p.pw + 2 * c2 + (q & 1);
}

}  // namespace b200mar
