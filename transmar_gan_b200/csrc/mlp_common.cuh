// mlp_common.cuh — parameter block and smem carve-up of the fused MLP+LN2+residual kernels.
#pragma once
#include "common.cuh"

namespace b200mar {

struct MlpParams {
  const float* a;       // [n_tok, 64]  post-attention residual stream
  const float* z;       // [n_tok, 64]  block input (only for zsum)
  const float* fc1_w;   // [128, 64]
  const float* fc1_b;   // [128]
  const float* fc2_w;   // [64, 128]
  const float* fc2_b;   // [64]
  const float* ln_w;    // [64]
  const float* ln_b;    // [64]
  float* out;           // a + LN2(mlp(a))           (may be null)
  float* zsum;          // out + z                    (may be null)
  float* a_out;         // mlp_tc with a_is_delta: the materialised a = z + delta (may be null)
  int a_is_delta;       // mlp_tc: `a` holds LN1(attn) only and the kernel forms a = z + a
  // backward only
  const float* dout;
  float* da;
  float* d_fc1_w; float* d_fc1_b; float* d_fc2_w; float* d_fc2_b; float* d_ln_w; float* d_ln_b;
  long long n_tok;
};

struct MlpSmem {
  static constexpr int oW1 = 0;                          // [128][68]
  static constexpr int oW2 = oW1 + kHid * kLd64;         // [64][132]
  static constexpr int oB1 = oW2 + kDim * kLd128;
  static constexpr int oB2 = oB1 + kHid;
  static constexpr int oLnW = oB2 + kDim;
  static constexpr int oLnB = oLnW + kDim;
  static constexpr int oA = oLnB + kDim;                 // [64][68]
  static constexpr int oH = oA + kTok * kLd64;           // [64][132]
  static constexpr int oY = oH + kTok * kLd128;          // [64][68]
  static constexpr int kFloatsFwd = oY + kTok * kLd64;
  // backward extras
  static constexpr int oPre = kFloatsFwd;                // [64][132] pre-activation / dpre
  static constexpr int oDy = oPre + kTok * kLd128;       // [64][68] dout tile / dm
  static constexpr int kFloatsBwd = oDy + kTok * kLd64;
};

__device__ __forceinline__ void load_mlp_weights(float* smem, const MlpParams& p) {
  using L = MlpSmem;
  load_matrix(smem + L::oW1, kLd64, p.fc1_w, kHid, kDim, kHid);
  load_matrix(smem + L::oW2, kLd128, p.fc2_w, kDim, kHid, kDim);
  for (int i = threadIdx.x; i < kHid; i += blockDim.x) smem[L::oB1 + i] = p.fc1_b[i];
  for (int i = threadIdx.x; i < kDim; i += blockDim.x) {
    smem[L::oB2 + i] = p.fc2_b[i];
    smem[L::oLnW + i] = p.ln_w[i];
    smem[L::oLnB + i] = p.ln_b[i];
  }
}

}  // namespace b200mar
