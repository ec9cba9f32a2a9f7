// ngram_common.cuh — parameter block / smem layout of the N-gram context kernels.
#pragma once
#include "common.cuh"

namespace b200mar {

struct NgramParams {
  const float* u;            // [B, wh, ww, 32] unigram embeddings (channels last)
  const float* qkv_w;        // [3*NA, 32]
  const float* qkv_b;        // [3*NA]
  const float* proj_w;       // [32, NA]
  const float* proj_b;       // [32]
  const float* logit_scale;  // [nh]
  const float* bias_table;   // [9, nh]
  const float* merge_w;      // [64, 64]
  const float* merge_b;      // [64]
  float* ctx;                // [B, wh, ww, 64]
  // backward only
  const float* dctx;
  float* du;
  float* d_qkv_w; float* d_qkv_b; float* d_proj_w; float* d_proj_b;
  float* d_logit_scale; float* d_bias_table; float* d_merge_w; float* d_merge_b;
  int B, wh, ww;
};

template <int NH>
struct NgramSmem {
  static constexpr int HD = 32 / NH;          // 5 (6 heads) or 8 (4 heads)
  static constexpr int NA = HD * NH;          // 30 or 32
  static constexpr int oWqkv = 0;             // [96][33]
  static constexpr int oBqkv = oWqkv + 96 * 33;
  static constexpr int oWp = oBqkv + 96;      // [32][33]
  static constexpr int oBp = oWp + 32 * 33;
  static constexpr int oWm = oBp + 32;        // [64][65]
  static constexpr int oBm = oWm + 64 * 65;
  static constexpr int oScale = oBm + 64;     // [8]
  static constexpr int oBias = oScale + 8;    // [NH][12]
  static constexpr int oScratch = oBias + 8 * 12;
  static constexpr int kScratchPerWarp = 8 * 32 + 8 * 96 + 4 * 64;   // tok / cat (4 windows) | qkv | om (4 windows)
  static constexpr int kFloatsFwd = oScratch + 8 * kScratchPerWarp;
};

// Neighbourhood token t (0..3 forward, 4..7 backward; within a direction t = 2*i + j,
// row-offset major, window_utils.py:65) of window (h0,w0) -> source window (r,c) of the
// unpadded unigram map.  Forward pad: row wh := row wh-2, column ww := column ww-2
// (window_utils.py:49-54); backward pad: row -1 := row 1, column -1 := column 1 (:55-60).
__device__ __forceinline__ void ngram_neighbour(int t, int h0, int w0, int wh, int ww, int& r,
                                                int& c) {
  const int i = (t >> 1) & 1, j = t & 1;
  if (t < 4) {
    r = h0 + i; if (r >= wh) r = wh - 2;
    c = w0 + j; if (c >= ww) c = ww - 2;
  } else {
    r = h0 + i - 1; if (r < 0) r = 1;
    c = w0 + j - 1; if (c < 0) c = 1;
  }
}
// relative position index inside the 2x2 n-gram window (attention.py:9-18 with win 2)
__device__ __forceinline__ int ngram_rel_index(int p, int q) {
  return ((p >> 1) - (q >> 1) + 1) * 3 + ((p & 1) - (q & 1) + 1);
}

template <int NH>
__device__ __forceinline__ void load_ngram_weights(float* smem, const NgramParams& p) {
  using L = NgramSmem<NH>;
  // All global loads of a thread are issued back to back (fixed trip counts for the 256-thread CTAs of these
  // kernels, 16- / 8-byte loads), then scattered into the padded shared-memory rows: one round of global latency
  // instead of ~30 dependent scalar rounds (the prologue was most of the 15 us fixed cost of a launch).
  const int tid = threadIdx.x;
  constexpr int NT = kThreads;
  constexpr int kQ = (3 * L::NA * 8 + NT - 1) / NT;            // qkv [3 NA][32]: float4 pieces
  constexpr int kP = (32 * L::NA / 2 + NT - 1) / NT;           // proj [32][NA]: float2 pieces (NA is even)
  constexpr int kM = 64 * 16 / NT;                             // merge [64][64]: float4 pieces
  static_assert(L::NA % 2 == 0 && (64 * 16) % NT == 0, "piece counts");
  float4 q[kQ], m[kM];
  float2 pr[kP];
  if (p.qkv_w) {                               // null when the qkv rows are precomputed per window
#pragma unroll
    for (int j = 0; j < kQ; ++j) {
      const int f = tid + j * NT;
      q[j] = f < 3 * L::NA * 8 ? __ldg(reinterpret_cast<const float4*>(p.qkv_w) + f) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
#pragma unroll
  for (int j = 0; j < kP; ++j) {
    const int f = tid + j * NT;
    pr[j] = f < 32 * L::NA / 2 ? __ldg(reinterpret_cast<const float2*>(p.proj_w) + f) : make_float2(0.f, 0.f);
  }
#pragma unroll
  for (int j = 0; j < kM; ++j) m[j] = __ldg(reinterpret_cast<const float4*>(p.merge_w) + tid + j * NT);
  if (p.qkv_w) {
#pragma unroll
    for (int j = 0; j < kQ; ++j) {
      const int f = tid + j * NT;                // rows >= 3 NA of the [96][33] tile are zero
      if (f < 96 * 8) {
        float* d = smem + L::oWqkv + (f >> 3) * 33 + ((f & 7) << 2);
        d[0] = q[j].x; d[1] = q[j].y; d[2] = q[j].z; d[3] = q[j].w;
      }
    }
    for (int i = tid; i < 96; i += NT) smem[L::oBqkv + i] = i < 3 * L::NA ? p.qkv_b[i] : 0.f;
  }
#pragma unroll
  for (int j = 0; j < kP; ++j) {
    const int f = tid + j * NT;
    if (f < 32 * L::NA / 2) {
      const int r = (2 * f) / L::NA, c = 2 * f - r * L::NA;
      smem[L::oWp + r * 33 + c] = pr[j].x;
      smem[L::oWp + r * 33 + c + 1] = pr[j].y;
    }
  }
  for (int i = tid; i < 32; i += NT) smem[L::oBp + i] = p.proj_b[i];
#pragma unroll
  for (int j = 0; j < kM; ++j) {
    const int f = tid + j * NT;
    float* d = smem + L::oWm + (f >> 4) * 65 + ((f & 15) << 2);
    d[0] = m[j].x; d[1] = m[j].y; d[2] = m[j].z; d[3] = m[j].w;
  }
  for (int i = tid; i < 64; i += NT) smem[L::oBm + i] = p.merge_b[i];
  if (tid < NH)
    smem[L::oScale + tid] = expf(fminf(p.logit_scale[tid], kLogitMax));
  for (int i = tid; i < NH * 9; i += NT) {
    const int h = i / 9, e = i - h * 9;
    smem[L::oBias + h * 12 + e] = p.bias_table[e * NH + h];
  }
}

}  // namespace b200mar
