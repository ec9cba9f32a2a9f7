// ngram_fwd.cu — N-gram context, forward (fp32).
//   unigram_fwd : grouped 8x8/stride-8 conv (components/window_utils.py:36-38,78) as one
//                 streaming pass over the token tensor, one warp per window.
//   ngram_ctx_fwd: seq_refl_win_pad fwd/bwd (:43-61) + 2x2 sliding WindowAttention(dim 32)
//                 (:63-73, attention.py:70-101) + AvgPool2d(2) + concat + 1x1 merge (:86-90),
//                 one warp per window, reflect padding folded into the index math.
#include "common.cuh"
#include "ngram_common.cuh"

namespace b200mar {

// ------------------------------------------------------------------------------- unigram
__global__ void __launch_bounds__(kThreads) unigram_fwd_kernel(
    const float* __restrict__ z, const float* __restrict__ w, const float* __restrict__ bias,
    float* __restrict__ u, int B, int ph, int pw) {
  __shared__ __align__(16) float sW[kTok * kDim];  // [t = 8i+j][c = 2g+k]
  for (int idx = threadIdx.x; idx < kTok * kDim; idx += blockDim.x) {
    const int t = idx >> 6, c = idx & 63;
    sW[idx] = w[(c >> 1) * 128 + (c & 1) * 64 + t];
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int wh = ph / kWin, ww = pw / kWin;
  const int nwin = B * wh * ww;
  const WinDiv wdiv = make_windiv(wh, ww);
  const float bl = bias[lane];
  for (int win = blockIdx.x * 8 + warp; win < nwin; win += gridDim.x * 8) {
    int w0, h0, b;
    win_pos(wdiv, win, w0, h0, b);
    const float* base = z + (((size_t)b * ph + h0 * kWin) * pw + w0 * kWin) * kDim + 2 * lane;
    float acc = 0.f;
#pragma unroll 1
    for (int i = 0; i < kWin; ++i) {
      float2 v[kWin];
#pragma unroll
      for (int j = 0; j < kWin; ++j)
        v[j] = *reinterpret_cast<const float2*>(base + ((size_t)i * pw + j) * kDim);
#pragma unroll
      for (int j = 0; j < kWin; ++j) {
        const float2 wv = *reinterpret_cast<const float2*>(sW + (i * kWin + j) * kDim + 2 * lane);
        acc = fmaf(v[j].x, wv.x, acc);
        acc = fmaf(v[j].y, wv.y, acc);
      }
    }
    u[(size_t)win * 32 + lane] = acc + bl;
  }
}

// Inference form: the same streaming pass also applies the n-gram attention's qkv Linear(32 -> 3*NA)
// to the window's unigram embedding, once per window (ngram_ctx_fwd re-derives it for each of the
// up to 8 neighbourhoods a window belongs to).  nq: [B*wh*ww][96] (rows >= 3*NA are zero).
// Four windows per warp and pass: the kernel was bound by shared-memory instruction issue (ncu: 31 % of the stalls
// MIO throttle, DRAM at 57 %), so each conv weight and each qkv weight read from shared memory now serves four
// windows, and four windows' rows (8 KB per warp) are in flight at a time.
template <bool S16>
__global__ void __launch_bounds__(kThreads, 2) unigram_qkv_fwd_kernel(
    const float* __restrict__ z, const float* __restrict__ w, const float* __restrict__ bias,
    const float* __restrict__ qkv_w, const float* __restrict__ qkv_b, int n_qkv,
    float* __restrict__ nq, float* __restrict__ u_out, int B, int ph, int pw) {
  __shared__ __align__(16) float sW[kTok * kDim];  // [t = 8i+j][c = 2g+k]
  __shared__ __align__(16) float sWq[32 * 96];     // [c][j]
  __shared__ __align__(16) float sU[8][32 * 4];    // per warp: [channel][window of the group]
  for (int idx = threadIdx.x; idx < kTok * kDim; idx += blockDim.x) {
    const int t = idx >> 6, c = idx & 63;
    sW[idx] = w[(c >> 1) * 128 + (c & 1) * 64 + t];
  }
  for (int idx = threadIdx.x; idx < 32 * 96; idx += blockDim.x) {
    const int c = idx / 96, j = idx - c * 96;
    sWq[idx] = j < n_qkv ? qkv_w[j * 32 + c] : 0.f;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int wh = ph / kWin, ww = pw / kWin;
  const int nwin = B * wh * ww;
  const WinDiv wdiv = make_windiv(wh, ww);
  const float bl = bias[lane];
  float bq[3];
#pragma unroll
  for (int g = 0; g < 3; ++g) bq[g] = lane + 32 * g < n_qkv ? qkv_b[lane + 32 * g] : 0.f;
  const int ngroup = (nwin + 3) >> 2;
  for (int grp = blockIdx.x * 8 + warp; grp < ngroup; grp += gridDim.x * 8) {
    size_t e0[4];                                     // element offset of (window k, row 0, token 0, channel 2*lane)
    bool ok[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int win = grp * 4 + k;
      ok[k] = win < nwin;
      int w0, h0, b;
      win_pos(wdiv, ok[k] ? win : 0, w0, h0, b);
      e0[k] = (((size_t)b * ph + h0 * kWin) * pw + w0 * kWin) * kDim + 2 * lane;
    }
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
    for (int i = 0; i < kWin; ++i) {
      float2 v[4][kWin];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
#pragma unroll
        for (int j = 0; j < kWin; ++j) {
          const size_t e = e0[k] + ((size_t)i * pw + j) * kDim;
          if (!ok[k]) {
            v[k][j] = make_float2(0.f, 0.f);
          } else if constexpr (S16) {
            const uint32_t u2 = *reinterpret_cast<const uint32_t*>(reinterpret_cast<const unsigned short*>(z) + e);
            v[k][j] = make_float2(__uint_as_float(u2 << 16), __uint_as_float(u2 & 0xFFFF0000u));
          } else {
            v[k][j] = *reinterpret_cast<const float2*>(z + e);
          }
        }
      }
#pragma unroll
      for (int j = 0; j < kWin; ++j) {
        const float2 wv = *reinterpret_cast<const float2*>(sW + (i * kWin + j) * kDim + 2 * lane);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          acc[k] = fmaf(v[k][j].x, wv.x, acc[k]);
          acc[k] = fmaf(v[k][j].y, wv.y, acc[k]);
        }
      }
    }
    *reinterpret_cast<float4*>(&sU[warp][lane * 4]) = make_float4(acc[0] + bl, acc[1] + bl, acc[2] + bl, acc[3] + bl);
    if (u_out) {                                      // training: the unigram embedding itself, saved for the backward
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (ok[k]) u_out[(size_t)(grp * 4 + k) * 32 + lane] = acc[k] + bl;
    }
    __syncwarp();
    float q[3][4];
#pragma unroll
    for (int g = 0; g < 3; ++g)
#pragma unroll
      for (int k = 0; k < 4; ++k) q[g][k] = bq[g];
#pragma unroll 8
    for (int c = 0; c < 32; ++c) {
      const float4 u4 = *reinterpret_cast<const float4*>(&sU[warp][c * 4]);
      const float uu[4] = {u4.x, u4.y, u4.z, u4.w};
      const float* wr = sWq + c * 96 + lane;
      const float w0 = wr[0], w1 = wr[32], w2 = wr[64];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        q[0][k] = fmaf(w0, uu[k], q[0][k]);
        q[1][k] = fmaf(w1, uu[k], q[1][k]);
        q[2][k] = fmaf(w2, uu[k], q[2][k]);
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (ok[k]) {
        float* o = nq + (size_t)(grp * 4 + k) * 96 + lane;
        o[0] = q[0][k]; o[32] = q[1][k]; o[64] = q[2][k];
      }
    }
    __syncwarp();
  }
}

// ------------------------------------------------------------------------------- context
// PRE = true: p.u holds the per-window qkv rows [nwin][96] of unigram_qkv_fwd_kernel instead of the
// unigram embeddings; the neighbourhood gather fetches them directly.
template <int NH, bool PRE>
__global__ void __launch_bounds__(kThreads) ngram_ctx_fwd_kernel(NgramParams p) {
  using L = NgramSmem<NH>;
  constexpr int HD = L::HD, NA = L::NA;
  extern __shared__ __align__(16) float smem[];
  load_ngram_weights<NH>(smem, p);
  __syncthreads();
  const float* sWqkv = smem + L::oWqkv;
  const float* sBqkv = smem + L::oBqkv;
  const float* sWp = smem + L::oWp;
  const float* sBp = smem + L::oBp;
  const float* sWm = smem + L::oWm;
  const float* sBm = smem + L::oBm;
  const float* sScale = smem + L::oScale;
  const float* sBias = smem + L::oBias;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // per-warp scratch.  The proj / merge matrix-vector products are done for FOUR windows at a time so that one pass
  // over the weights in shared memory serves four outputs (the kernel is bound by shared-memory instruction issue):
  // sOm4 / sCat4 hold the four windows' vectors interleaved, [channel][window], and are read as float4.
  float* sTok = smem + L::oScratch + warp * L::kScratchPerWarp;  // [8][32] (non-PRE, inside a window); then sCat4
  float* sCat4 = sTok;                                           // [64][4]
  float* sQkv = sTok + 8 * 32;                                   // [8][96]
  float* sOm4 = sQkv + 8 * 96;                                   // [64][4]  mean attention out per dir

  const int wh = p.wh, ww = p.ww;
  const int nwin = p.B * wh * ww;
  const WinDiv wdiv = make_windiv(wh, ww);
  // PRE: the 8 neighbourhood rows of the NEXT window are fetched into registers while this window
  // computes (24 loads in flight per lane hide the L2 latency of the gather)
  float nxt[24];
  auto fetch = [&](int win_) {
    int w0, h0, b;
    win_pos(wdiv, win_, w0, h0, b);
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      int r, c;
      ngram_neighbour(t, h0, w0, wh, ww, r, c);
      const float* src = p.u + ((size_t)(b * wh + r) * ww + c) * 96 + lane;
      nxt[3 * t] = __ldg(src); nxt[3 * t + 1] = __ldg(src + 32); nxt[3 * t + 2] = __ldg(src + 64);
    }
  };
  const int ngroup = (nwin + 3) >> 2, gstride = gridDim.x * 8;
  if (PRE && (int)(blockIdx.x * 8 + warp) < ngroup) fetch((blockIdx.x * 8 + warp) * 4);
  for (int grp = blockIdx.x * 8 + warp; grp < ngroup; grp += gstride) {
#pragma unroll 1
    for (int k = 0; k < 4; ++k) {
      const int win = grp * 4 + k;
      if (win >= nwin) {                                         // ragged last group: zero vectors
        sOm4[lane * 4 + k] = 0.f; sOm4[(lane + 32) * 4 + k] = 0.f;
        continue;
      }
      int w0, h0, b;
      win_pos(wdiv, win, w0, h0, b);
      if (PRE) {
        // ---- qkv rows of the 8 neighbourhood tokens (4 forward, 4 backward), precomputed per window
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          sQkv[t * 96 + lane] = nxt[3 * t];
          sQkv[t * 96 + lane + 32] = nxt[3 * t + 1];
          sQkv[t * 96 + lane + 64] = nxt[3 * t + 2];
        }
        __syncwarp();
        const int nw = k < 3 ? win + 1 : (grp + gstride) * 4;
        if (nw < nwin) fetch(nw);
      } else {
        // ---- the 8 neighbourhood tokens (4 forward, 4 backward), lane = channel
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          int r, c;
          ngram_neighbour(t, h0, w0, wh, ww, r, c);
          sTok[t * 32 + lane] = p.u[((size_t)(b * wh + r) * ww + c) * 32 + lane];
        }
        __syncwarp();
        // ---- qkv = Linear(32 -> 3*NA)
#pragma unroll
        for (int jg = 0; jg < 3; ++jg) {
          const int j = lane + 32 * jg;
          if (j < 3 * NA) {
            float acc[8];
            const float bj = sBqkv[j];
#pragma unroll
            for (int t = 0; t < 8; ++t) acc[t] = bj;
#pragma unroll 8
            for (int c = 0; c < 32; ++c) {
              const float wv = sWqkv[j * 33 + c];
#pragma unroll
              for (int t = 0; t < 8; ++t) acc[t] = fmaf(wv, sTok[t * 32 + c], acc[t]);
            }
#pragma unroll
            for (int t = 0; t < 8; ++t) sQkv[t * 96 + j] = acc[t];
          }
        }
        __syncwarp();
      }
      // ---- 4-token cosine attention per (direction, head); mean over the 4 query tokens.  A (direction, head)
      //      pair is shared by QG adjacent lanes, each owning 4 / QG of the query tokens (2 * NH * QG = 24 or 32
      //      busy lanes instead of 2 * NH); the partial means are combined with shuffles.
      {
        constexpr int QG = NH == 6 ? 2 : 4, QPL = 4 / QG;    // query groups per (dir, head), queries per lane
        const bool busy = lane < 2 * NH * QG;
        const int pair = busy ? lane / QG : 0, qg = lane % QG;
        const int dir = pair / NH, h = pair % NH;
        float om[HD];
#pragma unroll
        for (int d = 0; d < HD; ++d) om[d] = 0.f;
        if (busy) {
          float kn[4][HD], vv[4][HD];
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const float* row = sQkv + (dir * 4 + t) * 96;
            float sk = 0.f;
#pragma unroll
            for (int d = 0; d < HD; ++d) {
              kn[t][d] = row[NA + h * HD + d];
              vv[t][d] = row[2 * NA + h * HD + d];
              sk = fmaf(kn[t][d], kn[t][d], sk);
            }
            // 1 / max(||.||, eps): eps^2 = 1e-24 is a normal float, so max(sq, eps^2) under rsqrt is exact
            const float ik = PRE ? rsqrtf(fmaxf(sk, kNormEps * kNormEps)) : 1.0f / fmaxf(sqrtf(sk), kNormEps);
#pragma unroll
            for (int d = 0; d < HD; ++d) kn[t][d] *= ik;
          }
          const float scale = sScale[h];
#pragma unroll
          for (int qi = 0; qi < QPL; ++qi) {
            const int pq = qg * QPL + qi;
            const float* row = sQkv + (dir * 4 + pq) * 96;
            float qn[HD], sq = 0.f;
#pragma unroll
            for (int d = 0; d < HD; ++d) { qn[d] = row[h * HD + d]; sq = fmaf(qn[d], qn[d], sq); }
            const float iq = (PRE ? rsqrtf(fmaxf(sq, kNormEps * kNormEps)) : 1.0f / fmaxf(sqrtf(sq), kNormEps)) * scale;
            float s[4], mx = -INFINITY;
#pragma unroll
            for (int kq = 0; kq < 4; ++kq) {
              float dot = 0.f;
#pragma unroll
              for (int d = 0; d < HD; ++d) dot = fmaf(qn[d], kn[kq][d], dot);
              s[kq] = fmaf(dot, iq, sBias[h * 12 + ngram_rel_index(pq, kq)]);
              mx = fmaxf(mx, s[kq]);
            }
            float sum = 0.f;
#pragma unroll
            for (int kq = 0; kq < 4; ++kq) { s[kq] = PRE ? __expf(s[kq] - mx) : expf(s[kq] - mx); sum += s[kq]; }
            const float inv = PRE ? __fdividef(0.25f, sum) : 0.25f / sum;   // softmax normalisation and the AvgPool2d(2) mean
#pragma unroll
            for (int kq = 0; kq < 4; ++kq) {
              const float pv = s[kq] * inv;
#pragma unroll
              for (int d = 0; d < HD; ++d) om[d] = fmaf(pv, vv[kq][d], om[d]);
            }
          }
        }
#pragma unroll
        for (int d = 0; d < HD; ++d) {
#pragma unroll
          for (int o = 1; o < QG; o <<= 1) om[d] += __shfl_xor_sync(0xffffffffu, om[d], o);
        }
        if (busy && qg == 0) {
#pragma unroll
          for (int d = 0; d < HD; ++d) sOm4[(dir * 32 + h * HD + d) * 4 + k] = om[d];
        }
      }
      __syncwarp();
    }
    // ---- proj (commutes with the mean), four windows: cat[k][dir*32 + c] = bp[c] + Wp[c,:] . om[k][dir]
    {
      float acc[2][4];
#pragma unroll
      for (int dir = 0; dir < 2; ++dir)
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[dir][k] = sBp[lane];
#pragma unroll 6
      for (int j = 0; j < NA; ++j) {
        const float w = sWp[lane * 33 + j];
#pragma unroll
        for (int dir = 0; dir < 2; ++dir) {
          const float4 o4 = *reinterpret_cast<const float4*>(sOm4 + (dir * 32 + j) * 4);
          acc[dir][0] = fmaf(w, o4.x, acc[dir][0]); acc[dir][1] = fmaf(w, o4.y, acc[dir][1]);
          acc[dir][2] = fmaf(w, o4.z, acc[dir][2]); acc[dir][3] = fmaf(w, o4.w, acc[dir][3]);
        }
      }
      __syncwarp();                                              // (non-PRE: sTok is dead, sCat4 may overwrite it)
#pragma unroll
      for (int dir = 0; dir < 2; ++dir)
        *reinterpret_cast<float4*>(sCat4 + (dir * 32 + lane) * 4) =
            make_float4(acc[dir][0], acc[dir][1], acc[dir][2], acc[dir][3]);
    }
    __syncwarp();
    // ---- merge 1x1 conv 64 -> 64, four windows (two outputs per lane, two partial sums each: short FMA chains)
    {
      const float* w0 = sWm + lane * 65;
      const float* w1 = sWm + (lane + 32) * 65;
      float a0[4], a1[4], b0[4], b1[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) { a0[k] = sBm[lane]; a1[k] = 0.f; b0[k] = sBm[lane + 32]; b1[k] = 0.f; }
#pragma unroll 4
      for (int c = 0; c < 64; c += 2) {
        const float4 c0 = *reinterpret_cast<const float4*>(sCat4 + c * 4);
        const float4 c1 = *reinterpret_cast<const float4*>(sCat4 + c * 4 + 4);
        const float u0 = w0[c], u1 = w0[c + 1], v0 = w1[c], v1 = w1[c + 1];
        a0[0] = fmaf(u0, c0.x, a0[0]); a0[1] = fmaf(u0, c0.y, a0[1]); a0[2] = fmaf(u0, c0.z, a0[2]); a0[3] = fmaf(u0, c0.w, a0[3]);
        a1[0] = fmaf(u1, c1.x, a1[0]); a1[1] = fmaf(u1, c1.y, a1[1]); a1[2] = fmaf(u1, c1.z, a1[2]); a1[3] = fmaf(u1, c1.w, a1[3]);
        b0[0] = fmaf(v0, c0.x, b0[0]); b0[1] = fmaf(v0, c0.y, b0[1]); b0[2] = fmaf(v0, c0.z, b0[2]); b0[3] = fmaf(v0, c0.w, b0[3]);
        b1[0] = fmaf(v1, c1.x, b1[0]); b1[1] = fmaf(v1, c1.y, b1[1]); b1[2] = fmaf(v1, c1.z, b1[2]); b1[3] = fmaf(v1, c1.w, b1[3]);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int win = grp * 4 + k;
        if (win < nwin) {
          p.ctx[(size_t)win * 64 + lane] = a0[k] + a1[k];
          p.ctx[(size_t)win * 64 + lane + 32] = b0[k] + b1[k];
        }
      }
    }
    __syncwarp();
  }
}

template <int NH, bool PRE>
static int launch_ngram_ctx_fwd(const NgramParams& p, cudaStream_t st) {
  using L = NgramSmem<NH>;
  const size_t smem = L::kFloatsFwd * sizeof(float);
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(ngram_ctx_fwd_kernel<NH, PRE>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured.done();
  }
  const int nwin = p.B * p.wh * p.ww;
  int grid = ((nwin + 3) / 4 + 7) / 8;               // a warp takes groups of four windows
  if (grid > 3 * sm_count()) grid = 3 * sm_count();
  ngram_ctx_fwd_kernel<NH, PRE><<<grid, kThreads, smem, st>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace b200mar

extern "C" int b200mar_unigram_fwd(const float* z, const float* weight, const float* bias, float* u,
                                   int B, int ph, int pw, void* stream) {
  using namespace b200mar;
  if (ph % kWin || pw % kWin || ph < kWin || pw < kWin) return -1;
  const int nwin = B * (ph / kWin) * (pw / kWin);
  int grid = (nwin + 7) / 8;
  if (grid > 8 * sm_count()) grid = 8 * sm_count();
  unigram_fwd_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(z, weight, bias, u, B, ph, pw);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_ngram_ctx_fwd(const float* u, const float* qkv_w, const float* qkv_b,
                                     const float* proj_w, const float* proj_b,
                                     const float* logit_scale, const float* bias_table,
                                     const float* merge_w, const float* merge_b, float* ctx, int B,
                                     int wh, int ww, int nh, void* stream) {
  using namespace b200mar;
  if (wh < 2 || ww < 2) return -1;  // the reference raises for a single window row/column
  NgramParams p{};
  p.u = u; p.qkv_w = qkv_w; p.qkv_b = qkv_b; p.proj_w = proj_w; p.proj_b = proj_b;
  p.logit_scale = logit_scale; p.bias_table = bias_table; p.merge_w = merge_w; p.merge_b = merge_b;
  p.ctx = ctx; p.B = B; p.wh = wh; p.ww = ww;
  if (nh == 6) return launch_ngram_ctx_fwd<6, false>(p, (cudaStream_t)stream);
  if (nh == 4) return launch_ngram_ctx_fwd<4, false>(p, (cudaStream_t)stream);
  return -1;
}

extern "C" int b200mar_unigram_qkv_fwd(const float* z, const float* weight, const float* bias,
                                       const float* qkv_w, const float* qkv_b, float* nq, int B,
                                       int ph, int pw, int nh, void* stream) {
  using namespace b200mar;
  if (ph % kWin || pw % kWin || ph < kWin || pw < kWin || (nh != 6 && nh != 4)) return -1;
  const int nwin = B * (ph / kWin) * (pw / kWin);
  int grid = ((nwin + 3) / 4 + 7) / 8;               // a warp takes groups of four windows
  if (grid > 2 * sm_count()) grid = 2 * sm_count();
  unigram_qkv_fwd_kernel<false><<<grid, kThreads, 0, (cudaStream_t)stream>>>(
      z, weight, bias, qkv_w, qkv_b, 3 * (32 / nh) * nh, nq, nullptr, B, ph, pw);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

// Training form: the same pass also writes the unigram embedding u [B*wh*ww][32] that the backward kernels consume.
extern "C" int b200mar_unigram_qkv_u_fwd(const float* z, const float* weight, const float* bias,
                                         const float* qkv_w, const float* qkv_b, float* nq, float* u, int B,
                                         int ph, int pw, int nh, void* stream) {
  using namespace b200mar;
  if (ph % kWin || pw % kWin || ph < kWin || pw < kWin || (nh != 6 && nh != 4) || !u) return -1;
  const int nwin = B * (ph / kWin) * (pw / kWin);
  int grid = ((nwin + 3) / 4 + 7) / 8;
  if (grid > 2 * sm_count()) grid = 2 * sm_count();
  unigram_qkv_fwd_kernel<false><<<grid, kThreads, 0, (cudaStream_t)stream>>>(
      z, weight, bias, qkv_w, qkv_b, 3 * (32 / nh) * nh, nq, u, B, ph, pw);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

// Precision mode "bf16": the same pass over a BF16 stream z16 [B, ph*pw, 64]; nq stays fp32.
extern "C" int b200mar_unigram_qkv_fwd_s16(const void* z16, const float* weight, const float* bias,
                                           const float* qkv_w, const float* qkv_b, float* nq, int B,
                                           int ph, int pw, int nh, void* stream) {
  using namespace b200mar;
  if (ph % kWin || pw % kWin || ph < kWin || pw < kWin || (nh != 6 && nh != 4)) return -1;
  const int nwin = B * (ph / kWin) * (pw / kWin);
  int grid = ((nwin + 3) / 4 + 7) / 8;
  if (grid > 2 * sm_count()) grid = 2 * sm_count();
  unigram_qkv_fwd_kernel<true><<<grid, kThreads, 0, (cudaStream_t)stream>>>(
      (const float*)z16, weight, bias, qkv_w, qkv_b, 3 * (32 / nh) * nh, nq, nullptr, B, ph, pw);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_ngram_ctx_pre_fwd(const float* nq, const float* proj_w, const float* proj_b,
                                         const float* logit_scale, const float* bias_table,
                                         const float* merge_w, const float* merge_b, float* ctx,
                                         int B, int wh, int ww, int nh, void* stream) {
  using namespace b200mar;
  if (wh < 2 || ww < 2) return -1;
  NgramParams p{};
  p.u = nq; p.proj_w = proj_w; p.proj_b = proj_b;
  p.logit_scale = logit_scale; p.bias_table = bias_table; p.merge_w = merge_w; p.merge_b = merge_b;
  p.ctx = ctx; p.B = B; p.wh = wh; p.ww = ww;
  if (nh == 6) return launch_ngram_ctx_fwd<6, true>(p, (cudaStream_t)stream);
  if (nh == 4) return launch_ngram_ctx_fwd<4, true>(p, (cudaStream_t)stream);
  return -1;
}

// ------------------------------------------------------------------------------- no-allocation variant
// b200mar_ngram_ctx_infer_noalloc = b200mar_unigram_qkv_fwd + b200mar_ngram_ctx_pre_fwd with the scratch between
// the two kernels (nq, [windows][96] fp32) held in a file-scope device array instead of a caller-allocated
// buffer, for callers that cannot allocate (CUDA-graph capture, tight serving loops).  No counterpart in the
// reference.
//   * limit: kNoAllocMaxWindows windows per call (6 MiB; one 1024x1024 slice or six 416x416 slices at stage 1);
//     every GPU that loads the library holds the array until the process exits;
//   * the array is shared by every call and keeps the last call's contents.  Nothing has to start at zero: the
//     first kernel writes all 96 columns of every window row the second kernel reads.  Calls that overlap on
//     the GPU (different streams) RACE on it - serialise them;
//   * host code sees `g_nq_scratch` only as a stand-in symbol: its device address is looked up once per device,
//     with that device current, through cudaGetSymbolAddress.
namespace b200mar {
constexpr int kNoAllocMaxWindows = 16384;
static __device__ __align__(256) float g_nq_scratch[(size_t)kNoAllocMaxWindows * 96];

static int nq_scratch_ptr(float** out) {
  static float* cache[64] = {nullptr};
  int dev = -1;
  B200MAR_CHECK(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && cache[dev]) { *out = cache[dev]; return 0; }
  float* ptr = nullptr;
  B200MAR_CHECK(cudaGetSymbolAddress((void**)&ptr, g_nq_scratch));
  if (!ptr) { last_error_ref() = "cudaGetSymbolAddress(g_nq_scratch) returned null"; return -4; }
  if (dev >= 0 && dev < 64) cache[dev] = ptr;
  *out = ptr;
  return 0;
}
}  // namespace b200mar

extern "C" int b200mar_ngram_ctx_infer_noalloc_max_windows(void) { return b200mar::kNoAllocMaxWindows; }

extern "C" int b200mar_ngram_ctx_infer_noalloc(const float* z, const float* uni_w, const float* uni_b,
                                               const float* qkv_w, const float* qkv_b,
                                               const float* proj_w, const float* proj_b,
                                               const float* logit_scale, const float* bias_table,
                                               const float* merge_w, const float* merge_b, float* ctx,
                                               int B, int ph, int pw, int nh, void* stream) {
  using namespace b200mar;
  if (B <= 0 || ph % kWin || pw % kWin || ph < 2 * kWin || pw < 2 * kWin || (nh != 6 && nh != 4)) return -1;
  const long long nwin = (long long)B * (ph / kWin) * (pw / kWin);
  if (nwin > kNoAllocMaxWindows) {
    last_error_ref() = "b200mar_ngram_ctx_infer_noalloc: B*(ph/8)*(pw/8) exceeds the 16384-window static scratch "
                       "(use b200mar_unigram_qkv_fwd + b200mar_ngram_ctx_pre_fwd with a caller-owned buffer)";
    return -5;
  }
  float* nq = nullptr;
  if (int rc = nq_scratch_ptr(&nq)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  int grid = (((int)nwin + 3) / 4 + 7) / 8;          // a warp takes groups of four windows
  if (grid > 2 * sm_count()) grid = 2 * sm_count();
  unigram_qkv_fwd_kernel<false><<<grid, kThreads, 0, st>>>(z, uni_w, uni_b, qkv_w, qkv_b, 3 * (32 / nh) * nh, nq, nullptr, B, ph, pw);
  B200MAR_CHECK(cudaGetLastError());
  NgramParams p{};
  p.u = nq; p.proj_w = proj_w; p.proj_b = proj_b;
  p.logit_scale = logit_scale; p.bias_table = bias_table; p.merge_w = merge_w; p.merge_b = merge_b;
  p.ctx = ctx; p.B = B; p.wh = ph / kWin; p.ww = pw / kWin;
  if (nh == 6) return launch_ngram_ctx_fwd<6, true>(p, st);
  return launch_ngram_ctx_fwd<4, true>(p, st);
}
