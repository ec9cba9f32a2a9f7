// ngram_bwd.cu — backward of the N-gram context path (fp32).
//   unigram_bwd  : gradient of the grouped 8x8/stride-8 conv (window_utils.py:36-38).
//   ngram_ctx_bwd: gradient of reflect-pad + 2x2 sliding cosine attention + avg-pool + merge
//                  (window_utils.py:43-90), one warp per window with full recompute.
#include "common.cuh"
#include "ngram_common.cuh"

namespace b200mar {

// ------------------------------------------------------------------------------- unigram
// Warp w of every CTA owns window row i = w for all windows the CTA visits, so its 16 weight
// gradients (8 columns x 2 channels per lane) live in registers for the whole kernel.
__global__ void __launch_bounds__(kThreads) unigram_bwd_kernel(
    const float* __restrict__ z, const float* __restrict__ w, const float* __restrict__ du,
    float* __restrict__ dz, const float* __restrict__ dz_add, float* __restrict__ dw, float* __restrict__ db,
    int B, int ph, int pw) {
  const int lane = threadIdx.x & 31, i = threadIdx.x >> 5;   // i = window row owned by this warp
  const int wh = ph / kWin, ww = pw / kWin;
  const int nwin = B * wh * ww;
  const WinDiv wdiv = make_windiv(wh, ww);
  float2 wr[kWin], acc[kWin];
#pragma unroll
  for (int j = 0; j < kWin; ++j) {
    wr[j].x = w[lane * 128 + i * 8 + j];
    wr[j].y = w[lane * 128 + 64 + i * 8 + j];
    acc[j] = make_float2(0.f, 0.f);
  }
  float dbl = 0.f;
  for (int win = blockIdx.x; win < nwin; win += gridDim.x) {
    int w0, h0, b;
    win_pos(wdiv, win, w0, h0, b);
    const float g = du[(size_t)win * 32 + lane];
    if (i == 0) dbl += g;
    const size_t base = (((size_t)b * ph + h0 * kWin + i) * pw + w0 * kWin) * kDim + 2 * lane;
    float2 zv[kWin], dv[kWin];
#pragma unroll
    for (int j = 0; j < kWin; ++j) {
      zv[j] = *reinterpret_cast<const float2*>(z + base + j * kDim);
      dv[j] = *reinterpret_cast<const float2*>(dz + base + j * kDim);
      if (dz_add) {                                   // gradient of the within-stage sum, folded in here
        const float2 e = *reinterpret_cast<const float2*>(dz_add + base + j * kDim);
        dv[j].x += e.x; dv[j].y += e.y;
      }
    }
#pragma unroll
    for (int j = 0; j < kWin; ++j) {
      dv[j].x = fmaf(wr[j].x, g, dv[j].x);
      dv[j].y = fmaf(wr[j].y, g, dv[j].y);
      *reinterpret_cast<float2*>(dz + base + j * kDim) = dv[j];
      acc[j].x = fmaf(g, zv[j].x, acc[j].x);
      acc[j].y = fmaf(g, zv[j].y, acc[j].y);
    }
  }
  // flush: the 16 gradients of a thread are two runs of 8 consecutive floats -> four 16-byte vector reductions
  // (red.global.add.v4.f32) instead of 16 scalar ones: the flush is bound by the number of atomic operations
  // at L2 (every CTA hits the same 4096 addresses), measured 10 us per 148 CTAs with scalar atomics
  if ((reinterpret_cast<uintptr_t>(dw) & 15) == 0) {   // dp.FlatGradBucket aligns its views to 16 bytes
    float* q = dw + lane * 128 + i * 8;
    atomicAdd(reinterpret_cast<float4*>(q), make_float4(acc[0].x, acc[1].x, acc[2].x, acc[3].x));
    atomicAdd(reinterpret_cast<float4*>(q + 4), make_float4(acc[4].x, acc[5].x, acc[6].x, acc[7].x));
    atomicAdd(reinterpret_cast<float4*>(q + 64), make_float4(acc[0].y, acc[1].y, acc[2].y, acc[3].y));
    atomicAdd(reinterpret_cast<float4*>(q + 68), make_float4(acc[4].y, acc[5].y, acc[6].y, acc[7].y));
  } else {
#pragma unroll
    for (int j = 0; j < kWin; ++j) {
      atomicAdd(dw + lane * 128 + i * 8 + j, acc[j].x);
      atomicAdd(dw + lane * 128 + 64 + i * 8 + j, acc[j].y);
    }
  }
  if (i == 0) atomicAdd(db + lane, dbl);
}

// ------------------------------------------------------------------------------- context
template <int NH>
struct NgramSmemBwd {
  using F = NgramSmem<NH>;
  static constexpr int oDWqkv = F::oScratch;          // [96][33]
  static constexpr int oDWp = oDWqkv + 96 * 33;       // [32][33]
  static constexpr int oDWm = oDWp + 32 * 33;         // [64][65]
  static constexpr int oDBias = oDWm + 64 * 65;       // [8][12]
  static constexpr int oDScale = oDBias + 96;         // [8]
  static constexpr int oScratch = oDScale + 8;
  // per warp: tok[8][32] qkv[8][96] om[64] cat[64] g[64] dcat[64] dom[64] dqkv[8][96]
  static constexpr int kScratchPerWarp = 256 + 768 + 64 * 5 + 768;
  static constexpr int kFloats = oScratch + 8 * kScratchPerWarp;
};

template <int NH>
__global__ void __launch_bounds__(kThreads) ngram_ctx_bwd_kernel(NgramParams p) {
  using L = NgramSmem<NH>;
  using LB = NgramSmemBwd<NH>;
  constexpr int HD = L::HD, NA = L::NA;
  extern __shared__ __align__(16) float smem[];
  load_ngram_weights<NH>(smem, p);
  // zero the gradient accumulators AND the per-warp scratch (pad rows / columns that are never written take
  // part in the block-wide outer products below)
  for (int i = threadIdx.x; i < LB::kFloats - LB::oDWqkv; i += blockDim.x) smem[LB::oDWqkv + i] = 0.f;
  __syncthreads();
  const float* sWqkv = smem + L::oWqkv;
  const float* sBqkv = smem + L::oBqkv;
  const float* sWp = smem + L::oWp;
  const float* sBp = smem + L::oBp;
  const float* sWm = smem + L::oWm;
  const float* sScale = smem + L::oScale;
  const float* sBias = smem + L::oBias;
  float* sDBias = smem + LB::oDBias;
  float* sDScale = smem + LB::oDScale;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* sTok = smem + LB::oScratch + warp * LB::kScratchPerWarp;
  float* sQkv = sTok + 256;
  float* sOm = sQkv + 768;
  float* sCat = sOm + 64;
  float* sG = sCat + 64;
  float* sDcat = sG + 64;
  float* sDom = sDcat + 64;
  float* sDqkv = sDom + 64;

  float dbm[2] = {0.f, 0.f}, dbp = 0.f, dbq[3] = {0.f, 0.f, 0.f};
  // Weight gradients are sums over windows of outer products of per-window vectors.  Every warp leaves its
  // window's vectors in its scratch; then the 256 threads of the CTA each own 32 gradient elements in REGISTERS
  // and add the 8 windows of the pass (no shared-memory atomics: those were ~8000 per window and 85 % of the
  // kernel):  merge dW[o][c] += g[o] cat[c]           o = tid/4,  c = 16 (tid%4) .. +16
  //           proj  dW[c][j] += dcat[d][c] om[d][j]   c = tid/8,  j =  4 (tid%8) .. +4   (both directions d)
  //           qkv   dW[j][c] += dqkv[t][j] tok[t][c]  j = tid/8 + 32 i (i < 3),  c = 4 (tid%8) .. +4   (8 tokens t)
  float accM[16], accP[4], accQ[12];
  float dBias9[9], dScale1 = 0.f;                    // (dir, head) lanes: bias-table / logit-scale gradients
#pragma unroll
  for (int i = 0; i < 9; ++i) dBias9[i] = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) accM[i] = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) accP[i] = 0.f;
#pragma unroll
  for (int i = 0; i < 12; ++i) accQ[i] = 0.f;
  const float* sScr = smem + LB::oScratch;

  const int wh = p.wh, ww = p.ww;
  const int nwin = p.B * wh * ww;
  const WinDiv wdiv = make_windiv(wh, ww);
  for (int base = blockIdx.x * 8; base < nwin; base += gridDim.x * 8) {
   const int win = base + warp;
   if (win < nwin) {
    int w0, h0, b;
    win_pos(wdiv, win, w0, h0, b);
    int nbr[8];
    // ================= forward recompute
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      int r, c;
      ngram_neighbour(t, h0, w0, wh, ww, r, c);
      nbr[t] = ((b * wh + r) * ww + c) * 32;
      sTok[t * 32 + lane] = p.u[(size_t)nbr[t] + lane];
    }
    __syncwarp();
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
    // per-(dir, head) state kept in registers between the forward and backward halves
    float qn[4][HD], kn[4][HD], P[4][4], nq[4], nk[4];
    const int dir = lane / NH, h = lane % NH;
    const bool active = lane < 2 * NH;
    if (active) {
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float* row = sQkv + (dir * 4 + t) * 96;
        float sq = 0.f, sk = 0.f;
#pragma unroll
        for (int d = 0; d < HD; ++d) {
          qn[t][d] = row[h * HD + d];
          kn[t][d] = row[NA + h * HD + d];
          sq = fmaf(qn[t][d], qn[t][d], sq);
          sk = fmaf(kn[t][d], kn[t][d], sk);
        }
        nq[t] = sqrtf(sq); nk[t] = sqrtf(sk);
        const float iq = 1.0f / fmaxf(nq[t], kNormEps), ik = 1.0f / fmaxf(nk[t], kNormEps);
#pragma unroll
        for (int d = 0; d < HD; ++d) { qn[t][d] *= iq; kn[t][d] *= ik; }
      }
      float om[HD];
#pragma unroll
      for (int d = 0; d < HD; ++d) om[d] = 0.f;
      const float scale = sScale[h];
#pragma unroll
      for (int pq = 0; pq < 4; ++pq) {
        float mx = -INFINITY;
#pragma unroll
        for (int kq = 0; kq < 4; ++kq) {
          float dot = 0.f;
#pragma unroll
          for (int d = 0; d < HD; ++d) dot = fmaf(qn[pq][d], kn[kq][d], dot);
          P[pq][kq] = dot * scale + sBias[h * 12 + ngram_rel_index(pq, kq)];
          mx = fmaxf(mx, P[pq][kq]);
        }
        float sum = 0.f;
#pragma unroll
        for (int kq = 0; kq < 4; ++kq) { P[pq][kq] = expf(P[pq][kq] - mx); sum += P[pq][kq]; }
        const float inv = 1.0f / sum;
#pragma unroll
        for (int kq = 0; kq < 4; ++kq) {
          P[pq][kq] *= inv;
          const float* vrow = sQkv + (dir * 4 + kq) * 96 + 2 * NA + h * HD;
#pragma unroll
          for (int d = 0; d < HD; ++d) om[d] = fmaf(0.25f * P[pq][kq], vrow[d], om[d]);
        }
      }
#pragma unroll
      for (int d = 0; d < HD; ++d) sOm[dir * 32 + h * HD + d] = om[d];
    }
    __syncwarp();
#pragma unroll
    for (int dd = 0; dd < 2; ++dd) {
      float acc = sBp[lane];
#pragma unroll 6
      for (int j = 0; j < NA; ++j) acc = fmaf(sWp[lane * 33 + j], sOm[dd * 32 + j], acc);
      sCat[dd * 32 + lane] = acc;
    }
    // ================= backward
    // ---- B1: merge conv
#pragma unroll
    for (int og = 0; og < 2; ++og) {
      const float gv = p.dctx[(size_t)win * 64 + lane + 32 * og];
      sG[lane + 32 * og] = gv;
      dbm[og] += gv;
    }
    __syncwarp();
#pragma unroll
    for (int cg = 0; cg < 2; ++cg) {
      const int c = lane + 32 * cg;
      float acc = 0.f;
#pragma unroll 8
      for (int o = 0; o < 64; ++o) acc = fmaf(sWm[o * 65 + c], sG[o], acc);
      sDcat[c] = acc;
    }
    __syncwarp();
    // ---- B2: proj (applied to the token-mean, so its input gradient is dom)
    dbp += sDcat[lane] + sDcat[32 + lane];
    if (lane < NA) {
#pragma unroll
      for (int dd = 0; dd < 2; ++dd) {
        float acc = 0.f;
#pragma unroll 8
        for (int c = 0; c < 32; ++c) acc = fmaf(sWp[c * 33 + lane], sDcat[dd * 32 + c], acc);
        sDom[dd * 32 + lane] = acc;
      }
    }
    __syncwarp();
    // ---- B3: 4-token cosine attention backward per (dir, head)
    if (active) {
      const float scale = sScale[h];
      float dom[HD], dPq[4];
#pragma unroll
      for (int d = 0; d < HD; ++d) dom[d] = sDom[dir * 32 + h * HD + d];
#pragma unroll
      for (int kq = 0; kq < 4; ++kq) {
        const float* vrow = sQkv + (dir * 4 + kq) * 96 + 2 * NA + h * HD;
        float acc = 0.f, colsum = 0.f;
#pragma unroll
        for (int d = 0; d < HD; ++d) acc = fmaf(dom[d], vrow[d], acc);
        dPq[kq] = 0.25f * acc;                                   // dP[p][kq], identical for all p
#pragma unroll
        for (int pq = 0; pq < 4; ++pq) colsum += P[pq][kq];
#pragma unroll
        for (int d = 0; d < HD; ++d)                              // dv[kq][d]
          sDqkv[(dir * 4 + kq) * 96 + 2 * NA + h * HD + d] = 0.25f * colsum * dom[d];
      }
      float dS[4][4], dsc = 0.f;
#pragma unroll
      for (int pq = 0; pq < 4; ++pq) {
        float rowdot = 0.f;
#pragma unroll
        for (int kq = 0; kq < 4; ++kq) rowdot = fmaf(dPq[kq], P[pq][kq], rowdot);
#pragma unroll
        for (int kq = 0; kq < 4; ++kq) {
          dS[pq][kq] = P[pq][kq] * (dPq[kq] - rowdot);
          dBias9[ngram_rel_index(pq, kq)] += dS[pq][kq];        // static index after unrolling: registers
          float cs = 0.f;
#pragma unroll
          for (int d = 0; d < HD; ++d) cs = fmaf(qn[pq][d], kn[kq][d], cs);
          dsc = fmaf(dS[pq][kq], cs, dsc);
        }
      }
      dScale1 += dsc;
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        float dqn[HD], dkn[HD], dotq = 0.f, dotk = 0.f;
#pragma unroll
        for (int d = 0; d < HD; ++d) {
          float aq = 0.f, ak = 0.f;
#pragma unroll
          for (int o = 0; o < 4; ++o) {
            aq = fmaf(dS[t][o], kn[o][d], aq);    // dqn[t] = s * sum_q dS[t][q] kn[q]
            ak = fmaf(dS[o][t], qn[o][d], ak);    // dkn[t] = s * sum_p dS[p][t] qn[p]
          }
          dqn[d] = aq * scale; dkn[d] = ak * scale;
          dotq = fmaf(qn[t][d], dqn[d], dotq);
          dotk = fmaf(kn[t][d], dkn[d], dotk);
        }
        const float iq = 1.0f / fmaxf(nq[t], kNormEps), ik = 1.0f / fmaxf(nk[t], kNormEps);
        const float pq_ = nq[t] >= kNormEps ? 1.f : 0.f, pk_ = nk[t] >= kNormEps ? 1.f : 0.f;
        float* drow = sDqkv + (dir * 4 + t) * 96;
#pragma unroll
        for (int d = 0; d < HD; ++d) {
          drow[h * HD + d] = (dqn[d] - pq_ * qn[t][d] * dotq) * iq;
          drow[NA + h * HD + d] = (dkn[d] - pk_ * kn[t][d] * dotk) * ik;
        }
      }
    }
    __syncwarp();
    // ---- B4: qkv Linear
#pragma unroll
    for (int jg = 0; jg < 3; ++jg) {
      const int j = lane + 32 * jg;
      if (j < 3 * NA) {
        float s = 0.f;
#pragma unroll
        for (int t = 0; t < 8; ++t) s += sDqkv[t * 96 + j];
        dbq[jg] += s;
      }
    }
    {
      float dt[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) dt[t] = 0.f;
#pragma unroll 2
      for (int j = 0; j < 3 * NA; ++j) {
        const float wv = sWqkv[j * 33 + lane];
#pragma unroll
        for (int t = 0; t < 8; ++t) dt[t] = fmaf(wv, sDqkv[t * 96 + j], dt[t]);
      }
#pragma unroll
      for (int t = 0; t < 8; ++t) atomicAdd(p.du + (size_t)nbr[t] + lane, dt[t]);
    }
   }
    // ================= block-wide weight-gradient accumulation over the windows of this pass
    __syncthreads();
    {
      const int nw = min(8, nwin - base);
      const int tid = threadIdx.x;
      const int mo = tid >> 2, mc = (tid & 3) * 16;
      const int pc = tid >> 3, pj = (tid & 7) * 4;
      for (int w = 0; w < nw; ++w) {
        const float* scr = sScr + w * LB::kScratchPerWarp;
        const float* wTok = scr;
        const float* wOm = scr + 1024;
        const float* wCat = scr + 1088;
        const float* wG = scr + 1152;
        const float* wDcat = scr + 1216;
        const float* wDqkv = scr + 1344;
        const float gv = wG[mo];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 c4 = *reinterpret_cast<const float4*>(wCat + mc + 4 * i);
          accM[4 * i] = fmaf(gv, c4.x, accM[4 * i]); accM[4 * i + 1] = fmaf(gv, c4.y, accM[4 * i + 1]);
          accM[4 * i + 2] = fmaf(gv, c4.z, accM[4 * i + 2]); accM[4 * i + 3] = fmaf(gv, c4.w, accM[4 * i + 3]);
        }
        {
          const float d0 = wDcat[pc], d1 = wDcat[32 + pc];
          const float4 o0 = *reinterpret_cast<const float4*>(wOm + pj);
          const float4 o1 = *reinterpret_cast<const float4*>(wOm + 32 + pj);
          accP[0] = fmaf(d0, o0.x, fmaf(d1, o1.x, accP[0])); accP[1] = fmaf(d0, o0.y, fmaf(d1, o1.y, accP[1]));
          accP[2] = fmaf(d0, o0.z, fmaf(d1, o1.z, accP[2])); accP[3] = fmaf(d0, o0.w, fmaf(d1, o1.w, accP[3]));
        }
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          const float4 x4 = *reinterpret_cast<const float4*>(wTok + t * 32 + pj);
#pragma unroll
          for (int i = 0; i < 3; ++i) {
            const float gq = wDqkv[t * 96 + pc + 32 * i];
            accQ[4 * i] = fmaf(gq, x4.x, accQ[4 * i]); accQ[4 * i + 1] = fmaf(gq, x4.y, accQ[4 * i + 1]);
            accQ[4 * i + 2] = fmaf(gq, x4.z, accQ[4 * i + 2]); accQ[4 * i + 3] = fmaf(gq, x4.w, accQ[4 * i + 3]);
          }
        }
      }
    }
    __syncthreads();                                   // the scratch is rewritten by the next pass
  }

  // ================= flush
  if (lane < 2 * NH) {                               // both directions of a head and all 8 warps share an entry
#pragma unroll
    for (int i = 0; i < 9; ++i) atomicAdd(sDBias + (lane % NH) * 12 + i, dBias9[i]);
    atomicAdd(sDScale + lane % NH, dScale1);
  }
#pragma unroll
  for (int og = 0; og < 2; ++og) atomicAdd(p.d_merge_b + lane + 32 * og, dbm[og]);
  atomicAdd(p.d_proj_b + lane, dbp);
#pragma unroll
  for (int jg = 0; jg < 3; ++jg)
    if (lane + 32 * jg < 3 * NA) atomicAdd(p.d_qkv_b + lane + 32 * jg, dbq[jg]);
  {
    const int tid = threadIdx.x;
    const int mo = tid >> 2, mc = (tid & 3) * 16, pc = tid >> 3, pj = (tid & 7) * 4;
    // the merge / qkv gradients of a thread are runs of 16 / 4 consecutive floats: 16-byte vector reductions when
    // the destinations are 16-byte aligned (dp.FlatGradBucket views are), as in unigram_bwd_kernel - every CTA hits
    // the same addresses, so the flush is bound by the number of reduction operations at L2
    const bool vec = ((reinterpret_cast<uintptr_t>(p.d_merge_w) | reinterpret_cast<uintptr_t>(p.d_qkv_w)) & 15) == 0;
    if (vec) {
#pragma unroll
      for (int i = 0; i < 4; ++i)
        atomicAdd(reinterpret_cast<float4*>(p.d_merge_w + mo * 64 + mc + 4 * i),
                  make_float4(accM[4 * i], accM[4 * i + 1], accM[4 * i + 2], accM[4 * i + 3]));
    } else {
#pragma unroll
      for (int i = 0; i < 16; ++i) atomicAdd(p.d_merge_w + mo * 64 + mc + i, accM[i]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (pj + i < NA) atomicAdd(p.d_proj_w + pc * NA + pj + i, accP[i]);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int j = pc + 32 * i;
      if (j < 3 * NA) {
        if (vec) {
          atomicAdd(reinterpret_cast<float4*>(p.d_qkv_w + j * 32 + pj),
                    make_float4(accQ[4 * i], accQ[4 * i + 1], accQ[4 * i + 2], accQ[4 * i + 3]));
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) atomicAdd(p.d_qkv_w + j * 32 + pj + k, accQ[4 * i + k]);
        }
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < NH * 9; i += blockDim.x)
    atomicAdd(p.d_bias_table + (i % 9) * NH + i / 9, sDBias[(i / 9) * 12 + i % 9]);
  if (threadIdx.x < NH && p.logit_scale[threadIdx.x] <= kLogitMax)
    atomicAdd(p.d_logit_scale + threadIdx.x, sDScale[threadIdx.x] * sScale[threadIdx.x]);
}

template <int NH>
static int launch_ngram_ctx_bwd(const NgramParams& p, cudaStream_t st) {
  using LB = NgramSmemBwd<NH>;
  const size_t smem = LB::kFloats * sizeof(float);
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(ngram_ctx_bwd_kernel<NH>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured.done();
  }
  const int nwin = p.B * p.wh * p.ww;
  int grid = (nwin + 7) / 8;
  if (grid > sm_count()) grid = sm_count();
  B200MAR_CHECK(cudaMemsetAsync(p.du, 0, (size_t)nwin * 32 * sizeof(float), st));
  ngram_ctx_bwd_kernel<NH><<<grid, kThreads, smem, st>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace b200mar

extern "C" int b200mar_unigram_bwd_add(const float* z, const float* weight, const float* du, float* dz,
                                       const float* dz_add, float* dweight, float* dbias, int B, int ph,
                                       int pw, void* stream) {
  using namespace b200mar;
  if (ph % kWin || pw % kWin || ph < kWin || pw < kWin) return -1;
  const int nwin = B * (ph / kWin) * (pw / kWin);
  int grid = nwin;
  // every CTA flushes the whole 4096-entry weight gradient: 3 CTAs per SM saturate the streaming part, more only
  // add flush traffic (measured at batch 16 x 128^2: 90 / 68 / 65 / 65 / 72 / 76 us for 1 / 2 / 3 / 4 / 6 / 8)
  if (grid > 3 * sm_count()) grid = 3 * sm_count();
  unigram_bwd_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(z, weight, du, dz, dz_add, dweight, dbias,
                                                                  B, ph, pw);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_unigram_bwd(const float* z, const float* weight, const float* du, float* dz,
                                   float* dweight, float* dbias, int B, int ph, int pw,
                                   void* stream) {
  return b200mar_unigram_bwd_add(z, weight, du, dz, nullptr, dweight, dbias, B, ph, pw, stream);
}

extern "C" int b200mar_ngram_ctx_bwd(const float* u, const float* dctx, const float* qkv_w,
                                     const float* qkv_b, const float* proj_w, const float* proj_b,
                                     const float* logit_scale, const float* bias_table,
                                     const float* merge_w, const float* merge_b, float* du,
                                     float* d_qkv_w, float* d_qkv_b, float* d_proj_w,
                                     float* d_proj_b, float* d_logit_scale, float* d_bias_table,
                                     float* d_merge_w, float* d_merge_b, int B, int wh, int ww,
                                     int nh, void* stream) {
  using namespace b200mar;
  if (wh < 2 || ww < 2) return -1;
  NgramParams p{};
  p.u = u; p.dctx = dctx; p.qkv_w = qkv_w; p.qkv_b = qkv_b; p.proj_w = proj_w; p.proj_b = proj_b;
  p.logit_scale = logit_scale; p.bias_table = bias_table; p.merge_w = merge_w; p.merge_b = merge_b;
  p.du = du; p.d_qkv_w = d_qkv_w; p.d_qkv_b = d_qkv_b; p.d_proj_w = d_proj_w; p.d_proj_b = d_proj_b;
  p.d_logit_scale = d_logit_scale; p.d_bias_table = d_bias_table; p.d_merge_w = d_merge_w;
  p.d_merge_b = d_merge_b; p.B = B; p.wh = wh; p.ww = ww;
  if (nh == 6) return launch_ngram_ctx_bwd<6>(p, (cudaStream_t)stream);
  if (nh == 4) return launch_ngram_ctx_bwd<4>(p, (cudaStream_t)stream);
  return -1;
}
