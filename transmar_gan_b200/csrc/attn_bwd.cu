// attn_bwd.cu — backward of the fused N-gram window partition + scaled-cosine window
// attention + LN1 + residual (fp32 SIMT path).  Everything is recomputed from the block input
// z and the n-gram context; only per-CTA register / shared-memory accumulators and one atomic
// flush per CTA are used for the parameter gradients.
//
// Forward being differentiated (components/attention.py:70-101, transformer_blocks.py:117):
//   X = gather(z) + ctx ; [q|k|v] = X Wqkv^T + b ; qn = q/max(|q|,eps), kn likewise
//   S = s_h qn kn^T + bias_h + mask ; P = softmax(S) ; O_h = P v ; Yp = O Wp^T + bp
//   a = z + scatter(g1 * LN(Yp) + be1)
#include "common.cuh"
#include "attn_common.cuh"

namespace b200mar {

template <int NH>
__global__ void __launch_bounds__(kThreads, 1) attn_bwd_kernel(AttnParams p) {
  using L = AttnSmem<NH>;
  constexpr int HD = L::HD, A = L::A, HDP = L::HDP, LDY = L::kLdDY;
  constexpr int NE = (HD + 3) / 4;   // head-dim elements per thread in the (row, d%4) mapping
  extern __shared__ __align__(16) float smem[];
  float* sWqkv = smem + L::oWqkv;
  float* sWp = smem + L::oWp;
  float* sBqkv = smem + L::oBqkv;
  float* sBp = smem + L::oBp;
  float* sLnW = smem + L::oLnW;
  float* sScale = smem + L::oScale;
  float* sBias = smem + L::oBias;
  float* sXD = smem + L::bXD;
  float* sQ = smem + L::bQ;
  float* sK = smem + L::bK;
  float* sV = smem + L::bV;
  float* sNq = smem + L::bNq;
  float* sNk = smem + L::bNk;
  float* sBuf = smem + L::bBuf;
  float* sDY = smem + L::bDY;
  float* sO = sDY;                       // alias (dead before dY is written)
  float* sDa = sDY + kTok * kLd64;       // alias
  float* sDBias = smem + L::bDBias;
  int* sTok = reinterpret_cast<int*>(smem + L::bTok);
  int* sCwin = sTok + kTok;
  int* sReg = sCwin + kTok;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const Geom g = make_geom(p.B, p.ph, p.pw, p.shift);
  const int nwin = g.B * g.wh * g.ww;

  load_attn_weights<NH>(smem, p);
  for (int i = tid; i < NH * L::kBiasLd; i += kThreads) sDBias[i] = 0.f;
  __syncthreads();

  // persistent per-thread gradient accumulators
  float accWqkv[3][4][4], accWp[4][4];
#pragma unroll
  for (int c = 0; c < 3; ++c) zero_acc(accWqkv[c]);
  zero_acc(accWp);
  float dbqkv = 0.f;                                 // column tid (< 192)
  float dbp[2] = {0.f, 0.f}, dg[2] = {0.f, 0.f}, dbe[2] = {0.f, 0.f};
  float dscale[NH];
#pragma unroll
  for (int h = 0; h < NH; ++h) dscale[h] = 0.f;

  for (int win = blockIdx.x; win < nwin; win += gridDim.x) {
    // ================= phase 0: slot table, gather X and the incoming gradient
    if (tid < kTok) {
      const Slot s = slot_of(g, win, tid);
      sTok[tid] = (int)s.tok; sCwin[tid] = s.cwin; sReg[tid] = s.region;
    }
    __syncthreads();
    {
      const int sub = lane >> 4, c = (lane & 15) << 2;
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int m = warp * 8 + it * 2 + sub;
        const size_t tb = (size_t)sTok[m] * kDim + c;
        const float4 zv = *reinterpret_cast<const float4*>(p.z + tb);
        const float4 cv = *reinterpret_cast<const float4*>(p.ctx + (size_t)sCwin[m] * kDim + c);
        *reinterpret_cast<float4*>(sXD + m * kLd64 + c) =
            make_float4(zv.x + cv.x, zv.y + cv.y, zv.z + cv.z, zv.w + cv.w);
        *reinterpret_cast<float4*>(sDa + m * kLd64 + c) = *reinterpret_cast<const float4*>(p.da + tb);
      }
    }
    __syncthreads();
    // ================= phase 1: recompute q,k,v ; normalise q,k (keep the norms)
#pragma unroll 1
    for (int part = 0; part < 3; ++part) {
      float acc[4][4]; zero_acc(acc);
      gemm64<false, false>(acc, sXD, kLd64, sWqkv + part * A * kLd64, kLd64, kDim);
      float* dst = part == 0 ? sQ : (part == 1 ? sK : sV);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = gemm_col<false>(j);
        if (n < A) {
          const int h = n / HD, d = n - h * HD;
          const float bias = sBqkv[part * A + n];
#pragma unroll
          for (int i = 0; i < 4; ++i) dst[(h * kTok + gemm_row<false>(i)) * HDP + d] = acc[i][j] + bias;
        }
      }
    }
    __syncthreads();
    for (int it = tid; it < 2 * NH * kTok; it += kThreads) {
      const bool isq = it < NH * kTok;
      const int hm = it % (NH * kTok);
      float* v = (isq ? sQ : sK) + hm * HDP;
      float ss = 0.f;
#pragma unroll
      for (int d = 0; d < HD; ++d) ss = fmaf(v[d], v[d], ss);
      const float nrm = sqrtf(ss);
      (isq ? sNq : sNk)[hm] = nrm;
      const float inv = 1.0f / fmaxf(nrm, kNormEps);
#pragma unroll
      for (int d = 0; d < HD; ++d) v[d] *= inv;
    }
    __syncthreads();
    // ================= phase 2: recompute O (same arithmetic as the forward kernel)
    {
      const int dl = lane & 15, half = lane >> 4;
#pragma unroll 1
      for (int h = 0; h < NH; ++h) {
        const float scale = sScale[h];
        const float* q = sQ + h * kTok * HDP;
        const float* k = sK + h * kTok * HDP;
        const float* v = sV + h * kTok * HDP;
        const float* bias = sBias + h * L::kBiasLd;
#pragma unroll 1
        for (int rr = 0; rr < 8; ++rr) {
          const int r = warp * 8 + rr;
          float* myP = sBuf + r * kLd64;
          float s0 = 0.f, s1 = 0.f;
#pragma unroll
          for (int d = 0; d < HD; ++d) {
            const float qd = q[r * HDP + d];
            s0 = fmaf(qd, k[lane * HDP + d], s0);
            s1 = fmaf(qd, k[(lane + 32) * HDP + d], s1);
          }
          s0 = s0 * scale + bias[rel_index(r, lane)];
          s1 = s1 * scale + bias[rel_index(r, lane + 32)];
          if (p.shift > 0) {
            const int rg = sReg[r];
            if (sReg[lane] != rg) s0 += -100.0f;
            if (sReg[lane + 32] != rg) s1 += -100.0f;
          }
          const float mx = warp_max(fmaxf(s0, s1));
          const float e0 = expf(s0 - mx), e1 = expf(s1 - mx);
          const float inv = 1.0f / warp_sum(e0 + e1);
          __syncwarp();
          myP[lane] = e0 * inv;
          myP[lane + 32] = e1 * inv;
          __syncwarp();
          float o = 0.f;
          if (dl < HD) {
#pragma unroll 8
            for (int c2 = 0; c2 < 32; ++c2) {
              const int c = 2 * c2 + half;
              o = fmaf(myP[c], v[c * HDP + dl], o);
            }
          }
          o += __shfl_xor_sync(0xffffffffu, o, 16);
          if (half == 0 && dl < HD) sO[r * kLd64 + h * HD + dl] = o;
        }
      }
    }
    __syncthreads();
    // ================= phase 3: Yp = O Wp^T + bp -> sBuf ; LayerNorm backward -> sBuf = dYp
    {
      float acc[4][4]; zero_acc(acc);
      gemm64<false, false>(acc, sO, kLd64, sWp, kLd64, A);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = gemm_col<false>(j);
        const float bias = sBp[n];
#pragma unroll
        for (int i = 0; i < 4; ++i) sBuf[gemm_row<false>(i) * kLd64 + n] = acc[i][j] + bias;
      }
    }
    __syncthreads();
#pragma unroll 1
    for (int rr = 0; rr < 8; ++rr) {
      const int m = warp * 8 + rr;
      const float y0 = sBuf[m * kLd64 + lane], y1 = sBuf[m * kLd64 + lane + 32];
      const float mean = warp_sum(y0 + y1) * (1.0f / kDim);
      const float c0 = y0 - mean, c1 = y1 - mean;
      const float rstd = 1.0f / sqrtf(warp_sum(c0 * c0 + c1 * c1) * (1.0f / kDim) + kLnEps);
      const float n0 = c0 * rstd, n1 = c1 * rstd;
      const float g0 = sDa[m * kLd64 + lane], g1 = sDa[m * kLd64 + lane + 32];
      dg[0] = fmaf(g0, n0, dg[0]); dg[1] = fmaf(g1, n1, dg[1]);
      dbe[0] += g0; dbe[1] += g1;
      const float t0 = g0 * sLnW[lane], t1 = g1 * sLnW[lane + 32];
      const float m1 = warp_sum(t0 + t1) * (1.0f / kDim);
      const float m2 = warp_sum(t0 * n0 + t1 * n1) * (1.0f / kDim);
      const float d0 = rstd * (t0 - m1 - n0 * m2), d1 = rstd * (t1 - m1 - n1 * m2);
      sBuf[m * kLd64 + lane] = d0; sBuf[m * kLd64 + lane + 32] = d1;
      dbp[0] += d0; dbp[1] += d1;
    }
    __syncthreads();
    // ================= phase 4: dWp += dYp^T O ; dO = dYp Wp -> sXD
    gemm64<true, true>(accWp, sBuf, kLd64, sO, kLd64, kTok);
    {
      float acc[4][4]; zero_acc(acc);
      gemm64<false, true>(acc, sBuf, kLd64, sWp, kLd64, kDim);
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<float4*>(sXD + gemm_row<false>(i) * kLd64 + gemm_col<true>(0)) =
            make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    }
    __syncthreads();
    // zero the pad columns [3A, 192) of the dY tile (the region held O / da until now)
    if (3 * A < 192) {
      constexpr int PADC = 192 - 3 * A > 0 ? 192 - 3 * A : 1;
      for (int i = tid; i < kTok * PADC; i += kThreads)
        sDY[(i / PADC) * LDY + 3 * A + (i % PADC)] = 0.f;
    }
    // ================= phase 5: per head — P, dV, dS, dqn/dkn -> dq/dk
    {
      const int row4 = tid >> 2, dgp = tid & 3;
#pragma unroll 1
      for (int h = 0; h < NH; ++h) {
        const float scale = sScale[h];
        const float* q = sQ + h * kTok * HDP;
        const float* k = sK + h * kTok * HDP;
        const float* v = sV + h * kTok * HDP;
        const float* bias = sBias + h * L::kBiasLd;
        float* dbias = sDBias + h * L::kBiasLd;
        float p0[8], p1[8];
        // ---- a1: probabilities of this warp's 8 rows -> registers + sBuf
#pragma unroll
        for (int rr = 0; rr < 8; ++rr) {
          const int r = warp * 8 + rr;
          float s0 = 0.f, s1 = 0.f;
#pragma unroll
          for (int d = 0; d < HD; ++d) {
            const float qd = q[r * HDP + d];
            s0 = fmaf(qd, k[lane * HDP + d], s0);
            s1 = fmaf(qd, k[(lane + 32) * HDP + d], s1);
          }
          s0 = s0 * scale + bias[rel_index(r, lane)];
          s1 = s1 * scale + bias[rel_index(r, lane + 32)];
          if (p.shift > 0) {
            const int rg = sReg[r];
            if (sReg[lane] != rg) s0 += -100.0f;
            if (sReg[lane + 32] != rg) s1 += -100.0f;
          }
          const float mx = warp_max(fmaxf(s0, s1));
          const float e0 = expf(s0 - mx), e1 = expf(s1 - mx);
          const float inv = 1.0f / warp_sum(e0 + e1);
          p0[rr] = e0 * inv; p1[rr] = e1 * inv;
          sBuf[r * kLd64 + lane] = p0[rr];
          sBuf[r * kLd64 + lane + 32] = p1[rr];
        }
        __syncthreads();
        // ---- a2: dV[c][d] = sum_r P[r][c] dO[r][h*HD+d]      (thread: c = row4, d = dgp + 4e)
        {
          float dv[NE];
#pragma unroll
          for (int e = 0; e < NE; ++e) dv[e] = 0.f;
#pragma unroll 4
          for (int r = 0; r < kTok; ++r) {
            const float pv = sBuf[r * kLd64 + row4];
#pragma unroll
            for (int e = 0; e < NE; ++e) {
              const int d = dgp + 4 * e;
              if (d < HD) dv[e] = fmaf(pv, sXD[r * kLd64 + h * HD + d], dv[e]);
            }
          }
#pragma unroll
          for (int e = 0; e < NE; ++e) {
            const int d = dgp + 4 * e;
            if (d < HD) sDY[row4 * LDY + 2 * A + h * HD + d] = dv[e];
          }
        }
        __syncthreads();
        // ---- a3: dP, dS rows (warp-rows) -> sBuf ; bias-table and logit-scale gradients
#pragma unroll
        for (int rr = 0; rr < 8; ++rr) {
          const int r = warp * 8 + rr;
          float dp0 = 0.f, dp1 = 0.f, cs0 = 0.f, cs1 = 0.f;
#pragma unroll
          for (int d = 0; d < HD; ++d) {
            const float dod = sXD[r * kLd64 + h * HD + d];
            const float qd = q[r * HDP + d];
            dp0 = fmaf(dod, v[lane * HDP + d], dp0);
            dp1 = fmaf(dod, v[(lane + 32) * HDP + d], dp1);
            cs0 = fmaf(qd, k[lane * HDP + d], cs0);
            cs1 = fmaf(qd, k[(lane + 32) * HDP + d], cs1);
          }
          const float rowdot = warp_sum(dp0 * p0[rr] + dp1 * p1[rr]);
          const float ds0 = p0[rr] * (dp0 - rowdot), ds1 = p1[rr] * (dp1 - rowdot);
          dscale[h] = fmaf(ds0, cs0, fmaf(ds1, cs1, dscale[h]));
          atomicAdd(dbias + rel_index(r, lane), ds0);
          atomicAdd(dbias + rel_index(r, lane + 32), ds1);
          sBuf[r * kLd64 + lane] = ds0;
          sBuf[r * kLd64 + lane + 32] = ds1;
        }
        __syncthreads();
        // ---- a4: dqn = s * dS kn ; dkn = s * dS^T qn ; F.normalize Jacobian -> dq, dk in sDY
        {
          float dq[NE], dk[NE];
#pragma unroll
          for (int e = 0; e < NE; ++e) { dq[e] = 0.f; dk[e] = 0.f; }
#pragma unroll 4
          for (int c = 0; c < kTok; ++c) {
            const float dsr = sBuf[row4 * kLd64 + c];   // dS[row4][c]
            const float dsc = sBuf[c * kLd64 + row4];   // dS[c][row4]
#pragma unroll
            for (int e = 0; e < NE; ++e) {
              const int d = dgp + 4 * e;
              if (d < HD) {
                dq[e] = fmaf(dsr, k[c * HDP + d], dq[e]);
                dk[e] = fmaf(dsc, q[c * HDP + d], dk[e]);
              }
            }
          }
          float dotq = 0.f, dotk = 0.f;
#pragma unroll
          for (int e = 0; e < NE; ++e) {
            const int d = dgp + 4 * e;
            if (d < HD) {
              dq[e] *= scale; dk[e] *= scale;
              dotq = fmaf(q[row4 * HDP + d], dq[e], dotq);
              dotk = fmaf(k[row4 * HDP + d], dk[e], dotk);
            }
          }
          dotq += __shfl_xor_sync(0xffffffffu, dotq, 1);
          dotq += __shfl_xor_sync(0xffffffffu, dotq, 2);
          dotk += __shfl_xor_sync(0xffffffffu, dotk, 1);
          dotk += __shfl_xor_sync(0xffffffffu, dotk, 2);
          const float nq = sNq[h * kTok + row4], nk = sNk[h * kTok + row4];
          const float iq = 1.0f / fmaxf(nq, kNormEps), ik = 1.0f / fmaxf(nk, kNormEps);
          const float pq = nq >= kNormEps ? 1.f : 0.f, pk = nk >= kNormEps ? 1.f : 0.f;
#pragma unroll
          for (int e = 0; e < NE; ++e) {
            const int d = dgp + 4 * e;
            if (d < HD) {
              sDY[row4 * LDY + h * HD + d] = (dq[e] - pq * q[row4 * HDP + d] * dotq) * iq;
              sDY[row4 * LDY + A + h * HD + d] = (dk[e] - pk * k[row4 * HDP + d] * dotk) * ik;
            }
          }
        }
        __syncthreads();
      }
    }
    // ================= phase 6: dWqkv, dbqkv, dX -> dz / dctx
    {
      const int sub = lane >> 4, c = (lane & 15) << 2;
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int m = warp * 8 + it * 2 + sub;
        const float4 zv = *reinterpret_cast<const float4*>(p.z + (size_t)sTok[m] * kDim + c);
        const float4 cv = *reinterpret_cast<const float4*>(p.ctx + (size_t)sCwin[m] * kDim + c);
        *reinterpret_cast<float4*>(sXD + m * kLd64 + c) =
            make_float4(zv.x + cv.x, zv.y + cv.y, zv.z + cv.z, zv.w + cv.w);
      }
    }
    __syncthreads();
#pragma unroll
    for (int chunk = 0; chunk < 3; ++chunk)
      gemm64<true, true>(accWqkv[chunk], sDY + chunk * 64, LDY, sXD, kLd64, kTok);
    if (tid < 192) {
      float s = 0.f;
#pragma unroll 8
      for (int m = 0; m < kTok; ++m) s += sDY[m * LDY + tid];
      dbqkv += s;
    }
    {
      float acc[4][4]; zero_acc(acc);
      gemm64<false, true>(acc, sDY, LDY, sWqkv, kLd64, 192);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int m = gemm_row<false>(i), c = gemm_col<true>(0);
        const size_t tb = (size_t)sTok[m] * kDim + c;
        const float4 gv = *reinterpret_cast<const float4*>(p.da + tb);
        *reinterpret_cast<float4*>(p.dz + tb) =
            make_float4(gv.x + acc[i][0], gv.y + acc[i][1], gv.z + acc[i][2], gv.w + acc[i][3]);
        *reinterpret_cast<float4*>(sBuf + m * kLd64 + c) =
            make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
      }
    }
    __syncthreads();
    {
      // dctx: sum dX over the tokens that share a source window (4 quadrants of 16 slots)
      const int quad = tid >> 6, c = tid & 63;
      const int i0 = (quad >> 1) * 4, j0 = (quad & 1) * 4;
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) s += sBuf[((i0 + i) * kWin + j0 + j) * kLd64 + c];
      atomicAdd(p.dctx + (size_t)sCwin[i0 * kWin + j0] * kDim + c, s);
    }
    __syncthreads();
  }

  // ================= flush parameter gradients
#pragma unroll
  for (int chunk = 0; chunk < 3; ++chunk)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int n = chunk * 64 + gemm_row<true>(i);
      if (n < 3 * A) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          atomicAdd(p.d_qkv_w + n * kDim + gemm_col<true>(j), accWqkv[chunk][i][j]);
      }
    }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int col = gemm_col<true>(j);
      if (col < A) atomicAdd(p.d_proj_w + gemm_row<true>(i) * A + col, accWp[i][j]);
    }
  if (tid < 3 * A) atomicAdd(p.d_qkv_b + tid, dbqkv);
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    atomicAdd(p.d_proj_b + lane + 32 * e, dbp[e]);
    atomicAdd(p.d_ln_w + lane + 32 * e, dg[e]);
    atomicAdd(p.d_ln_b + lane + 32 * e, dbe[e]);
  }
#pragma unroll
  for (int h = 0; h < NH; ++h) {
    const float v = warp_sum(dscale[h]);
    // d/d(logit_scale) of exp(min(ls, ln100)): torch.clamp passes the gradient for ls <= max
    if (lane == 0 && p.logit_scale[h] <= kLogitMax) atomicAdd(p.d_logit_scale + h, v * sScale[h]);
  }
  __syncthreads();
  for (int i = tid; i < NH * 225; i += kThreads) {
    const int h = i / 225, e = i - h * 225;
    atomicAdd(p.d_bias_table + e * NH + h, sDBias[h * L::kBiasLd + e]);
  }
}

template <int NH>
static int launch_attn_bwd(const AttnParams& p, cudaStream_t st) {
  using L = AttnSmem<NH>;
  const size_t smem = L::kFloatsBwd * sizeof(float);
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(attn_bwd_kernel<NH>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured.done();
  }
  const int nwin = p.B * (p.ph / kWin) * (p.pw / kWin);
  const int grid = nwin < sm_count() ? nwin : sm_count();
  attn_bwd_kernel<NH><<<grid, kThreads, smem, st>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace b200mar

extern "C" int b200mar_attn_bwd(const float* z, const float* ctx, const float* da,
                                const float* qkv_w, const float* qkv_b, const float* proj_w,
                                const float* proj_b, const float* logit_scale,
                                const float* bias_table, const float* ln_w, const float* ln_b,
                                float* dz, float* dctx, float* d_qkv_w, float* d_qkv_b,
                                float* d_proj_w, float* d_proj_b, float* d_logit_scale,
                                float* d_bias_table, float* d_ln_w, float* d_ln_b, int B, int ph,
                                int pw, int nh, int shift, void* stream) {
  using namespace b200mar;
  if (ph % kWin || pw % kWin || ph < kWin || pw < kWin || (shift != 0 && shift != kShift)) return -1;
  if ((long long)B * ph * pw >= (1ll << 31)) return -1;
  AttnParams p{};
  p.z = z; p.ctx = ctx; p.da = da; p.qkv_w = qkv_w; p.qkv_b = qkv_b; p.proj_w = proj_w;
  p.proj_b = proj_b; p.logit_scale = logit_scale; p.bias_table = bias_table; p.ln_w = ln_w;
  p.ln_b = ln_b; p.dz = dz; p.dctx = dctx; p.d_qkv_w = d_qkv_w; p.d_qkv_b = d_qkv_b;
  p.d_proj_w = d_proj_w; p.d_proj_b = d_proj_b; p.d_logit_scale = d_logit_scale;
  p.d_bias_table = d_bias_table; p.d_ln_w = d_ln_w; p.d_ln_b = d_ln_b;
  p.B = B; p.ph = ph; p.pw = pw; p.shift = shift;
  cudaStream_t st = (cudaStream_t)stream;
  if (nh == 6) return launch_attn_bwd<6>(p, st);
  if (nh == 4) return launch_attn_bwd<4>(p, st);
  return -1;
}
