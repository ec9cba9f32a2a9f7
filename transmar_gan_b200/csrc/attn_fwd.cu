// attn_fwd.cu — fused N-gram window partition + scaled-cosine window attention + LN1 +
// residual, forward (fp32 SIMT path).  One persistent CTA streams (shifted) windows;
// the block's weights stay resident in shared memory.
//
// Reference semantics: components/window_utils.py:127-147 (context add, roll, partition),
// components/attention.py:70-101 (cosine attention), components/transformer_blocks.py:63-112
// (mask, unpartition, reverse roll) and :117 (x + norm1(.)).
#include "common.cuh"
#include "attn_common.cuh"

namespace b200mar {

template <int NH>
__global__ void __launch_bounds__(kThreads, 1)
attn_fwd_kernel(AttnParams p) {
  using L = AttnSmem<NH>;
  constexpr int HD = L::HD, A = L::A, HDP = L::HDP;
  extern __shared__ __align__(16) float smem[];
  float* sWqkv = smem + L::oWqkv;
  float* sWp = smem + L::oWp;
  float* sBqkv = smem + L::oBqkv;
  float* sBp = smem + L::oBp;
  float* sLnW = smem + L::oLnW;
  float* sLnB = smem + L::oLnB;
  float* sScale = smem + L::oScale;
  float* sBias = smem + L::oBias;
  float* sX = smem + L::oX;
  float* sQ = smem + L::oQ;
  float* sK = smem + L::oK;
  float* sV = smem + L::oV;
  float* sO = smem + L::oO;
  float* sP = smem + L::oP;
  int* sTok = reinterpret_cast<int*>(smem + L::oTok);
  int* sCwin = sTok + kTok;
  int* sReg = sCwin + kTok;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const Geom g = make_geom(p.B, p.ph, p.pw, p.shift);
  const int nwin = g.B * g.wh * g.ww;

  load_attn_weights<NH>(smem, p);
  __syncthreads();

  for (int win = blockIdx.x; win < nwin; win += gridDim.x) {
    // ---- slot table
    if (tid < kTok) {
      const Slot s = slot_of(g, win, tid);
      sTok[tid] = (int)s.tok; sCwin[tid] = s.cwin; sReg[tid] = s.region;
    }
    __syncthreads();
    // ---- gather X = z + ctx   (token-major, stride kLd64)
    {
      const int sub = lane >> 4, c = (lane & 15) << 2;
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int m = warp * 8 + it * 2 + sub;
        const float4 zv = *reinterpret_cast<const float4*>(p.z + (size_t)sTok[m] * kDim + c);
        const float4 cv = *reinterpret_cast<const float4*>(p.ctx + (size_t)sCwin[m] * kDim + c);
        *reinterpret_cast<float4*>(sX + m * kLd64 + c) =
            make_float4(zv.x + cv.x, zv.y + cv.y, zv.z + cv.z, zv.w + cv.w);
      }
    }
    __syncthreads();
    // ---- QKV projection: three 64x64 output blocks (q | k | v), raw values to sQ/sK/sV
#pragma unroll 1
    for (int part = 0; part < 3; ++part) {
      float acc[4][4]; zero_acc(acc);
      gemm64<false, false>(acc, sX, kLd64, sWqkv + part * A * kLd64, kLd64, kDim);
      float* dst = part == 0 ? sQ : (part == 1 ? sK : sV);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = gemm_col<false>(j);
        if (n < A) {
          const int h = n / HD, d = n - h * HD;
          const float bias = sBqkv[part * A + n];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int m = gemm_row<false>(i);
            dst[(h * kTok + m) * HDP + d] = acc[i][j] + bias;
          }
        }
      }
    }
    __syncthreads();
    // ---- F.normalize(q), F.normalize(k) in place
    for (int it = tid; it < 2 * NH * kTok; it += kThreads) {
      float* v = (it < NH * kTok ? sQ : sK) + (it % (NH * kTok)) * HDP;
      float ss = 0.f;
#pragma unroll
      for (int d = 0; d < HD; ++d) ss = fmaf(v[d], v[d], ss);
      const float inv = 1.0f / fmaxf(sqrtf(ss), kNormEps);
#pragma unroll
      for (int d = 0; d < HD; ++d) v[d] *= inv;
    }
    __syncthreads();
    // ---- per head: warp `warp` owns query rows 8*warp .. 8*warp+7
    {
      float* myP = sP + warp * kTok;
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
          // O[r, h*HD + d] = sum_c P[c] * V[c][d]; lanes (d = lane&15, half) split c even/odd
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
    // ---- output projection -> sX (reused as Yp, token-major)
    {
      float acc[4][4]; zero_acc(acc);
      gemm64<false, false>(acc, sO, kLd64, sWp, kLd64, A);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = gemm_col<false>(j);
        const float bias = sBp[n];
#pragma unroll
        for (int i = 0; i < 4; ++i) sX[gemm_row<false>(i) * kLd64 + n] = acc[i][j] + bias;
      }
    }
    __syncthreads();
    // ---- LayerNorm(norm1) + residual + scatter to source token positions
#pragma unroll 1
    for (int rr = 0; rr < 8; ++rr) {
      const int m = warp * 8 + rr;
      const float y0 = sX[m * kLd64 + lane], y1 = sX[m * kLd64 + lane + 32];
      const float mean = warp_sum(y0 + y1) * (1.0f / kDim);
      const float d0 = y0 - mean, d1 = y1 - mean;
      const float var = warp_sum(d0 * d0 + d1 * d1) * (1.0f / kDim);
      const float rstd = 1.0f / sqrtf(var + kLnEps);
      const size_t base = (size_t)sTok[m] * kDim;
      p.a[base + lane] = p.z[base + lane] + (d0 * rstd * sLnW[lane] + sLnB[lane]);
      p.a[base + lane + 32] = p.z[base + lane + 32] + (d1 * rstd * sLnW[lane + 32] + sLnB[lane + 32]);
    }
    __syncthreads();
  }
}

template <int NH>
static int launch_attn_fwd(const AttnParams& p, cudaStream_t st) {
  using L = AttnSmem<NH>;
  const size_t smem = L::kFloatsFwd * sizeof(float);
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(attn_fwd_kernel<NH>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured.done();
  }
  const int nwin = p.B * (p.ph / kWin) * (p.pw / kWin);
  const int grid = nwin < sm_count() ? nwin : sm_count();
  attn_fwd_kernel<NH><<<grid, kThreads, smem, st>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace b200mar

extern "C" int b200mar_attn_fwd(const float* z, const float* ctx, const float* qkv_w,
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
  if (nh == 6) return launch_attn_fwd<6>(p, st);
  if (nh == 4) return launch_attn_fwd<4>(p, st);
  return -1;
}
