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

// Two 256-thread HALVES per CTA, each streaming its own windows through its own working set while both share the
// resident weights: the kernel is a chain of latency-bound phases (per-row softmax with warp reductions, small tile
// GEMMs), and one 8-warp CTA per SM (what 165 KB of shared memory allowed) left two warps per scheduler.
template <int NH>
struct AttnFwdSmem {
  using L = AttnSmem<NH>;
  static constexpr int oWork = L::oX;                               // weights / bias / LN / scale / bias table below
  static constexpr int wX = 0;                                      // [64][68] X = z + ctx; later O; later Y
  static constexpr int wQ = wX + kTok * kLd64;
  static constexpr int wK = wQ + L::kHeadFloats;
  static constexpr int wV = wK + L::kHeadFloats;
  static constexpr int wP = wV + L::kHeadFloats;                    // [8 warps][2 rows][64]
  static constexpr int wTok = wP + 16 * kTok;
  static constexpr int kWorkFloats = wTok + 4 * kTok;
  static constexpr int kFloats = oWork + 2 * kWorkFloats;
  static_assert(kFloats * 4 <= 232448, "attn_fwd shared memory exceeds 227 KB");
};

__device__ __forceinline__ void half_sync(int half) {
  asm volatile("bar.sync %0, 256;" ::"r"(1 + half) : "memory");
}
// gemm64 of common.cuh for a 256-thread HALF of the CTA (t = thread index inside the half)
__device__ __forceinline__ void gemm64_half(float (&acc)[4][4], const float* __restrict__ A, int lda,
                                            const float* __restrict__ B, int ldb, int K, int t) {
  const int ty = t >> 4, tx = t & 15;
#pragma unroll 2
  for (int k = 0; k < K; k += 4) {
    float a[4][4], b[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 v = *reinterpret_cast<const float4*>(A + (ty + 16 * i) * lda + k);
      a[i][0] = v.x; a[i][1] = v.y; a[i][2] = v.z; a[i][3] = v.w;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float4 v = *reinterpret_cast<const float4*>(B + (tx + 16 * j) * ldb + k);
      b[j][0] = v.x; b[j][1] = v.y; b[j][2] = v.z; b[j][3] = v.w;
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk)
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i][kk], b[j][kk], acc[i][j]);
  }
}

template <int NH>
__global__ void __launch_bounds__(2 * kThreads, 1)
attn_fwd_kernel(AttnParams p) {
  using L = AttnSmem<NH>;
  using F = AttnFwdSmem<NH>;
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
  const int half = __shfl_sync(0xffffffffu, (int)threadIdx.x >> 8, 0);     // warp-uniform
  const int tid = threadIdx.x & 255, lane = tid & 31, warp = tid >> 5;
  float* work = smem + F::oWork + half * F::kWorkFloats;
  float* sX = work + F::wX;
  float* sQ = work + F::wQ;
  float* sK = work + F::wK;
  float* sV = work + F::wV;
  float* sO = sX;                                  // O overwrites X (dead after the QKV products)
  float* sP = work + F::wP;
  int* sTok = reinterpret_cast<int*>(work + F::wTok);
  int* sCwin = sTok + kTok;
  int* sReg = sCwin + kTok;
  const int ty = tid >> 4, tx = tid & 15;          // tile-product thread coordinates inside the half

  const Geom g = make_geom(p.B, p.ph, p.pw, p.shift);
  const int nwin = g.B * g.wh * g.ww;

  load_attn_weights<NH>(smem, p);
  __syncthreads();

  for (int win = blockIdx.x * 2 + half; win < nwin; win += gridDim.x * 2) {
    // ---- slot table
    if (tid < kTok) {
      const Slot s = slot_of(g, win, tid);
      sTok[tid] = (int)s.tok; sCwin[tid] = s.cwin; sReg[tid] = s.region;
    }
    half_sync(half);
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
    half_sync(half);
    // ---- QKV projection: three 64x64 output blocks (q | k | v), raw values to sQ/sK/sV
#pragma unroll 1
    for (int part = 0; part < 3; ++part) {
      float acc[4][4]; zero_acc(acc);
      gemm64_half(acc, sX, kLd64, sWqkv + part * A * kLd64, kLd64, kDim, tid);
      float* dst = part == 0 ? sQ : (part == 1 ? sK : sV);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = tx + 16 * j;
        if (n < A) {
          const int h = n / HD, d = n - h * HD;
          const float bias = sBqkv[part * A + n];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int m = ty + 16 * i;
            dst[(h * kTok + m) * HDP + d] = acc[i][j] + bias;
          }
        }
      }
    }
    half_sync(half);
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
    half_sync(half);
    // ---- per head: warp `warp` owns query rows 8*warp .. 8*warp+7, TWO rows per iteration: the K and V values a
    //      lane reads from shared memory serve both rows, and the two rows' warp reductions interleave
    {
      float* myP = sP + warp * 2 * kTok;
      const int dl = lane & 15, hf = lane >> 4;
#pragma unroll 1
      for (int h = 0; h < NH; ++h) {
        const float scale = sScale[h];
        const float* q = sQ + h * kTok * HDP;
        const float* k = sK + h * kTok * HDP;
        const float* v = sV + h * kTok * HDP;
        const float* bias = sBias + h * L::kBiasLd;
#pragma unroll 1
        for (int rr = 0; rr < 8; rr += 2) {
          const int r = warp * 8 + rr;                 // rows r and r + 1 (same window row: r is even)
          float s00 = 0.f, s01 = 0.f, s10 = 0.f, s11 = 0.f;
#pragma unroll
          for (int d = 0; d < HD; ++d) {
            const float k0 = k[lane * HDP + d], k1 = k[(lane + 32) * HDP + d];
            const float qa = q[r * HDP + d], qb = q[(r + 1) * HDP + d];
            s00 = fmaf(qa, k0, s00); s01 = fmaf(qa, k1, s01);
            s10 = fmaf(qb, k0, s10); s11 = fmaf(qb, k1, s11);
          }
          s00 = s00 * scale + bias[rel_index(r, lane)];
          s01 = s01 * scale + bias[rel_index(r, lane + 32)];
          s10 = s10 * scale + bias[rel_index(r + 1, lane)];
          s11 = s11 * scale + bias[rel_index(r + 1, lane + 32)];
          if (p.shift > 0) {
            const int ra = sReg[r], rb = sReg[r + 1], l0 = sReg[lane], l1 = sReg[lane + 32];
            if (l0 != ra) s00 += -100.0f;
            if (l1 != ra) s01 += -100.0f;
            if (l0 != rb) s10 += -100.0f;
            if (l1 != rb) s11 += -100.0f;
          }
          float ma = fmaxf(s00, s01), mb = fmaxf(s10, s11);
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            ma = fmaxf(ma, __shfl_xor_sync(0xffffffffu, ma, o));
            mb = fmaxf(mb, __shfl_xor_sync(0xffffffffu, mb, o));
          }
          const float e00 = expf(s00 - ma), e01 = expf(s01 - ma), e10 = expf(s10 - mb), e11 = expf(s11 - mb);
          float ta = e00 + e01, tb = e10 + e11;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            ta += __shfl_xor_sync(0xffffffffu, ta, o);
            tb += __shfl_xor_sync(0xffffffffu, tb, o);
          }
          const float ia = 1.0f / ta, ib = 1.0f / tb;
          __syncwarp();
          myP[lane] = e00 * ia; myP[lane + 32] = e01 * ia;
          myP[kTok + lane] = e10 * ib; myP[kTok + lane + 32] = e11 * ib;
          __syncwarp();
          // O[r, h*HD + d] = sum_c P[c] * V[c][d]; lanes (d = lane&15, half) split c even/odd
          float oa = 0.f, ob = 0.f;
          if (dl < HD) {
#pragma unroll 8
            for (int c2 = 0; c2 < 32; ++c2) {
              const int c = 2 * c2 + hf;
              const float vv = v[c * HDP + dl];
              oa = fmaf(myP[c], vv, oa);
              ob = fmaf(myP[kTok + c], vv, ob);
            }
          }
          oa += __shfl_xor_sync(0xffffffffu, oa, 16);
          ob += __shfl_xor_sync(0xffffffffu, ob, 16);
          if (hf == 0 && dl < HD) {
            sO[r * kLd64 + h * HD + dl] = oa;
            sO[(r + 1) * kLd64 + h * HD + dl] = ob;
          }
        }
      }
    }
    half_sync(half);
    // ---- output projection: O and Y share one buffer, so the products are held until every read is done
    {
      float acc[4][4]; zero_acc(acc);
      gemm64_half(acc, sO, kLd64, sWp, kLd64, A, tid);
      half_sync(half);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = tx + 16 * j;
        const float bias = sBp[n];
#pragma unroll
        for (int i = 0; i < 4; ++i) sX[(ty + 16 * i) * kLd64 + n] = acc[i][j] + bias;
      }
    }
    half_sync(half);
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
    half_sync(half);
  }
}

template <int NH>
static int launch_attn_fwd(const AttnParams& p, cudaStream_t st) {
  const size_t smem = AttnFwdSmem<NH>::kFloats * sizeof(float);
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(attn_fwd_kernel<NH>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured.done();
  }
  const int nwin = p.B * (p.ph / kWin) * (p.pw / kWin);
  const int pairs = (nwin + 1) / 2;
  const int grid = pairs < sm_count() ? pairs : sm_count();
  attn_fwd_kernel<NH><<<grid, 2 * kThreads, smem, st>>>(p);
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
