// glue_fwd.cu — the non-NSTB generator glue, forward, as fused
// channels-last kernels: everything between the transformer stages stays in the [B, T, 64] token
// layout (no NCHW <-> token flips, no materialised pooled / concatenated / permuted copies).
//
//   shallow_tok      ShallowExtractor conv3x3 1->64 written token-major        (encoder_decoder.py:23-27, ngswin.py:172)
//   cascade_proj     InterPool (2x2 max) of the cascade + concat + Linear       (pooling.py:13-17, ngswin.py:175-178, encoder_decoder.py:117,136)
//   bottleneck_cat   BottleneckPool + LeakyReLU skip + pixel-shuffle + concat   (pooling.py:32-37, encoder_decoder.py:159-164,191-192)
//   bottleneck_tail  depthwise 3x3 + GELU + Linear 84->64 + LN, + e1 (decoder in) (encoder_decoder.py:194-200, ngswin.py:189)
//   final_dot9       LN(norm) + global skip, then the 64->1 conv's per-tap dots (ngswin.py:190,204; reconstruction.py:14)
//   recon            tap gather (conv 64->1), conv3x3 1->1, tanh, crop           (reconstruction.py:14-19, ngswin.py:208)
#include "common.cuh"
#include "tc.cuh"
#include "tma.cuh"
#include "../../include/b200mar.h"

namespace b200mar {

__device__ __forceinline__ float leaky(float x) { return x > 0.f ? x : 0.01f * x; }
__device__ __forceinline__ float4 max4(float4 a, float4 b) {
  return make_float4(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z), fmaxf(a.w, b.w));
}
__device__ __forceinline__ float4 ld4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

// ------------------------------------------------------------------ shallow conv, token-major out
__global__ void __launch_bounds__(256) shallow_tok_kernel(const float* __restrict__ x,
                                                          const float* __restrict__ w,
                                                          const float* __restrict__ b,
                                                          float* __restrict__ c0, int B, int H, int W) {
  // 16 threads (4 channels each) per group of FOUR consecutive pixels of a row: the 3 x 6 input patch is loaded once
  // for the four outputs and the pixel index is decomposed once per group, by multiply-shift (the kernel was
  // instruction-bound at 45 % of its HBM floor: two integer divisions and nine guarded loads per float4 written)
  const int cq = threadIdx.x & 15, pl = threadIdx.x >> 4;
  float wr[4][9], br[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    br[j] = b[cq * 4 + j];
#pragma unroll
    for (int k = 0; k < 9; ++k) wr[j][k] = w[(cq * 4 + j) * 9 + k];
  }
  const unsigned gw = ((unsigned)W + 3u) / 4u;
  const unsigned ngrp = (unsigned)B * H * gw;          // < 2^31 (checked by the launcher): 32-bit index math
  const FastDiv fgw = make_fastdiv(gw), fH = make_fastdiv((unsigned)H);
  for (unsigned g = blockIdx.x * 16u + pl; g < ngrp; g += gridDim.x * 16u) {
    const unsigned rowi = fdiv(g, fgw);                // b * H + y
    const int x0 = (int)(g - rowi * gw) * 4;
    const unsigned bb = fdiv(rowi, fH);
    const int yy = (int)(rowi - bb * (unsigned)H);
    const float* img = x + (size_t)bb * H * W;
    float in[3][6];
#pragma unroll
    for (int dy = 0; dy < 3; ++dy) {
      const int y2 = yy + dy - 1;
      const bool rowok = y2 >= 0 && y2 < H;
#pragma unroll
      for (int dx = 0; dx < 6; ++dx) {
        const int x2 = x0 + dx - 1;
        in[dy][dx] = (rowok && x2 >= 0 && x2 < W) ? __ldg(img + (size_t)y2 * W + x2) : 0.f;
      }
    }
#pragma unroll
    for (int px = 0; px < 4; ++px) {
      if (x0 + px < W) {
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          o[j] = br[j];
#pragma unroll
          for (int dy = 0; dy < 3; ++dy)
#pragma unroll
            for (int dx = 0; dx < 3; ++dx) o[j] = fmaf(wr[j][dy * 3 + dx], in[dy][px + dx], o[j]);
        }
        *reinterpret_cast<float4*>(c0 + ((size_t)rowi * W + x0 + px) * 64 + cq * 4) = make_float4(o[0], o[1], o[2], o[3]);
      }
    }
  }
}

// ------------------------------------------------------------------ InterPool + concat + Linear
// out[b, t, :] = W [64, K] * cat(pool2x2(srcA)[t], pool2x2(srcB)[t] (optional), direct[t]) + bias,
// K = 128 (no srcB) or 192; optionally also writes pool2x2(srcA) (the next stage's / bottleneck's
// pooled shallow features).  64 tokens per tile, SIMT 64x64xK tile product.
struct CascadeParams {
  const float* srcA; const float* srcB; const float* direct; const float* w; const float* bias;
  float* out; float* poolA;
  int B, h, wd, K;
};
__global__ void __launch_bounds__(kThreads, 1) cascade_proj_kernel(CascadeParams p) {
  extern __shared__ __align__(16) float smem[];
  const int ld = p.K + 4;                                  // == 4 mod 32 for K = 128, 192
  float* sW = smem;                                        // [64][ld]
  float* sA = smem + 64 * ld;                              // [64][ld]
  const int tid = threadIdx.x;
  load_matrix(sW, ld, p.w, 64, p.K, 64);
  const long long ntok = (long long)p.B * p.h * p.wd;
  const long long ntile = (ntok + 63) / 64;
  const int W2 = 2 * p.wd;
  for (long long tile = blockIdx.x; tile < ntile; tile += gridDim.x) {
    __syncthreads();
    const long long t0 = tile * 64;
    for (int idx = tid; idx < 64 * (p.K / 4); idx += kThreads) {
      const int i = idx / (p.K / 4), c4 = (idx - i * (p.K / 4)) * 4;
      const long long t = t0 + i;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (t < ntok) {
        const int xx = (int)(t % p.wd), yy = (int)((t / p.wd) % p.h);
        const long long b = t / ((long long)p.wd * p.h);
        const long long hi = ((b * 2 * p.h + 2 * yy) * W2 + 2 * xx) * 64;      // top-left source token
        const int part = c4 >> 6, c = c4 & 63;
        const bool direct = part == (p.srcB ? 2 : 1);
        if (direct) {
          v = ld4(p.direct + t * 64 + c);
        } else {
          const float* s = (part == 0 ? p.srcA : p.srcB) + hi + c;
          v = max4(max4(ld4(s), ld4(s + 64)), max4(ld4(s + (long long)W2 * 64), ld4(s + (long long)W2 * 64 + 64)));
          if (part == 0 && p.poolA) *reinterpret_cast<float4*>(p.poolA + t * 64 + c) = v;
        }
      }
      *reinterpret_cast<float4*>(sA + i * ld + c4) = v;
    }
    __syncthreads();
    float acc[4][4]; zero_acc(acc);
    gemm64<false, false>(acc, sA, ld, sW, ld, p.K);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const long long t = t0 + gemm_row<false>(i);
      if (t < ntok) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int n = gemm_col<false>(j);
          p.out[t * 64 + n] = acc[i][j] + __ldg(p.bias + n);
        }
      }
    }
  }
}

// ------------------------------------------------------------------ bottleneck: skip + shuffle + concat
__global__ void __launch_bounds__(256) bottleneck_cat_kernel(
    const float* __restrict__ c0, const float* __restrict__ p0, const float* __restrict__ p00,
    const float* __restrict__ e1, const float* __restrict__ e2, const float* __restrict__ e3,
    float* __restrict__ cat, int B, int H, int W) {
  const unsigned n = (unsigned)B * H * W * 21u;         // < 2^31 (checked by the launcher)
  const FastDiv fW = make_fastdiv((unsigned)W), fH = make_fastdiv((unsigned)H);
  for (unsigned idx = blockIdx.x * 256u + threadIdx.x; idx < n; idx += gridDim.x * 256u) {
    const unsigned t = idx / 21u;                       // (compile-time divisor: multiply-shift already)
    const int g = (int)(idx - t * 21u);
    float4 v;
    if (g < 16) {
      const float4 a = ld4(e1 + (size_t)t * 64 + g * 4), s = ld4(c0 + (size_t)t * 64 + g * 4);
      v = make_float4(a.x + leaky(s.x), a.y + leaky(s.y), a.z + leaky(s.z), a.w + leaky(s.w));
    } else {
      const unsigned rowi = fdiv(t, fW);
      const unsigned bq = fdiv(rowi, fH);
      const int xx = (int)(t - rowi * W), yy = (int)(rowi - bq * (unsigned)H);
      const long long b = bq;
      float o[4];
      if (g < 20) {                       // scale 2: channel c <- source channel (c*2 + y%2)*2 + x%2
        const long long lo = ((b * (H / 2) + yy / 2) * (W / 2) + xx / 2) * 64;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int k = (((g - 16) * 4 + j) * 2 + (yy & 1)) * 2 + (xx & 1);
          o[j] = __ldg(e2 + lo + k) + leaky(__ldg(p0 + lo + k));
        }
      } else {                            // scale 4: channel c <- source channel (c*4 + y%4)*4 + x%4
        const long long lo = ((b * (H / 4) + yy / 4) * (W / 4) + xx / 4) * 64;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int k = ((j * 4 + (yy & 3)) * 4) + (xx & 3);
          o[j] = __ldg(e3 + lo + k) + leaky(__ldg(p00 + lo + k));
        }
      }
      v = make_float4(o[0], o[1], o[2], o[3]);
    }
    *reinterpret_cast<float4*>(cat + (size_t)t * 84 + g * 4) = v;
  }
}

// ------------------------------------------------------------------ bottleneck tail
// out = LN(pointwise(GELU(depthwise3x3(cat) + b_dw)) + b_pw) + e1 ; one 8x8 spatial tile per pass
struct TailParams {
  const float* cat; const float* dw_w; const float* dw_b; const float* pw_w; const float* pw_b;
  const float* ln_w; const float* ln_b; const float* e1; float* out;
  int B, H, W;
};
struct TailSmem {
  static constexpr int kLdA = 100;                       // K = 84 padded to 88; stride == 4 mod 32
  static constexpr int oW = 0;                           // [64][100] pointwise weight
  static constexpr int oDw = oW + 64 * kLdA;             // [9][84] depthwise weight, tap-major
  static constexpr int oDb = oDw + 9 * 84;               // [84]
  static constexpr int oHalo = oDb + 84;                 // [100][84]
  static constexpr int oA = oHalo + 100 * 84;            // [64][100]
  static constexpr int oY = oA + 64 * kLdA;              // [64][68]
  static constexpr int kFloats = oY + 64 * kLd64;
};
__global__ void __launch_bounds__(kThreads, 1) bottleneck_tail_kernel(TailParams p) {
  using L = TailSmem;
  extern __shared__ __align__(16) float smem[];
  float* sW = smem + L::oW;
  float* sDw = smem + L::oDw;
  float* sDb = smem + L::oDb;
  float* sHalo = smem + L::oHalo;
  float* sA = smem + L::oA;
  float* sY = smem + L::oY;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int idx = tid; idx < 64 * L::kLdA; idx += kThreads) {
    const int n = idx / L::kLdA, k = idx - n * L::kLdA;
    sW[idx] = k < 84 ? p.pw_w[n * 84 + k] : 0.f;
  }
  for (int idx = tid; idx < 9 * 84; idx += kThreads) sDw[idx] = p.dw_w[(idx % 84) * 9 + idx / 84];
  for (int idx = tid; idx < 84; idx += kThreads) sDb[idx] = p.dw_b[idx];
  for (int idx = tid; idx < 64 * L::kLdA; idx += kThreads) sA[idx] = 0.f;      // K pad columns stay zero
  const int th = p.H / 8, tw = p.W / 8;
  const int ntile = p.B * th * tw;
  for (int tile = blockIdx.x; tile < ntile; tile += gridDim.x) {
    __syncthreads();
    const int tx0 = (tile % tw) * 8, ty0 = ((tile / tw) % th) * 8, b = tile / (tw * th);
    // ---- 10x10 halo of the 84-channel concat (zero outside the image: conv padding 1)
    for (int idx = tid; idx < 100 * 21; idx += kThreads) {
      const int hp = idx / 21, g = idx - hp * 21;
      const int yy = ty0 + hp / 10 - 1, xx = tx0 + hp % 10 - 1;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (yy >= 0 && yy < p.H && xx >= 0 && xx < p.W)
        v = ld4(p.cat + (((long long)b * p.H + yy) * p.W + xx) * 84 + g * 4);
      *reinterpret_cast<float4*>(sHalo + hp * 84 + g * 4) = v;
    }
    __syncthreads();
    // ---- depthwise 3x3 + GELU -> A tile [64 tokens][84]
    for (int idx = tid; idx < 64 * 84; idx += kThreads) {
      const int t = idx / 84, c = idx - t * 84;
      const int ty = t >> 3, tx = t & 7;
      float a = sDb[c];
#pragma unroll
      for (int dy = 0; dy < 3; ++dy)
#pragma unroll
        for (int dx = 0; dx < 3; ++dx)
          a = fmaf(sDw[(dy * 3 + dx) * 84 + c], sHalo[((ty + dy) * 10 + tx + dx) * 84 + c], a);
      sA[t * L::kLdA + c] = gelu_erf(a);
    }
    __syncthreads();
    // ---- pointwise 84 -> 64
    {
      float acc[4][4]; zero_acc(acc);
      gemm64<false, false>(acc, sA, L::kLdA, sW, L::kLdA, 88);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = gemm_col<false>(j);
        const float bias = __ldg(p.pw_b + n);
#pragma unroll
        for (int i = 0; i < 4; ++i) sY[gemm_row<false>(i) * kLd64 + n] = acc[i][j] + bias;
      }
    }
    __syncthreads();
    // ---- LayerNorm + e1, one warp per token row
#pragma unroll 1
    for (int rr = 0; rr < 8; ++rr) {
      const int m = warp * 8 + rr;
      const float y0 = sY[m * kLd64 + lane], y1 = sY[m * kLd64 + lane + 32];
      const float mean = warp_sum(y0 + y1) * (1.0f / 64.f);
      const float d0 = y0 - mean, d1 = y1 - mean;
      const float rstd = 1.0f / sqrtf(warp_sum(d0 * d0 + d1 * d1) * (1.0f / 64.f) + kLnEps);
      const long long tok = (((long long)b * p.H + ty0 + (m >> 3)) * p.W + tx0 + (m & 7)) * 64;
      p.out[tok + lane] = d0 * rstd * __ldg(p.ln_w + lane) + __ldg(p.ln_b + lane) + __ldg(p.e1 + tok + lane);
      p.out[tok + lane + 32] = d1 * rstd * __ldg(p.ln_w + lane + 32) + __ldg(p.ln_b + lane + 32) +
                               __ldg(p.e1 + tok + lane + 32);
    }
  }
}

// ------------------------------------------------------------------ final LN + skip + per-tap dots
// t = LN(dec) + c0 ;  taps[tok][k] = sum_c w1[c][k] * t[c]  (k = dy*3+dx of the 64->1 conv)
__global__ void __launch_bounds__(256) final_dot9_kernel(const float* __restrict__ dec,
                                                         const float* __restrict__ c0,
                                                         const float* __restrict__ ln_w,
                                                         const float* __restrict__ ln_b,
                                                         const float* __restrict__ w1,
                                                         float* __restrict__ taps, long long ntok) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = ((long long)blockIdx.x * 256 + threadIdx.x) >> 5;
  const long long nwarp = ((long long)gridDim.x * 256) >> 5;
  float wa[9], wb[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) { wa[k] = w1[lane * 9 + k]; wb[k] = w1[(lane + 32) * 9 + k]; }
  const float g0 = ln_w[lane], g1 = ln_w[lane + 32], b0 = ln_b[lane], b1 = ln_b[lane + 32];
  for (long long t = warp0; t < ntok; t += nwarp) {
    const float y0 = __ldg(dec + t * 64 + lane), y1 = __ldg(dec + t * 64 + lane + 32);
    const float mean = warp_sum(y0 + y1) * (1.0f / 64.f);
    const float d0 = y0 - mean, d1 = y1 - mean;
    const float rstd = 1.0f / sqrtf(warp_sum(d0 * d0 + d1 * d1) * (1.0f / 64.f) + kLnEps);
    const float t0 = d0 * rstd * g0 + b0 + __ldg(c0 + t * 64 + lane);
    const float t1 = d1 * rstd * g1 + b1 + __ldg(c0 + t * 64 + lane + 32);
    float mine = 0.f;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const float s = warp_sum(fmaf(t0, wa[k], t1 * wb[k]));
      if (lane == k) mine = s;
    }
    if (lane < 9) taps[t * 9 + lane] = mine;
  }
}

// TMA form of the same pass: 128-token tiles of dec and c0 arrive as 128 B-swizzled bulk tensor loads,
// thread t owns token row t (LayerNorm and the nine 64-long dots run in registers, no shuffles); the
// rows are pulled into registers first so that the next tile's loads overlap the arithmetic.
constexpr int kDotWorkers = 3;
struct Dot9Smem {
  static constexpr int oW = 0;                        // fp32 [64 c][12] (9 taps, padded)
  static constexpr int oLnW = oW + 64 * 12 * 4;
  static constexpr int oLnB = oLnW + 256;
  static constexpr int oBar = oLnB + 256;
  static constexpr int oWork = ((oBar + 64 + 1023) / 1024) * 1024;
  static constexpr int kWorkBytes = 2 * 32768;        // dec tile, c0 tile
  static constexpr int kBytes = oWork + kDotWorkers * kWorkBytes;
};
__global__ void __launch_bounds__(128 * kDotWorkers, 1)
final_dot9_tma_kernel(const __grid_constant__ CUtensorMap mapDec, const __grid_constant__ CUtensorMap mapC0,
                      const float* __restrict__ ln_w, const float* __restrict__ ln_b,
                      const float* __restrict__ w1, float* __restrict__ taps, long long ntok) {
  using L = Dot9Smem;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  float* sW = reinterpret_cast<float*>(smem_raw + L::oW);
  float* sLnW = reinterpret_cast<float*>(smem_raw + L::oLnW);
  float* sLnB = reinterpret_cast<float*>(smem_raw + L::oLnB);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + L::oBar);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int wk = warp >> 2, wq = warp & 3;
  unsigned char* sWk = smem_raw + L::oWork + wk * L::kWorkBytes;
  uint64_t* bar = bars + wk;
  for (int i = tid; i < 64 * 12; i += 128 * kDotWorkers) {
    const int c = i / 12, k = i - c * 12;
    sW[i] = k < 9 ? w1[c * 9 + k] : 0.f;
  }
  if (tid < 64) { sLnW[tid] = ln_w[tid]; sLnB[tid] = ln_b[tid]; }
  if (tid == 0) {
    for (int i = 0; i < kDotWorkers; ++i) tc::mbar_init(bars + i, 1);
    tc::fence_barrier_init();
    tma::prefetch_desc(&mapDec); tma::prefetch_desc(&mapC0);
  }
  tc::fence_async_smem();
  __syncthreads();
  const int row = wq * 32 + lane;
  const bool leader = wq == 0 && lane == 0;
  const long long ntile = (ntok + 127) / 128;
  const long long stride = (long long)gridDim.x * kDotWorkers;
  auto issue = [&](long long t_) {
    tma::mbar_expect_tx(bar, 65536);
    tma::load_2d(&mapDec, bar, sWk, 0, (int)(t_ * 128));
    tma::load_2d(&mapDec, bar, sWk + 16384, 32, (int)(t_ * 128));
    tma::load_2d(&mapC0, bar, sWk + 32768, 0, (int)(t_ * 128));
    tma::load_2d(&mapC0, bar, sWk + 49152, 32, (int)(t_ * 128));
  };
  long long tile = (long long)blockIdx.x * kDotWorkers + wk;
  if (leader && tile < ntile) issue(tile);
  uint32_t ph = 0;
  for (; tile < ntile; tile += stride) {
    tc::mbar_wait(bar, ph, nullptr); ph ^= 1;
    float y[64], c[64];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const float4 a = *reinterpret_cast<const float4*>(sWk + tma::tile_off(row, i));
      const float4 b = *reinterpret_cast<const float4*>(sWk + 32768 + tma::tile_off(row, i));
      y[4 * i] = a.x; y[4 * i + 1] = a.y; y[4 * i + 2] = a.z; y[4 * i + 3] = a.w;
      c[4 * i] = b.x; c[4 * i + 1] = b.y; c[4 * i + 2] = b.z; c[4 * i + 3] = b.w;
    }
    tc::worker_sync(1 + wk);                           // rows are in registers: refill the buffers
    if (leader && tile + stride < ntile) issue(tile + stride);
    float mean4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 64; ++i) mean4[i & 3] += y[i];
    const float mean = ((mean4[0] + mean4[1]) + (mean4[2] + mean4[3])) * (1.0f / 64.f);
    float var4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 64; ++i) { y[i] -= mean; var4[i & 3] = fmaf(y[i], y[i], var4[i & 3]); }
    const float var = (var4[0] + var4[1]) + (var4[2] + var4[3]);
    const float rstd = 1.0f / sqrtf(var * (1.0f / 64.f) + kLnEps);
    float d[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) d[k] = 0.f;
#pragma unroll
    for (int i = 0; i < 64; ++i) {
      const float t = fmaf(y[i] * rstd, sLnW[i], sLnB[i]) + c[i];
      const float4 wa = *reinterpret_cast<const float4*>(sW + i * 12);
      const float4 wb = *reinterpret_cast<const float4*>(sW + i * 12 + 4);
      const float wc = sW[i * 12 + 8];
      d[0] = fmaf(t, wa.x, d[0]); d[1] = fmaf(t, wa.y, d[1]); d[2] = fmaf(t, wa.z, d[2]);
      d[3] = fmaf(t, wa.w, d[3]); d[4] = fmaf(t, wb.x, d[4]); d[5] = fmaf(t, wb.y, d[5]);
      d[6] = fmaf(t, wb.z, d[6]); d[7] = fmaf(t, wb.w, d[7]); d[8] = fmaf(t, wc, d[8]);
    }
    const long long tok = tile * 128 + row;
    if (tok < ntok) {
#pragma unroll
      for (int k = 0; k < 9; ++k) taps[tok * 9 + k] = d[k];
    }
  }
}

// ------------------------------------------------------------------ reconstruction tail
// out1 = b1 + sum_k taps[y+dy-1, x+dx-1][k] (zero outside), out = tanh(conv3x3(out1; w2) + b2), crop
__global__ void __launch_bounds__(256) recon_kernel(const float* __restrict__ taps,
                                                    const float* __restrict__ b1,
                                                    const float* __restrict__ w2,
                                                    const float* __restrict__ b2,
                                                    float* __restrict__ out, int B, int H, int W,
                                                    int H0, int W0) {
  __shared__ float s1[10][34];                            // out1 on the 8x32 tile + 1-pixel halo
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int tilesx = (W + 31) / 32, tilesy = (H + 7) / 8;
  const int ntile = B * tilesx * tilesy;
  float w[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) w[k] = w2[k];
  const float bias1 = b1[0], bias2 = b2[0];
  for (int tile = blockIdx.x; tile < ntile; tile += gridDim.x) {
    const int x0 = (tile % tilesx) * 32, y0 = ((tile / tilesx) % tilesy) * 8, b = tile / (tilesx * tilesy);
    const float* tp = taps + (long long)b * H * W * 9;
    __syncthreads();
    for (int idx = threadIdx.x; idx < 10 * 34; idx += 256) {
      const int hy = idx / 34, hx = idx - hy * 34;
      const int yy = y0 + hy - 1, xx = x0 + hx - 1;
      float v = 0.f;                                      // outside the image: second conv's zero padding
      if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
        v = bias1;
#pragma unroll
        for (int dy = 0; dy < 3; ++dy)
#pragma unroll
          for (int dx = 0; dx < 3; ++dx) {
            const int y2 = yy + dy - 1, x2 = xx + dx - 1;
            if (y2 >= 0 && y2 < H && x2 >= 0 && x2 < W) v += __ldg(tp + ((long long)y2 * W + x2) * 9 + dy * 3 + dx);
          }
      }
      s1[hy][hx] = v;
    }
    __syncthreads();
    const int yy = y0 + ty, xx = x0 + tx;
    if (yy < H0 && xx < W0) {
      float a = bias2;
#pragma unroll
      for (int dy = 0; dy < 3; ++dy)
#pragma unroll
        for (int dx = 0; dx < 3; ++dx) a = fmaf(w[dy * 3 + dx], s1[ty + dy][tx + dx], a);
      out[((long long)b * H0 + yy) * W0 + xx] = tanhf(a);
    }
  }
}

static int grid_for(long long work_items, int per_block) {
  long long g = (work_items + per_block - 1) / per_block;
  const long long cap = (long long)sm_count() * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace b200mar

using namespace b200mar;

extern "C" int b200mar_shallow_tok_fwd(const float* x, const float* w, const float* b, float* c0,
                                       int B, int H, int W, void* stream) {
  if (B <= 0 || H <= 0 || W <= 0 || (long long)B * H * W >= (1ll << 31)) return -1;
  shallow_tok_kernel<<<grid_for((long long)B * H * ((W + 3) / 4), 16 * 4), 256, 0, (cudaStream_t)stream>>>(x, w, b, c0, B, H, W);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_cascade_proj_fwd(const float* srcA, const float* srcB, const float* direct,
                                        const float* w, const float* bias, float* out, float* poolA,
                                        int B, int h, int wd, void* stream) {
  if (B <= 0 || h <= 0 || wd <= 0 || !srcA || !direct) return -1;
  CascadeParams p{srcA, srcB, direct, w, bias, out, poolA, B, h, wd, srcB ? 192 : 128};
  const size_t smem = (size_t)2 * 64 * (p.K + 4) * sizeof(float);
  B200MAR_CHECK(cudaFuncSetAttribute(cascade_proj_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 64 * 196 * 4));
  const long long ntile = ((long long)B * h * wd + 63) / 64;
  const int grid = (int)(ntile < 2LL * sm_count() ? ntile : 2LL * sm_count());
  cascade_proj_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_bottleneck_cat_fwd(const float* c0, const float* p0, const float* p00,
                                          const float* e1, const float* e2, const float* e3,
                                          float* cat, int B, int H, int W, void* stream) {
  if (B <= 0 || H % 4 || W % 4 || (long long)B * H * W * 21 >= (1ll << 31)) return -1;
  bottleneck_cat_kernel<<<grid_for((long long)B * H * W * 21, 256 * 4), 256, 0, (cudaStream_t)stream>>>(
      c0, p0, p00, e1, e2, e3, cat, B, H, W);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_bottleneck_tail_fwd(const float* cat, const float* dw_w, const float* dw_b,
                                           const float* pw_w, const float* pw_b, const float* ln_w,
                                           const float* ln_b, const float* e1, float* out,
                                           int B, int H, int W, void* stream) {
  if (B <= 0 || H % 8 || W % 8) return -1;
  TailParams p{cat, dw_w, dw_b, pw_w, pw_b, ln_w, ln_b, e1, out, B, H, W};
  const size_t smem = TailSmem::kFloats * sizeof(float);
  B200MAR_CHECK(cudaFuncSetAttribute(bottleneck_tail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int ntile = B * (H / 8) * (W / 8);
  const int grid = ntile < 2 * sm_count() ? ntile : 2 * sm_count();
  bottleneck_tail_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_final_recon_fwd(const float* dec, const float* c0, const float* ln_w,
                                       const float* ln_b, const float* w1, const float* b1,
                                       const float* w2, const float* b2, float* taps, float* out,
                                       int B, int H, int W, int H0, int W0, void* stream) {
  if (B <= 0 || H <= 0 || W <= 0 || H0 > H || W0 > W) return -1;
  const long long ntok = (long long)B * H * W;
  {
    CUtensorMap mD, mC;
    if (int rc = tma::encode_rows64(&mD, dec, ntok)) return rc;
    if (int rc = tma::encode_rows64(&mC, c0, ntok)) return rc;
    static DeviceOnce configured;
    if (configured.first()) {
      B200MAR_CHECK(cudaFuncSetAttribute(final_dot9_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)Dot9Smem::kBytes));
      configured.done();
    }
    const long long ntile9 = (ntok + 127) / 128;
    long long grid = (ntile9 + kDotWorkers - 1) / kDotWorkers;
    if (grid > sm_count()) grid = sm_count();
    final_dot9_tma_kernel<<<(int)grid, 128 * kDotWorkers, Dot9Smem::kBytes, (cudaStream_t)stream>>>(
        mD, mC, ln_w, ln_b, w1, taps, ntok);
    B200MAR_CHECK(cudaGetLastError());
  }
  const int ntile = B * ((W + 31) / 32) * ((H + 7) / 8);
  recon_kernel<<<grid_for(ntile, 1), 256, 0, (cudaStream_t)stream>>>(taps, b1, w2, b2, out, B, H, W, H0, W0);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
