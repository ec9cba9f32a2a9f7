// glue_bwd.cu — backward twins of the fused generator glue of glue_fwd.cu (the training
// half): with them the TRAINING forward takes the fused channels-last glue path too and the ~140 stock PyTorch /
// cuDNN launches of a step (layout flips, pools, small GEMMs, convolutions and their backward) disappear.
// fp32 FMA-pipe kernels: the glue is 2.8 % of the model's FLOPs and memory-bound; every parameter gradient is
// ACCUMULATED with atomics (the convention of every backward entry point of this library), activations' gradients
// are written.
//
//   final_recon_bwd   tanh, conv3x3 1->1, conv3x3 64->1 (per-tap form), global skip, LayerNorm(norm)
//                     (reconstruction.py:14-19, ngswin.py:190,204,208)
//   shallow_tok_bwd   ShallowExtractor conv3x3 1->64 (encoder_decoder.py:23-27)
//   cascade_proj_bwd  InterPool (2x2 max, first-maximum routing like nn.MaxPool2d) + concat + Linear
//                     (pooling.py:13-17, ngswin.py:175-178, encoder_decoder.py:117,136)
//   bottleneck_bwd    LN + Linear 84->64 + GELU + depthwise 3x3 + pixel-shuffle / concat / LeakyReLU skips
//                     (encoder_decoder.py:159-164,188-200, pooling.py:32-37)
#include "common.cuh"
#include "../../include/b200mar.h"

namespace b200mar {

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

static int grid_cap(long long blocks, int per_sm) {
  const long long cap = (long long)sm_count() * per_sm;
  return (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}

// ================================================================================================ final + recon
// Forward (glue_fwd.cu): v = LN(dec) + c0; taps[t][k] = sum_c w1[c][k] v[t][c]; o1[p] = b1 + sum_k taps[p + off_k][k]
// (off_k = (k/3 - 1, k%3 - 1), zero outside the image); o2[p] = b2 + sum_k w2[k] o1[p + off_k]; y = tanh(o2), cropped.
//
// recon_bwd_kernel (image space, one thread per pixel of the PADDED H x W grid):
//   g2[p] = dy[p] (1 - y[p]^2) inside the crop, else 0
//   g1[p] = sum_k w2[k] g2[p - off_k]                       -> written, [B, H, W]
//   dw2[k] += o1[p] g2[p - off_k],  db2 += g2[p],  db1 += g1[p]
__global__ void __launch_bounds__(256) recon_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                                                        const float* __restrict__ taps, const float* __restrict__ b1,
                                                        const float* __restrict__ w2, float* __restrict__ g1,
                                                        float* __restrict__ dw2, float* __restrict__ db2,
                                                        float* __restrict__ db1, int B, int H, int W, int H0, int W0) {
  __shared__ float red[8][12];
  float w[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) w[k] = w2[k];
  const float bias1 = b1[0];
  float acc[11];
#pragma unroll
  for (int i = 0; i < 11; ++i) acc[i] = 0.f;
  const long long npix = (long long)B * H * W;
  for (long long p = blockIdx.x * 256ll + threadIdx.x; p < npix; p += (long long)gridDim.x * 256) {
    const int xx = (int)(p % W), yy = (int)((p / W) % H);
    const long long b = p / ((long long)W * H);
    const float* dyb = dy + b * H0 * W0;
    const float* yb = y + b * H0 * W0;
    const float* tp = taps + b * H * W * 9;
    float o1 = bias1;                                       // o1 at this pixel (inside the image by construction)
    float gsum = 0.f;
    float g2n[9];
#pragma unroll
    for (int dyi = 0; dyi < 3; ++dyi)
#pragma unroll
      for (int dxi = 0; dxi < 3; ++dxi) {
        const int k = dyi * 3 + dxi;
        const int y2 = yy + dyi - 1, x2 = xx + dxi - 1;      // p + off_k
        if (y2 >= 0 && y2 < H && x2 >= 0 && x2 < W) o1 += __ldg(tp + ((long long)y2 * W + x2) * 9 + k);
        const int y3 = yy - (dyi - 1), x3 = xx - (dxi - 1);  // p - off_k
        float g = 0.f;
        if (y3 >= 0 && y3 < H0 && x3 >= 0 && x3 < W0) {
          const float yo = __ldg(yb + (long long)y3 * W0 + x3);
          g = __ldg(dyb + (long long)y3 * W0 + x3) * (1.0f - yo * yo);
        }
        g2n[k] = g;
        gsum = fmaf(w[k], g, gsum);
      }
    g1[p] = gsum;
#pragma unroll
    for (int k = 0; k < 9; ++k) acc[k] = fmaf(o1, g2n[k], acc[k]);
    acc[9] += g2n[4];                                       // off_4 = (0,0): g2 at this pixel
    acc[10] += gsum;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < 11; ++i) {
    const float s = warp_sum(acc[i]);
    if (lane == 0) red[warp][i] = s;
  }
  __syncthreads();
  if (threadIdx.x < 11) {
    float s = 0.f;
#pragma unroll
    for (int wv = 0; wv < 8; ++wv) s += red[wv][threadIdx.x];
    if (threadIdx.x < 9) atomicAdd(dw2 + threadIdx.x, s);
    else if (threadIdx.x == 9) atomicAdd(db2, s);
    else atomicAdd(db1, s);
  }
}

// final_bwd_kernel (token space, one warp per token, lane owns channels lane and lane + 32):
//   dtaps[t][k] = g1[t - off_k] (0 outside the image);  dv[c] = sum_k w1[c][k] dtaps[k]  -> dc0 (written)
//   dw1[c][k] += v[c] dtaps[k];  LayerNorm backward of dv -> ddec (written), d_ln_w += dv xhat, d_ln_b += dv
__global__ void __launch_bounds__(256) final_bwd_kernel(const float* __restrict__ dec, const float* __restrict__ c0,
                                                        const float* __restrict__ ln_w, const float* __restrict__ ln_b,
                                                        const float* __restrict__ w1, const float* __restrict__ g1,
                                                        float* __restrict__ ddec, float* __restrict__ dc0,
                                                        float* __restrict__ dw1, float* __restrict__ dlnw,
                                                        float* __restrict__ dlnb, int B, int H, int W) {
  __shared__ float red[8][64][12];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float wa[9], wb[9], aa[11], ab[11];
#pragma unroll
  for (int k = 0; k < 9; ++k) { wa[k] = w1[lane * 9 + k]; wb[k] = w1[(lane + 32) * 9 + k]; }
#pragma unroll
  for (int i = 0; i < 11; ++i) { aa[i] = 0.f; ab[i] = 0.f; }
  const float ga = ln_w[lane], gb = ln_w[lane + 32], ba = ln_b[lane], bb = ln_b[lane + 32];
  const long long ntok = (long long)B * H * W;
  const long long warp0 = ((long long)blockIdx.x * 256 + threadIdx.x) >> 5;
  const long long nwarp = ((long long)gridDim.x * 256) >> 5;
  for (long long t = warp0; t < ntok; t += nwarp) {
    const int xx = (int)(t % W), yy = (int)((t / W) % H);
    const float* gimg = g1 + (t - (long long)yy * W - xx);
    // lane k < 9 fetches dtaps[k]; broadcast by shuffle
    float mine = 0.f;
    if (lane < 9) {
      const int y3 = yy - (lane / 3 - 1), x3 = xx - (lane % 3 - 1);
      if (y3 >= 0 && y3 < H && x3 >= 0 && x3 < W) mine = __ldg(gimg + (long long)y3 * W + x3);
    }
    const float y0 = __ldg(dec + t * 64 + lane), y1 = __ldg(dec + t * 64 + lane + 32);
    const float mean = warp_sum(y0 + y1) * (1.0f / 64.f);
    const float d0 = y0 - mean, d1 = y1 - mean;
    const float rstd = 1.0f / sqrtf(warp_sum(d0 * d0 + d1 * d1) * (1.0f / 64.f) + kLnEps);
    const float xh0 = d0 * rstd, xh1 = d1 * rstd;
    const float v0 = xh0 * ga + ba + __ldg(c0 + t * 64 + lane);
    const float v1 = xh1 * gb + bb + __ldg(c0 + t * 64 + lane + 32);
    float dv0 = 0.f, dv1 = 0.f;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const float dt = __shfl_sync(0xffffffffu, mine, k);
      dv0 = fmaf(wa[k], dt, dv0);
      dv1 = fmaf(wb[k], dt, dv1);
      aa[k] = fmaf(v0, dt, aa[k]);
      ab[k] = fmaf(v1, dt, ab[k]);
    }
    dc0[t * 64 + lane] = dv0;
    dc0[t * 64 + lane + 32] = dv1;
    aa[9] = fmaf(dv0, xh0, aa[9]); ab[9] = fmaf(dv1, xh1, ab[9]);     // d ln_w
    aa[10] += dv0; ab[10] += dv1;                                     // d ln_b
    const float dx0 = dv0 * ga, dx1 = dv1 * gb;
    const float m1 = warp_sum(dx0 + dx1) * (1.0f / 64.f);
    const float m2 = warp_sum(dx0 * xh0 + dx1 * xh1) * (1.0f / 64.f);
    ddec[t * 64 + lane] = rstd * (dx0 - m1 - xh0 * m2);
    ddec[t * 64 + lane + 32] = rstd * (dx1 - m1 - xh1 * m2);
  }
#pragma unroll
  for (int i = 0; i < 11; ++i) { red[warp][lane][i] = aa[i]; red[warp][lane + 32][i] = ab[i]; }
  __syncthreads();
  for (int idx = threadIdx.x; idx < 64 * 11; idx += 256) {
    const int c = idx / 11, i = idx - c * 11;
    float s = 0.f;
#pragma unroll
    for (int wv = 0; wv < 8; ++wv) s += red[wv][c][i];
    if (i < 9) atomicAdd(dw1 + c * 9 + i, s);
    else if (i == 9) atomicAdd(dlnw + c, s);
    else atomicAdd(dlnb + c, s);
  }
}

// ================================================================================================ shallow conv
// c0[t][c] = b[c] + sum_k w[c][k] x[t + off_k]  ->  dw[c][k] += sum_t dc0[t][c] x[t + off_k], db[c] += sum_t dc0[t][c]
// thread = (4 channels, one of 16 pixel lanes) as in the forward; 36 + 4 accumulators per thread.
__global__ void __launch_bounds__(256) shallow_tok_wgrad_kernel(const float* __restrict__ x,
                                                                const float* __restrict__ dc0,
                                                                float* __restrict__ dw, float* __restrict__ db,
                                                                int B, int H, int W) {
  __shared__ float red[16][64][10];
  const int cq = threadIdx.x & 15, pl = threadIdx.x >> 4;
  float acc[4][10];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int k = 0; k < 10; ++k) acc[j][k] = 0.f;
  const unsigned npix = (unsigned)B * H * W;
  for (unsigned pix = blockIdx.x * 16u + pl; pix < npix; pix += gridDim.x * 16u) {
    const unsigned rowi = pix / (unsigned)W;
    const int xx = (int)(pix - rowi * W), yy = (int)(rowi % (unsigned)H);
    const float* img = x + (pix - (unsigned)yy * W - xx);
    float in[9];
#pragma unroll
    for (int dy = 0; dy < 3; ++dy)
#pragma unroll
      for (int dx = 0; dx < 3; ++dx) {
        const int y2 = yy + dy - 1, x2 = xx + dx - 1;
        in[dy * 3 + dx] = (y2 >= 0 && y2 < H && x2 >= 0 && x2 < W) ? __ldg(img + (long long)y2 * W + x2) : 0.f;
      }
    const float4 g = ldg4(dc0 + (size_t)pix * 64 + cq * 4);
    const float gv[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
      for (int k = 0; k < 9; ++k) acc[j][k] = fmaf(gv[j], in[k], acc[j][k]);
      acc[j][9] += gv[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int k = 0; k < 10; ++k) red[pl][cq * 4 + j][k] = acc[j][k];
  __syncthreads();
  for (int idx = threadIdx.x; idx < 640; idx += 256) {
    const int c = idx / 10, k = idx - c * 10;
    float s = 0.f;
#pragma unroll
    for (int l = 0; l < 16; ++l) s += red[l][c][k];
    if (k < 9) atomicAdd(dw + c * 9 + k, s); else atomicAdd(db + c, s);
  }
}

// dx[p] = sum_k sum_c dc0[p - off_k][c] w[c][k]; one warp per pixel, lane owns channels lane, lane + 32
__global__ void __launch_bounds__(256) shallow_tok_dgrad_kernel(const float* __restrict__ dc0,
                                                                const float* __restrict__ w,
                                                                float* __restrict__ dx, int B, int H, int W) {
  const int lane = threadIdx.x & 31;
  float wa[9], wb[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) { wa[k] = w[lane * 9 + k]; wb[k] = w[(lane + 32) * 9 + k]; }
  const long long npix = (long long)B * H * W;
  const long long warp0 = ((long long)blockIdx.x * 256 + threadIdx.x) >> 5;
  const long long nwarp = ((long long)gridDim.x * 256) >> 5;
  for (long long p = warp0; p < npix; p += nwarp) {
    const int xx = (int)(p % W), yy = (int)((p / W) % H);
    const long long base = p - (long long)yy * W - xx;
    float s = 0.f;
#pragma unroll
    for (int dy = 0; dy < 3; ++dy)
#pragma unroll
      for (int dxi = 0; dxi < 3; ++dxi) {
        const int y3 = yy - (dy - 1), x3 = xx - (dxi - 1);
        if (y3 >= 0 && y3 < H && x3 >= 0 && x3 < W) {
          const float* g = dc0 + (base + (long long)y3 * W + x3) * 64;
          s = fmaf(__ldg(g + lane), wa[dy * 3 + dxi], s);
          s = fmaf(__ldg(g + lane + 32), wb[dy * 3 + dxi], s);
        }
      }
    s = warp_sum(s);
    if (lane == 0) dx[p] = s;
  }
}

// ================================================================================================ cascade proj
// Forward: out[t] = W [64, K] cat(pool2x2(srcA)[t], pool2x2(srcB)[t] (optional), direct[t]) + bias, poolA = pool2x2(srcA).
// Backward per tile of 64 output tokens: the concat tile is recomputed (pooling included), then
//   dcat = dout W            (three 64x64x64 tile products)  -> d_direct written; pooled parts routed to the FIRST
//                             maximum of each 2x2 neighbourhood (nn.MaxPool2d semantics, pooling.py:13-17); the
//                             gradient arriving on the returned poolA (dpool, optional) joins part 0
//   dW += dout^T cat, db += column sums of dout   (per-CTA partials in registers, one atomic flush)
struct CascadeBwdParams {
  const float* srcA; const float* srcB; const float* direct; const float* w;
  const float* dout; const float* dpool;
  float* d_srcA; float* d_srcB; float* d_direct; float* dw; float* db;
  int B, h, wd, K;
};
__global__ void __launch_bounds__(kThreads, 1) cascade_proj_bwd_kernel(CascadeBwdParams p) {
  extern __shared__ __align__(16) float smem[];
  const int ld = p.K + 4;
  float* sW = smem;                       // [64 n][ld]
  float* sA = sW + 64 * ld;               // [64 tokens][ld]   the concat tile
  float* sD = sA + 64 * ld;               // [64 tokens][68]   dout tile
  float* sG = sD + 64 * kLd64;            // [64 tokens][68]   one 64-column chunk of dcat
  const int tid = threadIdx.x;
  load_matrix(sW, ld, p.w, 64, p.K, 64);
  const long long ntok = (long long)p.B * p.h * p.wd;
  const long long ntile = (ntok + 63) / 64;
  const int W2 = 2 * p.wd, nchunk = p.K / 64, direct_part = nchunk - 1;
  float accW[3][4][4];
#pragma unroll
  for (int q = 0; q < 3; ++q) zero_acc(accW[q]);
  float accB = 0.f;
  for (long long tile = blockIdx.x; tile < ntile; tile += gridDim.x) {
    __syncthreads();
    const long long t0 = tile * 64;
    for (int idx = tid; idx < 64 * (p.K / 4); idx += kThreads) {
      const int i = idx / (p.K / 4), c4 = (idx - i * (p.K / 4)) * 4;
      const long long t = t0 + i;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (t < ntok) {
        const int part = c4 >> 6, c = c4 & 63;
        if (part == direct_part) {
          v = ldg4(p.direct + t * 64 + c);
        } else {
          const int xx = (int)(t % p.wd), yy = (int)((t / p.wd) % p.h);
          const long long b = t / ((long long)p.wd * p.h);
          const float* s = (part == 0 ? p.srcA : p.srcB) + ((b * 2 * p.h + 2 * yy) * W2 + 2 * xx) * 64 + c;
          const float4 a0 = ldg4(s), a1 = ldg4(s + 64), a2 = ldg4(s + (long long)W2 * 64), a3 = ldg4(s + (long long)W2 * 64 + 64);
          v = make_float4(fmaxf(fmaxf(a0.x, a1.x), fmaxf(a2.x, a3.x)), fmaxf(fmaxf(a0.y, a1.y), fmaxf(a2.y, a3.y)),
                          fmaxf(fmaxf(a0.z, a1.z), fmaxf(a2.z, a3.z)), fmaxf(fmaxf(a0.w, a1.w), fmaxf(a2.w, a3.w)));
        }
      }
      *reinterpret_cast<float4*>(sA + i * ld + c4) = v;
    }
    for (int idx = tid; idx < 64 * 16; idx += kThreads) {
      const int i = idx >> 4, c4 = (idx & 15) * 4;
      const long long t = t0 + i;
      *reinterpret_cast<float4*>(sD + i * kLd64 + c4) = t < ntok ? ldg4(p.dout + t * 64 + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncthreads();
    if (tid < 64) {
      float s = 0.f;
#pragma unroll 8
      for (int i = 0; i < 64; ++i) s += sD[i * kLd64 + tid];
      accB += s;
    }
#pragma unroll 1
    for (int q = 0; q < nchunk; ++q) {
      // dW[n][q*64 + k] += sum_i dout[i][n] cat[i][q*64 + k]
      gemm64<true, true>(accW[q], sD, kLd64, sA + q * 64, ld, 64);
      // dcat[i][q*64 + k] = sum_n dout[i][n] W[n][q*64 + k]
      float acc[4][4]; zero_acc(acc);
      gemm64<false, true>(acc, sD, kLd64, sW + q * 64, ld, 64);
      if (q == direct_part) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const long long t = t0 + gemm_row<false>(i);
          if (t < ntok)
            *reinterpret_cast<float4*>(p.d_direct + t * 64 + gemm_col<true>(0)) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        }
      } else {
        __syncthreads();                  // sG free (previous chunk's routing pass done)
#pragma unroll
        for (int i = 0; i < 4; ++i)
          *reinterpret_cast<float4*>(sG + gemm_row<false>(i) * kLd64 + gemm_col<true>(0)) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        __syncthreads();
        const float* src = q == 0 ? p.srcA : p.srcB;
        float* dsrc = q == 0 ? p.d_srcA : p.d_srcB;
        for (int idx = tid; idx < 64 * 16; idx += kThreads) {
          const int i = idx >> 4, c = (idx & 15) * 4;
          const long long t = t0 + i;
          if (t >= ntok) continue;
          float4 g = *reinterpret_cast<const float4*>(sG + i * kLd64 + c);
          if (q == 0 && p.dpool) {
            const float4 e = ldg4(p.dpool + t * 64 + c);
            g.x += e.x; g.y += e.y; g.z += e.z; g.w += e.w;
          }
          const int xx = (int)(t % p.wd), yy = (int)((t / p.wd) % p.h);
          const long long b = t / ((long long)p.wd * p.h);
          const long long o = ((b * 2 * p.h + 2 * yy) * W2 + 2 * xx) * 64 + c;
          const long long offs[4] = {0, 64, (long long)W2 * 64, (long long)W2 * 64 + 64};
          float4 a[4];
#pragma unroll
          for (int s4 = 0; s4 < 4; ++s4) a[s4] = ldg4(src + o + offs[s4]);
          const float gv[4] = {g.x, g.y, g.z, g.w};
          float r[4][4];
#pragma unroll
          for (int ch = 0; ch < 4; ++ch) {
            const float v0 = (&a[0].x)[ch], v1 = (&a[1].x)[ch], v2 = (&a[2].x)[ch], v3 = (&a[3].x)[ch];
            int best = 0; float m = v0;
            if (v1 > m) { m = v1; best = 1; }
            if (v2 > m) { m = v2; best = 2; }
            if (v3 > m) { m = v3; best = 3; }
#pragma unroll
            for (int s4 = 0; s4 < 4; ++s4) r[s4][ch] = best == s4 ? gv[ch] : 0.f;
          }
#pragma unroll
          for (int s4 = 0; s4 < 4; ++s4)
            *reinterpret_cast<float4*>(dsrc + o + offs[s4]) = make_float4(r[s4][0], r[s4][1], r[s4][2], r[s4][3]);
        }
      }
    }
  }
  // flush the per-CTA partials
  for (int q = 0; q < nchunk; ++q)
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
        atomicAdd(p.dw + (size_t)gemm_row<true>(i) * p.K + q * 64 + gemm_col<true>(j), accW[q][i][j]);
  if (tid < 64) atomicAdd(p.db + tid, accB);
}

// ================================================================================================ bottleneck
// Forward: cat [T, 84] = concat(e1 + leaky(c0), shuffle2(e2 + leaky(p0)), shuffle4(e3 + leaky(p00)));
//          pre = depthwise3x3(cat) + b_dw; g = GELU(pre); y = W_pw g + b_pw; out = LN(y) + e1.
// bottleneck_tail_bwd_kernel (one 8x8 spatial tile per pass, everything recomputed from the saved cat):
//   LayerNorm backward of dout -> dy; d_ln_w += dout xhat, d_ln_b += dout; d_pw_b += dy; d_pw_w += dy^T g;
//   ddwc = (dy W_pw) * GELU'(pre)  -> written to a [T, 84] scratch for the depthwise pass
struct TailBwdSmem {
  // K = 84: the 64-column chunk 1 of a row reads columns 64..127, i.e. up to 28 floats into the NEXT row (stride 100)
  // - harmless: those products land in output columns k >= 100, which are never used; kSlack keeps the last row's
  // over-read inside the allocation.  105 KB in all: two CTAs per SM.
  static constexpr int kLd = 100, kSlack = 32;
  static constexpr int oW = 0;                           // [64 n][100] pointwise weight (k >= 84 zero)
  static constexpr int oDw = oW + 64 * kLd + kSlack;     // [9][84] depthwise weight, tap-major
  static constexpr int oDb = oDw + 9 * 84;               // [84]
  static constexpr int oHalo = oDb + 84;                 // [100][84] cat halo; ALIASED by the pre-activation tile
  static constexpr int oPre = oHalo;                     // [64][100] pre-activation (written after the halo is dead)
  static constexpr int oG = oHalo + 100 * 84;            // [64][100] GELU(pre)
  static constexpr int oY = oG + 64 * kLd + kSlack;      // [64][68]  y, then dy
  static constexpr int kFloats = oY + 64 * kLd64;
  static_assert(100 * 84 >= 64 * kLd + kSlack, "the pre-activation tile must fit in the halo buffer");
};
struct TailBwdParams {
  const float* cat; const float* dw_w; const float* dw_b; const float* pw_w; const float* pw_b;
  const float* ln_w; const float* dout;
  float* ddwc; float* d_pw_w; float* d_pw_b; float* d_ln_w; float* d_ln_b;
  int B, H, W;
};
__global__ void __launch_bounds__(kThreads, 2) bottleneck_tail_bwd_kernel(TailBwdParams p) {
  using L = TailBwdSmem;
  extern __shared__ __align__(16) float smem[];
  float* sW = smem + L::oW;
  float* sDw = smem + L::oDw;
  float* sDb = smem + L::oDb;
  float* sHalo = smem + L::oHalo;
  float* sPre = smem + L::oPre;
  float* sG = smem + L::oG;
  float* sY = smem + L::oY;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int idx = tid; idx < 64 * L::kLd + L::kSlack; idx += kThreads) {
    const int n = idx / L::kLd, k = idx - n * L::kLd;
    sW[idx] = (n < 64 && k < 84) ? p.pw_w[n * 84 + k] : 0.f;
    sG[idx] = 0.f;                                       // pad columns stay zero
  }
  for (int idx = tid; idx < 9 * 84; idx += kThreads) sDw[idx] = p.dw_w[(idx % 84) * 9 + idx / 84];
  for (int idx = tid; idx < 84; idx += kThreads) sDb[idx] = p.dw_b[idx];
  float accW[2][4][4];
  zero_acc(accW[0]); zero_acc(accW[1]);
  float accB = 0.f, aLw0 = 0.f, aLw1 = 0.f, aLb0 = 0.f, aLb1 = 0.f;
  const float g0 = p.ln_w[lane], g1 = p.ln_w[lane + 32];
  const int th = p.H / 8, tw = p.W / 8;
  const int ntile = p.B * th * tw;
  for (int tile = blockIdx.x; tile < ntile; tile += gridDim.x) {
    __syncthreads();
    const int tx0 = (tile % tw) * 8, ty0 = ((tile / tw) % th) * 8, b = tile / (tw * th);
    for (int idx = tid; idx < 100 * 21; idx += kThreads) {
      const int hp = idx / 21, g = idx - hp * 21;
      const int yy = ty0 + hp / 10 - 1, xx = tx0 + hp % 10 - 1;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (yy >= 0 && yy < p.H && xx >= 0 && xx < p.W)
        v = ldg4(p.cat + (((long long)b * p.H + yy) * p.W + xx) * 84 + g * 4);
      *reinterpret_cast<float4*>(sHalo + hp * 84 + g * 4) = v;
    }
    __syncthreads();
    float pre[21];                                        // 64 * 84 / 256 elements per thread, staged in registers:
#pragma unroll                                            // the pre-activation tile overwrites the halo buffer
    for (int e = 0; e < 21; ++e) {
      const int idx = tid + e * kThreads;
      const int t = idx / 84, c = idx - t * 84;
      const int ty = t >> 3, tx = t & 7;
      float a = sDb[c];
#pragma unroll
      for (int dy = 0; dy < 3; ++dy)
#pragma unroll
        for (int dx = 0; dx < 3; ++dx)
          a = fmaf(sDw[(dy * 3 + dx) * 84 + c], sHalo[((ty + dy) * 10 + tx + dx) * 84 + c], a);
      pre[e] = a;
    }
    __syncthreads();                                      // every halo read is done
#pragma unroll
    for (int e = 0; e < 21; ++e) {
      const int idx = tid + e * kThreads;
      const int t = idx / 84, c = idx - t * 84;
      sPre[t * L::kLd + c] = pre[e];
      sG[t * L::kLd + c] = gelu_erf(pre[e]);
    }
    __syncthreads();
    {                                                     // y = W_pw g + b_pw
      float acc[4][4]; zero_acc(acc);
      gemm64<false, false>(acc, sG, L::kLd, sW, L::kLd, 88);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = gemm_col<false>(j);
        const float bias = __ldg(p.pw_b + n);
#pragma unroll
        for (int i = 0; i < 4; ++i) sY[gemm_row<false>(i) * kLd64 + n] = acc[i][j] + bias;
      }
    }
    __syncthreads();
    // LayerNorm backward, one warp per token row; dy overwrites y in place
#pragma unroll 1
    for (int rr = 0; rr < 8; ++rr) {
      const int m = warp * 8 + rr;
      const float y0 = sY[m * kLd64 + lane], y1 = sY[m * kLd64 + lane + 32];
      const float mean = warp_sum(y0 + y1) * (1.0f / 64.f);
      const float d0 = y0 - mean, d1 = y1 - mean;
      const float rstd = 1.0f / sqrtf(warp_sum(d0 * d0 + d1 * d1) * (1.0f / 64.f) + kLnEps);
      const float xh0 = d0 * rstd, xh1 = d1 * rstd;
      const long long tok = (((long long)b * p.H + ty0 + (m >> 3)) * p.W + tx0 + (m & 7)) * 64;
      const float o0 = __ldg(p.dout + tok + lane), o1 = __ldg(p.dout + tok + lane + 32);
      aLw0 = fmaf(o0, xh0, aLw0); aLw1 = fmaf(o1, xh1, aLw1); aLb0 += o0; aLb1 += o1;
      const float dx0 = o0 * g0, dx1 = o1 * g1;
      const float m1 = warp_sum(dx0 + dx1) * (1.0f / 64.f);
      const float m2 = warp_sum(dx0 * xh0 + dx1 * xh1) * (1.0f / 64.f);
      sY[m * kLd64 + lane] = rstd * (dx0 - m1 - xh0 * m2);
      sY[m * kLd64 + lane + 32] = rstd * (dx1 - m1 - xh1 * m2);
    }
    __syncthreads();
    if (tid < 64) {
      float s2 = 0.f;
#pragma unroll 8
      for (int i = 0; i < 64; ++i) s2 += sY[i * kLd64 + tid];
      accB += s2;
    }
#pragma unroll 1
    for (int q = 0; q < 2; ++q) {
      gemm64<true, true>(accW[q], sY, kLd64, sG + q * 64, L::kLd, 64);       // d_pw_w[n][q*64 + k]
      float acc[4][4]; zero_acc(acc);
      gemm64<false, true>(acc, sY, kLd64, sW + q * 64, L::kLd, 64);          // dg[i][q*64 + k]
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int m = gemm_row<false>(i);
        const long long tok = ((long long)b * p.H + ty0 + (m >> 3)) * p.W + tx0 + (m & 7);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int k = q * 64 + gemm_col<true>(j);
          if (k < 84) p.ddwc[tok * 84 + k] = acc[i][j] * gelu_erf_grad(sPre[m * L::kLd + k]);
        }
      }
    }
  }
  for (int q = 0; q < 2; ++q)
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = q * 64 + gemm_col<true>(j);
        if (k < 84) atomicAdd(p.d_pw_w + gemm_row<true>(i) * 84 + k, accW[q][i][j]);
      }
  if (tid < 64) atomicAdd(p.d_pw_b + tid, accB);
  // LayerNorm affine gradients: lanes hold channels (lane, lane + 32); reduce over the 8 warps through smem
  __syncthreads();
  float* red = sY;                                        // [8 warps][4][32]
  red[(warp * 4 + 0) * 32 + lane] = aLw0; red[(warp * 4 + 1) * 32 + lane] = aLw1;
  red[(warp * 4 + 2) * 32 + lane] = aLb0; red[(warp * 4 + 3) * 32 + lane] = aLb1;
  __syncthreads();
  if (tid < 128) {
    const int which = tid >> 5, l = tid & 31;
    float s2 = 0.f;
#pragma unroll
    for (int wv = 0; wv < 8; ++wv) s2 += red[(wv * 4 + which) * 32 + l];
    float* dst = which < 2 ? p.d_ln_w : p.d_ln_b;
    atomicAdd(dst + l + 32 * (which & 1), s2);
  }
}

// bottleneck_dw_bwd_kernel (one 8x8 tile per pass, halos of ddwc and of cat):
//   dcat[t][c] = sum_k ddwc[t - off_k][c] w_dw[c][k];  d_dw_w[c][k] += sum_t ddwc[t][c] cat[t + off_k][c];  d_dw_b += ddwc
//   then the concat / pixel-shuffle / LeakyReLU skips backwards:
//     c < 64        : de1[t][c] = dcat + dout[t][c] (the tail's "+ e1"),  dc0[t][c] = dcat leaky'(c0[t][c])
//     64 <= c < 80  : source (lo = (y/2, x/2), k = ((c-64)*2 + y%2)*2 + x%2): de2[lo][k] = dcat, dp0[lo][k] = dcat leaky'(p0)
//     80 <= c < 84  : source (lo = (y/4, x/4), k = ((c-80)*4 + y%4)*4 + x%4): de3, dp00 likewise
struct DwBwdParams {
  const float* cat; const float* ddwc; const float* dw_w; const float* dout;
  const float* c0; const float* p0; const float* p00;
  float* dc0; float* dp0; float* dp00; float* de1; float* de2; float* de3;
  float* d_dw_w; float* d_dw_b;
  int B, H, W;
};
__global__ void __launch_bounds__(kThreads, 1) bottleneck_dw_bwd_kernel(DwBwdParams p) {
  extern __shared__ __align__(16) float smem[];
  float* sDw = smem;                      // [9][84]
  float* sC = sDw + 9 * 84;               // [100][84] cat halo
  float* sD = sC + 100 * 84;              // [100][84] ddwc halo
  float* sRed = sD + 100 * 84;            // [3][84][10]
  const int tid = threadIdx.x;
  for (int idx = tid; idx < 9 * 84; idx += kThreads) sDw[idx] = p.dw_w[(idx % 84) * 9 + idx / 84];
  float acc[10];
#pragma unroll
  for (int k = 0; k < 10; ++k) acc[k] = 0.f;
  const int th = p.H / 8, tw = p.W / 8;
  const int ntile = p.B * th * tw;
  const int cmine = tid % 84, grp = tid / 84;            // threads 0..251: (group of tokens, channel)
  for (int tile = blockIdx.x; tile < ntile; tile += gridDim.x) {
    __syncthreads();
    const int tx0 = (tile % tw) * 8, ty0 = ((tile / tw) % th) * 8, b = tile / (tw * th);
    for (int idx = tid; idx < 100 * 21; idx += kThreads) {
      const int hp = idx / 21, g = idx - hp * 21;
      const int yy = ty0 + hp / 10 - 1, xx = tx0 + hp % 10 - 1;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f), d = v;
      if (yy >= 0 && yy < p.H && xx >= 0 && xx < p.W) {
        const long long o = (((long long)b * p.H + yy) * p.W + xx) * 84 + g * 4;
        v = ldg4(p.cat + o);
        d = ldg4(p.ddwc + o);
      }
      *reinterpret_cast<float4*>(sC + hp * 84 + g * 4) = v;
      *reinterpret_cast<float4*>(sD + hp * 84 + g * 4) = d;
    }
    __syncthreads();
    if (grp < 3) {                                        // depthwise weight / bias gradients
      for (int t = grp; t < 64; t += 3) {
        const int ty = t >> 3, tx = t & 7;
        const float d = sD[((ty + 1) * 10 + tx + 1) * 84 + cmine];
#pragma unroll
        for (int dy = 0; dy < 3; ++dy)
#pragma unroll
          for (int dx = 0; dx < 3; ++dx)
            acc[dy * 3 + dx] = fmaf(d, sC[((ty + dy) * 10 + tx + dx) * 84 + cmine], acc[dy * 3 + dx]);
        acc[9] += d;
      }
    }
    for (int idx = tid; idx < 64 * 84; idx += kThreads) {
      const int t = idx / 84, c = idx - t * 84;
      const int ty = t >> 3, tx = t & 7;
      float g = 0.f;
#pragma unroll
      for (int dy = 0; dy < 3; ++dy)
#pragma unroll
        for (int dx = 0; dx < 3; ++dx)      // dcat[t] += ddwc[t - off_k] w[k], off_k = (dy-1, dx-1)
          g = fmaf(sD[((ty + 2 - dy) * 10 + tx + 2 - dx) * 84 + c], sDw[(dy * 3 + dx) * 84 + c], g);
      const int yy = ty0 + ty, xx = tx0 + tx;
      if (c < 64) {
        const long long o = (((long long)b * p.H + yy) * p.W + xx) * 64 + c;
        p.de1[o] = g + __ldg(p.dout + o);
        p.dc0[o] = __ldg(p.c0 + o) > 0.f ? g : 0.01f * g;
      } else if (c < 80) {
        const int k = (((c - 64) * 2 + (yy & 1)) * 2) + (xx & 1);
        const long long o = (((long long)b * (p.H / 2) + yy / 2) * (p.W / 2) + xx / 2) * 64 + k;
        p.de2[o] = g;
        p.dp0[o] = __ldg(p.p0 + o) > 0.f ? g : 0.01f * g;
      } else {
        const int k = (((c - 80) * 4 + (yy & 3)) * 4) + (xx & 3);
        const long long o = (((long long)b * (p.H / 4) + yy / 4) * (p.W / 4) + xx / 4) * 64 + k;
        p.de3[o] = g;
        p.dp00[o] = __ldg(p.p00 + o) > 0.f ? g : 0.01f * g;
      }
    }
  }
  __syncthreads();
  if (grp < 3)
#pragma unroll
    for (int k = 0; k < 10; ++k) sRed[(grp * 84 + cmine) * 10 + k] = acc[k];
  __syncthreads();
  for (int idx = tid; idx < 84 * 10; idx += kThreads) {
    const int c = idx / 10, k = idx - c * 10;
    const float s2 = sRed[(0 * 84 + c) * 10 + k] + sRed[(1 * 84 + c) * 10 + k] + sRed[(2 * 84 + c) * 10 + k];
    if (k < 9) atomicAdd(p.d_dw_w + c * 9 + k, s2); else atomicAdd(p.d_dw_b + c, s2);
  }
}

}  // namespace b200mar

using namespace b200mar;

extern "C" int b200mar_final_recon_bwd(const float* dy, const float* y, const float* taps, const float* dec,
                                       const float* c0, const float* ln_w, const float* ln_b, const float* w1,
                                       const float* b1, const float* w2, float* g1_scratch, float* ddec, float* dc0,
                                       float* d_ln_w, float* d_ln_b, float* d_w1, float* d_b1, float* d_w2,
                                       float* d_b2, int B, int H, int W, int H0, int W0, void* stream) {
  if (B <= 0 || H <= 0 || W <= 0 || H0 > H || W0 > W || (long long)B * H * W >= (1ll << 31)) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  const long long npix = (long long)B * H * W;
  recon_bwd_kernel<<<grid_cap((npix + 255) / 256, 4), 256, 0, st>>>(dy, y, taps, b1, w2, g1_scratch, d_w2, d_b2, d_b1,
                                                                   B, H, W, H0, W0);
  B200MAR_CHECK(cudaGetLastError());
  final_bwd_kernel<<<grid_cap((npix + 7) / 8, 4), 256, 0, st>>>(dec, c0, ln_w, ln_b, w1, g1_scratch, ddec, dc0, d_w1,
                                                               d_ln_w, d_ln_b, B, H, W);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_shallow_tok_bwd(const float* x, const float* w, const float* dc0, float* dw, float* db,
                                       float* dx, int B, int H, int W, void* stream) {
  if (B <= 0 || H <= 0 || W <= 0 || (long long)B * H * W >= (1ll << 31)) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  const long long npix = (long long)B * H * W;
  shallow_tok_wgrad_kernel<<<grid_cap((npix + 127) / 128, 2), 256, 0, st>>>(x, dc0, dw, db, B, H, W);
  B200MAR_CHECK(cudaGetLastError());
  if (dx) {
    shallow_tok_dgrad_kernel<<<grid_cap((npix + 7) / 8, 8), 256, 0, st>>>(dc0, w, dx, B, H, W);
    B200MAR_CHECK(cudaGetLastError());
  }
  return 0;
}

extern "C" int b200mar_cascade_proj_bwd(const float* srcA, const float* srcB, const float* direct, const float* w,
                                        const float* dout, const float* dpool, float* d_srcA, float* d_srcB,
                                        float* d_direct, float* dw, float* db, int B, int h, int wd, void* stream) {
  if (B <= 0 || h <= 0 || wd <= 0 || !srcA || !direct || !dout || !d_srcA || !d_direct || (srcB && !d_srcB)) return -1;
  CascadeBwdParams p{srcA, srcB, direct, w, dout, dpool, d_srcA, d_srcB, d_direct, dw, db, B, h, wd, srcB ? 192 : 128};
  const size_t smem = ((size_t)2 * 64 * (p.K + 4) + 2 * 64 * kLd64) * sizeof(float);
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(cascade_proj_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)(((size_t)2 * 64 * 196 + 2 * 64 * kLd64) * sizeof(float))));
    configured.done();
  }
  const long long ntile = ((long long)B * h * wd + 63) / 64;
  const int grid = (int)(ntile < sm_count() ? ntile : sm_count());
  cascade_proj_bwd_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_bottleneck_bwd(const float* cat, const float* dout, const float* c0, const float* p0,
                                      const float* p00, const float* dw_w, const float* dw_b, const float* pw_w,
                                      const float* pw_b, const float* ln_w, float* ddwc_scratch, float* dc0,
                                      float* dp0, float* dp00, float* de1, float* de2, float* de3, float* d_dw_w,
                                      float* d_dw_b, float* d_pw_w, float* d_pw_b, float* d_ln_w, float* d_ln_b,
                                      int B, int H, int W, void* stream) {
  if (B <= 0 || H % 8 || W % 8 || H <= 0 || W <= 0) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  const int ntile = B * (H / 8) * (W / 8);
  const int grid = ntile < 2 * sm_count() ? ntile : 2 * sm_count();
  static DeviceOnce configured;
  const size_t smem1 = TailBwdSmem::kFloats * sizeof(float);
  const size_t smem2 = (size_t)(9 * 84 + 2 * 100 * 84 + 3 * 84 * 10) * sizeof(float);
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(bottleneck_tail_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
    B200MAR_CHECK(cudaFuncSetAttribute(bottleneck_dw_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
    configured.done();
  }
  TailBwdParams t{cat, dw_w, dw_b, pw_w, pw_b, ln_w, dout, ddwc_scratch, d_pw_w, d_pw_b, d_ln_w, d_ln_b, B, H, W};
  bottleneck_tail_bwd_kernel<<<grid, kThreads, smem1, st>>>(t);
  B200MAR_CHECK(cudaGetLastError());
  DwBwdParams d{cat, ddwc_scratch, dw_w, dout, c0, p0, p00, dc0, dp0, dp00, de1, de2, de3, d_dw_w, d_dw_b, B, H, W};
  bottleneck_dw_bwd_kernel<<<grid, kThreads, smem2, st>>>(d);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
