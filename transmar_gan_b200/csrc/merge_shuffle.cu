// merge_shuffle.cu — PatchMerging forward (components/encoder_decoder.py:51-74),
// pixel_shuffle_permute + concat (components/encoder_decoder.py:159-164,191-192) and the
// integer index-map exports used by the bit-exactness tests.
#include "common.cuh"
#include "merge_common.cuh"

namespace b200mar {

// ------------------------------------------------------------------------ PatchMerging fwd
__global__ void __launch_bounds__(kThreads, 1) patch_merge_fwd_kernel(MergeParams p) {
  using L = MergeSmem;
  extern __shared__ __align__(16) float smem[];
  float* sW = smem + L::oW;
  float* sLnW = smem + L::oLnW;
  float* sLnB = smem + L::oLnB;
  float* sX = smem + L::oX;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  load_matrix(sW, kLd256, p.red_w, kDim, 256, kDim);
  for (int i = tid; i < 256; i += kThreads) { sLnW[i] = p.ln_w[i]; sLnB[i] = p.ln_b[i]; }
  __syncthreads();
  const int oh = p.ph / 2, ow = p.pw / 2;
  const long long n_out = (long long)p.B * oh * ow;
  const long long ntile = (n_out + kTok - 1) / kTok;
  for (long long tile = blockIdx.x; tile < ntile; tile += gridDim.x) {
    // ---- gather the 2x2 neighbourhood: X[m][q*64 + c], q = 2*row_parity + col_parity
    {
      const int sub = lane >> 4, c = (lane & 15) << 2;
#pragma unroll 4
      for (int it = 0; it < 16; ++it) {
        const int pair = (warp * 16 + it) * 2 + sub;   // 0..255 = (m, q)
        const int m = pair >> 2, q = pair & 3;
        const long long o = tile * kTok + m;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (o < n_out) v = *reinterpret_cast<const float4*>(p.x + merge_src(p, o, q) * kDim + c);
        *reinterpret_cast<float4*>(sX + m * kLd256 + q * kDim + c) = v;
      }
    }
    __syncthreads();
    // ---- LayerNorm over the 256 concatenated channels (in place)
#pragma unroll 1
    for (int rr = 0; rr < 8; ++rr) {
      float* row = sX + (warp * 8 + rr) * kLd256;
      float v[8], s = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) { v[i] = row[lane + 32 * i]; s += v[i]; }
      const float mean = warp_sum(s) * (1.0f / 256.f);
      float q = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) { v[i] -= mean; q = fmaf(v[i], v[i], q); }
      const float rstd = 1.0f / sqrtf(warp_sum(q) * (1.0f / 256.f) + kLnEps);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        row[lane + 32 * i] = v[i] * rstd * sLnW[lane + 32 * i] + sLnB[lane + 32 * i];
    }
    __syncthreads();
    // ---- reduction Linear(256 -> 64, no bias)
    {
      float acc[4][4]; zero_acc(acc);
      gemm64<false, false>(acc, sX, kLd256, sW, kLd256, 256);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const long long o = tile * kTok + gemm_row<false>(i);
        if (o < n_out) {
#pragma unroll
          for (int j = 0; j < 4; ++j) p.y[o * kDim + gemm_col<false>(j)] = acc[i][j];
        }
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------ pixel shuffle + cat
// One thread per 16-byte group of the [B, H*W, 84] concat tensor (21 groups per token): groups 0..15 are a
// straight float4 copy of the full-resolution skip, groups 16..19 / 20 gather four channels of the half- /
// quarter-resolution skips (source channel stride 4 / 16).  Forward copies x_i -> cat, inverse copies
// cat -> x_i (the permutation's transpose, i.e. its gradient).  Token coordinates in 32-bit arithmetic
// (the launcher checks B*H*W < 2^31).
__global__ void __launch_bounds__(256) pixel_shuffle_cat_kernel(float* __restrict__ x0, float* __restrict__ x1,
                                                                float* __restrict__ x2, float* __restrict__ cat,
                                                                int B, int H, int W, int inverse) {
  const long long total = (long long)B * H * W * 21;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const unsigned tok = (unsigned)(idx / 21);
    const int g = (int)(idx - (long long)tok * 21);
    float4* cp = reinterpret_cast<float4*>(cat + (size_t)tok * 84 + g * 4);
    if (g < 16) {
      float4* sp = reinterpret_cast<float4*>(x0 + (size_t)tok * 64 + g * 4);
      if (inverse) *sp = *cp; else *cp = *sp;
      continue;
    }
    const unsigned C = tok % (unsigned)W, row = tok / (unsigned)W;
    const unsigned R = row % (unsigned)H, b = row / (unsigned)H;
    float* src;
    int stride;
    if (g < 20) {              // s = 2: channel c = 4 (g-16) + k  <-  source channel (c*2 + a)*2 + c'
      const unsigned h = R >> 1, a = R & 1, w = C >> 1, cc = C & 1;
      src = x1 + ((size_t)(b * (H >> 1) + h) * (W >> 1) + w) * 64 + (((g - 16) * 4 * 2 + a) * 2 + cc);
      stride = 4;
    } else {                   // s = 4: channel c = k  <-  source channel (c*4 + a)*4 + c'
      const unsigned h = R >> 2, a = R & 3, w = C >> 2, cc = C & 3;
      src = x2 + ((size_t)(b * (H >> 2) + h) * (W >> 2) + w) * 64 + (a * 4 + cc);
      stride = 16;
    }
    if (inverse) {
      const float4 v = *cp;
      src[0] = v.x; src[stride] = v.y; src[2 * stride] = v.z; src[3 * stride] = v.w;
    } else {
      *cp = make_float4(src[0], src[stride], src[2 * stride], src[3 * stride]);
    }
  }
}

// ------------------------------------------------------------------------ index-map exports
__global__ void partition_index_kernel(long long* tok_idx, long long* ctx_idx, Geom g) {
  const int n = g.B * g.wh * g.ww * kTok;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const Slot s = slot_of(g, i >> 6, i & 63);
    tok_idx[i] = s.tok;
    ctx_idx[i] = s.cwin;
  }
}
__global__ void shift_mask_kernel(float* mask, Geom g) {
  const int n = g.wh * g.ww * kTok * kTok;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int q = i & 63, pq = (i >> 6) & 63, win = i >> 12;
    const int rp = slot_of(g, win, pq).region, rq = slot_of(g, win, q).region;
    mask[i] = (g.shift > 0 && rp != rq) ? -100.0f : 0.0f;
  }
}
__global__ void rel_index_kernel(long long* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < kTok * kTok) out[i] = rel_index(i >> 6, i & 63);
}

}  // namespace b200mar

extern "C" int b200mar_patch_merge_fwd(const float* x, const float* ln_w, const float* ln_b,
                                       const float* red_w, float* y, int B, int ph, int pw,
                                       void* stream) {
  using namespace b200mar;
  if (ph % 2 || pw % 2) return -1;
  MergeParams p{};
  p.x = x; p.ln_w = ln_w; p.ln_b = ln_b; p.red_w = red_w; p.y = y; p.B = B; p.ph = ph; p.pw = pw;
  const size_t smem = MergeSmem::kFloatsFwd * sizeof(float);
  static DeviceOnce configured;
  if (configured.first()) {
    B200MAR_CHECK(cudaFuncSetAttribute(patch_merge_fwd_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured.done();
  }
  const long long ntile = ((long long)B * (ph / 2) * (pw / 2) + kTok - 1) / kTok;
  const int grid = (int)(ntile < sm_count() ? ntile : sm_count());
  patch_merge_fwd_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(p);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_pixel_shuffle_cat(float* x0, float* x1, float* x2, float* cat, int B, int H,
                                         int W, int inverse, void* stream) {
  using namespace b200mar;
  if (H % 4 || W % 4 || (long long)B * H * W >= (1ll << 31)) return -1;
  const long long total = (long long)B * H * W * 21;      // 16-byte groups
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  pixel_shuffle_cat_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(x0, x1, x2, cat, B, H, W,
                                                                          inverse);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_partition_index(long long* tok_idx, long long* ctx_idx, int B, int ph,
                                       int pw, int shift, void* stream) {
  using namespace b200mar;
  if (ph % kWin || pw % kWin || (shift != 0 && shift != kShift)) return -1;
  const Geom g = make_geom(B, ph, pw, shift);
  const int n = g.B * g.wh * g.ww * kTok;
  partition_index_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(tok_idx, ctx_idx, g);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_shift_mask(float* mask, int ph, int pw, int shift, void* stream) {
  using namespace b200mar;
  if (ph % kWin || pw % kWin) return -1;
  const Geom g = make_geom(1, ph, pw, shift);
  const int n = g.wh * g.ww * kTok * kTok;
  shift_mask_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(mask, g);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_rel_index(long long* rel, void* stream) {
  using namespace b200mar;
  rel_index_kernel<<<16, 256, 0, (cudaStream_t)stream>>>(rel);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
