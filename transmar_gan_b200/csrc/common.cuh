// common.cuh — shared device helpers for the b200mar kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "b200mar kernels are written for sm_100a (B200) only"
#endif

namespace b200mar {

constexpr int kWin = 8;          // window side (models/generator/ngswin.py:46 window_size=8)
constexpr int kTok = 64;         // tokens per window
constexpr int kDim = 64;         // embed dim
constexpr int kHid = 128;        // MLP hidden (mlp_ratio 2)
constexpr int kShift = 4;        // window_size // 2
constexpr int kThreads = 256;    // CTA size of the tile kernels
constexpr int kLd64 = 68;        // smem row stride for 64-wide rows  (== 4 mod 32: LDS.128 conflict-free)
constexpr int kLd128 = 132;      // smem row stride for 128-wide rows
constexpr int kLd256 = 260;      // smem row stride for 256-wide rows
constexpr float kLnEps = 1e-5f;
constexpr float kLogitMax = 4.605170185988092f;  // ln(1/0.01), attention.py:84
constexpr float kNormEps = 1e-12f;               // F.normalize eps, attention.py:83

// ---------------------------------------------------------------- host-side error plumbing
const char*& last_error_ref();
int sm_count();                  // of the CURRENT device (cached per device)
// "once per device" guard for per-function attributes (cudaFuncSetAttribute is per device, and a process may
// drive several GPUs): `static DeviceOnce c; if (c.first()) { ...set attributes...; c.done(); }`
struct DeviceOnce {
  unsigned long long mask = 0;   // bit d set = configured on device d (d < 64; larger ordinals always reconfigure)
  int dev = -1;
  bool first() {
    if (cudaGetDevice(&dev) != cudaSuccess) dev = -1;
    if (dev < 0 || dev >= 64) return true;
    return !((__atomic_load_n(&mask, __ATOMIC_ACQUIRE) >> dev) & 1ull);
  }
  void done() {
    if (dev >= 0 && dev < 64) __atomic_fetch_or(&mask, 1ull << dev, __ATOMIC_RELEASE);
  }
};
#define B200MAR_CHECK(call)                                                   \
  do {                                                                        \
    cudaError_t _e = (call);                                                  \
    if (_e != cudaSuccess) { last_error_ref() = #call; return (int)_e; }      \
  } while (0)

// ---------------------------------------------------------------- window geometry
// Division by a loop-invariant divisor without the ~25-instruction IDIV sequence (3 of them on the XU pipe):
// Granlund-Montgomery round-up multiplier, exact for every 32-bit n and d >= 1.
struct FastDiv { uint32_t m, s1, s2; };
__host__ __device__ inline FastDiv make_fastdiv(uint32_t d) {
  uint32_t l = 0;
  while ((1ull << l) < d) ++l;
  FastDiv f;
  f.m = (uint32_t)((((1ull << l) - d) << 32) / d + 1);
  f.s1 = l < 1u ? l : 1u;
  f.s2 = l > 0u ? l - 1u : 0u;
  return f;
}
__device__ __forceinline__ uint32_t fdiv(uint32_t n, const FastDiv& f) {
  const uint32_t t = __umulhi(f.m, n);
  return (t + ((n - t) >> f.s1)) >> f.s2;
}
struct Geom {
  int B, ph, pw, wh, ww, shift;
  FastDiv fww, fwwh;             // division by ww and by ww * wh (window index -> (w, h, b))
};
__host__ __device__ inline Geom make_geom(int B, int ph, int pw, int shift) {
  Geom g; g.B = B; g.ph = ph; g.pw = pw; g.wh = ph / kWin; g.ww = pw / kWin; g.shift = shift;
  g.fww = make_fastdiv((uint32_t)g.ww); g.fwwh = make_fastdiv((uint32_t)(g.ww * g.wh));
  return g;
}
// window index -> (w, h, b), row-major over (b, h, w); no integer division
struct WinDiv {                  // for kernels that carry (wh, ww) instead of a Geom
  int ww, wwh;
  FastDiv fww, fwwh;
};
__host__ __device__ inline WinDiv make_windiv(int wh, int ww) {
  WinDiv d; d.ww = ww; d.wwh = ww * wh; d.fww = make_fastdiv((uint32_t)ww); d.fwwh = make_fastdiv((uint32_t)(ww * wh));
  return d;
}
__device__ __forceinline__ void win_pos(const WinDiv& d, int win, int& w, int& h, int& b) {
  b = (int)fdiv((uint32_t)win, d.fwwh);
  const int rem = win - b * d.wwh;
  h = (int)fdiv((uint32_t)rem, d.fww);
  w = rem - h * d.ww;
}
__device__ __forceinline__ void win_pos(const Geom& g, int win, int& w, int& h, int& b) {
  b = (int)fdiv((uint32_t)win, g.fwwh);
  const int rem = win - b * (g.ww * g.wh);
  h = (int)fdiv((uint32_t)rem, g.fww);
  w = rem - h * g.ww;
}
// Window `win` = (b,h,w) row-major, slot t = 8*i + j of the (shifted) window.
// Source row/col in the unshifted token grid: r=(8h+i+s)%ph, c=(8w+j+s)%pw
// (window_utils.py:134-145: context added per unshifted window, then roll(-s,-s), then partition).
struct Slot {
  long long tok;   // flat token index  b*ph*pw + r*pw + c
  int cwin;        // flat source window index b*wh*ww + (r/8)*ww + c/8
  int region;      // 9-region id of the shifted-frame position (transformer_blocks.py:63-76)
};
__device__ __forceinline__ Slot slot_of(const Geom& g, int w, int h, int b, int t) {
  const int R = h * kWin + (t >> 3);
  const int C = w * kWin + (t & 7);
  int r = R + g.shift; if (r >= g.ph) r -= g.ph;
  int c = C + g.shift; if (c >= g.pw) c -= g.pw;
  Slot s;
  s.tok = ((long long)b * g.ph + r) * g.pw + c;
  s.cwin = (b * g.wh + (r >> 3)) * g.ww + (c >> 3);
  const int rid = R < g.ph - kWin ? 0 : (R < g.ph - kShift ? 1 : 2);
  const int cid = C < g.pw - kWin ? 0 : (C < g.pw - kShift ? 1 : 2);
  s.region = rid * 3 + cid;
  return s;
}
__device__ __forceinline__ Slot slot_of(const Geom& g, int win, int t) {
  int w, h, b;
  win_pos(g, win, w, h, b);
  return slot_of(g, w, h, b, t);
}
// relative-position index of (query p, key q) inside an 8x8 window, attention.py:9-18
__device__ __forceinline__ int rel_index(int p, int q) {
  return ((p >> 3) - (q >> 3) + 7) * 15 + ((p & 7) - (q & 7) + 7);
}

// ---------------------------------------------------------------- small math
__device__ __forceinline__ float gelu_erf(float x) {
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}
__device__ __forceinline__ float gelu_erf_grad(float x) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
  const float pdf = 0.39894228040143267794f * expf(-0.5f * x * x);
  return cdf + x * pdf;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---------------------------------------------------------------- CTA-wide 64x64 GEMM tile
// acc[i][j] += sum_{r<K} A(m_i, r) * B(n_j, r)  for a 64x64 output block, 256 threads,
// thread (ty,tx) = (tid>>4, tid&15) owning 4 rows x 4 cols.
//   A_T = false: A stored [m][r] (row stride lda, float4 along r), rows m_i = ty + 16 i
//   A_T = true : A stored [r][m] (float4 along m),                 rows m_i = 4 ty + i
//   B_T = false: B stored [n][r],                                   cols n_j = tx + 16 j
//   B_T = true : B stored [r][n],                                   cols n_j = 4 tx + j
// Row strides must be multiples of 4 floats and base pointers 16-byte aligned; strides
// == 4 (mod 32) make the non-transposed float4 reads bank-conflict-free.  K % 4 == 0.
template <bool A_T> __device__ __forceinline__ int gemm_row(int i) {
  const int ty = threadIdx.x >> 4;
  return A_T ? 4 * ty + i : ty + 16 * i;
}
template <bool B_T> __device__ __forceinline__ int gemm_col(int j) {
  const int tx = threadIdx.x & 15;
  return B_T ? 4 * tx + j : tx + 16 * j;
}

template <bool A_T, bool B_T>
__device__ __forceinline__ void gemm64(float (&acc)[4][4], const float* __restrict__ A, int lda,
                                       const float* __restrict__ B, int ldb, int K) {
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
#pragma unroll 2
  for (int k = 0; k < K; k += 4) {
    float a[4][4], b[4][4];  // [row i][kk], [col j][kk]
    if (A_T) {
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const float4 v = *reinterpret_cast<const float4*>(A + (k + kk) * lda + 4 * ty);
        a[0][kk] = v.x; a[1][kk] = v.y; a[2][kk] = v.z; a[3][kk] = v.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 v = *reinterpret_cast<const float4*>(A + (ty + 16 * i) * lda + k);
        a[i][0] = v.x; a[i][1] = v.y; a[i][2] = v.z; a[i][3] = v.w;
      }
    }
    if (B_T) {
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const float4 v = *reinterpret_cast<const float4*>(B + (k + kk) * ldb + 4 * tx);
        b[0][kk] = v.x; b[1][kk] = v.y; b[2][kk] = v.z; b[3][kk] = v.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float4 v = *reinterpret_cast<const float4*>(B + (tx + 16 * j) * ldb + k);
        b[j][0] = v.x; b[j][1] = v.y; b[j][2] = v.z; b[j][3] = v.w;
      }
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk)
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i][kk], b[j][kk], acc[i][j]);
  }
}

__device__ __forceinline__ void zero_acc(float (&acc)[4][4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
}

// copy a [rows][cols] row-major global matrix into smem with row stride ld (cols % 4 == 0),
// zero-filling rows [rows, rows_pad).
__device__ __forceinline__ void load_matrix(float* dst, int ld, const float* __restrict__ src,
                                            int rows, int cols, int rows_pad) {
  const int c4 = cols >> 2;
  for (int idx = threadIdx.x; idx < rows_pad * c4; idx += blockDim.x) {
    const int r = idx / c4, c = (idx - r * c4) << 2;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < rows) v = *reinterpret_cast<const float4*>(src + (size_t)r * cols + c);
    *reinterpret_cast<float4*>(dst + r * ld + c) = v;
  }
}

}  // namespace b200mar
