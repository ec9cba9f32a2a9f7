// radon.cu — parallel-beam Radon forward projector and its backprojection for the physics-consistency loss
// (call sites training/train_combined.py:269-272 `Radon(PATCH_SIZE, angles)`,
// losses/gan_losses.py:222-253 `projector.forward([B,H,W]) -> [B, n_angles, det]`).
//
// torch-radon (external/torch-radon, an EMPTY submodule of the reference: sources absent, parity unpinned) is
// restated from its published algorithm:
//   forward  : ray-driven.  Detector k of angle th is the line  x cos th + y sin th = s_k,
//              s_k = (k - D/2 + 1/2) * spacing, in coordinates centred on the image (pixel (i,j) centre at
//              (j + 1/2 - W/2, i + 1/2 - H/2)).  The ray is clipped to the image rectangle, cut into
//              ceil(length - 0.001) equal steps (<= 1 pixel, up to that margin, which keeps an axis-aligned ray of an
//              N-pixel image at exactly N steps whatever the rounding of the clip), sampled at the step midpoints with bilinear interpolation (zero outside
//              the image) and the sum is scaled by the step length.
//   backward : pixel-driven backprojection, as torch-radon's autograd backward: every pixel sums, over the angles,
//              the sinogram linearly interpolated at its own detector coordinate.
// B200 shape of it: the 128x128 crop is 64 KB, so a CTA keeps the whole image in shared memory with a one-pixel
// ring of zeros, which removes every bounds check from the sampling loop (any image up to 200 KB; larger ones are
// read through L1/L2 with checks) and walks `apc` angles of it, consecutive threads on consecutive
// detectors (neighbouring rays touch neighbouring pixels); the backward keeps a sin/cos table in shared memory and
// gathers from the L2-resident sinogram - no atomics, bitwise repeatable.
#include "common.cuh"

namespace b200mar {

constexpr int kRadonMaxAngles = 1024;
constexpr size_t kRadonSmemImage = 200 * 1024;

__device__ __forceinline__ float bilinear0(const float* img, int H, int W, float u, float v) {
  // texture-style coordinates: pixel (i, j) has its centre at (u, v) = (j + 1/2, i + 1/2); zero outside
  const float xb = u - 0.5f, yb = v - 0.5f;
  const float xf = floorf(xb), yf = floorf(yb);
  const int x0 = (int)xf, y0 = (int)yf;
  const float fx = xb - xf, fy = yb - yf;
  const bool xin0 = x0 >= 0 && x0 < W, xin1 = x0 + 1 >= 0 && x0 + 1 < W;
  const bool yin0 = y0 >= 0 && y0 < H, yin1 = y0 + 1 >= 0 && y0 + 1 < H;
  const float* p = img + y0 * W + x0;
  const float v00 = (yin0 && xin0) ? p[0] : 0.f, v01 = (yin0 && xin1) ? p[1] : 0.f;
  const float v10 = (yin1 && xin0) ? p[W] : 0.f, v11 = (yin1 && xin1) ? p[W + 1] : 0.f;
  const float top = fmaf(fx, v01 - v00, v00), bot = fmaf(fx, v11 - v10, v10);
  return fmaf(fy, bot - top, top);
}

// The same sample from the shared-memory copy, which carries a one-pixel ring of zeros (row stride W + 2): a point
// of a ray clipped to the image has floor(u - 1/2) in [-1, W-1], so no bounds checks are needed.
__device__ __forceinline__ float bilinear_ring(const float* img, int ld, float u, float v) {
  const float xb = u + 0.5f, yb = v + 0.5f;             // (u - 1/2) + 1: ring offset
  const float xf = floorf(xb), yf = floorf(yb);
  const float fx = xb - xf, fy = yb - yf;
  const float* p = img + (int)yf * ld + (int)xf;
  const float v00 = p[0], v01 = p[1], v10 = p[ld], v11 = p[ld + 1];
  const float top = fmaf(fx, v01 - v00, v00), bot = fmaf(fx, v11 - v10, v10);
  return fmaf(fy, bot - top, top);
}

constexpr int kRadonThreads = 512;

template <bool SMEM>
__global__ void __launch_bounds__(kRadonThreads) radon_fwd_kernel(const float* __restrict__ img_g, const float* __restrict__ angles,
                                                        float* __restrict__ sino, int H, int W, int A, int D,
                                                        float spacing, int apc) {
  extern __shared__ float sImg[];
  const float* img = img_g + (size_t)blockIdx.y * H * W;
  const int ld = W + 2;
  if (SMEM) {
    for (int i = threadIdx.x; i < (H + 2) * ld; i += kRadonThreads) {
      const int y = i / ld - 1, x = i % ld - 1;
      sImg[i] = (y >= 0 && y < H && x >= 0 && x < W) ? __ldg(img + y * W + x) : 0.f;
    }
    __syncthreads();
  }
  const int a0 = blockIdx.x * apc, a1 = min(a0 + apc, A);
  const float hx = 0.5f * (float)W, hy = 0.5f * (float)H;
  const float T = 0.71f * (float)max(H, W);              // half ray length: beyond the corner for every angle
  float* out = sino + (size_t)blockIdx.y * A * D;
  for (int r = a0 * D + threadIdx.x; r < a1 * D; r += kRadonThreads) {
    const int a = r / D, k = r - a * D;
    float sn, cs;
    sincosf(angles[a], &sn, &cs);
    const float s = ((float)k - 0.5f * (float)D + 0.5f) * spacing;
    const float sx = s * cs - T * sn, sy = s * sn + T * cs;      // start; direction (2T sn, -2T cs)
    const float dx = 2.f * T * sn, dy = -2.f * T * cs;
    // slab clip of alpha in [0, 1] to [-hx, hx] x [-hy, hy]
    float lo = 0.f, hi = 1.f;
    if (fabsf(dx) > 1e-12f) {
      const float t0 = (-hx - sx) / dx, t1 = (hx - sx) / dx;
      lo = fmaxf(lo, fminf(t0, t1)); hi = fminf(hi, fmaxf(t0, t1));
    } else if (sx < -hx || sx > hx) hi = -1.f;
    if (fabsf(dy) > 1e-12f) {
      const float t0 = (-hy - sy) / dy, t1 = (hy - sy) / dy;
      lo = fmaxf(lo, fminf(t0, t1)); hi = fminf(hi, fmaxf(t0, t1));
    } else if (sy < -hy || sy > hy) hi = -1.f;
    float acc = 0.f, step = 0.f;
    if (hi > lo) {
      const float len = 2.f * T * (hi - lo);
      const int n = max((int)ceilf(len - 1e-3f), 1);      // an axis-aligned ray of a 128-pixel image: exactly 128 steps
      step = len / (float)n;
      const float vx = dx * (hi - lo) / (float)n, vy = dy * (hi - lo) / (float)n;
      float px = sx + dx * lo + hx + 0.5f * vx, py = sy + dy * lo + hy + 0.5f * vy;    // first midpoint, texture coords
      // clamp the clip's rounding (a midpoint a hair outside the rectangle) so that the ring copy needs no checks
      const float ux = (float)W, uy = (float)H;
      for (int j = 0; j < n; ++j) {
        if (SMEM) acc += bilinear_ring(sImg, ld, fminf(fmaxf(px, 0.f), ux), fminf(fmaxf(py, 0.f), uy));
        else acc += bilinear0(img, H, W, px, py);
        px += vx; py += vy;
      }
    }
    out[r] = acc * step;
  }
}

__global__ void __launch_bounds__(256) radon_bwd_kernel(const float* __restrict__ sino_g, const float* __restrict__ angles,
                                                        float* __restrict__ img, int H, int W, int A, int D,
                                                        float inv_spacing) {
  __shared__ float sCs[kRadonMaxAngles], sSn[kRadonMaxAngles];
  for (int a = threadIdx.x; a < A; a += 256) sincosf(angles[a], &sSn[a], &sCs[a]);
  __syncthreads();
  const int pix = blockIdx.x * 256 + threadIdx.x;
  if (pix >= H * W) return;
  const int i = pix / W, j = pix - i * W;
  const float x = (float)j + 0.5f - 0.5f * (float)W, y = (float)i + 0.5f - 0.5f * (float)H;
  const float* sino = sino_g + (size_t)blockIdx.y * A * D;
  const float c0 = 0.5f * (float)D - 0.5f;               // detector index of s = 0
  float acc = 0.f;
  for (int a = 0; a < A; ++a) {
    const float kb = (x * sCs[a] + y * sSn[a]) * inv_spacing + c0;
    const float kf = floorf(kb);
    const int k0 = (int)kf;
    const float f = kb - kf;
    const float* row = sino + a * D;
    const float v0 = (k0 >= 0 && k0 < D) ? __ldg(row + k0) : 0.f;
    const float v1 = (k0 + 1 >= 0 && k0 + 1 < D) ? __ldg(row + k0 + 1) : 0.f;
    acc += fmaf(f, v1 - v0, v0);
  }
  img[(size_t)blockIdx.y * H * W + pix] = acc;
}

}  // namespace b200mar

extern "C" int b200mar_radon_fwd(const float* img, const float* angles, float* sino, int nimg, int H, int W,
                                 int n_angles, int det_count, float det_spacing, void* stream) {
  using namespace b200mar;
  if (nimg <= 0 || nimg > 65535 || H < 1 || W < 1 || n_angles < 1 || n_angles > kRadonMaxAngles || det_count < 1 ||
      !(det_spacing > 0.f) || (long long)H * W > (1ll << 30) || (long long)n_angles * det_count > (1ll << 30))
    return -1;
  // angles per CTA: about three CTAs per SM over the whole batch (one wave), rounded up so that a CTA's rays are
  // a whole number of passes of its 512 threads when the detector count allows
  int want = (3 * sm_count() + nimg - 1) / nimg;
  int apc = (n_angles + want - 1) / want;
  if (apc < 1) apc = 1;
  {
    int g = det_count, t = kRadonThreads;                 // gcd
    while (t) { const int q = g % t; g = t; t = q; }
    const int unit = kRadonThreads / g;                   // apc * det_count % 512 == 0  <=>  apc % unit == 0
    const int rounded = (apc + unit - 1) / unit * unit;
    if (rounded <= 2 * apc) apc = rounded;
  }
  dim3 grid((n_angles + apc - 1) / apc, nimg);
  const size_t bytes = (size_t)(H + 2) * (W + 2) * sizeof(float);
  if (bytes <= kRadonSmemImage) {
    static bool attr_set = false;          // idempotent; a race sets the same value twice
    if (!attr_set) {
      B200MAR_CHECK(cudaFuncSetAttribute(radon_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kRadonSmemImage));
      attr_set = true;
    }
    radon_fwd_kernel<true><<<grid, kRadonThreads, bytes, (cudaStream_t)stream>>>(img, angles, sino, H, W, n_angles, det_count,
                                                                      det_spacing, apc);
  } else {
    radon_fwd_kernel<false><<<grid, kRadonThreads, 0, (cudaStream_t)stream>>>(img, angles, sino, H, W, n_angles, det_count,
                                                                   det_spacing, apc);
  }
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int b200mar_radon_bwd(const float* sino, const float* angles, float* img, int nimg, int H, int W,
                                 int n_angles, int det_count, float det_spacing, void* stream) {
  using namespace b200mar;
  if (nimg <= 0 || nimg > 65535 || H < 1 || W < 1 || n_angles < 1 || n_angles > kRadonMaxAngles || det_count < 1 ||
      !(det_spacing > 0.f) || (long long)H * W > (1ll << 30) || (long long)n_angles * det_count > (1ll << 30))
    return -1;
  dim3 grid((H * W + 255) / 256, nimg);
  radon_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(sino, angles, img, H, W, n_angles, det_count,
                                                           1.0f / det_spacing);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
