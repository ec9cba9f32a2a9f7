// metal_loss.cu — the three image-space terms of the generator loss in ONE pass over (fake, real, ct), values and
// dL/dfake together:
//
//   M    = ct > threshold                                             losses/gan_losses.py:48-64   extract_metal_mask
//   band = max_pool2d(M, 2r+1, stride 1, padding r)                   :67-85                        dilate_mask
//   w    = min(1 + beta * band, w_max)                                :88-112                       compute_weight_map
//   rec   = mean |w * (fake - real)|                                  :115-139                      compute_metal_aware_loss
//   edge  = mean  w * (|gx(fake) - gx(real)| + |gy(fake) - gy(real)|) :142-191   forward differences, last column /
//                                                                                 row replicated     (…_edge_loss)
//   metal = mean |M * (fake - real)|                                  :194-219                      metal_consistency_loss
//   total = lam_rec * rec + lam_edge * edge + lam_metal * metal       training/train_combined.py:456-485
//
// Stock PyTorch runs 6 pooling / elementwise forward passes plus their backward over the same three images; here a
// CTA stages a 32x32 tile (+ halo) in shared memory, dilates the mask separably, and every pixel gathers its own
// gradient: the replicate padding makes difference W-2 (H-2) count twice, with the weights of columns W-2 and W-1.
// HBM-bound: 3 reads + 1 write of 4 bytes per pixel.  sums[4] = {rec, edge, metal, total} accumulate with one
// atomic per CTA and quantity (the caller zeroes them).
#include "common.cuh"

namespace b200mar {

constexpr int kLT = 32;                  // tile side
constexpr int kLMaxRadius = 32;

__device__ __forceinline__ float sgnf(float x) { return (float)(x > 0.f) - (float)(x < 0.f); }

struct MetalLossArgs {
  const float* fake; const float* real; const float* ct;
  float* dfake; float* sums;
  int nimg, H, W, radius;
  float threshold, beta, w_max, lam_rec, lam_edge, lam_metal;
};

__global__ void __launch_bounds__(256) metal_loss_kernel(MetalLossArgs a) {
  extern __shared__ unsigned char smem_raw[];
  const int R = a.radius, halo = R + 1, E = kLT + 2 * halo, P = kLT + 2;      // P: tile + 1-pixel ring
  float* sF = reinterpret_cast<float*>(smem_raw);           // [P][P]  fake
  float* sR = sF + P * P;                                   // [P][P]  real
  float* sWt = sR + P * P;                                  // [P][P]  weight map
  unsigned char* sM = reinterpret_cast<unsigned char*>(sWt + P * P);   // [E][E]  mask
  unsigned char* sRow = sM + E * E;                         // [E][P]  mask dilated along x
  const int H = a.H, W = a.W;
  const int x0 = blockIdx.x * kLT, y0 = blockIdx.y * kLT;
  const size_t img = (size_t)blockIdx.z * H * W;
  const int tid = threadIdx.x;

  for (int i = tid; i < E * E; i += 256) {                  // mask with halo r+1 (outside the image: 0)
    const int y = y0 - halo + i / E, x = x0 - halo + i % E;
    unsigned char m = 0;
    if (y >= 0 && y < H && x >= 0 && x < W) m = a.ct[img + (size_t)y * W + x] > a.threshold;
    sM[i] = m;
  }
  for (int i = tid; i < P * P; i += 256) {                  // fake / real with a 1-pixel ring
    const int y = y0 - 1 + i / P, x = x0 - 1 + i % P;
    float f = 0.f, r = 0.f;
    if (y >= 0 && y < H && x >= 0 && x < W) {
      f = a.fake[img + (size_t)y * W + x];
      r = a.real[img + (size_t)y * W + x];
    }
    sF[i] = f; sR[i] = r;
  }
  __syncthreads();
  for (int i = tid; i < E * P; i += 256) {                  // dilation along x for every staged row
    const int yy = i / P, xx = i % P;                       // ring column xx <-> mask column xx + R .. xx + 2R
    const unsigned char* row = sM + yy * E + xx;
    unsigned char m = 0;
    for (int k = 0; k <= 2 * R; ++k) m |= row[k];
    sRow[i] = m;
  }
  __syncthreads();
  for (int i = tid; i < P * P; i += 256) {                  // dilation along y -> weight map on the ring
    const int yy = i / P, xx = i % P;
    unsigned char m = 0;
    for (int k = 0; k <= 2 * R; ++k) m |= sRow[(yy + k) * P + xx];
    sWt[i] = fminf(1.0f + a.beta * (float)m, a.w_max);
  }
  __syncthreads();

  const float invN = 1.0f / ((float)a.nimg * (float)H * (float)W);
  float s_rec = 0.f, s_edge = 0.f, s_met = 0.f;
  const int tx = tid & 31;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int ty = (tid >> 5) + 8 * k;
    const int y = y0 + ty, x = x0 + tx;
    if (y >= H || x >= W) continue;
    const int c = (ty + 1) * P + tx + 1;                    // centre in ring coordinates
    const float f = sF[c], r = sR[c], w = sWt[c];
    const float m = (float)sM[(ty + halo) * E + tx + halo];
    const float d = f - r;
    s_rec += fabsf(w * d);
    s_met += fabsf(m * d);
    float g = a.lam_rec * w * sgnf(w * d) + a.lam_metal * m * sgnf(m * d);
    // x differences: Dx(j) = (f[j+1]-f[j]) - (r[j+1]-r[j]), j in [0, W-2]; column W-1 repeats Dx(W-2)
    float ge = 0.f;
    {
      const float dr = (x + 1 < W) ? (sF[c + 1] - f) - (sR[c + 1] - r) : 0.f;       // Dx(x)
      const float dl = (x >= 1) ? (f - sF[c - 1]) - (r - sR[c - 1]) : 0.f;          // Dx(x-1)
      s_edge += w * fabsf(x + 1 < W ? dr : dl);
      if (x >= 1) ge += (sWt[c - 1] + (x == W - 1 ? w : 0.f)) * sgnf(dl);           // +Wx(x-1) s(x-1)
      if (x + 1 < W) ge -= (w + (x == W - 2 ? sWt[c + 1] : 0.f)) * sgnf(dr);        // -Wx(x)   s(x)
    }
    {
      const float dd = (y + 1 < H) ? (sF[c + P] - f) - (sR[c + P] - r) : 0.f;
      const float du = (y >= 1) ? (f - sF[c - P]) - (r - sR[c - P]) : 0.f;
      s_edge += w * fabsf(y + 1 < H ? dd : du);
      if (y >= 1) ge += (sWt[c - P] + (y == H - 1 ? w : 0.f)) * sgnf(du);
      if (y + 1 < H) ge -= (w + (y == H - 2 ? sWt[c + P] : 0.f)) * sgnf(dd);
    }
    a.dfake[img + (size_t)y * W + x] = (g + a.lam_edge * ge) * invN;
  }

  // CTA reduction of the three sums, one atomic per quantity
  __shared__ float sRed[3][8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s_rec += __shfl_xor_sync(0xffffffffu, s_rec, o);
    s_edge += __shfl_xor_sync(0xffffffffu, s_edge, o);
    s_met += __shfl_xor_sync(0xffffffffu, s_met, o);
  }
  if (tx == 0) { sRed[0][tid >> 5] = s_rec; sRed[1][tid >> 5] = s_edge; sRed[2][tid >> 5] = s_met; }
  __syncthreads();
  if (tid < 3) {
    float v = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) v += sRed[tid][i];
    v *= invN;
    atomicAdd(a.sums + tid, v);
    const float lam = tid == 0 ? a.lam_rec : (tid == 1 ? a.lam_edge : a.lam_metal);
    atomicAdd(a.sums + 3, lam * v);
  }
}

}  // namespace b200mar

extern "C" int b200mar_metal_losses(const float* fake, const float* real, const float* ct, float* dfake, float* sums4,
                                    int nimg, int H, int W, int radius, float threshold, float beta, float w_max,
                                    float lam_rec, float lam_edge, float lam_metal, void* stream) {
  using namespace b200mar;
  if (nimg <= 0 || H < 2 || W < 2 || radius < 0 || radius > kLMaxRadius || nimg > 65535) return -1;
  MetalLossArgs a{fake, real, ct, dfake, sums4, nimg, H, W, radius, threshold, beta, w_max, lam_rec, lam_edge, lam_metal};
  const int halo = radius + 1, E = kLT + 2 * halo, P = kLT + 2;
  const size_t smem = (size_t)3 * P * P * sizeof(float) + (size_t)E * E + (size_t)E * P;
  dim3 grid((W + kLT - 1) / kLT, (H + kLT - 1) / kLT, nimg);
  metal_loss_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(a);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
