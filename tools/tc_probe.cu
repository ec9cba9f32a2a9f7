// tc_probe.cu — single-CTA micro-kernel that exercises the tcgen05 operand conventions the
// attention kernel relies on, so each can be checked against a plain matmul in the tests:
//   mode 0: D = A * B^T, A [128,K] and B [N,K] both K-major from shared memory
//   mode 1: D = A * Bm,  Bm [K,N] MN-major (N contiguous) from shared memory
//   mode 2: as mode 1, but A is first moved into TMEM (tcgen05.st) and used as a TMEM operand
// `col0` shifts the accumulator / TMEM-A base column to test unaligned column addressing.
// Development tooling, NOT part of libb200mar.so: tools/gpu_probe.py builds it into tools/libtcprobe.so
// (nvcc -I transmar_gan_b200/csrc) on demand.
#include "common.cuh"
#include "tc.cuh"
#include <cuda_fp16.h>
#include <cuda_bf16.h>

namespace b200mar {

B200MAR_DEFINE_TC_ERROR_GETTER(tc_error_probe)
const char*& last_error_ref() { static thread_local const char* m = "no error"; return m; }
int sm_count() { return 148; }

__global__ void __launch_bounds__(128, 1) tc_probe_kernel(const float* __restrict__ A,
                                                          const float* __restrict__ B,
                                                          float* __restrict__ D, int N, int K,
                                                          int mode, int col0) {
  extern __shared__ __align__(128) float smem[];
  float* sA = smem;                    // [128][K] chunk-major
  float* sB = smem + 128 * K;          // K-major: [N][K] chunk-major ; MN-major: see below
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int t = threadIdx.x, lane = t & 31, wq = t >> 5;
  for (int idx = t; idx < 128 * K; idx += 128) {
    const int r = idx / K, k = idx % K;
    sA[tc::cm_off(128, r, k)] = A[idx];
  }
  int raw_lbo = 0, raw_sbo = 0;
  if ((mode & 15) == 3) {
    raw_lbo = col0 & 0xFFFF; raw_sbo = (col0 >> 16) & 0xFFFF; col0 = 0;
    for (int idx = t; idx < 4096; idx += 128) sB[idx] = B[idx];
  } else if ((mode & 15) == 0 || (mode & 15) == 2) {
    const int Rp = N + ((mode >> 4) == 1 ? 1 : 0);            // variant 1: padded k-chunk stride
    for (int idx = t; idx < N * K; idx += 128) {
      const int n = idx / K, k = idx % K;
      sB[tc::cm_off(Rp, n, k)] = B[idx];
    }
  } else {
    // MN-major, no swizzle: element (n, k) at (n%4) + (k%8)*4 + (k/8)*32 + (n/4)*(K*4) floats:
    // 8 k-rows of 16 bytes (4 consecutive n) form a 128-byte core matrix; LBO = 128 B between
    // successive 8-k groups, SBO = K*16 B between successive 4-n groups.
    const int variant = mode >> 4;
    for (int idx = t; idx < K * N; idx += 128) {
      const int k = idx / N, n = idx % N;
      if (variant < 2) sB[(n & 3) + (k & 7) * 4 + (k >> 3) * 32 + (n >> 2) * (K * 4)] = B[idx];
      else             sB[(n & 3) + (k & 7) * 4 + (n >> 2) * 32 + (k >> 3) * (N * 8)] = B[idx];
    }

  }
  const int variant = mode >> 4;
  mode &= 15;
  if (t == 0) { tc::mbar_init(&bar, 1); tc::fence_barrier_init(); }
  if (t < 32) tc::tmem_alloc<512>(&slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tbase = slot;
  const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
  const uint32_t dcol = col0, acol = 256 + col0;
  if (mode == 2) {
    // A row of this thread -> TMEM columns [acol, acol+K)
    const int row = wq * 32 + lane;
    for (int c = 0; c < K; c += 16) {
      uint32_t r[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(A[row * K + c + i]);
      asm volatile(
          "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
          "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
          ::"r"(tbase + lane_addr + acol + c), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]),
            "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]),
            "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    tc::tc_fence_before();
    __syncthreads();
  }
  if (t == 0) {
    tc::tc_fence_after();
    const uint32_t aS = tc::smem_u32(sA), bS = tc::smem_u32(sB);
    uint32_t idesc = tc::idesc_tf32(128, N);
    if (mode == 1 || mode == 3) idesc |= (1u << 16);          // b_major = MN
    if (mode == 3) {
      tc::mma_tf32(tbase + dcol, tc::smem_desc(aS, 2048, 128), tc::smem_desc(bS, raw_lbo * 16, raw_sbo * 16),
                   idesc, 0u);
      K = 0;   // single MMA only
    }
    for (int k = 0; k < K; k += 8) {
      uint64_t bd;
      const int Rp = N + (variant == 1 ? 1 : 0);
      if (mode == 0 || mode == 2) bd = tc::smem_desc(bS + (k >> 2) * (Rp * 16), Rp * 16, 128);
      else if (variant == 0) bd = tc::smem_desc(bS + (k >> 3) * 128, 128, K * 16);
      else if (variant == 1) bd = tc::smem_desc(bS + (k >> 3) * 128, K * 16, 128);
      else if (variant == 2) bd = tc::smem_desc(bS + (k >> 3) * (N * 32), N * 32, 128);
      else                   bd = tc::smem_desc(bS + (k >> 3) * (N * 32), 128, N * 32);
      if (mode == 2) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
            ::"r"(tbase + dcol), "r"(tbase + acol + k), "l"(bd), "r"(idesc), "r"(k > 0 ? 1u : 0u)
            : "memory");
      } else {
        const uint64_t ad = tc::smem_desc(aS + (k >> 2) * 2048, 2048, 128);
        tc::mma_tf32(tbase + dcol, ad, bd, idesc, k > 0 ? 1u : 0u);
      }
    }
    tc::mma_commit(&bar);
  }
  tc::mbar_wait(&bar, 0, &g_tc_error);
  tc::tc_fence_after();
  const int row = wq * 32 + lane;
  for (int c = 0; c < N; c += 16) {
    float v[16];
    tc::tmem_ld16(tbase + lane_addr + dcol + c, v);
#pragma unroll
    for (int i = 0; i < 16; ++i) D[row * N + c + i] = v[i];
  }
  tc::tc_fence_before();
  __syncthreads();
  if (t < 32) tc::tmem_dealloc<512>(slot);
}

// fp16-operand variants: mode 4 = both operands K-major fp16 in shared memory (chunk-major, 8
// elements per 16-byte chunk), mode 5 = A packed two-per-column in TMEM.
__global__ void __launch_bounds__(128, 1) tc_probe_f16_kernel(const float* __restrict__ A,
                                                              const float* __restrict__ B,
                                                              float* __restrict__ D, int N, int K,
                                                              int mode, int col0) {
  extern __shared__ __align__(128) float smem[];
  __half* sA = reinterpret_cast<__half*>(smem);          // [128][K]: (k/8)*(128*8) + r*8 + k%8
  __half* sB = sA + 128 * K;                             // [N][K]  : (k/8)*(N*8) + n*8 + k%8
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int t = threadIdx.x, lane = t & 31, wq = t >> 5;
  for (int idx = t; idx < 128 * K; idx += 128) {
    const int r = idx / K, k = idx % K;
    sA[(k >> 3) * 1024 + r * 8 + (k & 7)] = __float2half_rn(A[idx]);
  }
  for (int idx = t; idx < N * K; idx += 128) {
    const int n = idx / K, k = idx % K;
    sB[(k >> 3) * (N * 8) + n * 8 + (k & 7)] = __float2half_rn(B[idx]);
  }
  if (t == 0) { tc::mbar_init(&bar, 1); tc::fence_barrier_init(); }
  if (t < 32) tc::tmem_alloc<512>(&slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tbase = slot;
  const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
  const uint32_t dcol = col0, acol = 256 + col0;
  const int row = wq * 32 + lane;
  if (mode == 5) {
    for (int c = 0; c < K / 2; c += 8) {                  // 8 columns = 16 fp16 per store
      float v[16];
#pragma unroll
      for (int i = 0; i < 8; ++i)
        v[i] = __uint_as_float(tc::pack_f16(A[row * K + 2 * (c + i)], A[row * K + 2 * (c + i) + 1]));
      asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                   ::"r"(tbase + lane_addr + acol + c), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]),
                     "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
    }
    tc::tmem_wait_st();
    tc::tc_fence_before();
    __syncthreads();
  }
  if (wq == 0) {
    tc::tc_fence_after();
    if (tc::elect_one()) {
      const uint32_t idesc = tc::idesc_f16(128, N);
      const uint64_t da = tc::smem_desc(tc::smem_u32(sA), 2048, 128);
      const uint64_t db = tc::smem_desc(tc::smem_u32(sB), N * 16, 128);
      for (int k = 0; k < K; k += 16) {
        if (mode == 5) tc::mma_f16_ts(tbase + dcol, tbase + acol + k / 2, db + (k >> 3) * N, idesc, k > 0);
        else           tc::mma_f16(tbase + dcol, da + (k >> 3) * 128, db + (k >> 3) * N, idesc, k > 0);
      }
      tc::mma_commit(&bar);
    }
    __syncwarp();
  }
  tc::mbar_wait(&bar, 0, &g_tc_error);
  tc::tc_fence_after();
  for (int c = 0; c < N; c += 16) {
    float v[16];
    tc::tmem_ld16(tbase + lane_addr + dcol + c, v);
#pragma unroll
    for (int i = 0; i < 16; ++i) D[row * N + c + i] = v[i];
  }
  tc::tc_fence_before();
  __syncthreads();
  if (t < 32) tc::tmem_dealloc<512>(slot);
}

// bf16 / MN-major variants (the layouts the backward kernels rely on):
//   mode 6: D[m][n] = sum_k A[k][m] B[k][n], A [K][128] and B [K][N] both chunk-major tiles of K rows read
//           as MN-major operands (the transposes of the tiles);
//   mode 7: D = A B with A [128][K] packed two-per-column in TMEM and B [K][N] MN-major in shared memory.
__global__ void __launch_bounds__(128, 1) tc_probe_bf16_kernel(const float* __restrict__ A,
                                                               const float* __restrict__ B,
                                                               float* __restrict__ D, int N, int K,
                                                               int mode, int col0) {
  extern __shared__ __align__(128) float smem[];
  unsigned char* sA = reinterpret_cast<unsigned char*>(smem);       // [K rows][128 ch] chunk-major bf16
  unsigned char* sB = sA + K * 128 * 2;                               // [K rows][N ch]
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int t = threadIdx.x, lane = t & 31, wq = t >> 5;
  if (mode == 6)
    for (int idx = t; idx < K * 128; idx += 128) {
      const int k = idx / 128, m = idx % 128;
      *reinterpret_cast<__nv_bfloat16*>(sA + (m >> 3) * (K * 16) + k * 16 + (m & 7) * 2) = __float2bfloat16_rn(A[idx]);
    }
  for (int idx = t; idx < K * N; idx += 128) {
    const int k = idx / N, n = idx % N;
    *reinterpret_cast<__nv_bfloat16*>(sB + (n >> 3) * (K * 16) + k * 16 + (n & 7) * 2) = __float2bfloat16_rn(B[idx]);
  }
  if (t == 0) { tc::mbar_init(&bar, 1); tc::fence_barrier_init(); }
  if (t < 32) tc::tmem_alloc<512>(&slot);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tbase = slot;
  const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
  const uint32_t dcol = col0, acol = 256 + col0;
  const int row = wq * 32 + lane;
  if (mode == 7) {
    for (int c = 0; c < K / 2; c += 8) {
      float v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i)
        v[i] = __uint_as_float(tc::pack_bf16(A[row * K + 2 * (c + i)], A[row * K + 2 * (c + i) + 1]));
      asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                   ::"r"(tbase + lane_addr + acol + c), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]),
                     "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
    }
    tc::tmem_wait_st();
    tc::tc_fence_before();
    __syncthreads();
  }
  if (wq == 0) {
    tc::tc_fence_after();
    if (tc::elect_one()) {
      const uint32_t idesc = tc::idesc_bf16(128, N, mode == 6, true);
      const uint64_t da = tc::desc_mnmajor16(tc::smem_u32(sA), K);
      const uint64_t db = tc::desc_mnmajor16(tc::smem_u32(sB), K);
      for (int k = 0; k < K; k += 16) {
        if (mode == 7) tc::mma_f16_ts(tbase + dcol, tbase + acol + k / 2, db + (k >> 4) * 16, idesc, k > 0);
        else           tc::mma_f16(tbase + dcol, da + (k >> 4) * 16, db + (k >> 4) * 16, idesc, k > 0);
      }
      tc::mma_commit(&bar);
    }
    __syncwarp();
  }
  tc::mbar_wait(&bar, 0, &g_tc_error);
  tc::tc_fence_after();
  for (int c = 0; c < N; c += 16) {
    float v[16];
    tc::tmem_ld16(tbase + lane_addr + dcol + c, v);
#pragma unroll
    for (int i = 0; i < 16; ++i) D[row * N + c + i] = v[i];
  }
  tc::tc_fence_before();
  __syncthreads();
  if (t < 32) tc::tmem_dealloc<512>(slot);
}

}  // namespace b200mar

extern "C" int b200mar_tc_probe(const float* A, const float* B, float* D, int N, int K, int mode,
                                int col0, void* stream) {
  using namespace b200mar;
  if (mode == 6 || mode == 7) {
    if (N % 16 || N < 16 || N > 256 || K % 16 || K < 16 || K > 128 || col0 < 0 || col0 + N > 256) return -1;
    const size_t smem16 = (size_t)(128 * K + N * K) * 2;
    B200MAR_CHECK(cudaFuncSetAttribute(tc_probe_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       200 * 1024));
    tc_probe_bf16_kernel<<<1, 128, smem16, (cudaStream_t)stream>>>(A, B, D, N, K, mode, col0);
    B200MAR_CHECK(cudaGetLastError());
    return 0;
  }
  if (mode == 4 || mode == 5) {
    if (N % 16 || N < 16 || N > 256 || K % 16 || K < 16 || K > 128 || col0 < 0 || col0 + N > 256) return -1;
    const size_t smem16 = (size_t)(128 * K + N * K) * 2;
    B200MAR_CHECK(cudaFuncSetAttribute(tc_probe_f16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       200 * 1024));
    tc_probe_f16_kernel<<<1, 128, smem16, (cudaStream_t)stream>>>(A, B, D, N, K, mode, col0);
    B200MAR_CHECK(cudaGetLastError());
    return 0;
  }
  if (N % 16 || N < 16 || N > 256 || K % 8 || K < 8 || K > 128 || mode < 0 || (mode & 15) > 3) return -1;
  if ((mode & 15) != 3 && (col0 < 0 || col0 + N > 256 || col0 + K > 256)) return -1;
  const size_t smem = (size_t)(128 * K + (N + 1) * K + ((mode & 15) == 3 ? 4096 : 0)) * sizeof(float);
  B200MAR_CHECK(cudaFuncSetAttribute(tc_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     200 * 1024));
  tc_probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(A, B, D, N, K, mode, col0);
  B200MAR_CHECK(cudaGetLastError());
  return 0;
}
