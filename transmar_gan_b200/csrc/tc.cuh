// tc.cuh — tcgen05 / TMEM / mbarrier primitives (inline PTX, sm_100a) used by the tensor-core
// kernels.  Conventions:
//   * operands are fp32 in shared memory, consumed as TF32 (kind::tf32, K = 8 per MMA);
//   * operand tiles use the canonical NO-SWIZZLE K-major layout ("chunk-major"):
//       element (row r, k) of an R-row tile lives at byte  (k/4)*(R*16) + r*16 + (k%4)*4
//     i.e. 8x16-byte core matrices are 128 contiguous bytes, SBO = 128 B (next 8 rows),
//     LBO = R*16 B (next 16-byte k-chunk).  A thread that owns row r writes one float4 per
//     k-chunk and consecutive rows are consecutive 16-byte slots -> conflict-free;
//   * accumulators are M = 128 rows (TMEM lane = row) x N fp32 columns, cta_group::1;
//   * every mbarrier wait is bounded (2^22 try_wait rounds, seconds): on timeout a global error flag is
//     raised and the kernel TRAPS, so a descriptor bug can neither hang the GPU box nor be computed through.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <cuda_fp16.h>
#include "tc_ldst.cuh"

namespace b200mar {

// per-translation-unit error flag raised by a timed-out mbarrier wait (no -rdc, so each .cu
// that includes this header owns one; lib.cu sums them in b200mar_tc_error()).
static __device__ int g_tc_error = 0;
#define B200MAR_DEFINE_TC_ERROR_GETTER(name)                                       \
  int name() {                                                                     \
    int v = 0;                                                                     \
    if (cudaMemcpyFromSymbol(&v, g_tc_error, sizeof(int)) != cudaSuccess) return -1; \
    return v;                                                                      \
  }

namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// Bounded wait on phase parity; returns false (and raises *err) on timeout.
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity, int* err) {
  const uint32_t addr = smem_u32(bar);
#pragma unroll 1
  for (uint32_t spin = 0; spin < (1u << 22); ++spin) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    if (done) return true;
  }
  // A barrier that never completes is a bug (bad descriptor, lost commit): raise the flag for
  // b200mar_tc_error() and kill the kernel - continuing would consume unsynchronised TMEM / shared memory and
  // silently corrupt outputs or gradients.  The trap surfaces as a CUDA error on the next runtime call.
  if (err) atomicExch(err, 1);
  __threadfence_system();
  __trap();
  return false;
}

// ---------------------------------------------------------------- fences
__device__ __forceinline__ void fence_async_smem() {   // generic-proxy smem writes -> async proxy (MMA)
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// ---------------------------------------------------------------- TMEM allocation (one warp)
template <int COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t* slot_in_smem) {
  static_assert(COLS == 32 || COLS == 64 || COLS == 128 || COLS == 256 || COLS == 512, "pow2 >= 32");
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
               ::"r"(smem_u32(slot_in_smem)), "n"(COLS) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}

// ---------------------------------------------------------------- descriptors
// shared-memory matrix descriptor, no swizzle, K-major (cute::UMMA::SmemDescriptor bit layout):
// [0,14) start>>4 | [16,30) LBO>>4 | [32,46) SBO>>4 | [46,48) version=1 | [61,64) layout=0
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// instruction descriptor, kind::tf32, fp32 accumulate, both operands K-major
// (cute::UMMA::InstrDescriptor): c_format[4,6)=1 a_format[7,10)=2 b_format[10,13)=2
// n_dim[17,23)=N>>3 m_dim[24,29)=M>>4
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// instruction descriptor, kind::f16 with fp16 operands (a_format = b_format = 0), fp32 accumulate
__host__ __device__ constexpr uint32_t idesc_f16(int M, int N) {
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// instruction descriptor, kind::f16 with BF16 operands (a_format = b_format = 1), fp32 accumulate;
// a_mn / b_mn select an MN-major operand (bit 15 / 16: the transpose of a chunk-major tile, see below)
__host__ __device__ constexpr uint32_t idesc_bf16(int M, int N, bool a_mn = false, bool b_mn = false) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (a_mn ? (1u << 15) : 0u) | (b_mn ? (1u << 16) : 0u) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// two fp32 -> packed bf16x2 (lo = a, hi = b), round-to-nearest-even
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}
// A 16-bit chunk-major tile of R rows: element (row r, channel c) at byte (c/8)*(R*16) + r*16 + (c%8)*2.
//   * as a K-major operand (M/N index = row, K = channel): smem_desc(base, LBO = R*16, SBO = 128),
//     one K = 16 step advances the start by 2*R*16 bytes;
//   * as an MN-major operand (M/N index = channel, K = row; i.e. the TRANSPOSE of the tile):
//     smem_desc(base, LBO = 128, SBO = R*16), one K = 16 step (16 rows) advances the start by 256 bytes.
__device__ __forceinline__ uint64_t desc_kmajor16(uint32_t saddr, int R) { return smem_desc(saddr, R * 16, 128); }
__device__ __forceinline__ uint64_t desc_mnmajor16(uint32_t saddr, int R) { return smem_desc(saddr, 128, R * 16); }
// kind::f16: one K = 16 step, operands K-major fp16 (8 elements per 16-byte k-chunk)
__device__ __forceinline__ void mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                        uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// kind::f16 with the A operand in TMEM: 128 lanes x 8 columns per K = 16 step, two fp16 per column
__device__ __forceinline__ void mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc,
                                           uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// two fp32 -> packed fp16x2 (lo = a, hi = b), round-to-nearest-even
__device__ __forceinline__ uint32_t pack_f16(float a, float b) {
  uint32_t r;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}
// lo part of an fp16 hi + lo pair: fp16x2(a - hi.lo16, b - hi.hi16) for hi = pack_f16(a, b).  The differences are
// exact in fp32; |lo| <= 2^-11 |value|, and below 2^-14 the fp16 subnormals still resolve 2^-24 absolute.
__device__ __forceinline__ uint32_t lo_f16(float a, float b, uint32_t hi) {
  const float2 h = __half22float2(*reinterpret_cast<const __half2*>(&hi));
  return pack_f16(a - h.x, b - h.y);
}
// second term of a BF16 pair: bf16x2(a - hi.lo16, b - hi.hi16) for hi = pack_bf16(a, b); 16 significant bits in all
__device__ __forceinline__ uint32_t lo_bf16(float a, float b, uint32_t hi) {
  return pack_bf16(a - __uint_as_float(hi << 16), b - __uint_as_float(hi & 0xFFFF0000u));
}
// D[tmem] (+)= A[smem] * B[smem]^T, one K = 8 step.  Issued by ONE thread.
__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                         uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// same, with the A operand (128 lanes x 8 fp32 columns per K step) read from TMEM
__device__ __forceinline__ void mma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc,
                                            uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// K-major operand pair with explicit k-chunk strides (bytes between successive 16-byte k-chunks):
// K/8 MMAs, D (+)= A[128 x K] * B[N x K]^T.
__device__ __forceinline__ void mma_ss_tf32(uint32_t tmem_d, uint32_t a_saddr, uint32_t lboA,
                                            uint32_t b_saddr, uint32_t lboB, int K, uint32_t idesc,
                                            bool accumulate_first) {
#pragma unroll
  for (int k = 0; k < K; k += 8) {
    const uint64_t ad = smem_desc(a_saddr + (k >> 2) * lboA, lboA, 128);
    const uint64_t bd = smem_desc(b_saddr + (k >> 2) * lboB, lboB, 128);
    mma_tf32(tmem_d, ad, bd, idesc, (k > 0 || accumulate_first) ? 1u : 0u);
  }
}
// A from TMEM columns [tmem_a, tmem_a + K), B K-major in shared memory
__device__ __forceinline__ void mma_ts_tf32(uint32_t tmem_d, uint32_t tmem_a, uint32_t b_saddr,
                                            uint32_t lboB, int K, uint32_t idesc, bool accumulate_first) {
#pragma unroll
  for (int k = 0; k < K; k += 8) {
    const uint64_t bd = smem_desc(b_saddr + (k >> 2) * lboB, lboB, 128);
    mma_tf32_ts(tmem_d, tmem_a + k, bd, idesc, (k > 0 || accumulate_first) ? 1u : 0u);
  }
}
// Full-K product over a chunk-major operand pair: K/8 MMAs, each advancing both start
// addresses by two 16-byte k-chunks (2*LBO).  rowsA/rowsB = tile heights (LBO = rows*16).
__device__ __forceinline__ void mma_tile_tf32(uint32_t tmem_d, uint32_t a_saddr, int rowsA,
                                              uint32_t b_saddr, int rowsB, int K, uint32_t idesc,
                                              bool accumulate_first) {
  const uint32_t lboA = rowsA * 16, lboB = rowsB * 16;
  for (int k = 0; k < K; k += 8) {
    const uint64_t ad = smem_desc(a_saddr + (k >> 2) * lboA, lboA, 128);
    const uint64_t bd = smem_desc(b_saddr + (k >> 2) * lboB, lboB, 128);
    mma_tf32(tmem_d, ad, bd, idesc, (k > 0 || accumulate_first) ? 1u : 0u);
  }
}
// make all previously issued MMAs arrive on an mbarrier when they complete
// fp32 -> tf32, round to nearest (the MMA itself ignores the low 13 mantissa bits, i.e. truncates)
__device__ __forceinline__ float round_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__device__ __forceinline__ float4 round_tf32(float4 v) {
  return make_float4(round_tf32(v.x), round_tf32(v.y), round_tf32(v.z), round_tf32(v.w));
}

__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
               ::"r"(smem_u32(bar)) : "memory");
}

// ---------------------------------------------------------------- TMEM -> registers
// 32 lanes x 32-bit, 16 consecutive columns per thread (thread's lane = 32*(warp%4) + laneid)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  // load + wait in ONE asm statement so no consumer can be scheduled between them
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n\t"
      "tcgen05.wait::ld.sync.aligned;"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// one lane of a converged warp (the MMA issuer): warp-uniform branch + elect keeps the operands
// of the following tcgen05.mma in uniform registers (no per-lane waterfall loop)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
// named barrier for one 128-thread worker (ids 1..15; 0 is __syncthreads)
__device__ __forceinline__ void worker_sync(int id) {
  asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory");
}

// named barrier of one 128-thread worker with an OR reduction of a per-thread predicate
__device__ __forceinline__ bool worker_sync_or(int id, bool pred) {
  uint32_t r;
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "setp.ne.u32 q, %2, 0;\n\t"
      "bar.red.or.pred p, %1, 128, q;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(r) : "r"(id), "r"((uint32_t)pred) : "memory");
  return r != 0;
}
// chunk-major operand addressing (floats): element (r, k) of an R-row tile
__device__ __forceinline__ int cm_off(int R, int r, int k) { return (k >> 2) * (R * 4) + r * 4 + (k & 3); }

}  // namespace tc
}  // namespace b200mar
