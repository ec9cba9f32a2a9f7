// tma.cuh — Tensor Memory Accelerator helpers (inline PTX, sm_100a): 2-D tiled tensor maps over the
// token-major fp32 residual stream [n_tok][64], bulk tensor loads signalled on an mbarrier, bulk
// tensor stores and reduce-add stores (bulk_group completion).
//
// Tile convention: a 128-token tile is moved as TWO boxes of [128 rows][32 floats] (inner extent
// 128 B = the SWIZZLE_128B span).  In shared memory box h (channels 32h..32h+31) is a 16 KB block,
// row r at byte r*128, and the 16-byte chunk c of that row sits at chunk slot c ^ (r & 7).  A thread
// that owns row r and walks the chunks with a compile-time index therefore hits 8 distinct
// 16-byte bank groups per quarter-warp: conflict-free whole-row access with static register indices.
// Tile buffers must be 1024-byte aligned.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200mar {
namespace tma {

// byte offset of 16-byte chunk i (0..15) of row r inside a [2 boxes][128 rows][128 B] swizzled tile
__device__ __forceinline__ uint32_t tile_off(int r, int i) {
  return (uint32_t)((i >> 3) * 16384 + r * 128 + (((i & 7) ^ (r & 7)) << 4));
}

__device__ __forceinline__ void prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(
                   (uint32_t)__cvta_generic_to_shared(bar)),
               "r"(bytes)
               : "memory");
}

// global -> shared box load; completes `bytes of the box` on the mbarrier
__device__ __forceinline__ void load_2d(const CUtensorMap* m, uint64_t* bar, void* smem_dst, int c0,
                                        int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(c0), "r"(c1)
      : "memory");
}

// shared -> global box store (bulk_group)
__device__ __forceinline__ void store_2d(const CUtensorMap* m, const void* smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"((uint32_t)__cvta_generic_to_shared(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}

// shared -> global box reduction: global[box] += shared[box] (fp32 add, one add per element)
__device__ __forceinline__ void reduce_add_2d(const CUtensorMap* m, const void* smem_src, int c0,
                                              int c1) {
  asm volatile(
      "cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
          reinterpret_cast<uint64_t>(m)),
      "r"((uint32_t)__cvta_generic_to_shared(smem_src)), "r"(c0), "r"(c1)
      : "memory");
}

// 3-D form (coordinates fastest-first)
__device__ __forceinline__ void load_3d(const CUtensorMap* m, uint64_t* bar, void* smem_dst, int c0,
                                        int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(c0), "r"(c1),
      "r"(c2)
      : "memory");
}

// 4-D forms over the token grid [B][ph][pw][64] (coordinates fastest-first: channel, x, y, b)
__device__ __forceinline__ void load_4d(const CUtensorMap* m, uint64_t* bar, void* smem_dst, int c0,
                                        int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(c0), "r"(c1),
      "r"(c2), "r"(c3)
      : "memory");
}
// L2 prefetch of a box (no shared-memory destination, no completion signal)
__device__ __forceinline__ void prefetch_4d(const CUtensorMap* m, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.prefetch.tensor.4d.L2.global.tile [%0, {%1, %2, %3, %4}];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void store_4d(const CUtensorMap* m, const void* smem_src, int c0, int c1,
                                         int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
          reinterpret_cast<uint64_t>(m)),
      "r"((uint32_t)__cvta_generic_to_shared(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

__device__ __forceinline__ void commit_group() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all committed bulk groups of this thread have finished READING shared memory
__device__ __forceinline__ void wait_read_all() {
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
// all but the most recent committed bulk group of this thread have finished reading shared memory
__device__ __forceinline__ void wait_read_1() {
  asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
}
// all committed bulk groups of this thread are complete (global writes performed)
__device__ __forceinline__ void wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ---- host: tensor map over a token-major fp32 matrix [n_rows][64], box [128][32], 128 B swizzle.
// cuTensorMapEncodeTiled is fetched through the runtime (no link-time libcuda dependency, so the
// library also loads on a machine without a driver).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess || !p)
      return nullptr;
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// ---- host: tensor map over the token grid [B][ph][pw][64] fp32; one box = half the channels of one
// 8x8 window: [1][8 rows][8 tokens][32 floats] -> 64 smem rows (window slot t = 8*i + j) of 128 B,
// 128 B swizzle (16-byte chunk c of slot t at chunk slot c ^ (t & 7)).
inline int encode_windows(CUtensorMap* out, const float* base, int B, int ph, int pw) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return -2;
  const cuuint64_t gdim[4] = {64, (cuuint64_t)pw, (cuuint64_t)ph, (cuuint64_t)B};
  const cuuint64_t gstride[3] = {256, (cuuint64_t)pw * 256, (cuuint64_t)ph * pw * 256};
  const cuuint32_t box[4] = {32, 8, 8, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(base), gdim, gstride,
                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -3;
}

// ---- host: the same over a BF16 token grid [B][ph][pw][64] (precision mode "bf16"): one box = ALL 64 channels of
// one 8x8 window: [1][8 rows][8 tokens][64 bf16 = 128 B] -> 64 smem rows of 128 B, 128 B swizzle.
inline int encode_windows16(CUtensorMap* out, const void* base, int B, int ph, int pw) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return -2;
  const cuuint64_t gdim[4] = {64, (cuuint64_t)pw, (cuuint64_t)ph, (cuuint64_t)B};
  const cuuint64_t gstride[3] = {128, (cuuint64_t)pw * 128, (cuuint64_t)ph * pw * 128};
  const cuuint32_t box[4] = {64, 8, 8, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gdim, gstride,
                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -3;
}

// ---- host: a token-major BF16 matrix [n_rows][64]: box [128 rows][64 bf16 = 128 B], 128 B swizzle (16 KB)
inline int encode_rows64_16(CUtensorMap* out, const void* base, long long n_rows) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return -2;
  const cuuint64_t gdim[2] = {64, (cuuint64_t)n_rows};
  const cuuint64_t gstride[1] = {128};
  const cuuint32_t box[2] = {64, 128};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride,
                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -3;
}

// ---- host: a token-major BF16 matrix [n_rows][C] (C a multiple of 64): box [128 rows][64 bf16 = 128 B] at a channel
// offset that is a multiple of 64, 128 B swizzle (16 KB) - the store side of kernels that stage a bf16 export tile
inline int encode_rows16_box64(CUtensorMap* out, const void* base, long long n_rows, int C) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return -2;
  const cuuint64_t gdim[2] = {(cuuint64_t)C, (cuuint64_t)n_rows};
  const cuuint64_t gstride[1] = {(cuuint64_t)C * 2};
  const cuuint32_t box[2] = {64, 128};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride,
                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -3;
}

// ---- host: general form over a token grid [B][H][W][64] fp32: box = [1][by rows][bx tokens][32 floats]
// taking every `step`-th token in x and y (step = 2: one parity class of the 2x2 PatchMerging
// neighbourhood), 128 B swizzle.  The box lands as by*bx rows of 128 B, row = y*bx + x.
inline int encode_grid(CUtensorMap* out, const float* base, int B, int H, int W, int bx, int by,
                       int step) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return -2;
  const cuuint64_t gdim[4] = {64, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  const cuuint64_t gstride[3] = {256, (cuuint64_t)W * 256, (cuuint64_t)H * W * 256};
  const cuuint32_t box[4] = {32, (cuuint32_t)(bx * step), (cuuint32_t)(by * step), 1};
  const cuuint32_t estr[4] = {1, (cuuint32_t)step, (cuuint32_t)step, 1};
  const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(base), gdim, gstride,
                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -3;
}

// ---- host: un-swizzled halo box over a channels-last grid [B][H][W][C] fp32 (C * 4 a multiple of
// 16): box = [1][bh][bw][C]; out-of-bounds positions (negative or past the edge) are zero-filled,
// which is exactly a convolution's zero padding.
inline int encode_halo(CUtensorMap* out, const float* base, int B, int H, int W, int C, int bw, int bh) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return -2;
  const cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  const cuuint64_t gstride[3] = {(cuuint64_t)C * 4, (cuuint64_t)W * C * 4, (cuuint64_t)H * W * C * 4};
  const cuuint32_t box[4] = {(cuuint32_t)C, (cuuint32_t)bw, (cuuint32_t)bh, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(base), gdim, gstride,
                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -3;
}

// ---- host: a row-major 16-bit matrix [n_rows][C] (C a multiple of 8) viewed as [C/8 chunks][n_rows][8]:
// one box = [C/8][128 rows][8 elements] lands in shared memory CHUNK-MAJOR (chunk c at c*2048, row r at
// r*16), i.e. as the no-swizzle tcgen05 operand tile of tc.cuh; rows past n_rows are zero-filled.
inline int encode_rows16_chunkmajor(CUtensorMap* out, const void* base, long long n_rows, int C) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return -2;
  const cuuint64_t gdim[3] = {8, (cuuint64_t)n_rows, (cuuint64_t)(C / 8)};
  const cuuint64_t gstride[2] = {(cuuint64_t)C * 2, 16};
  const cuuint32_t box[3] = {8, 128, (cuuint32_t)(C / 8)};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), gdim, gstride,
                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -3;
}

inline int encode_rows64(CUtensorMap* out, const float* base, long long n_rows) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return -2;
  const cuuint64_t gdim[2] = {64, (cuuint64_t)n_rows};
  const cuuint64_t gstride[1] = {256};
  const cuuint32_t box[2] = {32, 128};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride,
                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -3;
}

}  // namespace tma
}  // namespace b200mar
