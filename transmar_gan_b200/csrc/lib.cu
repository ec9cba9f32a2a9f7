// lib.cu — library-level plumbing of libb200mar.so: ABI version, error string, device probe.
#include "common.cuh"
#include "../../include/b200mar.h"

namespace b200mar {
const char*& last_error_ref() {
  static thread_local const char* msg = "no error";
  return msg;
}
int sm_count() {
  static int cache[64] = {0};                 // per device ordinal: a process may drive several GPUs
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev >= 0 && dev < 64 && cache[dev] > 0) return cache[dev];
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  if (dev >= 0 && dev < 64) cache[dev] = n;
  return n;
}
int tc_error_mlp();
int tc_error_attn();
int tc_error_merge();
int tc_error_bottleneck();
int tc_error_cascade();
int tc_error_bwd();
int tc_error_attn_bwd();
}  // namespace b200mar

// 1 if any tensor-core kernel of this process hit its bounded mbarrier wait (descriptor-bug guard)
extern "C" int b200mar_tc_error(void) {
  const int a = b200mar::tc_error_mlp(), b = 0, c = b200mar::tc_error_attn(),
            d = b200mar::tc_error_merge(), e = b200mar::tc_error_bottleneck(),
            f = b200mar::tc_error_cascade() | b200mar::tc_error_bwd() | b200mar::tc_error_attn_bwd();
  if (a < 0 || b < 0 || c < 0 || d < 0 || e < 0 || f < 0) return -1;
  return (a | b | c | d | e | f) ? 1 : 0;
}

extern "C" int b200mar_abi_version(void) { return 7; }
extern "C" const char* b200mar_last_error(void) { return b200mar::last_error_ref(); }
extern "C" int b200mar_device_ok(int dev) {
  int major = 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
  return major == 10 ? 1 : 0;
}
