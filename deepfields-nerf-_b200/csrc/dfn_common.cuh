// Shared host/device helpers for libdeepfields_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#define DFN_OK 0
#define DFN_ERR_ARG 1
#define DFN_ERR_LAUNCH 2

void dfn_set_error(const char* fmt, ...);

#define DFN_CHECK_ARG(cond, ...)                 \
  do {                                           \
    if (!(cond)) {                               \
      dfn_set_error(__VA_ARGS__);                \
      return DFN_ERR_ARG;                        \
    }                                            \
  } while (0)

#define DFN_CHECK_LAUNCH(name)                                                    \
  do {                                                                            \
    cudaError_t e__ = cudaGetLastError();                                         \
    if (e__ != cudaSuccess) {                                                     \
      dfn_set_error("%s: launch failed: %s", name, cudaGetErrorString(e__));      \
      return DFN_ERR_LAUNCH;                                                      \
    }                                                                             \
  } while (0)

#define DFN_CUDA(call, name)                                                      \
  do {                                                                            \
    cudaError_t e__ = (call);                                                     \
    if (e__ != cudaSuccess) {                                                     \
      dfn_set_error("%s: %s", name, cudaGetErrorString(e__));                     \
      return DFN_ERR_LAUNCH;                                                      \
    }                                                                             \
  } while (0)

// SM count of the CURRENT device (cached per device: one process may drive several GPUs)
static inline int dfn_num_sms() {
  static int cache[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  const int slot = dev >= 0 && dev < 64 ? dev : 0;
  if (cache[slot] == 0) {
    int n = 0;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    cache[slot] = n > 0 ? n : 148;
  }
  return cache[slot];
}

// SM count the MLP launchers size their grids to: the device's, or the calling thread's budget (dfn_set_sm_limit, dfn_error.cu)
int dfn_mlp_sms();

// ---- device helpers ---------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ uint32_t pack_bf16x2_relu(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
// 16-byte global store as ONE instruction.  Written as asm because the compiler otherwise splits a uint4 store whose
// four words do not sit in an aligned register quad into 4+4+8-byte stores (measured: 2x on the training forward).
__device__ __forceinline__ void st_global_v4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// the same with an L2 eviction-priority hint (policy from make_evict_first_policy): write-once streams (activation
// stash, dz scratch) should not displace the weight images the same kernel keeps re-reading from L2
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ uint64_t make_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint64_t make_evict_last_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void st_global_v4_hint(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.v4.b32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d), "l"(pol) : "memory");
}
__device__ __forceinline__ float bf16lo_to_f32(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16hi_to_f32(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ float round_bf16(float x) {
  return __bfloat162float(__float2bfloat16_rn(x));
}
