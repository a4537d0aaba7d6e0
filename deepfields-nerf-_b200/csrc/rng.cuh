// Counter-based random numbers for the in-kernel draws of the train step (the reference draws with torch.rand /
// torch.randn_like inside sampling.py:20, train.py:53 and model.py:162; parity tests inject recorded draws instead).
// Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11): key = seed, counter = (item index,
// training step, stream id), so every draw is a pure function of (seed, step, what, which) - no state to carry, CUDA-graph
// replays advance through the step counter in device memory, ranks differ through the seed.
#pragma once
#include <stdint.h>

namespace rng {

constexpr uint32_t kStreamJitter = 1, kStreamPdf = 2, kStreamNoiseCoarse = 3, kStreamNoiseFine = 4;

__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}

// uniform in [0, 1): 24 random bits, like torch.rand for float32
__device__ __forceinline__ float u01(uint32_t x) { return static_cast<float>(x >> 8) * 5.9604644775390625e-08f; }

// four uniforms of item `idx` (64-bit) in stream `stream` at training step `step`
__device__ __forceinline__ float4 uniform4(unsigned long long seed, uint32_t step, uint32_t stream, unsigned long long idx) {
  const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(idx), static_cast<uint32_t>(idx >> 32), step, stream),
                                make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32)));
  return make_float4(u01(r.x), u01(r.y), u01(r.z), u01(r.w));
}

// one standard normal of item `idx` (Box-Muller on two of the four words; u1 in (0, 1])
__device__ __forceinline__ float normal1(unsigned long long seed, uint32_t step, uint32_t stream, unsigned long long idx) {
  const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(idx), static_cast<uint32_t>(idx >> 32), step, stream),
                                make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32)));
  const float u1 = (static_cast<float>(r.x >> 8) + 1.0f) * 5.9604644775390625e-08f;
  const float u2 = u01(r.y);
  return sqrtf(-2.0f * __logf(u1)) * __cosf(6.283185307179586f * u2);
}

}  // namespace rng
