// Device helpers shared by the Instant-NGP kernels (ngp.cu: inference + the unfused training pieces; ngp_train.cu: the fused
// training backward): level table, tcnn's grid indexing / trilinear weights, SH degree 4, mma.sync building blocks.
#pragma once
#include "dfn_common.cuh"
#include "../../include/deepfields_b200.h"
#include <cuda_fp16.h>
#include <math.h>

namespace ngp {

constexpr int kLevels = DFN_NGP_LEVELS;
constexpr uint32_t kPrimeY = 2654435761u, kPrimeZ = 805459861u;

struct LevelDev {
  float scale;
  uint32_t res, size, offset, hashed;
  uint32_t magic;  // floor(2^32 / size): branch-free i % size on the dense levels
};
struct GridDev {
  LevelDev lv[kLevels];
};

inline GridDev to_dev(const dfn_hashgrid* g) {
  GridDev d;
  for (int l = 0; l < kLevels; ++l) {
    d.lv[l].scale = g->scale[l];
    d.lv[l].res = g->res[l];
    d.lv[l].size = g->size[l];
    d.lv[l].offset = g->offset[l];
    d.lv[l].hashed = ((unsigned long long)g->res[l] * g->res[l] * g->res[l] > g->size[l]) ? 1u : 0u;
    d.lv[l].magic = (uint32_t)((1ull << 32) / g->size[l]);
  }
  return d;
}

// tcnn grid_index<3>: dense x + y*res + z*res^2 while the stride fits, else the coherent prime hash; then % size
// (size is a power of two on hashed levels).  Both forms are branch-free: lanes of one warp work on different levels.
__device__ __forceinline__ uint32_t index_hashed(uint32_t x, uint32_t y, uint32_t z, const LevelDev& L) {
  return (x ^ (y * kPrimeY) ^ (z * kPrimeZ)) & (L.size - 1u);
}
__device__ __forceinline__ uint32_t index_dense(uint32_t x, uint32_t y, uint32_t z, const LevelDev& L) {
  const uint32_t i = x + y * L.res + z * L.res * L.res;
  uint32_t r = i - __umulhi(i, L.magic) * L.size;  // quotient estimate is exact or one short
  return r >= L.size ? r - L.size : r;
}
__device__ __forceinline__ uint32_t grid_index(uint32_t x, uint32_t y, uint32_t z, const LevelDev& L) {
  return L.hashed ? index_hashed(x, y, z, L) : index_dense(x, y, z, L);
}

struct Cell {
  uint32_t gx, gy, gz;
  float fx, fy, fz;
};
// tcnn pos_fract: pos = fma(scale, x, 0.5); cell = floor(pos); frac = pos - cell
__device__ __forceinline__ Cell locate(float px, float py, float pz, float scale) {
  Cell c;
  float x = __fmaf_rn(scale, px, 0.5f), y = __fmaf_rn(scale, py, 0.5f), z = __fmaf_rn(scale, pz, 0.5f);
  float tx = floorf(x), ty = floorf(y), tz = floorf(z);
  c.gx = (uint32_t)(int)tx, c.gy = (uint32_t)(int)ty, c.gz = (uint32_t)(int)tz;
  c.fx = x - tx, c.fy = y - ty, c.fz = z - tz;
  return c;
}
__device__ __forceinline__ float corner_weight(const Cell& c, int idx) {
  float w = (idx & 1) ? c.fx : 1.0f - c.fx;
  w = __fmul_rn(w, (idx & 2) ? c.fy : 1.0f - c.fy);
  return __fmul_rn(w, (idx & 4) ? c.fz : 1.0f - c.fz);
}

// the 8 corner entries of one level (kGeneric: the table pointer may be shared memory).  kMode 1: the level is hashed for
// every lane of the warp, 0: dense for every lane, 2: mixed (select).
template <bool kGeneric, int kMode>
__device__ __forceinline__ void gather_level(const __half2* __restrict__ tab, const LevelDev& L, const Cell& c, uint32_t (&raw)[8]) {
  uint32_t hy[2], hz[2];
  if (kMode == 1) {
    hy[0] = c.gy * kPrimeY, hy[1] = hy[0] + kPrimeY;
    hz[0] = c.gz * kPrimeZ, hz[1] = hz[0] + kPrimeZ;
  }
  // The corners come in x pairs (x, x+1).  Whenever the two entries are the halves of one aligned 8-byte word (always on
  // a dense level with even x, and on a hashed level with even x because (x+1)^h == (x^h)^1) one 8-byte load fetches
  // both: a quarter fewer L1 wavefronts on average, same values, same accumulation order.
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t y = c.gy + (k & 1), z = c.gz + (k >> 1);
    uint32_t i0, i1;
    if (kMode == 1) {
      const uint32_t h = hy[k & 1] ^ hz[k >> 1];
      i0 = (c.gx ^ h) & (L.size - 1u), i1 = ((c.gx + 1u) ^ h) & (L.size - 1u);
    } else if (kMode == 0) {
      i0 = index_dense(c.gx, y, z, L), i1 = index_dense(c.gx + 1u, y, z, L);
    } else {
      i0 = grid_index(c.gx, y, z, L), i1 = grid_index(c.gx + 1u, y, z, L);
    }
    const uint2* p8 = reinterpret_cast<const uint2*>(tab + (i0 & ~1u));
    const uint2 w = kGeneric ? *p8 : __ldg(p8);
    raw[2 * k] = (i0 & 1u) ? w.y : w.x;
    uint32_t other = (i0 & 1u) ? w.x : w.y;
    if (i1 != (i0 ^ 1u)) {
      const uint32_t* p4 = reinterpret_cast<const uint32_t*>(tab + i1);
      other = kGeneric ? *p4 : __ldg(p4);
    }
    raw[2 * k + 1] = other;
  }
}
template <bool kGeneric>
__device__ __forceinline__ void gather_level_uniform(const __half2* __restrict__ tab, const LevelDev& L, const Cell& c, uint32_t (&raw)[8]) {
  const bool all_hashed = __all_sync(0xffffffffu, L.hashed != 0), all_dense = __all_sync(0xffffffffu, L.hashed == 0);
  if (all_hashed) gather_level<kGeneric, 1>(tab, L, c, raw);
  else if (all_dense) gather_level<kGeneric, 0>(tab, L, c, raw);
  else gather_level<kGeneric, 2>(tab, L, c, raw);
}
// trilinear blend, fp32 accumulate in corner order, one rounding to fp16; packed half2 (low = feature 0)
__device__ __forceinline__ uint32_t blend_level(const Cell& c, const uint32_t (&raw)[8]) {
  float a0 = 0.f, a1 = 0.f;
  const float wxy[4] = {__fmul_rn(1.0f - c.fx, 1.0f - c.fy), __fmul_rn(c.fx, 1.0f - c.fy), __fmul_rn(1.0f - c.fx, c.fy),
                        __fmul_rn(c.fx, c.fy)};
#pragma unroll
  for (int idx = 0; idx < 8; ++idx) {
    const float w = __fmul_rn(wxy[idx & 3], (idx & 4) ? c.fz : 1.0f - c.fz);  // == corner_weight(c, idx)
    const __half2 v = *reinterpret_cast<const __half2*>(&raw[idx]);
    a0 = __fmaf_rn(w, __low2float(v), a0);
    a1 = __fmaf_rn(w, __high2float(v), a1);
  }
  __half2 r = __floats2half2_rn(a0, a1);
  return *reinterpret_cast<uint32_t*>(&r);
}
__device__ __forceinline__ uint32_t encode_level(const __half2* __restrict__ tab, const LevelDev& L, float px, float py, float pz) {
  const Cell c = locate(px, py, pz, L.scale);
  uint32_t raw[8];
  gather_level<false, 2>(tab, L, c, raw);
  return blend_level(c, raw);
}

// tcnn spherical_harmonics.h, degree 4; directions arrive in [0,1] and are mapped to [-1,1]
__device__ __forceinline__ void sh4(float dx, float dy, float dz, float (&o)[16]) {
  const float x = __fmaf_rn(dx, 2.f, -1.f), y = __fmaf_rn(dy, 2.f, -1.f), z = __fmaf_rn(dz, 2.f, -1.f);
  const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
  o[0] = 0.28209479177387814f;
  o[1] = -0.48860251190291987f * y;
  o[2] = 0.48860251190291987f * z;
  o[3] = -0.48860251190291987f * x;
  o[4] = 1.0925484305920792f * xy;
  o[5] = -1.0925484305920792f * yz;
  o[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
  o[7] = -1.0925484305920792f * xz;
  o[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
  o[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
  o[10] = 2.8906114426405538f * xy * z;
  o[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
  o[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
  o[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
  o[14] = 1.4453057213202769f * z * (x2 - y2);
  o[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
}

// index of the view direction of point p when `group` consecutive points share one (samples of a ray): p / group
__device__ __forceinline__ long long dir_index(long long p, int group) {
  if (group == 1) return p;
  return (p >> 32) == 0 ? static_cast<long long>(static_cast<uint32_t>(p) / static_cast<uint32_t>(group)) : p / group;
}

// ---- mma.sync building blocks ----------------------------------------------------------------------------------------
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_h2(float lo, float hi) {
  __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint32_t pack_h2_relu(float lo, float hi) { return pack_h2(fmaxf(lo, 0.f), fmaxf(hi, 0.f)); }

// one bias-free layer on a 16-row tile: c[j] = A[16,K] * W[8j..8j+8, K]^T.  W in shared memory, row stride K+8 halves
// (bank = 4g + t: conflict free).
template <int K, int NT>
__device__ __forceinline__ void layer(const uint32_t (&a)[K / 16][4], const __half* W, int g, int t, float (&c)[NT][4]) {
  constexpr int LD = K + 8;
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    c[j][0] = c[j][1] = c[j][2] = c[j][3] = 0.f;
#pragma unroll
    for (int kb = 0; kb < K / 16; ++kb) {
      const uint32_t* w = reinterpret_cast<const uint32_t*>(W + (8 * j + g) * LD + 16 * kb + 2 * t);
      mma16816(c[j], a[kb], w[0], w[4]);
    }
  }
}
// accumulator fragments of 2*KB n-tiles -> ReLU -> fp16 A fragments of the next layer (KB k-blocks)
template <int KB>
__device__ __forceinline__ void relu_repack(const float (&c)[2 * KB][4], uint32_t (&a)[KB][4]) {
#pragma unroll
  for (int kb = 0; kb < KB; ++kb) {
    a[kb][0] = pack_h2_relu(c[2 * kb][0], c[2 * kb][1]);
    a[kb][1] = pack_h2_relu(c[2 * kb][2], c[2 * kb][3]);
    a[kb][2] = pack_h2_relu(c[2 * kb + 1][0], c[2 * kb + 1][1]);
    a[kb][3] = pack_h2_relu(c[2 * kb + 1][2], c[2 * kb + 1][3]);
  }
}

// shared-memory image of the weights (halves): dmlp W0 [64][40], W1 [16][72]; cmlp W0 [64][40], W1 [64][72], W2 [16][72]
constexpr int kOffD0 = 0, kOffD1 = kOffD0 + 64 * 40, kOffC0 = kOffD1 + 16 * 72, kOffC1 = kOffC0 + 64 * 40,
              kOffC2 = kOffC1 + 64 * 72, kWHalves = kOffC2 + 16 * 72;

__device__ __forceinline__ void stage_weights(__half* dst, const __half* __restrict__ src, int rows, int K) {
  for (int i = threadIdx.x; i < rows * K; i += blockDim.x) dst[(i / K) * (K + 8) + (i % K)] = src[i];
}


}  // namespace ngp
