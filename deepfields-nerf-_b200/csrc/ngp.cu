// Instant-NGP variant (scripts/instantngp.py:7-104): multiresolution hash-grid encode + its scatter-add backward,
// SH degree-4 direction encoding, and the two bias-free tiny MLPs (32->64->16 ReLU/ReLU, 32->64->64->3 ReLU/Sigmoid).
//
// The arithmetic the reference delegates to tiny-cuda-nn (un-vendored, commit 74ea173d) is restated from its published
// algorithm (grid.h / common_device.h / spherical_harmonics.h / fully_fused_mlp.cu): parity is UNPINNED.
// Documented choices where tcnn cannot be checked: the interpolation and the MLP accumulate in fp32 (tcnn: fp16) and
// round to fp16 once per level / per layer.
//
// Kernels
//   ngp_forward_kernel     inference: ONE launch, nothing but pos/dir in and rgb/features out touches HBM.  A warp owns
//                          32 points; lane (g = lane/4, t = lane%4) encodes levels {t, t+4, t+8, t+12} of points
//                          {g, g+8, g+16, g+24}, which IS the A-fragment layout of mma.m16n8k16, so the encoding never
//                          leaves registers; each layer's accumulator fragments are re-packed into the next layer's A
//                          fragments in registers too.  Weights (fp16) and the two smallest dense levels sit in shared
//                          memory.  These 19 kFLOP/point MLPs are far below the tensor roofline, the kernel is bound by the
//                          128 table gathers per point; that is why they run on mma.sync here rather than tcgen05.
//   hashgrid_forward_kernel / hashgrid_backward_kernel   training path: thread per (point, level), level fastest, so the
//                          [P,32] fp16 rows are written / read fully coalesced; backward = 8 float2 red.add per thread.
//   sh4_kernel             direction encoding for the training path.
#include "ngp_common.cuh"

namespace {
using namespace ngp;

// ---- training-path kernels -------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) hashgrid_forward_kernel(const float* __restrict__ pos, long long P,
                                                               const __half2* __restrict__ table,
                                                               const __grid_constant__ GridDev grid, uint32_t* __restrict__ out) {
  __shared__ LevelDev lv[kLevels];
  if (threadIdx.x < kLevels) lv[threadIdx.x] = grid.lv[threadIdx.x];
  __syncthreads();
  const long long total = P * kLevels;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long p = i / kLevels;
    const int l = (int)(i % kLevels);
    const LevelDev L = lv[l];
    out[i] = encode_level(table + L.offset, L, __ldg(pos + 3 * p), __ldg(pos + 3 * p + 1), __ldg(pos + 3 * p + 2));
  }
}

__global__ void __launch_bounds__(256) hashgrid_backward_kernel(const float* __restrict__ pos, long long P,
                                                                const uint32_t* __restrict__ d_out,
                                                                const __grid_constant__ GridDev grid, float grad_scale,
                                                                int first_level, float2* __restrict__ d_table) {
  __shared__ LevelDev lv[kLevels];
  if (threadIdx.x < kLevels) lv[threadIdx.x] = grid.lv[threadIdx.x];
  __syncthreads();
  const long long total = P * kLevels;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long p = i / kLevels;
    const int l = (int)(i % kLevels);
    if (l < first_level) continue;  // those levels are reduced in shared memory by hashgrid_backward_coarse_kernel
    const LevelDev L = lv[l];
    const uint32_t raw = __ldg(d_out + i);
    const __half2 dv = *reinterpret_cast<const __half2*>(&raw);
    const float d0 = __low2float(dv) * grad_scale, d1 = __high2float(dv) * grad_scale;
    if (d0 == 0.f && d1 == 0.f) continue;
    const Cell c = locate(__ldg(pos + 3 * p), __ldg(pos + 3 * p + 1), __ldg(pos + 3 * p + 2), L.scale);
    float2* base = d_table + L.offset;
#pragma unroll
    for (int k = 0; k < 4; ++k) {  // x pairs: one 16-byte vector reduction when both entries share an aligned float4
      const uint32_t y = c.gy + (k & 1), z = c.gz + (k >> 1);
      const uint32_t i0 = grid_index(c.gx, y, z, L), i1 = grid_index(c.gx + 1u, y, z, L);
      const float w0 = corner_weight(c, 2 * k), w1 = corner_weight(c, 2 * k + 1);
      if (i1 == (i0 ^ 1u)) {
        const bool odd = i0 & 1u;
        const float wa = odd ? w1 : w0, wb = odd ? w0 : w1;  // wa: entry i0 & ~1, wb: the next one
        atomicAdd(reinterpret_cast<float4*>(base + (i0 & ~1u)), make_float4(wa * d0, wa * d1, wb * d0, wb * d1));
      } else {
        atomicAdd(base + i0, make_float2(w0 * d0, w0 * d1));  // red.global.add.v2.f32
        atomicAdd(base + i1, make_float2(w1 * d0, w1 * d1));
      }
    }
  }
}

// The smallest dense levels (level 0: 4 096 entries, level 1: 13 824) take 8 atomics per point each: in global memory
// that is thousands of reductions per address (measured: level 0 alone 17 % of the backward).  One persistent CTA per SM
// accumulates them in shared memory and flushes once with 16-byte vector reductions.
__global__ void __launch_bounds__(1024, 1) hashgrid_backward_coarse_kernel(const float* __restrict__ pos, long long P,
                                                                           const uint32_t* __restrict__ d_out,
                                                                           const __grid_constant__ GridDev grid, float grad_scale,
                                                                           int n_lv, uint32_t n_entries, float2* __restrict__ d_table) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* acc = reinterpret_cast<float2*>(smem_raw);
  __shared__ LevelDev lv[kLevels];
  if (threadIdx.x < kLevels) lv[threadIdx.x] = grid.lv[threadIdx.x];
  for (uint32_t i = threadIdx.x; i < n_entries; i += blockDim.x) acc[i] = make_float2(0.f, 0.f);
  __syncthreads();
  const long long total = P * n_lv;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long p = i / n_lv;
    const int l = (int)(i % n_lv);
    const LevelDev L = lv[l];
    const uint32_t raw = __ldg(d_out + p * kLevels + l);
    const __half2 dv = *reinterpret_cast<const __half2*>(&raw);
    const float d0 = __low2float(dv) * grad_scale, d1 = __high2float(dv) * grad_scale;
    if (d0 == 0.f && d1 == 0.f) continue;
    const Cell c = locate(__ldg(pos + 3 * p), __ldg(pos + 3 * p + 1), __ldg(pos + 3 * p + 2), L.scale);
#pragma unroll
    for (int idx = 0; idx < 8; ++idx) {
      const uint32_t j = L.offset + index_dense(c.gx + (idx & 1), c.gy + ((idx >> 1) & 1), c.gz + ((idx >> 2) & 1), L);
      const float w = corner_weight(c, idx);
      atomicAdd(&acc[j].x, w * d0);
      atomicAdd(&acc[j].y, w * d1);
    }
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < n_entries / 2; i += blockDim.x) {  // level sizes are multiples of 8 entries
    const float4 v = reinterpret_cast<const float4*>(acc)[i];
    if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f) atomicAdd(reinterpret_cast<float4*>(d_table) + i, v);
  }
}

__global__ void __launch_bounds__(256) sh4_kernel(const float* __restrict__ dirs, long long P, __half* __restrict__ out) {
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < P; p += (long long)gridDim.x * blockDim.x) {
    float o[16];
    sh4(__ldg(dirs + 3 * p), __ldg(dirs + 3 * p + 1), __ldg(dirs + 3 * p + 2), o);
    uint32_t w[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      __half2 h = __floats2half2_rn(o[2 * k], o[2 * k + 1]);
      w[k] = *reinterpret_cast<uint32_t*>(&h);
    }
    st_global_v4(out + 16 * p, w[0], w[1], w[2], w[3]);
    st_global_v4(out + 16 * p + 8, w[4], w[5], w[6], w[7]);
  }
}

#ifndef DFN_NGP_THREADS
#define DFN_NGP_THREADS 256  // bring-up override: -DDFN_NGP_THREADS=320 trades registers (96) for warps (20 per SM)
#endif
constexpr int kNgpThreads = DFN_NGP_THREADS, kNgpWarps = kNgpThreads / 32;

__global__ void __launch_bounds__(kNgpThreads, 2)
ngp_forward_kernel(const float* __restrict__ pos, const float* __restrict__ dirs, int dirs_stride, int dirs_group, long long P,
                   const __half2* __restrict__ table, const __grid_constant__ GridDev grid,
                   const __half* __restrict__ dmlp_w, const __half* __restrict__ cmlp_w, int smem_levels,
                   __half* __restrict__ features, __half* __restrict__ rgb, float* __restrict__ rgb32, float* __restrict__ sigma32) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __half* W = reinterpret_cast<__half*>(smem_raw);
  LevelDev* lv = reinterpret_cast<LevelDev*>(smem_raw + kWHalves * 2);
  uint32_t* coarse = reinterpret_cast<uint32_t*>(smem_raw + kWHalves * 2 + kLevels * sizeof(LevelDev));

  stage_weights(W + kOffD0, dmlp_w, 64, 32);
  stage_weights(W + kOffD1, dmlp_w + 64 * 32, 16, 64);
  stage_weights(W + kOffC0, cmlp_w, 64, 32);
  stage_weights(W + kOffC1, cmlp_w + 64 * 32, 64, 64);
  stage_weights(W + kOffC2, cmlp_w + 64 * 32 + 64 * 64, 16, 64);
  if (threadIdx.x < kLevels) lv[threadIdx.x] = grid.lv[threadIdx.x];
  // the smallest dense levels (level 0: 16 KB, level 1: 54 KB) are gathered from shared memory
  uint32_t coarse_entries = 0;
  for (int l = 0; l < smem_levels; ++l) coarse_entries += grid.lv[l].size;
  for (uint32_t i = threadIdx.x; i < coarse_entries; i += blockDim.x) coarse[i] = reinterpret_cast<const uint32_t*>(table)[i];
  __syncthreads();

  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const long long n_tiles = (P + 31) / 32;
  const long long warp_global = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long warp_stride = (long long)gridDim.x * (blockDim.x >> 5);

  for (long long tile = warp_global; tile < n_tiles; tile += warp_stride) {
    const long long p0 = tile * 32;
    // ---- hash-grid encode straight into A fragments: enc[m][kb][r] ----
    uint32_t enc[2][2][4];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
      for (int h = 0; h < 2; ++h) {  // rows g (+8)
        long long p = p0 + 16 * m + 8 * h + g;
        p = p < P ? p : P - 1;
        const float px = __ldg(pos + 3 * p), py = __ldg(pos + 3 * p + 1), pz = __ldg(pos + 3 * p + 2);
        // the lane's four levels 8kb + 4q + t of this point: all 32 gathers are issued before the first blend
        Cell cell[4];
        uint32_t raw[4][8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int l = 8 * (j >> 1) + 4 * (j & 1) + t;
          const LevelDev L = lv[l];
          cell[j] = locate(px, py, pz, L.scale);
          if (j == 0) {  // only levels 0..3 can be shared-memory resident
            const __half2* tab = (l < smem_levels) ? reinterpret_cast<const __half2*>(coarse) + L.offset : table + L.offset;
            gather_level_uniform<true>(tab, L, cell[j], raw[j]);
          } else {
            gather_level_uniform<false>(table + L.offset, L, cell[j], raw[j]);
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) enc[m][j >> 1][2 * (j & 1) + h] = blend_level(cell[j], raw[j]);
      }

#pragma unroll
    for (int m = 0; m < 2; ++m) {
      const long long pr0 = p0 + 16 * m + g, pr1 = pr0 + 8;
      uint32_t cin[2][4];  // colour MLP input: k-block 0 = the 16 features, k-block 1 = SH
      {
        // ---- density MLP 32 -> 64 -> 16 ----
        float c1[8][4];
        layer<32, 8>(enc[m], W + kOffD0, g, t, c1);
        uint32_t a1[4][4];
        relu_repack<4>(c1, a1);
        float c2[2][4];
        layer<64, 2>(a1, W + kOffD1, g, t, c2);
        uint32_t f[1][4];
        relu_repack<1>(c2, f);  // output activation ReLU (instantngp.py:47); sigma = feature 0 (:79)
#pragma unroll
        for (int r = 0; r < 4; ++r) cin[0][r] = f[0][r];
        if (sigma32 != nullptr && t == 0) {   // sigma = feature 0 (fp16 value, stored as fp32 for the compositing kernels)
          if (pr0 < P) sigma32[pr0] = __low2float(*reinterpret_cast<const __half2*>(&f[0][0]));
          if (pr1 < P) sigma32[pr1] = __low2float(*reinterpret_cast<const __half2*>(&f[0][1]));
        }
        if (features != nullptr) {
          if (pr0 < P) {
            *reinterpret_cast<uint32_t*>(features + 16 * pr0 + 2 * t) = f[0][0];
            *reinterpret_cast<uint32_t*>(features + 16 * pr0 + 8 + 2 * t) = f[0][2];
          }
          if (pr1 < P) {
            *reinterpret_cast<uint32_t*>(features + 16 * pr1 + 2 * t) = f[0][1];
            *reinterpret_cast<uint32_t*>(features + 16 * pr1 + 8 + 2 * t) = f[0][3];
          }
        }
      }
      if (rgb == nullptr && rgb32 == nullptr) continue;
      // ---- SH degree 4 of the two rows, columns 2t, 2t+1, 2t+8, 2t+9 ----
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        long long p = (h ? pr1 : pr0);
        p = p < P ? p : P - 1;
        const float* d = dirs + (long long)dirs_stride * dir_index(p, dirs_group);
        float o[16];
        sh4(__ldg(d), __ldg(d + 1), __ldg(d + 2), o);
        float s0 = o[0], s1 = o[1], s2 = o[8], s3 = o[9];
        if (t == 1) s0 = o[2], s1 = o[3], s2 = o[10], s3 = o[11];
        if (t == 2) s0 = o[4], s1 = o[5], s2 = o[12], s3 = o[13];
        if (t == 3) s0 = o[6], s1 = o[7], s2 = o[14], s3 = o[15];
        cin[1][h] = pack_h2(s0, s1);
        cin[1][2 + h] = pack_h2(s2, s3);
      }
      // ---- colour MLP 32 -> 64 -> 64 -> 3 (padded to 16 rows; only n-tile 0 is needed) ----
      uint32_t a2[4][4];
      {
        float c1[8][4];
        layer<32, 8>(cin, W + kOffC0, g, t, c1);
        uint32_t a1[4][4];
        relu_repack<4>(c1, a1);
        float c2[8][4];
        layer<64, 8>(a1, W + kOffC1, g, t, c2);
        relu_repack<4>(c2, a2);
      }
      float c3[1][4];
      layer<64, 1>(a2, W + kOffC2, g, t, c3);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const long long p = h ? pr1 : pr0;
        if (p >= P) continue;
        const float v0 = 1.0f / (1.0f + __expf(-c3[0][2 * h])), v1 = 1.0f / (1.0f + __expf(-c3[0][2 * h + 1]));
        if (t == 0) {
          if (rgb) { rgb[3 * p] = __float2half_rn(v0); rgb[3 * p + 1] = __float2half_rn(v1); }
          if (rgb32) { rgb32[3 * p] = __half2float(__float2half_rn(v0)); rgb32[3 * p + 1] = __half2float(__float2half_rn(v1)); }
        } else if (t == 1) {
          if (rgb) rgb[3 * p + 2] = __float2half_rn(v0);
          if (rgb32) rgb32[3 * p + 2] = __half2float(__float2half_rn(v0));
        }
      }
    }
  }
}

}  // namespace

// ---- C ABI -------------------------------------------------------------------------------------------------------------
extern "C" long long dfn_hashgrid_init(dfn_hashgrid* g, int log2_hashmap_size) {
  if (g == nullptr || log2_hashmap_size < 4 || log2_hashmap_size > 24) {
    dfn_set_error("dfn_hashgrid_init: log2_hashmap_size %d outside [4, 24]", log2_hashmap_size);
    return -1;
  }
  // instantngp.py:23-33: 16 levels, 2 features, base resolution 16, per-level scale 1.5
  // tcnn grid_scale = exp2(level * log2(1.5)) * 16 - 1, evaluated by tcnn with the device's exp2f (not reproducible
  // offline); 16 * 1.5^level - 1 is exactly representable in fp32 for every level here, so the exact value is used.
  unsigned long long off = 0;
  float growth = 1.0f;
  for (int l = 0; l < kLevels; ++l, growth *= 1.5f) {
    const float scale = growth * 16.0f - 1.0f;
    const unsigned res = (unsigned)ceilf(scale) + 1u;                 // tcnn grid_resolution
    unsigned long long n = (unsigned long long)res * res * res;
    n = (n + 7ull) / 8ull * 8ull;
    if (n > (1ull << log2_hashmap_size)) n = 1ull << log2_hashmap_size;
    g->scale[l] = scale;
    g->res[l] = res;
    g->size[l] = (unsigned)n;
    g->offset[l] = (unsigned)off;
    off += n;
  }
  g->n_entries = (long long)off;
  g->log2_hashmap_size = log2_hashmap_size;
  return (long long)off;
}

static int check_grid(const dfn_hashgrid* g, const char* who) {
  DFN_CHECK_ARG(g != nullptr && g->n_entries > 0, "%s: grid descriptor not initialised (dfn_hashgrid_init)", who);
  return DFN_OK;
}
static int blocks_for(long long work, int per_sm) {
  long long b = (work + 255) / 256;
  long long cap = (long long)dfn_num_sms() * per_sm;
  return (int)(b < cap ? (b < 1 ? 1 : b) : cap);
}

extern "C" int dfn_hashgrid_forward(const float* pos, long long P, const void* table_f16, const dfn_hashgrid* g, void* out_f16,
                                    void* stream) {
  if (int rc = check_grid(g, "dfn_hashgrid_forward")) return rc;
  DFN_CHECK_ARG(P >= 0, "dfn_hashgrid_forward: P < 0");
  if (P == 0) return DFN_OK;
  DFN_CHECK_ARG(pos && table_f16 && out_f16, "dfn_hashgrid_forward: null pointer");
  DFN_CHECK_ARG(((uintptr_t)table_f16 & 7) == 0, "dfn_hashgrid_forward: table must be 8-byte aligned");
  hashgrid_forward_kernel<<<blocks_for(P * kLevels, 8), 256, 0, (cudaStream_t)stream>>>(
      pos, P, (const __half2*)table_f16, to_dev(g), (uint32_t*)out_f16);
  DFN_CHECK_LAUNCH("dfn_hashgrid_forward");
  return DFN_OK;
}

extern "C" int dfn_hashgrid_backward(const float* pos, long long P, const void* d_out_f16, const dfn_hashgrid* g, float grad_scale,
                                     float* d_table, void* stream) {
  if (int rc = check_grid(g, "dfn_hashgrid_backward")) return rc;
  DFN_CHECK_ARG(P >= 0, "dfn_hashgrid_backward: P < 0");
  if (P == 0) return DFN_OK;
  DFN_CHECK_ARG(pos && d_out_f16 && d_table, "dfn_hashgrid_backward: null pointer");
  DFN_CHECK_ARG(((uintptr_t)d_table & 15) == 0, "dfn_hashgrid_backward: d_table must be 16-byte aligned");
  // dense levels whose fp32 gradient fits one SM's shared memory (level 0 and 1 of the reference configuration)
  int n_lv = 0;
  unsigned coarse_entries = 0;
  while (n_lv < kLevels && (unsigned long long)g->res[n_lv] * g->res[n_lv] * g->res[n_lv] <= g->size[n_lv] &&
         8ull * (coarse_entries + g->size[n_lv]) <= 160 * 1024)
    coarse_entries += g->size[n_lv++];
  if (n_lv > 0 && P >= 4096) {
    // (set on every launch: the attribute belongs to the function in the CURRENT device's context)
    DFN_CUDA(cudaFuncSetAttribute(hashgrid_backward_coarse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024),
             "dfn_hashgrid_backward");
    long long nb = (P * n_lv + 1023) / 1024;
    if (nb > dfn_num_sms()) nb = dfn_num_sms();
    hashgrid_backward_coarse_kernel<<<(int)nb, 1024, 8ull * coarse_entries, (cudaStream_t)stream>>>(
        pos, P, (const uint32_t*)d_out_f16, to_dev(g), grad_scale, n_lv, coarse_entries, (float2*)d_table);
    DFN_CHECK_LAUNCH("dfn_hashgrid_backward (coarse levels)");
  } else {
    n_lv = 0;
  }
  hashgrid_backward_kernel<<<blocks_for(P * kLevels, 8), 256, 0, (cudaStream_t)stream>>>(
      pos, P, (const uint32_t*)d_out_f16, to_dev(g), grad_scale, n_lv, (float2*)d_table);
  DFN_CHECK_LAUNCH("dfn_hashgrid_backward");
  return DFN_OK;
}

extern "C" int dfn_sh4_encode(const float* dirs, long long P, void* out_f16, void* stream) {
  DFN_CHECK_ARG(P >= 0, "dfn_sh4_encode: P < 0");
  if (P == 0) return DFN_OK;
  DFN_CHECK_ARG(dirs && out_f16, "dfn_sh4_encode: null pointer");
  sh4_kernel<<<blocks_for(P, 8), 256, 0, (cudaStream_t)stream>>>(dirs, P, (__half*)out_f16);
  DFN_CHECK_LAUNCH("dfn_sh4_encode");
  return DFN_OK;
}

extern "C" int dfn_ngp_forward(const float* pos, const float* dirs, int dirs_stride, long long P, const void* table_f16,
                               const dfn_hashgrid* g, const void* dmlp_w_f16, const void* cmlp_w_f16, void* features_f16,
                               void* rgb_f16, void* stream) {
  if (int rc = check_grid(g, "dfn_ngp_forward")) return rc;
  DFN_CHECK_ARG(P >= 0, "dfn_ngp_forward: P < 0");
  if (P == 0) return DFN_OK;
  DFN_CHECK_ARG(pos && table_f16 && dmlp_w_f16 && cmlp_w_f16, "dfn_ngp_forward: null pointer");
  DFN_CHECK_ARG(((uintptr_t)table_f16 & 7) == 0, "dfn_ngp_forward: table must be 8-byte aligned");
  DFN_CHECK_ARG(rgb_f16 == nullptr || dirs != nullptr, "dfn_ngp_forward: rgb requested without view directions");
  DFN_CHECK_ARG(dirs_stride == 0 || dirs_stride == 3, "dfn_ngp_forward: dirs_stride must be 3 (per point) or 0 (one view)");
  DFN_CHECK_ARG(features_f16 || rgb_f16, "dfn_ngp_forward: no output requested");
  // dense levels that fit the shared-memory budget (two CTAs per SM): level 0 and 1 of the reference configuration
  int smem_levels = 0;
  size_t coarse_bytes = 0;
  while (smem_levels < kLevels && (unsigned long long)g->res[smem_levels] * g->res[smem_levels] * g->res[smem_levels] <= g->size[smem_levels] &&
         coarse_bytes + 4ull * g->size[smem_levels] <= 80 * 1024)
    coarse_bytes += 4ull * g->size[smem_levels++];
  const size_t smem = kWHalves * 2 + kLevels * sizeof(LevelDev) + coarse_bytes;
  DFN_CUDA(cudaFuncSetAttribute(ngp_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024), "dfn_ngp_forward");
  const long long tiles = (P + 31) / 32;
  long long blocks = (tiles + kNgpWarps - 1) / kNgpWarps;
  const long long cap = 2ll * dfn_num_sms();
  if (blocks > cap) blocks = cap;
  ngp_forward_kernel<<<(int)blocks, kNgpThreads, smem, (cudaStream_t)stream>>>(pos, dirs, dirs_stride, 1, P, (const __half2*)table_f16,
                                                                        to_dev(g), (const __half*)dmlp_w_f16,
                                                                        (const __half*)cmlp_w_f16, smem_levels,
                                                                        (__half*)features_f16, (__half*)rgb_f16, nullptr, nullptr);
  DFN_CHECK_LAUNCH("dfn_ngp_forward");
  return DFN_OK;
}

// the same query with fp32 outputs for the compositing kernels of the train step: rgb [P,3] and sigma [P] hold the fp16 values
extern "C" int dfn_ngp_forward_f32(const float* pos, const float* dirs, int dirs_stride, int dirs_group, long long P, const void* table_f16,
                                   const dfn_hashgrid* g, const void* dmlp_w_f16, const void* cmlp_w_f16, float* rgb, float* sigma,
                                   void* stream) {
  if (int rc = check_grid(g, "dfn_ngp_forward_f32")) return rc;
  DFN_CHECK_ARG(P >= 0, "dfn_ngp_forward_f32: P < 0");
  if (P == 0) return DFN_OK;
  DFN_CHECK_ARG(pos && dirs && table_f16 && dmlp_w_f16 && cmlp_w_f16 && rgb && sigma, "dfn_ngp_forward_f32: null pointer");
  DFN_CHECK_ARG(((uintptr_t)table_f16 & 7) == 0, "dfn_ngp_forward_f32: table must be 8-byte aligned");
  DFN_CHECK_ARG(dirs_stride == 0 || dirs_stride == 3, "dfn_ngp_forward_f32: dirs_stride must be 3 (per point) or 0 (one view)");
  DFN_CHECK_ARG(dirs_group >= 1, "dfn_ngp_forward_f32: dirs_group must be >= 1");
  int smem_levels = 0;
  size_t coarse_bytes = 0;
  while (smem_levels < kLevels && (unsigned long long)g->res[smem_levels] * g->res[smem_levels] * g->res[smem_levels] <= g->size[smem_levels] &&
         coarse_bytes + 4ull * g->size[smem_levels] <= 80 * 1024)
    coarse_bytes += 4ull * g->size[smem_levels++];
  const size_t smem = kWHalves * 2 + kLevels * sizeof(LevelDev) + coarse_bytes;
  DFN_CUDA(cudaFuncSetAttribute(ngp_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024), "dfn_ngp_forward_f32");
  const long long tiles = (P + 31) / 32;
  long long blocks = (tiles + kNgpWarps - 1) / kNgpWarps;
  const long long cap = 2ll * dfn_num_sms();
  if (blocks > cap) blocks = cap;
  ngp_forward_kernel<<<(int)blocks, kNgpThreads, smem, (cudaStream_t)stream>>>(pos, dirs, dirs_stride, dirs_group, P, (const __half2*)table_f16,
                                                                        to_dev(g), (const __half*)dmlp_w_f16,
                                                                        (const __half*)cmlp_w_f16, smem_levels, nullptr, nullptr, rgb,
                                                                        sigma);
  DFN_CHECK_LAUNCH("dfn_ngp_forward_f32");
  return DFN_OK;
}
