// Small HBM-bound kernels around the MLP: materialised positional encoding (reference
// scripts/model.py:27-53), dense-grid coordinates (scripts/extract_vol.py:57-82), Otsu histogram and
// u8 packing (scripts/extract_vol.py:123-206), fused AdamW on the flat parameter buffer
// (scripts/train.py:244-259).
#include "dfn_common.cuh"
#include "rng.cuh"
#include <cuda_fp16.h>

namespace {

// one thread per (point, dimension, frequency): sin and cos of the same angle come from ONE sincosf, and no thread does
// 64-bit division (the batch is cut into pieces of < 2^31 / D points by the launcher)
__global__ void posenc_kernel(const float* __restrict__ x, const float* __restrict__ freqs, unsigned P, int L,
                              int include_input, float* __restrict__ out) {
  const unsigned per = 3u * L + (include_input ? 3u : 0u);   // work items per point: 3L angles (+ 3 copies)
  const unsigned D = 6u * L + (include_input ? 3u : 0u);
  const unsigned total = P * per;
  for (unsigned idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const unsigned p = idx / per, c = idx - p * per;
    float* o = out + static_cast<size_t>(p) * D;
    if (include_input && c < 3u) {
      o[c] = x[static_cast<size_t>(p) * 3 + c];
    } else {
      const unsigned j = c - (include_input ? 3u : 0u);       // j = d * L + k
      const unsigned d = j / L, k = j - d * L;
      float sn, cs;
      sincosf(__fmul_rn(x[static_cast<size_t>(p) * 3 + d], freqs[k]), &sn, &cs);
      const unsigned off = include_input ? 3u : 0u;
      o[off + j] = sn;
      o[off + 3u * L + j] = cs;
    }
  }
}

__global__ void grid_points_kernel(const float* __restrict__ lin, int R, long long start, long long count,
                                   float* __restrict__ out) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < count;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long g = start + i;
    const int iz = static_cast<int>(g % R), iy = static_cast<int>((g / R) % R), ix = static_cast<int>(g / (static_cast<long long>(R) * R));
    out[i * 3] = lin[ix];
    out[i * 3 + 1] = lin[iy];
    out[i * 3 + 2] = lin[iz];
  }
}

// 256-bin histogram of values over [lo,hi] (numpy/skimage binning: last bin closed)
__global__ void histogram_kernel(const float* __restrict__ v, long long n, float lo, float hi, int nbins,
                                 unsigned long long* __restrict__ hist) {
  extern __shared__ unsigned int sh[];
  for (int i = threadIdx.x; i < nbins; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const double scale = static_cast<double>(nbins) / (static_cast<double>(hi) - static_cast<double>(lo));
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float x = v[i];
    if (x >= lo && x <= hi) {
      int b = static_cast<int>((static_cast<double>(x) - lo) * scale);
      b = b >= nbins ? nbins - 1 : b;
      atomicAdd(&sh[b], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nbins; i += blockDim.x)
    if (sh[i]) atomicAdd(&hist[i], static_cast<unsigned long long>(sh[i]));
}

__global__ void pack_volume_kernel(const float* __restrict__ sigma, const float* __restrict__ rgb, long long n,
                                   float thr, unsigned char* __restrict__ occ, unsigned char* __restrict__ col) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    occ[i] = sigma[i] > thr ? 1 : 0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      // (rgb * 255).byte(): fp32 multiply then truncation
      const float s = __fmul_rn(rgb[i * 3 + c], 255.0f);
      col[i * 3 + c] = static_cast<unsigned char>(static_cast<int>(s));
    }
  }
}

// torch.optim.AdamW single-tensor update on a flat fp32 buffer; lr/beta1 come from a device table so
// the launch is CUDA-graph replayable (scripts/train.py:244-259 AdamW + OneCycleLR(cycle_momentum)).
__global__ void adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                             float* __restrict__ v, long long n, const float* __restrict__ lr_table,
                             const float* __restrict__ beta1_table, const int* __restrict__ step_ptr, int table_len,
                             float beta2, float eps, float wd, float grad_scale) {
  const int step0 = *step_ptr;  // 0-based optimiser step
  const int ti = step0 < table_len ? step0 : table_len - 1;
  const float lr = lr_table[ti], b1 = beta1_table[ti];
  const float t = static_cast<float>(step0 + 1);
  // bias corrections as torch computes them (1 - beta^step); beta1 varies per step under cycle_momentum,
  // torch uses the current beta1 ** step.
  const float bc1 = 1.0f - powf(b1, t);
  const float bc2 = 1.0f - powf(beta2, t);
  const float step_size = lr / bc1;
  const float bc2_sqrt = sqrtf(bc2);
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float gi = g[i] * grad_scale;
    float pi = p[i] * (1.0f - lr * wd);
    const float mi = b1 * m[i] + (1.0f - b1) * gi;
    const float vi = beta2 * v[i] + (1.0f - beta2) * gi * gi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    pi -= step_size * (mi / denom);
    p[i] = pi; m[i] = mi; v[i] = vi;
  }
}

__global__ void step_inc_kernel(int* step_ptr) { *step_ptr += 1; }

// flag[0] = 1 if any gradient is inf / NaN (flag must be zero on entry): the guard torch.amp.GradScaler puts in front of
// optimizer.step (scripts/train.py:262-323) as one pass over the flat gradient
__global__ void nonfinite_flag_kernel(const float* __restrict__ g, long long n, int* __restrict__ flag) {
  bool bad = false;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float v = g[i];
    bad = bad || !(fabsf(v) <= 3.4028234e38f);
  }
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(flag, 1);
}

// adamw_kernel + an fp16 copy of the updated parameters (the operands the Instant-NGP kernels read) + skip when *skip != 0
__global__ void adamw_shadow_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                    float* __restrict__ v, long long n, const float* __restrict__ lr_table,
                                    const float* __restrict__ beta1_table, const int* __restrict__ step_ptr, int table_len,
                                    float beta2, float eps, float wd, float grad_scale, __half* __restrict__ shadow,
                                    const int* __restrict__ skip) {
  if (skip && *skip != 0) return;
  const int step0 = *step_ptr;
  const int ti = step0 < table_len ? step0 : table_len - 1;
  const float lr = lr_table[ti], b1 = beta1_table[ti];
  const float t = static_cast<float>(step0 + 1);
  const float bc1 = 1.0f - powf(b1, t);
  const float bc2 = 1.0f - powf(beta2, t);
  const float step_size = lr / bc1;
  const float bc2_sqrt = sqrtf(bc2);
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float gi = g[i] * grad_scale;
    float pi = p[i] * (1.0f - lr * wd);
    const float mi = b1 * m[i] + (1.0f - b1) * gi;
    const float vi = beta2 * v[i] + (1.0f - beta2) * gi * gi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    pi -= step_size * (mi / denom);
    p[i] = pi; m[i] = mi; v[i] = vi;
    shadow[i] = __float2half_rn(pi);
  }
}
__global__ void step_inc_unless_kernel(int* step_ptr, const int* skip) { if (!skip || *skip == 0) *step_ptr += 1; }

// the uniforms / normals the in-kernel draws use, written out (known-answer tests of csrc/rng.cuh)
__global__ void philox_dump_kernel(unsigned long long seed, const int* __restrict__ step_ptr, unsigned stream, long long n,
                                   float* __restrict__ uniforms, float* __restrict__ normals) {
  const uint32_t step = static_cast<uint32_t>(*step_ptr);
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float4 r = rng::uniform4(seed, step, stream, static_cast<unsigned long long>(i));
    if (uniforms) reinterpret_cast<float4*>(uniforms)[i] = r;
    if (normals) normals[i] = rng::normal1(seed, step, stream, static_cast<unsigned long long>(i));
  }
}

inline int grid1d(long long n, int block) {
  long long g = (n + block - 1) / block;
  const long long cap = static_cast<long long>(dfn_num_sms()) * 8;
  return static_cast<int>(g < 1 ? 1 : (g < cap ? g : cap));
}

}  // namespace

extern "C" int dfn_posenc(const float* x, const float* freqs, long long P, int L, int include_input, float* out,
                          void* stream) {
  DFN_CHECK_ARG(x && freqs && out, "posenc: null pointer");
  DFN_CHECK_ARG(L >= 0 && L <= 32, "posenc: L=%d out of range", L);
  const int D = (include_input ? 3 : 0) + 6 * L;
  if (P == 0 || D == 0) return DFN_OK;
  const long long per = 3LL * L + (include_input ? 3 : 0);
  const long long piece = (1LL << 31) / per;     // points per launch: the kernel indexes its work items with 32 bits
  for (long long p0 = 0; p0 < P; p0 += piece) {
    const long long n = P - p0 < piece ? P - p0 : piece;
    posenc_kernel<<<grid1d(n * per, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x + p0 * 3, freqs, static_cast<unsigned>(n), L,
                                                                                         include_input, out + p0 * D);
    DFN_CHECK_LAUNCH("posenc");
  }
  return DFN_OK;
}

extern "C" int dfn_grid_points(const float* lin, int R, long long start, long long count, float* out, void* stream) {
  DFN_CHECK_ARG(lin && out && R > 0, "grid_points: bad args");
  DFN_CHECK_ARG(start >= 0 && start + count <= static_cast<long long>(R) * R * R, "grid_points: range outside R^3");
  if (count == 0) return DFN_OK;
  grid_points_kernel<<<grid1d(count, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(lin, R, start, count, out);
  DFN_CHECK_LAUNCH("grid_points");
  return DFN_OK;
}

extern "C" int dfn_histogram(const float* v, long long n, float lo, float hi, int nbins, unsigned long long* hist,
                             void* stream) {
  DFN_CHECK_ARG(v && hist && nbins > 0 && nbins <= 4096 && hi > lo, "histogram: bad args");
  if (n == 0) return DFN_OK;
  histogram_kernel<<<grid1d(n, 256), 256, nbins * sizeof(unsigned int), static_cast<cudaStream_t>(stream)>>>(
      v, n, lo, hi, nbins, hist);
  DFN_CHECK_LAUNCH("histogram");
  return DFN_OK;
}

extern "C" int dfn_pack_volume(const float* sigma, const float* rgb, long long n, float thr, unsigned char* occ,
                               unsigned char* col, void* stream) {
  DFN_CHECK_ARG(sigma && rgb && occ && col, "pack_volume: null pointer");
  if (n == 0) return DFN_OK;
  pack_volume_kernel<<<grid1d(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(sigma, rgb, n, thr, occ, col);
  DFN_CHECK_LAUNCH("pack_volume");
  return DFN_OK;
}

extern "C" int dfn_adamw_step(float* p, const float* g, float* m, float* v, long long n, const float* lr_table,
                              const float* beta1_table, int* step_ptr, int table_len, float beta2, float eps,
                              float weight_decay, float grad_scale, void* stream) {
  DFN_CHECK_ARG(p && g && m && v && lr_table && beta1_table && step_ptr && table_len > 0, "adamw_step: bad args");
  if (n == 0) return DFN_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  adamw_kernel<<<grid1d(n, 256), 256, 0, st>>>(p, g, m, v, n, lr_table, beta1_table, step_ptr, table_len, beta2, eps,
                                               weight_decay, grad_scale);
  step_inc_kernel<<<1, 1, 0, st>>>(step_ptr);
  DFN_CHECK_LAUNCH("adamw_step");
  return DFN_OK;
}

extern "C" int dfn_philox_dump(unsigned long long seed, const int* step_ptr, unsigned rng_stream, long long n_items, float* uniforms,
                               float* normals, void* stream) {
  DFN_CHECK_ARG(step_ptr && (uniforms || normals), "philox_dump: null pointer");
  if (n_items == 0) return DFN_OK;
  philox_dump_kernel<<<grid1d(n_items, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(seed, step_ptr, rng_stream, n_items, uniforms, normals);
  DFN_CHECK_LAUNCH("philox_dump");
  return DFN_OK;
}

// AdamW(+OneCycleLR tables) on a flat fp32 buffer that also refreshes an fp16 copy of the parameters (what the Instant-NGP
// kernels read: tcnn casts its fp32 master parameters to fp16 per call).  skip_flag (nullable, device int): the update and the
// step increment are skipped when it is non-zero - set it with dfn_nonfinite_flag, the guard GradScaler.step applies.
extern "C" int dfn_adamw_step_shadow(float* p, const float* g, float* m, float* v, long long n, const float* lr_table,
                                     const float* beta1_table, int* step_ptr, int table_len, float beta2, float eps,
                                     float weight_decay, float grad_scale, void* shadow_f16, const int* skip_flag, void* stream) {
  DFN_CHECK_ARG(p && g && m && v && lr_table && beta1_table && step_ptr && shadow_f16 && table_len > 0, "adamw_step_shadow: bad args");
  if (n == 0) return DFN_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  adamw_shadow_kernel<<<grid1d(n, 256), 256, 0, st>>>(p, g, m, v, n, lr_table, beta1_table, step_ptr, table_len, beta2, eps,
                                                      weight_decay, grad_scale, static_cast<__half*>(shadow_f16), skip_flag);
  step_inc_unless_kernel<<<1, 1, 0, st>>>(step_ptr, skip_flag);
  DFN_CHECK_LAUNCH("adamw_step_shadow");
  return DFN_OK;
}

extern "C" int dfn_nonfinite_flag(const float* g, long long n, int* flag, void* stream) {
  DFN_CHECK_ARG(g && flag, "nonfinite_flag: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DFN_CUDA(cudaMemsetAsync(flag, 0, sizeof(int), st), "nonfinite_flag");
  if (n == 0) return DFN_OK;
  nonfinite_flag_kernel<<<grid1d(n, 256), 256, 0, st>>>(g, n, flag);
  DFN_CHECK_LAUNCH("nonfinite_flag");
  return DFN_OK;
}
