// Ray sampling kernels: stratified depths (reference scripts/sampling.py:10-32), inverse-CDF
// hierarchical sampling with a fused per-ray pdf/cdf build and searchsorted(right=True)
// (scripts/train.py:31-75), the coarse+fine depth merge (scripts/sampling.py:40) and point emission
// pts = o + d*z (scripts/sampling.py:24-31, :43-48).  All arithmetic that must be bit-identical to the
// reference's separate torch ops uses explicit round-to-nearest intrinsics (no FMA contraction).
#include "dfn_common.cuh"
#include <math_constants.h>

namespace {

// ------------------------------------------------------------------------------------------------
// stratified z: one thread per (ray, sample)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float lerp_nf(float near, float far, float t) {
  return __fadd_rn(__fmul_rn(near, __fsub_rn(1.0f, t)), __fmul_rn(far, t));
}

__global__ void stratified_kernel(const float* __restrict__ near, const float* __restrict__ far,
                                  const float* __restrict__ t_vals, const float* __restrict__ t_rand, int B, int S,
                                  float* __restrict__ z_out) {
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= static_cast<long long>(B) * S) return;
  const int ray = static_cast<int>(idx / S), i = static_cast<int>(idx % S);
  const float n = near[ray], f = far[ray];
  const float zi = lerp_nf(n, f, t_vals[i]);
  float lo = zi, hi = zi;
  if (i > 0) lo = __fmul_rn(0.5f, __fadd_rn(zi, lerp_nf(n, f, t_vals[i - 1])));
  if (i + 1 < S) hi = __fmul_rn(0.5f, __fadd_rn(lerp_nf(n, f, t_vals[i + 1]), zi));
  z_out[idx] = __fadd_rn(lo, __fmul_rn(__fsub_rn(hi, lo), t_rand[idx]));
}

// pts / dirs emission: one thread per point
__global__ void points_kernel(const float* __restrict__ o, const float* __restrict__ d, const float* __restrict__ z,
                              int B, int S, float* __restrict__ pts, float* __restrict__ dirs) {
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= static_cast<long long>(B) * S) return;
  const int ray = static_cast<int>(idx / S);
  const float zz = z[idx];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float dc = d[ray * 3 + c];
    if (pts) pts[idx * 3 + c] = __fadd_rn(o[ray * 3 + c], __fmul_rn(dc, zz));
    if (dirs) dirs[idx * 3 + c] = dc;
  }
}

// ------------------------------------------------------------------------------------------------
// sample_pdf: one warp per ray.  bins [B,NB], weights [B,NB-1] (or a ready cdf [B,NB]), u [B,NS]
// ------------------------------------------------------------------------------------------------
constexpr int kPdfWarps = 4;
constexpr int kMaxBins = 256;

template <bool HAVE_CDF>
__global__ void __launch_bounds__(kPdfWarps * 32) sample_pdf_kernel(
    const float* __restrict__ bins, const float* __restrict__ weights, const float* __restrict__ cdf_in,
    const float* __restrict__ u, int B, int NB, int NS, long long w_stride, int w_offset, int bins_mid,
    float* __restrict__ samples, long long* __restrict__ inds_out, float* __restrict__ cdf_out) {
  __shared__ float s_cdf[kPdfWarps][kMaxBins];
  __shared__ float s_bins[kPdfWarps][kMaxBins];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* cdf = s_cdf[warp];
  float* bn = s_bins[warp];
  for (int ray = blockIdx.x * kPdfWarps + warp; ray < B; ray += gridDim.x * kPdfWarps) {
    if (bins_mid) {  // bins = mid-points of z [B,NB+1]  (scripts/train.py:157)
      const float* zr = bins + static_cast<size_t>(ray) * (NB + 1);
      for (int i = lane; i < NB; i += 32) bn[i] = __fmul_rn(0.5f, __fadd_rn(zr[i], zr[i + 1]));
    } else {
      for (int i = lane; i < NB; i += 32) bn[i] = bins[static_cast<size_t>(ray) * NB + i];
    }
    if (HAVE_CDF) {
      for (int i = lane; i < NB; i += 32) cdf[i] = cdf_in[static_cast<size_t>(ray) * NB + i];
    } else {
      // pdf = (w + 1e-5) / sum(w + 1e-5); cdf = [0, cumsum(pdf)] with a double running sum rounded to
      // fp32 per prefix (the accumulation torch's CPU cumsum performs).
      const float* wr = weights + static_cast<size_t>(ray) * w_stride + w_offset;
      const int NW = NB - 1;
      float part = 0.f;
      for (int i = lane; i < NW; i += 32) part += __fadd_rn(wr[i], 1e-5f);
      const float tot = warp_sum(part);
      for (int i = lane; i < NW; i += 32) cdf[i + 1] = __fdiv_rn(__fadd_rn(wr[i], 1e-5f), tot);
      __syncwarp();
      if (lane == 0) {
        double run = 0.0;
        cdf[0] = 0.f;
        for (int i = 0; i < NW; ++i) {
          run += static_cast<double>(cdf[i + 1]);
          cdf[i + 1] = static_cast<float>(run);
        }
      }
    }
    __syncwarp();
    if (cdf_out)
      for (int i = lane; i < NB; i += 32) cdf_out[static_cast<size_t>(ray) * NB + i] = cdf[i];
    for (int s = lane; s < NS; s += 32) {
      const float uu = u[static_cast<size_t>(ray) * NS + s];
      // searchsorted(right=True): first index with cdf[idx] > u
      int lo = 0, hi = NB;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (cdf[mid] <= uu) lo = mid + 1; else hi = mid;
      }
      const int below = max(lo - 1, 0), above = min(lo, NB - 1);
      const float c0 = cdf[below], c1 = cdf[above];
      float denom = __fsub_rn(c1, c0);
      if (denom < 1e-5f) denom = 1.0f;
      const float t = __fdiv_rn(__fsub_rn(uu, c0), denom);
      const float b0 = bn[below], b1 = bn[above];
      samples[static_cast<size_t>(ray) * NS + s] = __fadd_rn(b0, __fmul_rn(t, __fsub_rn(b1, b0)));
      if (inds_out) inds_out[static_cast<size_t>(ray) * NS + s] = lo;
    }
    __syncwarp();
  }
}

// ------------------------------------------------------------------------------------------------
// merge: sort(cat(z_coarse, z_fine)) per ray; bitonic network over 256 padded keys in shared memory
// ------------------------------------------------------------------------------------------------
constexpr int kSortWarps = 4;
constexpr int kSortN = 256;

__global__ void __launch_bounds__(kSortWarps * 32) merge_sort_kernel(const float* __restrict__ za, int NA,
                                                                      const float* __restrict__ zb, int NBn, int B,
                                                                      float* __restrict__ out) {
  __shared__ float s_keys[kSortWarps][kSortN];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* k = s_keys[warp];
  const int N = NA + NBn;
  for (int ray = blockIdx.x * kSortWarps + warp; ray < B; ray += gridDim.x * kSortWarps) {
    for (int i = lane; i < kSortN; i += 32) {
      float v = CUDART_INF_F;
      if (i < NA) v = za[static_cast<size_t>(ray) * NA + i];
      else if (i < N) v = zb[static_cast<size_t>(ray) * NBn + (i - NA)];
      k[i] = v;
    }
    __syncwarp();
    for (int size = 2; size <= kSortN; size <<= 1) {
      for (int stride = size >> 1; stride > 0; stride >>= 1) {
        for (int t = lane; t < kSortN / 2; t += 32) {
          const int i = 2 * t - (t & (stride - 1));
          const int j = i + stride;
          const bool up = ((i & size) == 0);
          const float a = k[i], b = k[j];
          if ((a > b) == up) { k[i] = b; k[j] = a; }
        }
        __syncwarp();
      }
    }
    for (int i = lane; i < N; i += 32) out[static_cast<size_t>(ray) * N + i] = k[i];
    __syncwarp();
  }
}

inline int grid_for(long long items, int per_block) {
  long long g = (items + per_block - 1) / per_block;
  const long long cap = static_cast<long long>(dfn_num_sms()) * 16;
  return static_cast<int>(g < cap ? (g > 0 ? g : 1) : cap);
}

}  // namespace

extern "C" int dfn_stratified_sample(const float* near, const float* far, const float* t_vals, const float* t_rand,
                                     int B, int S, float* z_out, void* stream) {
  DFN_CHECK_ARG(near && far && t_vals && t_rand && z_out, "stratified_sample: null pointer");
  DFN_CHECK_ARG(B >= 0 && S >= 1, "stratified_sample: bad shape B=%d S=%d", B, S);
  const long long n = static_cast<long long>(B) * S;
  if (n == 0) return DFN_OK;
  stratified_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      near, far, t_vals, t_rand, B, S, z_out);
  DFN_CHECK_LAUNCH("stratified_sample");
  return DFN_OK;
}

extern "C" int dfn_points_on_rays(const float* rays_o, const float* rays_d, const float* z, int B, int S, float* pts,
                                  float* dirs, void* stream) {
  DFN_CHECK_ARG(rays_o && rays_d && z, "points_on_rays: null pointer");
  const long long n = static_cast<long long>(B) * S;
  if (n == 0) return DFN_OK;
  points_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      rays_o, rays_d, z, B, S, pts, dirs);
  DFN_CHECK_LAUNCH("points_on_rays");
  return DFN_OK;
}

extern "C" int dfn_sample_pdf(const float* bins, int bins_mid, const float* weights, long long w_stride,
                              int w_offset, const float* u, int B, int NB, int NS, float* samples, long long* inds_out,
                              float* cdf_out, void* stream) {
  DFN_CHECK_ARG(bins && weights && u && samples, "sample_pdf: null pointer");
  DFN_CHECK_ARG(NB >= 2 && NB <= kMaxBins, "sample_pdf: NB=%d outside [2,%d]", NB, kMaxBins);
  if (B == 0 || NS == 0) return DFN_OK;
  sample_pdf_kernel<false><<<grid_for(B, kPdfWarps), kPdfWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
      bins, weights, nullptr, u, B, NB, NS, w_stride, w_offset, bins_mid, samples, inds_out, cdf_out);
  DFN_CHECK_LAUNCH("sample_pdf");
  return DFN_OK;
}

extern "C" int dfn_sample_pdf_with_cdf(const float* bins, const float* cdf, const float* u, int B, int NB, int NS,
                                       float* samples, long long* inds_out, void* stream) {
  DFN_CHECK_ARG(bins && cdf && u && samples, "sample_pdf_with_cdf: null pointer");
  DFN_CHECK_ARG(NB >= 2 && NB <= kMaxBins, "sample_pdf_with_cdf: NB=%d outside [2,%d]", NB, kMaxBins);
  if (B == 0 || NS == 0) return DFN_OK;
  sample_pdf_kernel<true><<<grid_for(B, kPdfWarps), kPdfWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
      bins, nullptr, cdf, u, B, NB, NS, 0, 0, 0, samples, inds_out, nullptr);
  DFN_CHECK_LAUNCH("sample_pdf_with_cdf");
  return DFN_OK;
}

extern "C" int dfn_merge_sort_z(const float* z_coarse, int NA, const float* z_fine, int NBn, int B, float* z_out,
                                void* stream) {
  DFN_CHECK_ARG((z_coarse || NA == 0) && (z_fine || NBn == 0) && z_out, "merge_sort_z: null pointer");
  DFN_CHECK_ARG(NA >= 0 && NBn >= 0 && NA + NBn <= kSortN, "merge_sort_z: %d+%d keys exceed %d", NA, NBn, kSortN);
  if (B == 0) return DFN_OK;
  merge_sort_kernel<<<grid_for(B, kSortWarps), kSortWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
      z_coarse, NA, z_fine, NBn, B, z_out);
  DFN_CHECK_LAUNCH("merge_sort_z");
  return DFN_OK;
}
