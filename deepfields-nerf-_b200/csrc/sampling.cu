// Ray sampling kernels: stratified depths (reference scripts/sampling.py:10-32), inverse-CDF
// hierarchical sampling with a fused per-ray pdf/cdf build and searchsorted(right=True)
// (scripts/train.py:31-75), the coarse+fine depth merge (scripts/sampling.py:40) and point emission
// pts = o + d*z (scripts/sampling.py:24-31, :43-48).  All arithmetic that must be bit-identical to the
// reference's separate torch ops uses explicit round-to-nearest intrinsics (no FMA contraction).
#include "dfn_common.cuh"
#include "rng.cuh"
#include <math_constants.h>

namespace {

// ------------------------------------------------------------------------------------------------
// stratified z: one thread per (ray, sample)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float lerp_nf(float near, float far, float t) {
  return __fadd_rn(__fmul_rn(near, __fsub_rn(1.0f, t)), __fmul_rn(far, t));
}

__device__ __forceinline__ float strat_z(float n, float f, const float* __restrict__ t_vals, int i, int S, float tr) {
  const float zi = lerp_nf(n, f, t_vals[i]);
  float lo = zi, hi = zi;
  if (i > 0) lo = __fmul_rn(0.5f, __fadd_rn(zi, lerp_nf(n, f, t_vals[i - 1])));
  if (i + 1 < S) hi = __fmul_rn(0.5f, __fadd_rn(lerp_nf(n, f, t_vals[i + 1]), zi));
  return __fadd_rn(lo, __fmul_rn(__fsub_rn(hi, lo), tr));
}

// VEC = 4: S % 4 == 0, one thread per four consecutive samples of a ray (16-byte loads / stores)
// t_rand == nullptr: the jitter is drawn in the kernel (Philox; item = group of four samples)
template <int VEC>
__global__ void stratified_kernel(const float* __restrict__ near, const float* __restrict__ far,
                                  const float* __restrict__ t_vals, const float* __restrict__ t_rand, int B, int S,
                                  float* __restrict__ z_out, unsigned long long seed, const int* __restrict__ step_ptr) {
  const uint32_t step = t_rand ? 0u : static_cast<uint32_t>(*step_ptr);
  const long long n_items = static_cast<long long>(B) * S / VEC;
  const unsigned per_ray = static_cast<unsigned>(S / VEC);
  for (long long it = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; it < n_items;
       it += static_cast<long long>(gridDim.x) * blockDim.x) {
    long long ray;
    int i0;
    if ((it >> 32) == 0) {   // a 64-bit division costs ~100 instructions
      const unsigned r = static_cast<unsigned>(it) / per_ray;
      ray = r; i0 = static_cast<int>(static_cast<unsigned>(it) - r * per_ray) * VEC;
    } else {
      ray = it / per_ray; i0 = static_cast<int>(it % per_ray) * VEC;
    }
    const float n = __ldg(near + ray), f = __ldg(far + ray);
    if (VEC == 4) {
      const float4 tr = t_rand ? *reinterpret_cast<const float4*>(t_rand + it * 4)
                               : rng::uniform4(seed, step, rng::kStreamJitter, static_cast<unsigned long long>(it));
      float4 z;
      z.x = strat_z(n, f, t_vals, i0, S, tr.x);
      z.y = strat_z(n, f, t_vals, i0 + 1, S, tr.y);
      z.z = strat_z(n, f, t_vals, i0 + 2, S, tr.z);
      z.w = strat_z(n, f, t_vals, i0 + 3, S, tr.w);
      *reinterpret_cast<float4*>(z_out + it * 4) = z;
    } else {
      z_out[it] = strat_z(n, f, t_vals, i0, S, t_rand ? t_rand[it] : rng::uniform4(seed, step, rng::kStreamJitter, static_cast<unsigned long long>(it)).x);
    }
  }
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

// cdf[0] = 0, cdf[i+1] = fl32(sum_{j<=i} pdf[j]) with the running sum kept in double and rounded to fp32 per prefix - the
// accumulation torch's CPU cumsum performs.  On entry cdf[1..NW] holds the pdf.  The sum runs as a warp
// scan: every lane adds a contiguous run, the run totals are scanned with shuffles, the exclusive prefix of a lane is its
// left neighbour's inclusive one (a shuffle, not a subtraction).
// PRECONDITION of the bit-for-bit claim: the weights are compositing weights (sum <= 1, what scripts/train.py:158 passes).
// Then pdf = (w + 1e-5) / sum lies in [2^-17, 1], 24-bit mantissas spread over 17 binades plus 8 bits of carry fit the 53
// bits of a double, every partial sum is EXACT whatever the association, and the rounded prefixes are bit-identical to the
// sequential loop's.  The entry points accept any non-negative weights; with large or unnormalised ones (1e-5 / sum below
// ~2^-21) the partial sums round, the association matters, and the prefixes are only within an fp32 ulp of torch's.
__device__ __forceinline__ void pdf_to_cdf_warp(float* cdf, int NW, int lane) {
  const int per = (NW + 31) >> 5;              // <= 8
  const int i0 = lane * per;
  double loc[8];
  double run = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int i = i0 + k;
    if (k < per && i < NW) run += static_cast<double>(cdf[i + 1]);
    loc[k] = run;
  }
  double incl = run;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const double t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  double excl = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) excl = 0.0;
  __syncwarp();
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int i = i0 + k;
    if (k < per && i < NW) cdf[i + 1] = static_cast<float>(excl + loc[k]);
  }
  if (lane == 0) cdf[0] = 0.f;
}

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
      pdf_to_cdf_warp(cdf, NW, lane);
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
// sample_pdf + fine_points' sort in ONE kernel (the train step / render path): the NS new depths never leave the SM.
// One warp per ray: pdf -> cdf (warp scan) -> searchsorted + lerp (KPL draws per lane) -> bitonic sort of the new depths
// in registers (compare-exchange through shuffles) -> merge with the NA coarse depths by rank (two binary searches; the
// coarse depths of the stratified sampler are sorted by construction - if a ray's are not, its keys take the generic
// shared-memory bitonic network instead) -> one coalesced write of the NA + NS merged depths.
// Values are bit-identical to sample_pdf_kernel followed by merge_sort_kernel: same arithmetic, and a sorted sequence of
// the same multiset of floats is unique.
// ------------------------------------------------------------------------------------------------
constexpr int kSmWarps = 4;
constexpr int kSmMaxNA = 128;

template <int KPL>
__device__ __forceinline__ void bitonic_sort_regs(float (&v)[KPL], int lane) {
  constexpr int N = 32 * KPL;
#pragma unroll
  for (int k = 2; k <= N; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j >= KPL) {
        const int lj = j / KPL;
        const bool lower = (lane & lj) == 0;
#pragma unroll
        for (int i = 0; i < KPL; ++i) {
          const float other = __shfl_xor_sync(0xffffffffu, v[i], lj);
          const bool up = (((lane * KPL + i) & k) == 0);
          const bool take_min = (up == lower);
          const bool swap = take_min ? (other < v[i]) : (other > v[i]);
          v[i] = swap ? other : v[i];
        }
      } else {
#pragma unroll
        for (int i = 0; i < KPL; ++i) {
          const int p2 = i ^ j;
          if (p2 > i) {
            const bool up = (((lane * KPL + i) & k) == 0);
            const float a = v[i], b = v[p2];
            const bool swap = up ? (a > b) : (a < b);
            v[i] = swap ? b : a;
            v[p2] = swap ? a : b;
          }
        }
      }
    }
  }
}

// Sorted union of NA keys in shared memory (zc; `sorted` tells whether they are ascending) and NS keys in registers
// (v: key lane + 32 j in v[j], +inf beyond NS) -> so[0 .. NA + NS)
template <int KPL>
__device__ __forceinline__ void sort_merge_warp(float (&v)[KPL], const float* zc, int NA, int NS, bool sorted, float* sf, float* so,
                                                int lane) {
  const int N = NA + NS;
  if (sorted) {
      bitonic_sort_regs<KPL>(v, lane);                 // element e = lane * KPL + j, ascending; the +inf padding sorts last
#pragma unroll
      for (int j = 0; j < KPL; ++j) sf[lane * KPL + j] = v[j];
      __syncwarp();
      // rank merge, coarse first among equal values:  pos(coarse i) = i + #{fine < zc[i]},  pos(fine e) = e + #{coarse <= f}
#pragma unroll
      for (int j = 0; j < KPL; ++j) {
        const int e = lane * KPL + j;
        if (e < NS) {
          const float f = v[j];
          int lo = 0, hi = NA;
          while (lo < hi) { const int mid = (lo + hi) >> 1; if (zc[mid] <= f) lo = mid + 1; else hi = mid; }
          so[e + lo] = f;
        }
      }
      for (int i = lane; i < NA; i += 32) {
        const float c = zc[i];
        int lo = 0, hi = NS;
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (sf[mid] < c) lo = mid + 1; else hi = mid; }
        so[i + lo] = c;
      }
    } else {
      // generic path: bitonic network over all keys in shared memory (padded to a power of two with +inf)
      int M = 2;
      while (M < N) M <<= 1;
      for (int i = lane; i < M; i += 32) so[i] = i < NA ? zc[i] : CUDART_INF_F;
#pragma unroll
      for (int j = 0; j < KPL; ++j) if (lane + 32 * j < NS) so[NA + lane + 32 * j] = v[j];
      __syncwarp();
      for (int size = 2; size <= M; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
          for (int t = lane; t < M / 2; t += 32) {
            const int i = 2 * t - (t & (stride - 1));
            const int jj = i + stride;
            const bool up = ((i & size) == 0);
            const float a = so[i], b = so[jj];
            if ((a > b) == up) { so[i] = b; so[jj] = a; }
          }
          __syncwarp();
        }
    }
}

template <int KPL>
__global__ void __launch_bounds__(kSmWarps * 32) sample_merge_kernel(
    const float* __restrict__ z, const float* __restrict__ weights, const float* __restrict__ u, int B, int NA, int NS,
    float* __restrict__ z_all, float* __restrict__ z_new, unsigned long long seed, const int* __restrict__ step_ptr) {
  constexpr int NF = 32 * KPL;                        // padded number of new depths
  const uint32_t step = u ? 0u : static_cast<uint32_t>(*step_ptr);   // u == nullptr: the draws come from Philox
  __shared__ float s_cdf[kSmWarps][kSmMaxNA];
  __shared__ float s_bins[kSmWarps][kSmMaxNA];
  __shared__ float s_zc[kSmWarps][kSmMaxNA];
  __shared__ float s_f[kSmWarps][NF];
  __shared__ float s_out[kSmWarps][NF <= 128 ? 256 : 512];   // the generic network pads NA + NS keys to a power of two
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* cdf = s_cdf[warp];
  float* bn = s_bins[warp];
  float* zc = s_zc[warp];
  float* sf = s_f[warp];
  float* so = s_out[warp];
  const int NB = NA - 1, NW = NA - 2, N = NA + NS;
  for (int ray = blockIdx.x * kSmWarps + warp; ray < B; ray += gridDim.x * kSmWarps) {
    const float* zr = z + static_cast<size_t>(ray) * NA;
    const float* wr = weights + static_cast<size_t>(ray) * NA + 1;      // weights[..., 1:-1]  (scripts/train.py:158)
    for (int i = lane; i < NA; i += 32) zc[i] = zr[i];
    __syncwarp();
    for (int i = lane; i < NB; i += 32) bn[i] = __fmul_rn(0.5f, __fadd_rn(zc[i], zc[i + 1]));   // scripts/train.py:157
    float part = 0.f;
    for (int i = lane; i < NW; i += 32) part += __fadd_rn(wr[i], 1e-5f);
    const float tot = warp_sum(part);
    for (int i = lane; i < NW; i += 32) cdf[i + 1] = __fdiv_rn(__fadd_rn(wr[i], 1e-5f), tot);
    bool sorted = true;
    for (int i = lane; i + 1 < NA; i += 32) sorted = sorted && !(zc[i] > zc[i + 1]);
    sorted = __all_sync(0xffffffffu, sorted);
    __syncwarp();
    pdf_to_cdf_warp(cdf, NW, lane);
    __syncwarp();
    float v[KPL];
    float ur[KPL];
    if (!u) {
#pragma unroll
      for (int j4 = 0; j4 < KPL / 4; ++j4) {   // item = (ray, lane, group of four draws)
        const float4 r = rng::uniform4(seed, step, rng::kStreamPdf, (static_cast<unsigned long long>(ray) * 32 + lane) * (KPL / 4) + j4);
        ur[4 * j4] = r.x; ur[4 * j4 + 1] = r.y; ur[4 * j4 + 2] = r.z; ur[4 * j4 + 3] = r.w;
      }
    }
#pragma unroll
    for (int j = 0; j < KPL; ++j) {
      const int sidx = lane + 32 * j;
      v[j] = CUDART_INF_F;
      if (sidx < NS) {
        const float uu = u ? u[static_cast<size_t>(ray) * NS + sidx] : ur[j];
        int lo = 0, hi = NB;
        while (lo < hi) {   // searchsorted(right=True): first index with cdf[idx] > u
          const int mid = (lo + hi) >> 1;
          if (cdf[mid] <= uu) lo = mid + 1; else hi = mid;
        }
        const int below = max(lo - 1, 0), above = min(lo, NB - 1);
        const float c0 = cdf[below], c1 = cdf[above];
        float denom = __fsub_rn(c1, c0);
        if (denom < 1e-5f) denom = 1.0f;
        const float t = __fdiv_rn(__fsub_rn(uu, c0), denom);
        const float b0 = bn[below], b1 = bn[above];
        v[j] = __fadd_rn(b0, __fmul_rn(t, __fsub_rn(b1, b0)));
        if (z_new) z_new[static_cast<size_t>(ray) * NS + sidx] = v[j];
      }
    }
    sort_merge_warp<KPL>(v, zc, NA, NS, sorted, sf, so, lane);
    __syncwarp();
    for (int i = lane; i < N; i += 32) z_all[static_cast<size_t>(ray) * N + i] = so[i];
    __syncwarp();
  }
}

// ------------------------------------------------------------------------------------------------
// merge: sort(cat(z_coarse, z_fine)) per ray (scripts/sampling.py:40) for any inputs: the second list is sorted in
// registers, the first one is merged by rank when it is ascending (the stratified depths are) and goes through the generic
// shared-memory network otherwise
// ------------------------------------------------------------------------------------------------
constexpr int kSortN = 256;

template <int KPL>
__global__ void __launch_bounds__(kSmWarps * 32) merge_sort_kernel(const float* __restrict__ za, int NA,
                                                                   const float* __restrict__ zb, int NBn, int B,
                                                                   float* __restrict__ out) {
  constexpr int NF = 32 * KPL;
  __shared__ float s_zc[kSmWarps][kSmMaxNA];
  __shared__ float s_f[kSmWarps][NF];
  __shared__ float s_out[kSmWarps][NF <= 128 ? 256 : 512];   // the generic network pads NA + NS keys to a power of two
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* zc = s_zc[warp];
  const int N = NA + NBn;
  for (int ray = blockIdx.x * kSmWarps + warp; ray < B; ray += gridDim.x * kSmWarps) {
    for (int i = lane; i < NA; i += 32) zc[i] = za[static_cast<size_t>(ray) * NA + i];
    float v[KPL];
#pragma unroll
    for (int j = 0; j < KPL; ++j) {
      const int i = lane + 32 * j;
      v[j] = i < NBn ? zb[static_cast<size_t>(ray) * NBn + i] : CUDART_INF_F;
    }
    __syncwarp();
    bool sorted = true;
    for (int i = lane; i + 1 < NA; i += 32) sorted = sorted && !(zc[i] > zc[i + 1]);
    sorted = __all_sync(0xffffffffu, sorted);
    sort_merge_warp<KPL>(v, zc, NA, NBn, sorted, s_f[warp], s_out[warp], lane);
    __syncwarp();
    for (int i = lane; i < N; i += 32) out[static_cast<size_t>(ray) * N + i] = s_out[warp][i];
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
  const bool vec = (S % 4 == 0) && ((reinterpret_cast<uintptr_t>(t_rand) | reinterpret_cast<uintptr_t>(z_out)) % 16 == 0);
  if (vec) stratified_kernel<4><<<grid_for(n / 4, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(near, far, t_vals, t_rand, B, S, z_out, 0ull, nullptr);
  else stratified_kernel<1><<<grid_for(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(near, far, t_vals, t_rand, B, S, z_out, 0ull, nullptr);
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
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // the list that goes to registers is the longer one (<= 256 keys); the other one must fit the 128-key shared buffer
  const bool swap = NA > NBn;
  const float* a = swap ? z_fine : z_coarse;
  const float* b = swap ? z_coarse : z_fine;
  const int na = swap ? NBn : NA, nb = swap ? NA : NBn;
  if (nb <= 128) merge_sort_kernel<4><<<grid_for(B, kSmWarps), kSmWarps * 32, 0, st>>>(a, na, b, nb, B, z_out);
  else merge_sort_kernel<8><<<grid_for(B, kSmWarps), kSmWarps * 32, 0, st>>>(a, na, b, nb, B, z_out);
  DFN_CHECK_LAUNCH("merge_sort_z");
  return DFN_OK;
}

// sample_pdf on (mid-points of z, weights[:, 1:-1]) followed by sort(cat(z, samples)) - the two steps between the coarse and
// the fine network of scripts/train.py:157-163 - as one launch.  z [B,NA], weights [B,NA] (the full coarse weights),
// u [B,NS]; z_all [B,NA+NS]; z_new [B,NS] or NULL.
extern "C" int dfn_sample_merge(const float* z, const float* weights, const float* u, int B, int NA, int NS, float* z_all,
                                float* z_new, void* stream) {
  DFN_CHECK_ARG(z && weights && u && z_all, "sample_merge: null pointer");
  DFN_CHECK_ARG(NA >= 3 && NA <= kSmMaxNA && NS >= 1 && NS <= 256, "sample_merge: NA=%d outside [3,%d] or NS=%d outside [1,256]", NA, kSmMaxNA, NS);
  if (B == 0) return DFN_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (NS <= 128) sample_merge_kernel<4><<<grid_for(B, kSmWarps), kSmWarps * 32, 0, st>>>(z, weights, u, B, NA, NS, z_all, z_new, 0ull, nullptr);
  else sample_merge_kernel<8><<<grid_for(B, kSmWarps), kSmWarps * 32, 0, st>>>(z, weights, u, B, NA, NS, z_all, z_new, 0ull, nullptr);
  DFN_CHECK_LAUNCH("sample_merge");
  return DFN_OK;
}

// The same two entry points with the draws made IN the kernel (Philox4x32-10, csrc/rng.cuh): draw = f(seed, *step_ptr, item),
// step_ptr = the optimiser's step counter in device memory, so CUDA-graph replays advance and ranks differ through the seed.
extern "C" int dfn_stratified_sample_rng(const float* near, const float* far, const float* t_vals, unsigned long long seed,
                                         const int* step_ptr, int B, int S, float* z_out, void* stream) {
  DFN_CHECK_ARG(near && far && t_vals && step_ptr && z_out, "stratified_sample_rng: null pointer");
  DFN_CHECK_ARG(B >= 0 && S >= 1, "stratified_sample_rng: bad shape B=%d S=%d", B, S);
  const long long n = static_cast<long long>(B) * S;
  if (n == 0) return DFN_OK;
  if (S % 4 == 0 && reinterpret_cast<uintptr_t>(z_out) % 16 == 0)
    stratified_kernel<4><<<grid_for(n / 4, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(near, far, t_vals, nullptr, B, S, z_out, seed, step_ptr);
  else
    stratified_kernel<1><<<grid_for(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(near, far, t_vals, nullptr, B, S, z_out, seed, step_ptr);
  DFN_CHECK_LAUNCH("stratified_sample_rng");
  return DFN_OK;
}

extern "C" int dfn_sample_merge_rng(const float* z, const float* weights, unsigned long long seed, const int* step_ptr, int B,
                                    int NA, int NS, float* z_all, float* z_new, void* stream) {
  DFN_CHECK_ARG(z && weights && step_ptr && z_all, "sample_merge_rng: null pointer");
  DFN_CHECK_ARG(NA >= 3 && NA <= kSmMaxNA && NS >= 1 && NS <= 256, "sample_merge_rng: NA=%d outside [3,%d] or NS=%d outside [1,256]", NA, kSmMaxNA, NS);
  if (B == 0) return DFN_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (NS <= 128) sample_merge_kernel<4><<<grid_for(B, kSmWarps), kSmWarps * 32, 0, st>>>(z, weights, nullptr, B, NA, NS, z_all, z_new, seed, step_ptr);
  else sample_merge_kernel<8><<<grid_for(B, kSmWarps), kSmWarps * 32, 0, st>>>(z, weights, nullptr, B, NA, NS, z_all, z_new, seed, step_ptr);
  DFN_CHECK_LAUNCH("sample_merge_rng");
  return DFN_OK;
}
