// Ray generation and ray / cube intersection (reference scripts/precompute_rays.py:10-119, :206-228): the step
// in front of the training path.  The reference builds every ray of every camera in fp64 numpy, intersects it
// with the [-1,1]^3 cube in fp32 and keeps the hits (boolean-mask compaction, order preserved) in an .npz.
// Here: one pass counts the hits per 1024-ray chunk, a one-block scan turns the counts into offsets, a second
// pass recomputes the rays (cheap) and writes the kept ones in order.  Nothing but the outputs touches HBM:
// 56 B per kept ray + 3 B of image per generated ray.
#include "dfn_common.cuh"
#include "compact.cuh"

namespace {

using dfn_compact::kChunk;   // rays per CTA
using dfn_compact::scan_counts_kernel;

struct RayGenArgs {
  double kinv[9];          // inverse intrinsics, row-major (np.linalg.inv(K), precompute_rays.py:80)
  const double* poses;     // [M][4][4] camera-to-world, row-major
  long long n_rays;        // M * H * W
  int H, W;
  int colmap;              // F = diag(1,-1,-1) (precompute_rays.py:79)
  int normalize;
  float cube_min, cube_max;
};

// np.minimum / np.maximum propagate NaN (fminf / fmaxf do not)
__device__ __forceinline__ float np_min(float a, float b) { return (a != a || b != b) ? __int_as_float(0x7fc00000) : fminf(a, b); }
__device__ __forceinline__ float np_max(float a, float b) { return (a != a || b != b) ? __int_as_float(0x7fc00000) : fmaxf(a, b); }

// precompute_rays.py:76-119, ray n of the flattened [M,H,W] order; fp64 arithmetic, fp32 results
__device__ __forceinline__ void make_ray(const RayGenArgs& a, long long n, float (&o)[3], float (&d)[3]) {
  const long long hw = static_cast<long long>(a.H) * a.W;
  const long long m = n / hw;
  const int pix = static_cast<int>(n - m * hw);
  const double v = static_cast<double>(pix / a.W), u = static_cast<double>(pix % a.W);
  double c[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) c[i] = fma(a.kinv[i * 3 + 1], v, a.kinv[i * 3] * u) + a.kinv[i * 3 + 2];
  if (a.colmap) { c[1] = -c[1]; c[2] = -c[2]; }
  if (a.normalize) {
    double nrm = sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
    nrm = nrm < 1e-8 ? 1e-8 : nrm;
#pragma unroll
    for (int i = 0; i < 3; ++i) c[i] = c[i] / nrm;
  }
  const double* P = a.poses + m * 16;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    d[i] = static_cast<float>(fma(P[i * 4 + 2], c[2], fma(P[i * 4 + 1], c[1], P[i * 4] * c[0])));
    o[i] = static_cast<float>(P[i * 4 + 3]);
  }
}

// precompute_rays.py:10-64 on fp32 inputs; returns the hit flag (taken before t_near is clamped, :58-61)
__device__ __forceinline__ bool cube_hit(const float (&o)[3], const float (&d_in)[3], float cmin, float cmax, float& t_near,
                                         float& t_far) {
  const float eps = 1e-8f;
  float tn = 0.f, tf = 0.f;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float d = d_in[i];
    if (fabsf(d) < eps) d = __fmul_rn(d > 0.f ? 1.f : (d < 0.f ? -1.f : d), eps);   // np.sign(d) * epsilon
    const float t0 = __fdiv_rn(__fsub_rn(cmin, o[i]), d), t1 = __fdiv_rn(__fsub_rn(cmax, o[i]), d);
    const float lo = np_min(t0, t1), hi = np_max(t0, t1);
    tn = i == 0 ? lo : np_max(tn, lo);
    tf = i == 0 ? hi : np_min(tf, hi);
  }
  const bool hit = (tn <= tf) && (tf >= 0.f);
  t_near = np_max(tn, 0.f);
  t_far = tf;
  return hit;
}

__global__ void generate_rays_kernel(const RayGenArgs a, float* __restrict__ rays_o, float* __restrict__ rays_d) {
  for (long long n = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; n < a.n_rays;
       n += static_cast<long long>(gridDim.x) * blockDim.x) {
    float o[3], d[3];
    make_ray(a, n, o, d);
#pragma unroll
    for (int i = 0; i < 3; ++i) { rays_o[n * 3 + i] = o[i]; rays_d[n * 3 + i] = d[i]; }
  }
}

__global__ void cube_intersect_kernel(const float* __restrict__ rays_o, const float* __restrict__ rays_d, long long n_rays,
                                      float cmin, float cmax, uint8_t* __restrict__ hit, float* __restrict__ t_near,
                                      float* __restrict__ t_far) {
  for (long long n = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; n < n_rays;
       n += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float o[3] = {rays_o[n * 3], rays_o[n * 3 + 1], rays_o[n * 3 + 2]};
    const float d[3] = {rays_d[n * 3], rays_d[n * 3 + 1], rays_d[n * 3 + 2]};
    float tn, tf;
    hit[n] = cube_hit(o, d, cmin, cmax, tn, tf) ? 1 : 0;
    t_near[n] = tn;
    t_far[n] = tf;
  }
}

// pass 1: hits per chunk
// (also keeps the per-warp hit ballots: pass 2 recomputes only the rays it keeps - the fp64 arithmetic, not HBM, is what
// this path costs)
__global__ void __launch_bounds__(kChunk) count_hits_kernel(const RayGenArgs a, unsigned int* __restrict__ counts,
                                                            unsigned int* __restrict__ ballots) {
  const long long n = static_cast<long long>(blockIdx.x) * kChunk + threadIdx.x;
  bool hit = false;
  if (n < a.n_rays) {
    float o[3], d[3], tn, tf;
    make_ray(a, n, o, d);
    hit = cube_hit(o, d, a.cube_min, a.cube_max, tn, tf);
  }
  const unsigned int ballot = __ballot_sync(0xffffffffu, hit);
  if ((threadIdx.x & 31) == 0) ballots[n >> 5] = ballot;
  const int c = __syncthreads_count(hit);
  if (threadIdx.x == 0) counts[blockIdx.x] = static_cast<unsigned int>(c);
}

// pass 2: recompute, keep the hits in order
__global__ void __launch_bounds__(kChunk) compact_rays_kernel(const RayGenArgs a, const uint8_t* __restrict__ images,
                                                              const long long* __restrict__ offsets,
                                                              const unsigned int* __restrict__ ballots, float* __restrict__ rays_o,
                                                              float* __restrict__ rays_d, float* __restrict__ rgbs,
                                                              float* __restrict__ t_near, float* __restrict__ t_far) {
  __shared__ int warp_counts[kChunk / 32];
  const int lane = threadIdx.x & 31;
  const long long n = static_cast<long long>(blockIdx.x) * kChunk + threadIdx.x;
  const unsigned int ballot = ballots[n >> 5];   // (the ballot array is padded to whole chunks)
  const bool hit = (ballot >> lane) & 1u;
  const int rank = dfn_compact::rank_in_chunk(ballot, warp_counts);
  if (hit) {
    float o[3], d[3], tn, tf;
    make_ray(a, n, o, d);
    cube_hit(o, d, a.cube_min, a.cube_max, tn, tf);
    const long long dst = offsets[blockIdx.x] + rank;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      rays_o[dst * 3 + i] = o[i];
      rays_d[dst * 3 + i] = d[i];
      if (rgbs) rgbs[dst * 3 + i] = __fdiv_rn(static_cast<float>(images[n * 3 + i]), 255.0f);   // precompute_rays.py:131-132
    }
    t_near[dst] = tn;
    t_far[dst] = tf;
  }
}

inline int grid1d(long long n, int block) {
  long long g = (n + block - 1) / block;
  const long long cap = static_cast<long long>(dfn_num_sms()) * 8;
  return static_cast<int>(g < 1 ? 1 : (g < cap ? g : cap));
}

int fill_args(RayGenArgs& a, const double* kinv, const double* poses, int M, int H, int W, int colmap, int normalize, float cube_min,
              float cube_max) {
  DFN_CHECK_ARG(kinv && poses && M > 0 && H > 0 && W > 0, "rays: bad camera arguments");
  for (int i = 0; i < 9; ++i) a.kinv[i] = kinv[i];
  a.poses = poses;
  a.n_rays = static_cast<long long>(M) * H * W;
  a.H = H; a.W = W; a.colmap = colmap; a.normalize = normalize;
  a.cube_min = cube_min; a.cube_max = cube_max;
  return DFN_OK;
}

}  // namespace

extern "C" int dfn_generate_rays(const double* kinv_host, const double* poses, int M, int H, int W, int colmap, int normalize,
                                 float* rays_o, float* rays_d, void* stream) {
  RayGenArgs a{};
  if (int rc = fill_args(a, kinv_host, poses, M, H, W, colmap, normalize, -1.f, 1.f)) return rc;
  DFN_CHECK_ARG(rays_o && rays_d, "generate_rays: null output");
  generate_rays_kernel<<<grid1d(a.n_rays, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(a, rays_o, rays_d);
  DFN_CHECK_LAUNCH("generate_rays");
  return DFN_OK;
}

extern "C" int dfn_ray_cube_intersect(const float* rays_o, const float* rays_d, long long n_rays, float cube_min, float cube_max,
                                      unsigned char* intersects, float* t_near, float* t_far, void* stream) {
  DFN_CHECK_ARG(n_rays >= 0, "ray_cube_intersect: negative count");
  if (n_rays == 0) return DFN_OK;
  DFN_CHECK_ARG(rays_o && rays_d && intersects && t_near && t_far, "ray_cube_intersect: null pointer");
  cube_intersect_kernel<<<grid1d(n_rays, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(rays_o, rays_d, n_rays, cube_min, cube_max,
                                                                                            intersects, t_near, t_far);
  DFN_CHECK_LAUNCH("ray_cube_intersect");
  return DFN_OK;
}

extern "C" long long dfn_precompute_rays_workspace_bytes(long long n_rays) { return dfn_compact::workspace_bytes(n_rays); }

extern "C" int dfn_precompute_rays(const double* kinv_host, const double* poses, int M, int H, int W, int colmap, int normalize,
                                   const unsigned char* images, float cube_min, float cube_max, float* rays_o, float* rays_d,
                                   float* rgbs, float* t_near, float* t_far, long long* count, void* workspace, void* stream) {
  RayGenArgs a{};
  if (int rc = fill_args(a, kinv_host, poses, M, H, W, colmap, normalize, cube_min, cube_max)) return rc;
  DFN_CHECK_ARG(rays_o && rays_d && t_near && t_far && count && workspace, "precompute_rays: null pointer");
  DFN_CHECK_ARG((rgbs == nullptr) == (images == nullptr), "precompute_rays: images and rgbs go together");
  const long long chunks = (a.n_rays + kChunk - 1) / kChunk;
  DFN_CHECK_ARG(chunks < (1LL << 31), "precompute_rays: too many rays for one launch");
  const dfn_compact::Workspace w = dfn_compact::carve(workspace, a.n_rays);
  unsigned int* counts = w.counts;
  long long* offsets = w.offsets;
  unsigned int* ballots = w.ballots;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  count_hits_kernel<<<static_cast<unsigned int>(chunks), kChunk, 0, st>>>(a, counts, ballots);
  DFN_CHECK_LAUNCH("precompute_rays(count)");
  scan_counts_kernel<<<1, 1024, 0, st>>>(counts, chunks, offsets, count);
  DFN_CHECK_LAUNCH("precompute_rays(scan)");
  compact_rays_kernel<<<static_cast<unsigned int>(chunks), kChunk, 0, st>>>(a, images, offsets, ballots, rays_o, rays_d, rgbs, t_near, t_far);
  DFN_CHECK_LAUNCH("precompute_rays(compact)");
  return DFN_OK;
}
