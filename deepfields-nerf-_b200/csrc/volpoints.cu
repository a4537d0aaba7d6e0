// Occupied voxels of an extracted volume as a coloured point set (reference scripts/post_process_vol.py:37-50
// slice_aabb, :77-107 main): argwhere(occupancy) in C order, colours = rgb_volume / rgb_volume.max() gathered at
// those voxels, coordinates (index / (R-1)) * 2 - 1 in fp64, optionally sliced by an axis-aligned box.  The step
// after extract_vol; same two-pass ordered compaction as raygen.cu.
#include "compact.cuh"

namespace {
using dfn_compact::kChunk;
using dfn_compact::scan_counts_kernel;

struct VolArgs {
  const uint8_t* occ;      // [R^3]
  const uint8_t* rgb;      // [R^3,3]
  long long n;             // R^3
  int R;
  int slice;               // apply the box
  double lo[3], hi[3];     // aabb_min / aabb_max (post_process_vol.py:39-40)
};

// post_process_vol.py:95: (points / (res - 1)) * 2 - 1, elementwise fp64
__device__ __forceinline__ double voxel_coord(int i, int R) {
  return __dsub_rn(__dmul_rn(__ddiv_rn(static_cast<double>(i), static_cast<double>(R - 1)), 2.0), 1.0);
}

__device__ __forceinline__ void voxel_index(long long n, int R, int (&idx)[3]) {
  const unsigned int RR = static_cast<unsigned int>(R) * R;   // R <= 2048: n < 2^33 may exceed 32 bits -> 64-bit for ix only
  const long long ix = n / RR;
  const unsigned int rem = static_cast<unsigned int>(n - ix * RR);
  idx[0] = static_cast<int>(ix);
  idx[1] = static_cast<int>(rem / R);
  idx[2] = static_cast<int>(rem - idx[1] * R);
}

constexpr int kVoxThreads = kChunk / 4;   // every thread handles 4 consecutive voxels (one 4-byte + three 4-byte loads)

__device__ __forceinline__ bool keep_voxel(const VolArgs& a, long long n, unsigned int occ_byte) {
  bool keep = occ_byte != 0;
  if (keep && a.slice) {
    int idx[3];
    voxel_index(n, a.R, idx);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const double x = voxel_coord(idx[c], a.R);
      keep = keep && x >= a.lo[c] && x <= a.hi[c];   // slice_aabb, post_process_vol.py:42-45 (closed box)
    }
  }
  return keep;
}

// pass 1: kept voxels per chunk, keep bits (bit v % 32 of word v / 32), maximum of the colour bytes (post_process_vol.py:82)
__global__ void __launch_bounds__(kVoxThreads) count_voxels_kernel(const VolArgs a, unsigned int* __restrict__ counts,
                                                                   unsigned int* __restrict__ ballots,
                                                                   unsigned int* __restrict__ rgb_max) {
  const int lane = threadIdx.x & 31;
  const long long v0 = static_cast<long long>(blockIdx.x) * kChunk + 4 * threadIdx.x;
  unsigned int bits = 0, mx = 0;
  if (v0 + 3 < a.n) {
    const uint32_t o4 = *reinterpret_cast<const uint32_t*>(a.occ + v0);
    const uint32_t* c4 = reinterpret_cast<const uint32_t*>(a.rgb + v0 * 3);
    const uint32_t c[3] = {c4[0], c4[1], c4[2]};
#pragma unroll
    for (int i = 0; i < 3; ++i) mx = max(mx, max(max(c[i] & 0xFFu, (c[i] >> 8) & 0xFFu), max((c[i] >> 16) & 0xFFu, c[i] >> 24)));
#pragma unroll
    for (int i = 0; i < 4; ++i) bits |= (keep_voxel(a, v0 + i, (o4 >> (8 * i)) & 0xFFu) ? 1u : 0u) << i;
  } else {
    for (int i = 0; i < 4 && v0 + i < a.n; ++i) {
      const long long n = v0 + i;
      mx = max(mx, static_cast<unsigned int>(max(max(a.rgb[n * 3], a.rgb[n * 3 + 1]), a.rgb[n * 3 + 2])));
      bits |= (keep_voxel(a, n, a.occ[n]) ? 1u : 0u) << i;
    }
  }
  unsigned int word = bits << (4 * (lane & 7));   // 8 threads make one 32-voxel word
  word |= __shfl_xor_sync(0xffffffffu, word, 1);
  word |= __shfl_xor_sync(0xffffffffu, word, 2);
  word |= __shfl_xor_sync(0xffffffffu, word, 4);
  if ((lane & 7) == 0) ballots[static_cast<long long>(blockIdx.x) * (kChunk / 32) + (threadIdx.x >> 3)] = word;
  mx = __reduce_max_sync(0xffffffffu, mx);
  __shared__ unsigned int smax;
  if (threadIdx.x == 0) smax = 0;
  __syncthreads();
  if (lane == 0) atomicMax(&smax, mx);
  const int c = __syncthreads_count(bits & 1u) + __syncthreads_count(bits & 2u) + __syncthreads_count(bits & 4u) + __syncthreads_count(bits & 8u);
  if (threadIdx.x == 0) {
    counts[blockIdx.x] = static_cast<unsigned int>(c);
    if (smax > *reinterpret_cast<volatile unsigned int*>(rgb_max)) atomicMax(rgb_max, smax);   // monotone: the stale read only skips no-ops
  }
}

// pass 2: coordinates and colours of the kept voxels, in order
__global__ void __launch_bounds__(kVoxThreads) emit_voxels_kernel(const VolArgs a, const long long* __restrict__ offsets,
                                                                  const unsigned int* __restrict__ ballots,
                                                                  const unsigned int* __restrict__ rgb_max, double* __restrict__ points,
                                                                  float* __restrict__ rgbs) {
  __shared__ int warp_sums[kVoxThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long v0 = static_cast<long long>(blockIdx.x) * kChunk + 4 * threadIdx.x;
  const unsigned int word = ballots[static_cast<long long>(blockIdx.x) * (kChunk / 32) + (threadIdx.x >> 3)];
  const unsigned int bits = (word >> (4 * (lane & 7))) & 0xFu;
  // exclusive prefix of the per-thread counts inside the chunk
  const int cnt = __popc(bits);
  int incl = cnt;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int y = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl += y;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  int before = incl - cnt;
  for (int w = 0; w < warp; ++w) before += warp_sums[w];
  if (bits == 0) return;
  const float denom = static_cast<float>(*rgb_max);   // u8 tensor / u8 scalar tensor -> fp32 true division (:82)
  long long dst = offsets[blockIdx.x] + before;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (!((bits >> i) & 1u)) continue;
    const long long n = v0 + i;
    int idx[3];
    voxel_index(n, a.R, idx);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      points[dst * 3 + c] = voxel_coord(idx[c], a.R);
      rgbs[dst * 3 + c] = __fdiv_rn(static_cast<float>(a.rgb[n * 3 + c]), denom);
    }
    ++dst;
  }
}

}  // namespace

extern "C" long long dfn_volume_points_workspace_bytes(long long n_voxels) { return dfn_compact::workspace_bytes(n_voxels) + 16; }

namespace {
int fill_vol_args(VolArgs& a, const unsigned char* occupancy, const unsigned char* rgb_volume, int R, const double* aabb_min,
                  const double* aabb_max) {
  DFN_CHECK_ARG(occupancy && rgb_volume, "volume_points: null pointer");
  DFN_CHECK_ARG(R >= 2 && R <= 2048, "volume_points: resolution must be in [2, 2048]");
  DFN_CHECK_ARG((aabb_min == nullptr) == (aabb_max == nullptr), "volume_points: aabb_min and aabb_max go together");
  a.occ = occupancy; a.rgb = rgb_volume; a.R = R; a.n = static_cast<long long>(R) * R * R;
  a.slice = aabb_min != nullptr;
  for (int c = 0; c < 3; ++c) { a.lo[c] = a.slice ? aabb_min[c] : 0.0; a.hi[c] = a.slice ? aabb_max[c] : 0.0; }   // HOST pointers
  return DFN_OK;
}
unsigned int* rgb_max_of(void* workspace, long long n) {
  return reinterpret_cast<unsigned int*>(static_cast<uint8_t*>(workspace) + dfn_compact::workspace_bytes(n));
}
}  // namespace

// pass 1 + scan: *count = number of points pass 2 will write (so that the caller can size its outputs)
extern "C" int dfn_volume_points_count(const unsigned char* occupancy, const unsigned char* rgb_volume, int R, const double* aabb_min,
                                       const double* aabb_max, long long* count, void* workspace, void* stream) {
  VolArgs a{};
  if (int rc = fill_vol_args(a, occupancy, rgb_volume, R, aabb_min, aabb_max)) return rc;
  DFN_CHECK_ARG(count && workspace, "volume_points: null pointer");
  const long long chunks = (a.n + kChunk - 1) / kChunk;
  const dfn_compact::Workspace w = dfn_compact::carve(workspace, a.n);
  unsigned int* rgb_max = rgb_max_of(workspace, a.n);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DFN_CUDA(cudaMemsetAsync(rgb_max, 0, 4, st), "volume_points");
  count_voxels_kernel<<<static_cast<unsigned int>(chunks), kVoxThreads, 0, st>>>(a, w.counts, w.ballots, rgb_max);
  DFN_CHECK_LAUNCH("volume_points(count)");
  scan_counts_kernel<<<1, 1024, 0, st>>>(w.counts, chunks, w.offsets, count);
  DFN_CHECK_LAUNCH("volume_points(scan)");
  return DFN_OK;
}

// pass 2 on the workspace dfn_volume_points_count filled for the same volume
extern "C" int dfn_volume_points_emit(const unsigned char* occupancy, const unsigned char* rgb_volume, int R, double* points,
                                      float* rgbs, const void* workspace, void* stream) {
  VolArgs a{};
  if (int rc = fill_vol_args(a, occupancy, rgb_volume, R, nullptr, nullptr)) return rc;
  DFN_CHECK_ARG(points && rgbs && workspace, "volume_points: null pointer");
  const long long chunks = (a.n + kChunk - 1) / kChunk;
  const dfn_compact::Workspace w = dfn_compact::carve(const_cast<void*>(workspace), a.n);
  emit_voxels_kernel<<<static_cast<unsigned int>(chunks), kVoxThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      a, w.offsets, w.ballots, rgb_max_of(const_cast<void*>(workspace), a.n), points, rgbs);
  DFN_CHECK_LAUNCH("volume_points(emit)");
  return DFN_OK;
}

extern "C" int dfn_volume_points(const unsigned char* occupancy, const unsigned char* rgb_volume, int R, const double* aabb_min,
                                 const double* aabb_max, double* points, float* rgbs, long long* count, void* workspace,
                                 void* stream) {
  if (int rc = dfn_volume_points_count(occupancy, rgb_volume, R, aabb_min, aabb_max, count, workspace, stream)) return rc;
  return dfn_volume_points_emit(occupancy, rgb_volume, R, points, rgbs, workspace, stream);
}
