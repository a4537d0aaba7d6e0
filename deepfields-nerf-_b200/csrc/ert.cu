// Early ray termination for the inference render (BASELINE config 5; the reference has no inference path, its
// compositing is scripts/train.py:78-106): the fine network is queried front to back in depth segments, and after
// every segment the rays whose transmittance has dropped below eps leave the active list, so that the expensive
// part - the MLP - is not evaluated behind opaque surfaces.  Two small HBM-bound kernels around the MLP launches:
//   ert_update   one warp per active ray: scatter the segment's rgb / sigma into the full [B,S] arrays (unevaluated
//                samples stay sigma = 0 and add nothing in the final compositing pass), multiply the ray's running
//                transmittance by prod(1 - alpha + 1e-10) of the segment (train.py:89-95), flag the ray alive or not
//   compact      ordered compaction of the active list by those flags (compact.cuh)
#include "compact.cuh"
#include <mutex>
#include "../../include/deepfields_b200.h"

namespace {
using dfn_compact::kChunk;
using dfn_compact::scan_counts_kernel;

__global__ void ert_update_kernel(const int* __restrict__ active, int n, const float* __restrict__ seg_rgb,
                                  const float* __restrict__ seg_sigma, const float* __restrict__ z, int S, int s0, int C,
                                  float dist_last, float eps, float* __restrict__ rgb_full, float* __restrict__ sigma_full,
                                  float* __restrict__ T_run, unsigned char* __restrict__ alive) {
  const int lane = threadIdx.x & 31;
  const int k = static_cast<int>((static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5);
  if (k >= n) return;
  const long long ray = active ? active[k] : k;
  float prod = 1.f;
  for (int j = lane; j < C; j += 32) {
    const int i = s0 + j;
    const float sg = seg_sigma[static_cast<long long>(k) * C + j];
    const float dist = i + 1 < S ? __fsub_rn(z[ray * S + i + 1], z[ray * S + i]) : dist_last;
    const float alpha = __fsub_rn(1.f, expf(-__fmul_rn(sg, dist)));
    prod *= __fadd_rn(__fsub_rn(1.f, alpha), 1e-10f);
    sigma_full[ray * S + i] = sg;
#pragma unroll
    for (int c = 0; c < 3; ++c) rgb_full[(ray * S + i) * 3 + c] = seg_rgb[(static_cast<long long>(k) * C + j) * 3 + c];
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) prod *= __shfl_xor_sync(0xffffffffu, prod, off);
  if (lane == 0) {
    const float T = T_run[ray] * prod;
    T_run[ray] = T;
    alive[k] = T > eps ? 1 : 0;
  }
}

__global__ void __launch_bounds__(kChunk) count_flags_kernel(const unsigned char* __restrict__ flags, long long n,
                                                             unsigned int* __restrict__ counts, unsigned int* __restrict__ ballots) {
  const long long i = static_cast<long long>(blockIdx.x) * kChunk + threadIdx.x;
  const bool keep = i < n && flags[i] != 0;
  const unsigned int ballot = __ballot_sync(0xffffffffu, keep);
  if ((threadIdx.x & 31) == 0) ballots[i >> 5] = ballot;
  const int c = __syncthreads_count(keep);
  if (threadIdx.x == 0) counts[blockIdx.x] = static_cast<unsigned int>(c);
}

__global__ void __launch_bounds__(kChunk) emit_indices_kernel(const int* __restrict__ idx_in, const long long* __restrict__ offsets,
                                                              const unsigned int* __restrict__ ballots, int* __restrict__ idx_out) {
  __shared__ int warp_counts[kChunk / 32];
  const long long i = static_cast<long long>(blockIdx.x) * kChunk + threadIdx.x;
  const unsigned int ballot = ballots[i >> 5];
  const int rank = dfn_compact::rank_in_chunk(ballot, warp_counts);
  if ((ballot >> (threadIdx.x & 31)) & 1u) idx_out[offsets[blockIdx.x] + rank] = idx_in ? idx_in[i] : static_cast<int>(i);
}

}  // namespace

extern "C" int dfn_ert_update(const int* active, int n, const float* seg_rgb, const float* seg_sigma, const float* z, int S, int s0,
                              int C, float dist_last, float eps, float* rgb_full, float* sigma_full, float* T_run,
                              unsigned char* alive, void* stream) {
  DFN_CHECK_ARG(n >= 0 && S > 0 && C > 0 && s0 >= 0 && s0 + C <= S, "ert_update: segment outside the ray");
  if (n == 0) return DFN_OK;
  DFN_CHECK_ARG(seg_rgb && seg_sigma && z && rgb_full && sigma_full && T_run && alive, "ert_update: null pointer");
  const long long threads = static_cast<long long>(n) * 32;
  ert_update_kernel<<<static_cast<unsigned int>((threads + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      active, n, seg_rgb, seg_sigma, z, S, s0, C, dist_last, eps, rgb_full, sigma_full, T_run, alive);
  DFN_CHECK_LAUNCH("ert_update");
  return DFN_OK;
}

extern "C" long long dfn_compact_indices_workspace_bytes(long long n) { return dfn_compact::workspace_bytes(n); }

extern "C" int dfn_compact_indices(const unsigned char* flags, const int* idx_in, long long n, int* idx_out, long long* count,
                                   void* workspace, void* stream) {
  DFN_CHECK_ARG(n >= 0 && n < (1LL << 31), "compact_indices: bad length");
  DFN_CHECK_ARG(count && workspace, "compact_indices: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (n == 0) { DFN_CUDA(cudaMemsetAsync(count, 0, 8, st), "compact_indices"); return DFN_OK; }
  DFN_CHECK_ARG(flags && idx_out, "compact_indices: null pointer");
  const long long chunks = (n + kChunk - 1) / kChunk;
  const dfn_compact::Workspace w = dfn_compact::carve(workspace, n);
  count_flags_kernel<<<static_cast<unsigned int>(chunks), kChunk, 0, st>>>(flags, n, w.counts, w.ballots);
  DFN_CHECK_LAUNCH("compact_indices(count)");
  scan_counts_kernel<<<1, 1024, 0, st>>>(w.counts, chunks, w.offsets, count);
  DFN_CHECK_LAUNCH("compact_indices(scan)");
  emit_indices_kernel<<<static_cast<unsigned int>(chunks), kChunk, 0, st>>>(idx_in, w.offsets, w.ballots, idx_out);
  DFN_CHECK_LAUNCH("compact_indices(emit)");
  return DFN_OK;
}

// ---- dfn_compact_indices_static: the same compaction for callers that have no device memory to hand over as a workspace ----
// dfn_compact_indices needs ~0.14 B of scratch per flag (chunk counts, chunk offsets, warp ballots) and leaves the allocation to
// the caller.  Some hosts of this library cannot make one: a render thread living inside an engine that reserved the whole
// device for its own arena at start-up (a later cudaMalloc simply fails there), a frame loop that must not take the
// device-wide synchronisation cudaMalloc / cudaFree imply, or a plug-in whose contract hands it input and output buffers and
// nothing else.  For them the scratch lives in the library image itself: one fixed array per GPU, present from the moment the
// library is loaded on that GPU until the process exits, large enough for DFN_COMPACT_STATIC_MAX_N flags.
namespace {
constexpr long long kStaticMaxN = 1LL << 22;                         // = DFN_COMPACT_STATIC_MAX_N (static_assert below)
constexpr long long kStaticChunks = kStaticMaxN / kChunk;            // 4096
constexpr long long kStaticBytes = ((kStaticChunks * 4 + 16 + kStaticChunks * 8 + 16 + kStaticChunks * (kChunk / 8) + 255) / 256) * 256;
static_assert(kStaticBytes == 573696, "static scratch size quoted in the header / DESIGN.md");
static_assert(kStaticMaxN == DFN_COMPACT_STATIC_MAX_N, "header and kernel disagree on the limit");

// The scratch.  In device code this name is the array; in HOST code it is only a handle for the runtime's symbol table (its
// host address is meaningless to a kernel), so the launcher below asks the runtime for the device address once per GPU.
static __device__ __align__(256) unsigned char g_compact_scratch[kStaticBytes];

constexpr int kMaxDevices = 64;
std::mutex g_scratch_mutex;
void* g_scratch_ptr[kMaxDevices] = {nullptr};   // device address of g_compact_scratch on GPU i, filled on first use
}  // namespace

extern "C" long long dfn_compact_indices_static_max_n(void) { return kStaticMaxN; }

extern "C" int dfn_compact_indices_static(const unsigned char* flags, const int* idx_in, long long n, int* idx_out, long long* count,
                                          void* stream) {
  DFN_CHECK_ARG(n >= 0, "compact_indices_static: negative length");
  DFN_CHECK_ARG(n <= kStaticMaxN,
                "compact_indices_static: %lld flags exceed the built-in scratch (at most %lld per call); split the list or use "
                "dfn_compact_indices with a caller-owned workspace", n, kStaticMaxN);
  DFN_CHECK_ARG(count, "compact_indices_static: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (n == 0) { DFN_CUDA(cudaMemsetAsync(count, 0, 8, st), "compact_indices_static"); return DFN_OK; }
  DFN_CHECK_ARG(flags && idx_out, "compact_indices_static: null pointer");

  // set-up, once per GPU, with that GPU current: resolve the array's device address (every GPU has its own copy)
  int dev = -1;
  DFN_CUDA(cudaGetDevice(&dev), "compact_indices_static(cudaGetDevice)");
  DFN_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "compact_indices_static: device ordinal %d not supported", dev);
  void* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (g_scratch_ptr[dev] == nullptr) {
      void* p = nullptr;
      DFN_CUDA(cudaGetSymbolAddress(&p, g_compact_scratch), "compact_indices_static(cudaGetSymbolAddress)");
      DFN_CHECK_ARG(p != nullptr, "compact_indices_static: the runtime returned no address for the scratch array");
      g_scratch_ptr[dev] = p;
    }
    scratch = g_scratch_ptr[dev];
  }

  // from here on identical to dfn_compact_indices.  Nothing in the scratch has to start at zero: every count, offset and
  // ballot word this call reads was written earlier in this call (whatever the previous call left is overwritten or unread).
  const long long chunks = (n + kChunk - 1) / kChunk;
  const dfn_compact::Workspace w = dfn_compact::carve(scratch, n);
  count_flags_kernel<<<static_cast<unsigned int>(chunks), kChunk, 0, st>>>(flags, n, w.counts, w.ballots);
  DFN_CHECK_LAUNCH("compact_indices_static(count)");
  scan_counts_kernel<<<1, 1024, 0, st>>>(w.counts, chunks, w.offsets, count);
  DFN_CHECK_LAUNCH("compact_indices_static(scan)");
  emit_indices_kernel<<<static_cast<unsigned int>(chunks), kChunk, 0, st>>>(idx_in, w.offsets, w.ballots, idx_out);
  DFN_CHECK_LAUNCH("compact_indices_static(emit)");
  return DFN_OK;
}
