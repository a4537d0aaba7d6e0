// Ordered stream compaction shared by raygen.cu and volpoints.cu: pass 1 of the caller counts the kept elements per
// 1024-element chunk (and keeps the warp ballots), scan_counts_kernel turns the counts into offsets, pass 2 of the
// caller writes the kept elements in order.
#pragma once
#include "dfn_common.cuh"

namespace dfn_compact {

constexpr int kChunk = 1024;   // elements per CTA

// workspace layout: counts (u32 per chunk) | offsets (i64 per chunk) | ballots (u32 per warp), each 16-byte aligned
struct Workspace {
  unsigned int* counts;
  long long* offsets;
  unsigned int* ballots;
};
inline long long workspace_bytes(long long n) {
  const long long chunks = (n + kChunk - 1) / kChunk;
  return chunks * 4 + 16 + chunks * 8 + 16 + chunks * (kChunk / 8);
}
inline Workspace carve(void* workspace, long long n) {
  const long long chunks = (n + kChunk - 1) / kChunk;
  Workspace w;
  w.counts = static_cast<unsigned int*>(workspace);
  w.offsets = reinterpret_cast<long long*>(static_cast<uint8_t*>(workspace) + ((chunks * 4 + 15) / 16) * 16);
  w.ballots = reinterpret_cast<unsigned int*>(reinterpret_cast<uint8_t*>(w.offsets) + ((chunks * 8 + 15) / 16) * 16);
  return w;
}

// exclusive scan of the chunk counts (one block; the carry runs through the 1024-wide tiles), total -> *total_out
static __global__ void __launch_bounds__(1024) scan_counts_kernel(const unsigned int* __restrict__ counts, long long n_chunks,
                                                            long long* __restrict__ offsets, long long* __restrict__ total_out) {
  __shared__ long long warp_sums[32];
  __shared__ long long carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (long long base = 0; base < n_chunks; base += 1024) {
    const long long i = base + threadIdx.x;
    const long long x = i < n_chunks ? counts[i] : 0;
    long long s = x;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const long long y = __shfl_up_sync(0xffffffffu, s, off);
      if (lane >= off) s += y;
    }
    if (lane == 31) warp_sums[warp] = s;
    __syncthreads();
    if (warp == 0) {
      long long w = warp_sums[lane];
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const long long y = __shfl_up_sync(0xffffffffu, w, off);
        if (lane >= off) w += y;
      }
      warp_sums[lane] = w;   // inclusive
    }
    __syncthreads();
    const long long before = carry + (warp > 0 ? warp_sums[warp - 1] : 0) + (s - x);
    if (i < n_chunks) offsets[i] = before;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + x;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total_out = carry;
}


// position of a kept element inside its chunk (all threads of the CTA call this; warp_counts: kChunk / 32 ints of smem)
__device__ __forceinline__ int rank_in_chunk(unsigned int ballot, int* warp_counts) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) warp_counts[warp] = __popc(ballot);
  __syncthreads();
  if (warp == 0) {
    int w = warp_counts[lane];
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, w, off);
      if (lane >= off) w += y;
    }
    warp_counts[lane] = w;   // inclusive
  }
  __syncthreads();
  return (warp > 0 ? warp_counts[warp - 1] : 0) + __popc(ballot & ((1u << lane) - 1u));
}

}  // namespace dfn_compact
