// Bring-up microbenchmarks (not on the product path): issue-rate of tcgen05.mma under different
// commit / wait patterns and the TMEM -> register drain rate of the epilogue warps.  Results feed the
// pipeline design notes in DESIGN.md.
#include "../../deepfields-nerf-_b200/csrc/dfn_common.cuh"
#include "../../deepfields-nerf-_b200/csrc/umma.cuh"

namespace {
using namespace umma;

constexpr int kMaxDepth = 8;

struct BenchArgs {
  int mode;    // 0: K-major MMA stream; 1: TMEM drain by 8 warps; 2: MN-major MMA stream; 3: MMA + drain together
  int iters;   // K=256 GEMMs (16 MMAs each) per CTA
  int group;   // MMAs per tcgen05.commit
  int depth;   // committed groups in flight before the issuer waits
  int N;
  long long* cycles;  // [grid]
};

__device__ __forceinline__ uint64_t mk_desc(uint32_t lo, uint32_t hi) {
  uint64_t d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
  return d;
}

// mode 8: the K loop fully unrolled, descriptors = constant high word + incremented low word
template <int N, int GROUP>
__device__ __forceinline__ void tight_stream(uint32_t sa, uint32_t sb, uint32_t tbase, uint32_t bar0, int iters) {
  constexpr uint32_t idesc = instr_desc_bf16(128, N, 0, 0);
  constexpr uint32_t hi = (128u >> 4) | (1u << 14);
  const uint32_t a_lo = ((sa >> 4) & 0x3FFFu) | ((2048u >> 4) << 16);
  const uint32_t b_lo = ((sb >> 4) & 0x3FFFu) | ((static_cast<uint32_t>(N) * 16u >> 4) << 16);
  uint32_t ph0 = 0, ph1 = 0;
  int pend0 = 0, pend1 = 0;  // committed, not yet waited
  for (int it = 0; it < iters; ++it) {
    const uint32_t tacc = tbase + (it & 1) * 256;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const int slot = GROUP == 16 ? (it & 1) : ((k / GROUP) & 1);
      if (k % GROUP == 0) {
        if (slot == 0 && pend0) { mbar_wait(bar0, ph0); ph0 ^= 1; pend0 = 0; tc_fence_after(); }
        if (slot == 1 && pend1) { mbar_wait(bar0 + 8, ph1); ph1 ^= 1; pend1 = 0; tc_fence_after(); }
      }
      mma_bf16(tacc, mk_desc(a_lo + k * 256u, hi), mk_desc(b_lo + k * (2u * N * 16u >> 4), hi), idesc, k > 0 ? 1u : 0u);
      if (k % GROUP == GROUP - 1) {
        if (slot == 0) { mma_commit(bar0); pend0 = 1; } else { mma_commit(bar0 + 8); pend1 = 1; }
      }
    }
  }
  if (pend0) mbar_wait(bar0, ph0);
  if (pend1) mbar_wait(bar0 + 8, ph1);
}

// mode 9: converged warp, one asm block per group of 2 MMAs + commit, election inside the block
template <int N>
__device__ __forceinline__ void elect_stream(uint32_t sa, uint32_t sb, uint32_t tbase, uint32_t bar0, int iters) {
  constexpr uint32_t idesc = instr_desc_bf16(128, N, 0, 0);
  constexpr uint32_t hi = (128u >> 4) | (1u << 14);
  const uint32_t a_lo = ((sa >> 4) & 0x3FFFu) | ((2048u >> 4) << 16);
  const uint32_t b_lo = ((sb >> 4) & 0x3FFFu) | ((static_cast<uint32_t>(N) * 16u >> 4) << 16);
  uint32_t ph0 = 0, ph1 = 0;
  int pend0 = 0, pend1 = 0;
  for (int it = 0; it < iters; ++it) {
    const uint32_t tacc = tbase + (it & 1) * 256;
#pragma unroll
    for (int k = 0; k < 16; k += 2) {
      const int slot = (k >> 1) & 1;
      const uint32_t bar = slot == 0 ? bar0 : bar0 + 8;
      const uint32_t par = slot == 0 ? ph0 : ph1;
      const uint32_t need = slot == 0 ? pend0 : pend1;
      asm volatile(
          "{\n\t"
          ".reg .pred q, p, pe;\n\t"
          ".reg .b32 al, bl;\n\t"
          ".reg .b64 da, db;\n\t"
          "setp.eq.u32 p, %8, 0;\n\t"
          "@p bra EW_GO;\n\t"
          "EW_WAIT:\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 q, [%0], %1;\n\t"
          "@!q bra EW_WAIT;\n\t"
          "EW_GO:\n\t"
          "tcgen05.fence::after_thread_sync;\n\t"
          "elect.sync _|pe, 0xffffffff;\n\t"
          "setp.ne.b32 p, %7, 0;\n\t"
          "mov.b64 da, {%3, %5};\n\t"
          "mov.b64 db, {%4, %5};\n\t"
          "@pe tcgen05.mma.cta_group::1.kind::f16 [%2], da, db, %6, p;\n\t"
          "add.u32 al, %3, 256;\n\t"
          "add.u32 bl, %4, %9;\n\t"
          "mov.b64 da, {al, %5};\n\t"
          "mov.b64 db, {bl, %5};\n\t"
          "@pe tcgen05.mma.cta_group::1.kind::f16 [%2], da, db, %6, 1;\n\t"
          "@pe tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t"
          "}"
          ::"r"(bar), "r"(par), "r"(tacc), "r"(a_lo + k * 256u), "r"(b_lo + k * (2u * N * 16u >> 4)), "r"(hi), "r"(idesc),
            "r"(k > 0 ? 1u : 0u), "r"(need), "n"(2 * N * 16 >> 4)
          : "memory");
      if (slot == 0) { if (pend0) ph0 ^= 1; pend0 = 1; } else { if (pend1) ph1 ^= 1; pend1 = 1; }
    }
  }
  if (pend0) mbar_wait(bar0, ph0);
  if (pend1) mbar_wait(bar0 + 8, ph1);
}

__global__ void __launch_bounds__(320, 1) umma_bench_kernel(const BenchArgs b) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[kMaxDepth + 2];
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sa = smem_u32(smem), sb = sa + 65536;
  if (threadIdx.x == 0) {
    for (int i = 0; i < kMaxDepth + 2; ++i) mbar_init(smem_u32(&bars[i]), 1);
    fence_mbar_init();
  }
  // operands: pseudo-random bf16 in [-2,2)
  for (uint32_t i = threadIdx.x; i < (65536u + 131072u) / 4; i += blockDim.x) {
    uint32_t h = i * 2654435761u + blockIdx.x * 40503u;
    h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
    const uint32_t lo = 0x3F00u | (h & 0x80FFu), hi = 0x3F00u | ((h >> 16) & 0x80FFu);
    reinterpret_cast<uint32_t*>(smem)[i] = lo | (hi << 16);
  }
  fence_proxy_async();
  if (warp == 1) tmem_alloc(smem_u32(&tmem_ptr), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = *reinterpret_cast<volatile uint32_t*>(&tmem_ptr);
  const uint32_t bar_go = smem_u32(&bars[kMaxDepth]);
  long long t0 = 0, t1 = 0;

  const bool do_mma = b.mode == 0 || b.mode == 2 || (b.mode >= 3 && b.mode != 8 && b.mode != 9);
  const bool do_drain = b.mode == 1 || b.mode == 3;
  if (warp == 1 && b.mode == 9) {
    t0 = clock64();
    const uint32_t bar0 = smem_u32(&bars[0]);
    if (b.N == 256) elect_stream<256>(sa, sb, tbase, bar0, b.iters); else elect_stream<64>(sa, sb, tbase, bar0, b.iters);
    t1 = clock64();
    if (lane == 0) b.cycles[blockIdx.x] = t1 - t0;
  } else if (warp == 1 && lane == 0 && b.mode == 8) {
    t0 = clock64();
    const uint32_t bar0 = smem_u32(&bars[0]);
    if (b.N == 256) {
      if (b.group == 16) tight_stream<256, 16>(sa, sb, tbase, bar0, b.iters);
      else if (b.group == 4) tight_stream<256, 4>(sa, sb, tbase, bar0, b.iters);
      else if (b.group == 2) tight_stream<256, 2>(sa, sb, tbase, bar0, b.iters);
      else tight_stream<256, 1>(sa, sb, tbase, bar0, b.iters);
    } else {
      if (b.group == 16) tight_stream<64, 16>(sa, sb, tbase, bar0, b.iters);
      else if (b.group == 4) tight_stream<64, 4>(sa, sb, tbase, bar0, b.iters);
      else if (b.group == 2) tight_stream<64, 2>(sa, sb, tbase, bar0, b.iters);
      else tight_stream<64, 1>(sa, sb, tbase, bar0, b.iters);
    }
    t1 = clock64();
    b.cycles[blockIdx.x] = t1 - t0;
  } else if (warp == 1 && b.mode == 6) {
    // CUTLASS-style: the whole warp runs the loop converged, one elected lane issues
    const uint32_t idesc = instr_desc_bf16(128, b.N, 0, 0);
    t0 = clock64();
    int g = 0, in_group = 0;
    for (int it = 0; it < b.iters; ++it) {
      const uint32_t tacc = tbase + (it & 1) * 256;
      for (int k = 0; k < 16; ++k) {
        if (in_group == 0 && g >= b.depth) {
          mbar_wait(smem_u32(&bars[g % b.depth]), ((g / b.depth) - 1) & 1);
          tc_fence_after();
        }
        if (elect_one())
          mma_bf16(tacc, smem_desc(sa + k * 4096u, 2048u, 128u), smem_desc(sb + k * 2u * b.N * 16u, b.N * 16u, 128u), idesc, k > 0);
        __syncwarp();
        if (++in_group == b.group) {
          if (elect_one()) mma_commit(smem_u32(&bars[g % b.depth]));
          __syncwarp();
          in_group = 0; ++g;
        }
      }
    }
    for (int w = (g > b.depth ? g - b.depth : 0); w < g; ++w) mbar_wait(smem_u32(&bars[w % b.depth]), (w / b.depth) & 1);
    t1 = clock64();
    if (lane == 0) b.cycles[blockIdx.x] = t1 - t0;
  } else if (warp == 1 && lane == 0 && do_mma) {
    const bool mn = b.mode == 2;
    const uint32_t idesc = instr_desc_bf16(b.mode == 7 ? 64 : 128, b.N, mn, mn);
    const uint32_t a_lbo = mn ? 128u : 2048u, a_sbo = mn ? 2048u : 128u;
    const uint32_t b_lbo = mn ? 128u : b.N * 16u, b_sbo = mn ? 2048u : 128u;
    const uint32_t a_k = mn ? 256u : 4096u, b_k = mn ? 256u : 2u * b.N * 16u;
    t0 = clock64();
    int g = 0, in_group = 0;
    for (int it = 0; it < b.iters; ++it) {
      const uint32_t tacc = tbase + (it & 1) * 256;
      for (int k = 0; k < 16; ++k) {
        if (in_group == 0 && g >= b.depth) {
          mbar_wait(smem_u32(&bars[g % b.depth]), ((g / b.depth) - 1) & 1);
          tc_fence_after();
        }
        const int kk = mn ? (k & 7) : k;
        if (b.mode == 4) {
          // SWIZZLE_128B K-major: [k-block of 64][rows][128 B], 8-row atoms of 1024 B; +32 B per K=16 step
          const uint32_t ao = (k >> 2) * 16384u + (k & 3) * 32u, bo = (k >> 2) * (b.N * 128u) + (k & 3) * 32u;
          const uint64_t sw = static_cast<uint64_t>(2) << 61;
          mma_bf16(tacc, smem_desc(sa + ao, 16, 1024) | sw, smem_desc(sb + bo, 16, 1024) | sw, idesc, k > 0);
        } else if (b.mode == 5) {
          // independent accumulators: alternate the TMEM target every MMA
          mma_bf16(tbase + (k & 1) * 256, smem_desc(sa + kk * a_k, a_lbo, a_sbo), smem_desc(sb + kk * b_k, b_lbo, b_sbo), idesc, k > 1);
        } else {
        mma_bf16(tacc, smem_desc(sa + kk * a_k, a_lbo, a_sbo), smem_desc(sb + kk * b_k, b_lbo, b_sbo), idesc, k > 0);
        }
        if (++in_group == b.group) {
          mma_commit(smem_u32(&bars[g % b.depth]));
          in_group = 0; ++g;
        }
      }
    }
    if (in_group) { mma_commit(smem_u32(&bars[g % b.depth])); ++g; }
    // drain: wait for the last `depth` groups
    for (int w = (g > b.depth ? g - b.depth : 0); w < g; ++w) mbar_wait(smem_u32(&bars[w % b.depth]), (w / b.depth) & 1);
    t1 = clock64();
    b.cycles[blockIdx.x] = t1 - t0;
    (void)bar_go;
  }
  if (warp >= 2 && do_drain) {
    const int q = warp & 3, hh = (warp - 2) >> 2;
    const uint32_t taddr = tbase + (static_cast<uint32_t>(q * 32) << 16) + hh * 128;
    uint32_t acc = 0;
    t0 = clock64();
    for (int it = 0; it < b.iters; ++it) {
      const uint32_t tb = taddr + (it & 1) * 256;
      uint32_t v0[32], v1[32];
      tmem_ld32(tb, v0);
      tmem_ld_wait();
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {
        uint32_t (&cur)[32] = (cc & 1) ? v1 : v0;
        uint32_t (&nxt)[32] = (cc & 1) ? v0 : v1;
        if (cc < 3) tmem_ld32(tb + (cc + 1) * 32, nxt);
#pragma unroll
        for (int i = 0; i < 32; ++i) acc ^= cur[i];
        if (cc < 3) tmem_ld_wait();
      }
    }
    t1 = clock64();
    if (acc == 0x12345678u) b.cycles[0] = 0;  // keep the loads alive
    if (b.mode == 1 && threadIdx.x == 64) b.cycles[blockIdx.x] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tbase, 512);
}

}  // namespace

extern "C" int dfn_umma_bench(int mode, int iters, int group, int depth, int N, int grid, long long* cycles_out,
                              void* stream) {
  DFN_CHECK_ARG(cycles_out && iters > 0 && group >= 1 && group <= 16 && depth >= 1 && depth <= kMaxDepth, "umma_bench: bad args");
  DFN_CHECK_ARG(N >= 16 && N <= 256 && N % 16 == 0 && grid >= 1, "umma_bench: bad N/grid");
  BenchArgs b{mode, iters, group, depth, N, cycles_out};
  const uint32_t smem = 65536 + 131072;
  DFN_CUDA(cudaFuncSetAttribute(umma_bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem), "umma_bench");
  umma_bench_kernel<<<grid, 320, smem, static_cast<cudaStream_t>(stream)>>>(b);
  DFN_CHECK_LAUNCH("umma_bench");
  return DFN_OK;
}
