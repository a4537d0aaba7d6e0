// Bring-up microbenchmark (not on the product path): how fast can the epilogue warps of one CTA per SM push
// 64 KB activation tiles to global memory, alone and next to a tcgen05.mma stream / fences.
//   mode bit 0: warp 1 runs an MMA stream (N = 256) concurrently
//   mode bit 1: fence.proxy.async + mbarrier arrive after every tile (what publish() does)
//   mode bit 2: stores go through shared memory + one 64 KB bulk copy (TMA) per tile instead of STG.128
//   mode bit 3: st.global.cs (streaming) instead of default stores
#include "../../deepfields-nerf-_b200/csrc/dfn_common.cuh"
#include "../../deepfields-nerf-_b200/csrc/umma.cuh"

namespace {
using namespace umma;

struct SArgs { int mode, iters; uint8_t* buf; long long n_tiles; long long* cycles; };

__global__ void __launch_bounds__(320, 1) store_bench_kernel(const SArgs b) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[4];
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sa = smem_u32(smem), sb = sa + 65536;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bars[0]), 1); mbar_init(smem_u32(&bars[1]), 1); mbar_init(smem_u32(&bars[2]), 256); bars[3] = 0; fence_mbar_init(); }
  for (uint32_t i = threadIdx.x; i < (65536u + 131072u) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3F003F00u + i;
  fence_proxy_async();
  if (warp == 1) tmem_alloc(smem_u32(&tmem_ptr), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = *reinterpret_cast<volatile uint32_t*>(&tmem_ptr);
  if (warp == 1 && (b.mode & 1)) {
    // MMA stream: 2 MMAs per commit, two commit barriers alternating
    constexpr uint32_t idesc = instr_desc_bf16(128, 256, 0, 0);
    const uint32_t bar0 = smem_u32(&bars[0]);
    uint32_t ph[2] = {0, 0}; int pend[2] = {0, 0};
    const long long t0 = clock64();
    int done = 0;
    for (int it = 0; it < b.iters * 16; ++it, ++done) {      // 16 K=256 GEMMs per stored tile -> tensor pipe never idles
      for (int k = 0; k < 16; k += 2) {
        const int s = (k >> 1) & 1;
        if (pend[s]) { mbar_wait(bar0 + 8 * s, ph[s]); ph[s] ^= 1; tc_fence_after(); }
        if (elect_one()) {
          mma_bf16(tbase + (it & 1) * 256, smem_desc(sa + k * 4096u, 2048u, 128u), smem_desc(sb + k * 8192u, 4096u, 128u), idesc, k > 0);
          mma_bf16(tbase + (it & 1) * 256, smem_desc(sa + (k + 1) * 4096u, 2048u, 128u), smem_desc(sb + (k + 1) * 8192u, 4096u, 128u), idesc, 1);
          mma_commit(bar0 + 8 * s);
        }
        __syncwarp();
        pend[s] = 1;
      }
      if (*reinterpret_cast<volatile int*>(&bars[3]) != 0) break;   // the store warps are done
    }
    if (pend[0]) mbar_wait(bar0, ph[0]);
    if (pend[1]) mbar_wait(bar0 + 8, ph[1]);
    if (lane == 0) b.cycles[gridDim.x + blockIdx.x] = (clock64() - t0) / (done > 0 ? done * 16 : 1);   // cycles per MMA
  }
  if (warp == 0 && lane == 0 && (b.mode & 128)) {
    // weight-stream stand-in: 16 KB bulk loads from an L2-resident 1.25 MB region into a 3-stage ring, back to back
    __shared__ __align__(8) uint64_t lb[3];
    for (int i = 0; i < 3; ++i) mbar_init(smem_u32(&lb[i]), 1);
    fence_mbar_init();
    const uint8_t* wsrc = b.buf + ((b.n_tiles - 32) << 16);
    uint32_t ph[3] = {0, 0, 0};
    long long n = 0;
    const long long t0 = clock64();
    for (int i = 0; i < 3; ++i) { mbar_arrive_expect_tx(smem_u32(&lb[i]), 16384); bulk_g2s(sb + 65536 + i * 16384, wsrc + (i % 76) * 16384, 16384, smem_u32(&lb[i])); }
    while (*reinterpret_cast<volatile int*>(&bars[3]) == 0) {
      const int i = static_cast<int>(n % 3);
      mbar_wait(smem_u32(&lb[i]), ph[i]); ph[i] ^= 1;
      mbar_arrive_expect_tx(smem_u32(&lb[i]), 16384);
      bulk_g2s(sb + 65536 + i * 16384, wsrc + ((n + 3) % 76) * 16384, 16384, smem_u32(&lb[i]));
      ++n;
    }
    for (int i = 0; i < 3; ++i) mbar_wait(smem_u32(&lb[(n + i) % 3]), ph[(n + i) % 3]);
    b.cycles[2 * gridDim.x + blockIdx.x] = (clock64() - t0) / (n > 0 ? n : 1);   // cycles per 16 KB stage
  }
  if (warp >= 2) {
    const int wt = threadIdx.x - 64, row = (warp & 3) * 32 + lane, hh = (warp - 2) >> 2;
    const long long t0 = clock64();
    for (int it = 0; it < b.iters; ++it) {
      const long long tile = (static_cast<long long>(it) * gridDim.x + blockIdx.x) % b.n_tiles;
      uint8_t* dst = b.buf + (tile << 16);
      if (b.mode & 4) {
#pragma unroll
        for (int g = 0; g < 16; ++g)
          *reinterpret_cast<uint4*>(smem + (hh * 16 + g) * 2048 + row * 16) = make_uint4(it, g, wt, 0);
        fence_proxy_async();
        named_bar_sync(1, 256);
        if (wt == 0) { bulk_s2g(dst, sa, 65536); bulk_commit(); bulk_wait_read_all(); }
        named_bar_sync(1, 256);
      } else {
        uint32_t acc = 0;
        if (b.mode & 32) {   // drain a 128-column accumulator half like the epilogue does
          const uint32_t taddr = tbase + (static_cast<uint32_t>((warp & 3) * 32) << 16) + hh * 128 + (it & 1) * 256;
#pragma unroll
          for (int cc = 0; cc < 4; ++cc) {
            uint32_t v[32];
            tmem_ld32(taddr + cc * 32, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) acc ^= v[i];
          }
        }
        if (b.mode & 16) {   // the act-tile writes of the epilogue
#pragma unroll
          for (int g = 0; g < 16; ++g)
            *reinterpret_cast<uint4*>(smem + (hh * 16 + g) * 2048 + row * 16) = make_uint4(it, g, wt, acc);
        }
        if (!(b.mode & 64)) {
#pragma unroll
          for (int g = 0; g < 16; ++g) {
            uint4* p = reinterpret_cast<uint4*>(dst + (hh * 16 + g) * 2048 + row * 16);
            if (b.mode & 8) __stcs(p, make_uint4(it, g, wt, acc)); else *p = make_uint4(it, g, wt, acc);
          }
        }
      }
      if (b.mode & 2) { fence_proxy_async(); mbar_arrive(smem_u32(&bars[2])); }
    }
    if ((b.mode & 4) && wt == 0) bulk_wait_all();
    if (wt == 0) { b.cycles[blockIdx.x] = clock64() - t0; *reinterpret_cast<volatile int*>(&bars[3]) = 1; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tbase, 512);
}
}  // namespace

extern "C" int dfn_store_bench(int mode, int iters, int grid, void* buf, long long buf_bytes, long long* cycles_out, void* stream) {
  DFN_CHECK_ARG(buf && cycles_out && iters > 0 && grid > 0 && buf_bytes >= 65536, "store_bench: bad args");
  SArgs a{mode, iters, static_cast<uint8_t*>(buf), buf_bytes >> 16, cycles_out};
  const uint32_t smem = 65536 + 131072 + 8192;
  DFN_CUDA(cudaFuncSetAttribute(store_bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem), "store_bench");
  store_bench_kernel<<<grid, 320, smem, static_cast<cudaStream_t>(stream)>>>(a);
  DFN_CHECK_LAUNCH("store_bench");
  return DFN_OK;
}

// ---- cta_group::2 issue-rate probe: the leader of a CTA pair streams M=256 N=256 K=16 MMAs -----------------
namespace {
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(320, 1) pair_mma_kernel(int iters, int N, long long* cycles, uint8_t* buf,
                                                                                        long long n_tiles, int mode) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const uint32_t sa = smem_u32(smem), sb = sa + 65536;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bars[0]), 1); mbar_init(smem_u32(&bars[1]), 1); fence_mbar_init(); }
  for (uint32_t i = threadIdx.x; i < (65536u + 65536u) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3F003F00u + i;
  fence_proxy_async();
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_ptr)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  tc_fence_after();
  const uint32_t tbase = *reinterpret_cast<volatile uint32_t*>(&tmem_ptr);
  __shared__ volatile int mma_done;
  if (threadIdx.x == 0) mma_done = 0;
  __syncthreads();
  if (warp >= 2 && (mode & 3)) {
    // 8 warps of each CTA stream 64 KB tiles to global (mode bit 0) and / or read them back (bit 1) until the MMAs finish
    const int row = (warp & 3) * 32 + lane, hh = (warp - 2) >> 2;
    uint32_t acc = 0;
    for (long long it = 0; !mma_done && !(rank == 1 && it >= 16LL * iters); ++it) {
      const long long tile = (it * gridDim.x + blockIdx.x) % n_tiles;
      uint8_t* dst = buf + (tile << 16);
#pragma unroll
      for (int g = 0; g < 16; ++g) {
        uint4* p = reinterpret_cast<uint4*>(dst + (hh * 16 + g) * 2048 + row * 16);
        if (mode & 2) { const uint4 v = *p; acc ^= v.x ^ v.y ^ v.z ^ v.w; }
        if (mode & 1) *p = make_uint4(static_cast<uint32_t>(it), g, row, acc);
      }
    }
    if (acc == 0x12345u) cycles[0] = 1;
  }
  if (warp == 0 && rank == 0) {
    const uint32_t idesc = instr_desc_bf16(256, N, 0, 0);
    const uint32_t cb = (N / 2) * 16;   // bytes per K chunk of this CTA's half of B
    const uint32_t bar0 = smem_u32(&bars[0]);
    uint32_t ph[2] = {0, 0}; int pend[2] = {0, 0};
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      for (int k = 0; k < 16; k += 2) {
        const int s = (k >> 1) & 1;
        if (pend[s]) { mbar_wait(bar0 + 8 * s, ph[s]); ph[s] ^= 1; tc_fence_after(); }
        if (elect_one()) {
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const uint64_t da = smem_desc(sa + (k + i) * 4096u, 2048u, 128u), db = smem_desc(sb + (k + i) * 2 * cb, cb, 128u);
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tbase + (it & 1) * 256), "l"(da), "l"(db),
                         "r"(idesc), "r"(k + i > 0 ? 1u : 0u) : "memory");
          }
          asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar0 + 8 * s),
                       "h"(static_cast<uint16_t>(1)) : "memory");
        }
        __syncwarp();
        pend[s] = 1;
      }
    }
    if (pend[0]) mbar_wait(bar0, ph[0]);
    if (pend[1]) mbar_wait(bar0 + 8, ph[1]);
    if (lane == 0) cycles[blockIdx.x >> 1] = (clock64() - t0) / (iters * 16);
    mma_done = 1;
  }
  tc_fence_before();
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tbase), "r"(512u) : "memory");
}
}  // namespace

extern "C" int dfn_pair_mma_bench(int iters, int N, int grid, long long* cycles_out, void* buf, long long buf_bytes, int mode,
                                  void* stream) {
  DFN_CHECK_ARG(cycles_out && iters > 0 && grid > 0 && grid % 2 == 0 && N >= 32 && N <= 256 && N % 32 == 0, "pair_mma_bench: bad args");
  const uint32_t smem = 65536 + 65536;
  DFN_CUDA(cudaFuncSetAttribute(pair_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem), "pair_mma_bench");
  pair_mma_kernel<<<grid, 320, smem, static_cast<cudaStream_t>(stream)>>>(iters, N, cycles_out, static_cast<uint8_t*>(buf),
                                                                          buf_bytes >> 16, mode);
  DFN_CHECK_LAUNCH("pair_mma_bench");
  return DFN_OK;
}
