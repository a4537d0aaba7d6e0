// Bring-up kernel: one CTA computes D[128 x N] = A * B^T-style product with tcgen05.mma from operand
// images already in the chunk layout, with every descriptor field supplied by the caller.  The GPU test
// suite uses it to pin the shared-memory descriptor conventions (K-major and MN-major, LBO/SBO roles)
// that the MLP kernels rely on.
#include "dfn_common.cuh"
#include "umma.cuh"

namespace {
using namespace umma;

struct SelfArgs {
  const uint8_t* a; const uint8_t* b;
  uint32_t a_bytes, b_bytes;
  int N, K;
  uint32_t a_mn, b_mn, a_lbo, a_sbo, b_lbo, b_sbo, a_kstep, b_kstep;
  float* d;
};

__global__ void __launch_bounds__(128, 1) umma_selftest_kernel(const SelfArgs s) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sa = smem_u32(smem), sb = sa + ((s.a_bytes + 1023u) & ~1023u);
  const uint32_t bar_load = smem_u32(&bars[0]), bar_done = smem_u32(&bars[1]);
  if (threadIdx.x == 0) { mbar_init(bar_load, 1); mbar_init(bar_done, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_ptr), 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = *reinterpret_cast<volatile uint32_t*>(&tmem_ptr);
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(bar_load, s.a_bytes + s.b_bytes);
    for (uint32_t off = 0; off < s.a_bytes; off += 32768) {
      const uint32_t nb = s.a_bytes - off < 32768 ? s.a_bytes - off : 32768;
      bulk_g2s(sa + off, s.a + off, nb, bar_load);
    }
    for (uint32_t off = 0; off < s.b_bytes; off += 32768) {
      const uint32_t nb = s.b_bytes - off < 32768 ? s.b_bytes - off : 32768;
      bulk_g2s(sb + off, s.b + off, nb, bar_load);
    }
    mbar_wait(bar_load, 0);
    tc_fence_after();
    const uint32_t idesc = instr_desc_bf16(128, s.N, s.a_mn, s.b_mn);
    for (int k = 0; k < s.K / 16; ++k) {
      const uint64_t ad = smem_desc(sa + k * s.a_kstep, s.a_lbo, s.a_sbo);
      const uint64_t bd = smem_desc(sb + k * s.b_kstep, s.b_lbo, s.b_sbo);
      mma_bf16(tbase, ad, bd, idesc, k > 0 ? 1u : 0u);
    }
    mma_commit(bar_done);
  }
  __syncwarp();
  mbar_wait(bar_done, 0);
  tc_fence_after();
  const int row = warp * 32 + lane;
  for (int c0 = 0; c0 < s.N; c0 += 16) {
    uint32_t v[16];
    tmem_ld16(tbase + (static_cast<uint32_t>(warp * 32) << 16) + c0, v);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 16; ++i) s.d[static_cast<size_t>(row) * s.N + c0 + i] = __uint_as_float(v[i]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 256);
}

}  // namespace

extern "C" int dfn_umma_selftest(const void* a_tile, const void* b_tile, int N, int K, int a_mn_major, int b_mn_major,
                                 unsigned a_lbo, unsigned a_sbo, unsigned b_lbo, unsigned b_sbo, unsigned a_kstep,
                                 unsigned b_kstep, float* d_out, void* stream) {
  DFN_CHECK_ARG(a_tile && b_tile && d_out, "umma_selftest: null pointer");
  DFN_CHECK_ARG(N >= 16 && N <= 256 && N % 16 == 0 && K >= 16 && K <= 256 && K % 16 == 0, "umma_selftest: bad N/K");
  SelfArgs s{};
  s.a = static_cast<const uint8_t*>(a_tile); s.b = static_cast<const uint8_t*>(b_tile);
  s.a_bytes = 128u * K * 2u; s.b_bytes = static_cast<uint32_t>(N) * K * 2u;
  s.N = N; s.K = K; s.a_mn = a_mn_major; s.b_mn = b_mn_major;
  s.a_lbo = a_lbo; s.a_sbo = a_sbo; s.b_lbo = b_lbo; s.b_sbo = b_sbo; s.a_kstep = a_kstep; s.b_kstep = b_kstep;
  s.d = d_out;
  const uint32_t smem = ((s.a_bytes + 1023u) & ~1023u) + s.b_bytes;
  DFN_CUDA(cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem), "umma_selftest");
  umma_selftest_kernel<<<1, 128, smem, static_cast<cudaStream_t>(stream)>>>(s);
  DFN_CHECK_LAUNCH("umma_selftest");
  return DFN_OK;
}
