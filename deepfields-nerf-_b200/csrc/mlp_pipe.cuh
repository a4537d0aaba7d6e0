// The pipeline shared by the forward and data-gradient kernels (mlp_fwd.cu, mlp_bwd.cu): CTA pairs
// (tcgen05 cta_group::2), a weight ring fed by tensor-map TMA loads, and the unrolled MMA issue path.
// A kernel instantiates the templates with a policy class PP that provides
//   kNumJobs, kStages, kCPS (K chunks per ring stage), kRingStageBytes, kOffRing, bar_full_i / bar_empty_i / bar_accfull_i / bar_aready_i      (shared memory map)
//   job_N(j), job_chunks(j) (K chunks of 8, bias slice included), job_off(j), job_half_chunk(j), job_half_bytes(j)
//   struct ABases, bases(sbase, slot), a_chunk_lo<J>(kc, bases)                                  (where the A operand lives)
#pragma once
#include "mlp_common.cuh"
#include <cuda.h>
#include <cudaTypedefs.h>
#include <utility>
#include <stdlib.h>

namespace mlp_pipe {
using namespace mlp;
using namespace umma;


// The packed weight image seen as a 2-D byte tensor [bytes/128][128]; one map per box height (= stage size / 128):
// 128 rows = 16 KB (a full N=256 stage), 32 rows = the 2-chunk tail stage, 64 / 48 rows = the N=128 job's stages.
struct WeightMaps { CUtensorMap m[7]; };
constexpr unsigned kMapRows[7] = {128, 64, 48, 32, 16, 8, 256};   // box heights (x 128 bytes) a ring stage can have
__host__ __device__ constexpr int map_index(uint32_t stage_bytes) {
  return stage_bytes == 32768 ? 6 : stage_bytes == 16384 ? 0 : stage_bytes == 8192 ? 1 : stage_bytes == 6144 ? 2 : stage_bytes == 4096 ? 3 : stage_bytes == 2048 ? 4 : 5;
}

// ---- cluster helpers --------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of the same shared-memory location in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// wait with cluster-scope acquire: the arrivals come from the peer CTA
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  long long t0 = 0;
  for (;;) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) return;
    if (t0 == 0) t0 = clock64();
    else if (clock64() - t0 > 4000000000LL) {
      printf("dfn: cluster mbarrier timeout (block %d thread %d bar 0x%x parity %u)\n", blockIdx.x, threadIdx.x, bar, parity);
      __trap();
    }
  }
}

// tile of (iteration, slot) for this CTA: super-tile index * 2 + rank
__device__ __forceinline__ int tile_of(int it, int t, int rank) {
  const int n_pairs = gridDim.x >> 1, pair = blockIdx.x >> 1;
  return 2 * ((it * 2 + t) * n_pairs + pair) + rank;
}

// ---- weight ring + MMA issue, one instantiation per job so every shape is a compile-time constant -----
// A generic issue loop with run-time shapes costs ~300 cycles per tcgen05.mma in address arithmetic alone
// (N = 256 executes in 128).  Everything below is therefore unrolled with static ring-stage indices, and a
// whole stage (wait, MMAs, commit) is one asm block (tools/umma_bench.py measures both forms).
// A ring stage holds kCPS K-chunks of this CTA's half of the weight block (N/2 rows each).
template <class PP> constexpr int job_stages(int j) { return (PP::job_chunks(j) + PP::kCPS - 1) / PP::kCPS; }
template <class PP> constexpr int stage_chunks(int j, int k) { return (PP::job_chunks(j) - k * PP::kCPS) < PP::kCPS ? (PP::job_chunks(j) - k * PP::kCPS) : PP::kCPS; }
template <class PP> constexpr int stage_prefix(int j, int t) {  // ring uses before (job j, slot t) within one iteration
  int n = 0;
  for (int i = 0; i < j; ++i) n += 2 * job_stages<PP>(i);
  return n + t * job_stages<PP>(j);
}

template <class PP>
struct ProdState {
  uint32_t use[2][PP::kStages];   // parity of the number of fills so far per (slot, stage)
  uint64_t pol;               // L2 evict_last policy for the weight loads (the image is re-read by every tile)
  int last[PP::kStages];          // slot that used the stage last (-1: never)
  uint32_t row0;              // first 128-byte row of the replica of the weight image this pair streams (weight_replicas())
};

// tensor-map TMA load of `rows` x 128 bytes starting at row y of the weight image; completion (bytes) is signalled on
// `mbar`, a shared::cluster address that may belong to the peer (leader) CTA
__device__ __forceinline__ void tma_load_rows(uint32_t dst_smem, const CUtensorMap* map, uint32_t y, uint32_t mbar, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3}], [%4], %5;" ::"r"(dst_smem),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(0u), "r"(y), "r"(mbar), "l"(pol)
      : "memory");
}

template <class PP, int J>
__device__ __forceinline__ void produce_job(const WeightMaps& maps, uint32_t rank, uint32_t ring, uint32_t bar0, ProdState<PP>& ps,
                                            const bool (&has)[2]) {
  constexpr uint32_t cb = PP::job_half_chunk(J);
  const uint32_t src = PP::job_off(J) + rank * PP::job_half_bytes(J);   // byte offset in the weight image
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    if (!has[t]) continue;
#pragma unroll
    for (int k = 0; k < job_stages<PP>(J); ++k) {
      constexpr uint32_t kFull = PP::kCPS * cb;
      const uint32_t off = k * kFull;
      const uint32_t nb = static_cast<uint32_t>(stage_chunks<PP>(J, k)) * cb;
      const int stage = (stage_prefix<PP>(J, t) + k) % PP::kStages;
      // the previous user of this stage (either slot) must have released it (multicast commit of the leader)
      if (ps.last[stage] == 0) mbar_wait(bar0 + 8u * PP::bar_empty_i(0, stage), ps.use[0][stage] ^ 1);
      else if (ps.last[stage] == 1) mbar_wait(bar0 + 8u * PP::bar_empty_i(1, stage), ps.use[1][stage] ^ 1);
      ps.use[t][stage] ^= 1;
      ps.last[stage] = t;
      // both halves complete on the LEADER's full barrier; its producer expects the bytes of the pair
      const uint32_t full = bar0 + 8u * PP::bar_full_i(t, stage);
      const int mi = map_index(nb);
      if (rank == 0) mbar_arrive_expect_tx(full, 2 * nb);
      tma_load_rows(ring + stage * PP::kRingStageBytes, &maps.m[mi], ps.row0 + ((src + off) >> 7), rank == 0 ? full : map_to_cta(full, 0), ps.pol);
    }
  }
}
template <class PP, int... Js>
__device__ __forceinline__ void produce_all(const WeightMaps& maps, uint32_t rank, uint32_t ring, uint32_t bar0, ProdState<PP>& ps,
                                            const bool (&has)[2], std::integer_sequence<int, Js...>) {
  (produce_job<PP, Js>(maps, rank, ring, bar0, ps, has), ...);
}

// One ring stage as a single asm block: bounded wait for the weights of BOTH CTAs (the leader's full barrier counts
// its own producer and the peer's relay; a protocol bug traps instead of hanging), NMMA x tcgen05.mma.cta_group::2
// (K = 16 each) and the multicast commit that releases the stage in both CTAs.  MMA i < NMMA-1 reads A at
// a_lo + i*256 (two chunks further), the last one at a_last; B advances by B_STEP descriptor units per MMA.
#define DFN_STAGE_HEAD                                                        \
  "{\n\t"                                                                     \
  ".reg .pred q, p, pe;\n\t"                                                  \
  ".reg .b32 cnt, al, bl;\n\t"                                                \
  ".reg .b64 da, db;\n\t"                                                     \
  "mov.u32 cnt, 0;\n\t"                                                       \
  "SW_WAIT:\n\t"                                                              \
  "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 q, [%0], %1;\n\t" \
  "@q bra SW_GO;\n\t"                                                         \
  "add.u32 cnt, cnt, 1;\n\t"                                                  \
  "setp.lt.u32 p, cnt, 8000000;\n\t"                                          \
  "@p bra SW_WAIT;\n\t"                                                       \
  "trap;\n\t"                                                                 \
  "SW_GO:\n\t"                                                                \
  "tcgen05.fence::after_thread_sync;\n\t"                                     \
  "elect.sync _|pe, 0xffffffff;\n\t"                                          \
  "setp.ne.b32 p, %7, 0;\n\t"
#define DFN_STAGE_FIRST                                         \
  "mov.b64 da, {%3, %5};\n\t"                                   \
  "mov.b64 db, {%4, %5};\n\t"                                   \
  "@pe tcgen05.mma.cta_group::2.kind::f16 [%2], da, db, %6, p;\n\t"
#define DFN_STAGE_NEXT(i)                                       \
  "add.u32 al, %3, " #i "*256;\n\t"                             \
  "add.u32 bl, %4, " #i "*%10;\n\t"                             \
  "mov.b64 da, {al, %5};\n\t"                                   \
  "mov.b64 db, {bl, %5};\n\t"                                   \
  "@pe tcgen05.mma.cta_group::2.kind::f16 [%2], da, db, %6, 1;\n\t"
#define DFN_STAGE_LAST(i)                                       \
  "add.u32 bl, %4, " #i "*%10;\n\t"                             \
  "mov.b64 da, {%9, %5};\n\t"                                   \
  "mov.b64 db, {bl, %5};\n\t"                                   \
  "@pe tcgen05.mma.cta_group::2.kind::f16 [%2], da, db, %6, 1;\n\t"
#define DFN_STAGE_ONLY                                          \
  "mov.b64 da, {%9, %5};\n\t"                                   \
  "mov.b64 db, {%4, %5};\n\t"                                   \
  "@pe tcgen05.mma.cta_group::2.kind::f16 [%2], da, db, %6, p;\n\t"
#define DFN_COMMIT(b) \
  "@pe tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [" b "], %12;\n\t"
#define DFN_STAGE_TAIL DFN_COMMIT("%8") "}"
// last stage of a job: the second commit tells the epilogues of both CTAs that the accumulator is complete
#define DFN_STAGE_TAIL_ACC DFN_COMMIT("%8") DFN_COMMIT("%11") "}"
#define DFN_STAGE_OPERANDS                                                                                     \
  ::"r"(full_bar), "r"(parity), "r"(tacc), "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(acc_first), "r"(empty_bar), \
      "r"(a_last), "n"(B_STEP), "r"(acc_bar), "h"(static_cast<uint16_t>(3))                                       \
      : "memory"

// Executed by the whole (converged) issuer warp of the leader CTA.
template <int NMMA, int B_STEP, bool ACC>
__device__ __forceinline__ void stage_mma(uint32_t full_bar, uint32_t parity, uint32_t tacc, uint32_t a_lo, uint32_t a_last,
                                          uint32_t b_lo, uint32_t hi, uint32_t idesc, uint32_t acc_first, uint32_t empty_bar,
                                          uint32_t acc_bar) {
  static_assert((NMMA >= 1 && NMMA <= 4) || NMMA == 8, "1..4 or 8 MMAs per stage");
  if constexpr (ACC) {
    if constexpr (NMMA == 1) asm volatile(DFN_STAGE_HEAD DFN_STAGE_ONLY DFN_STAGE_TAIL_ACC DFN_STAGE_OPERANDS);
    else if constexpr (NMMA == 2) asm volatile(DFN_STAGE_HEAD DFN_STAGE_FIRST DFN_STAGE_LAST(1) DFN_STAGE_TAIL_ACC DFN_STAGE_OPERANDS);
    else if constexpr (NMMA == 3)
      asm volatile(DFN_STAGE_HEAD DFN_STAGE_FIRST DFN_STAGE_NEXT(1) DFN_STAGE_LAST(2) DFN_STAGE_TAIL_ACC DFN_STAGE_OPERANDS);
    else if constexpr (NMMA == 8)
      asm volatile(DFN_STAGE_HEAD DFN_STAGE_FIRST DFN_STAGE_NEXT(1) DFN_STAGE_NEXT(2) DFN_STAGE_NEXT(3) DFN_STAGE_NEXT(4) DFN_STAGE_NEXT(5)
                       DFN_STAGE_NEXT(6) DFN_STAGE_LAST(7) DFN_STAGE_TAIL_ACC DFN_STAGE_OPERANDS);
    else
      asm volatile(DFN_STAGE_HEAD DFN_STAGE_FIRST DFN_STAGE_NEXT(1) DFN_STAGE_NEXT(2) DFN_STAGE_LAST(3) DFN_STAGE_TAIL_ACC
                       DFN_STAGE_OPERANDS);
  } else {
    if constexpr (NMMA == 1) asm volatile(DFN_STAGE_HEAD DFN_STAGE_ONLY DFN_STAGE_TAIL DFN_STAGE_OPERANDS);
    else if constexpr (NMMA == 2) asm volatile(DFN_STAGE_HEAD DFN_STAGE_FIRST DFN_STAGE_LAST(1) DFN_STAGE_TAIL DFN_STAGE_OPERANDS);
    else if constexpr (NMMA == 3)
      asm volatile(DFN_STAGE_HEAD DFN_STAGE_FIRST DFN_STAGE_NEXT(1) DFN_STAGE_LAST(2) DFN_STAGE_TAIL DFN_STAGE_OPERANDS);
    else if constexpr (NMMA == 8)
      asm volatile(DFN_STAGE_HEAD DFN_STAGE_FIRST DFN_STAGE_NEXT(1) DFN_STAGE_NEXT(2) DFN_STAGE_NEXT(3) DFN_STAGE_NEXT(4) DFN_STAGE_NEXT(5)
                       DFN_STAGE_NEXT(6) DFN_STAGE_LAST(7) DFN_STAGE_TAIL DFN_STAGE_OPERANDS);
    else
      asm volatile(DFN_STAGE_HEAD DFN_STAGE_FIRST DFN_STAGE_NEXT(1) DFN_STAGE_NEXT(2) DFN_STAGE_LAST(3) DFN_STAGE_TAIL
                       DFN_STAGE_OPERANDS);
  }
}

struct SlotCtx {
  uint32_t ph0, ph1, ph2, ph3, ph4, ph5;   // phases of this slot's full[stage] barriers (scalars: must stay in registers)
  uint32_t pa;              // phase of this slot's aready barrier
  bool has;                 // the slot has a tile in the current iteration
};
struct IssueCtx {
  uint32_t sbase, bar0, tmem_base, rank;
  SlotCtx s0, s1;
  int lane;
#ifdef DFN_FWD_STATS
  long long w_full, w_aready;   // bring-up build (-DDFN_FWD_STATS): cycles the issuer waits for weights / for the epilogues
#endif
};
template <int T>
__device__ __forceinline__ SlotCtx& slot_of(IssueCtx& c) {
  if constexpr (T == 0) return c.s0; else return c.s1;
}
template <int S>
__device__ __forceinline__ uint32_t& ph_of(SlotCtx& c) {
  static_assert(S >= 0 && S < 6, "at most 6 ring stages");
  if constexpr (S == 0) return c.ph0;
  else if constexpr (S == 1) return c.ph1;
  else if constexpr (S == 2) return c.ph2;
  else if constexpr (S == 3) return c.ph3;
  else if constexpr (S == 4) return c.ph4;
  else return c.ph5;
}

template <class PP, int J, int T, int K>
__device__ __forceinline__ void issue_stage(IssueCtx& c, const typename PP::ABases& ab, uint32_t tacc) {
  constexpr int N = PP::job_N(J);
  constexpr uint32_t idesc = instr_desc_bf16(256, N, 0, 0);   // M = 256 over the CTA pair
  constexpr uint32_t cb = PP::job_half_chunk(J);              // bytes per K chunk of this CTA's half of B
  constexpr uint32_t hi = desc_hi(128);
  constexpr int B_STEP = 2 * (cb >> 4);                       // descriptor address units per MMA (two K chunks)
  constexpr int kc0 = K * PP::kCPS;
  constexpr int stage = (stage_prefix<PP>(J, T) + K) % PP::kStages;
  constexpr int nchunk = stage_chunks<PP>(J, K);
  constexpr bool last = K == job_stages<PP>(J) - 1;
  static_assert(nchunk % 2 == 0, "a K=16 MMA consumes two chunks");
  const uint32_t full = c.bar0 + 8u * PP::bar_full_i(T, stage), empty = c.bar0 + 8u * PP::bar_empty_i(T, stage);
  uint32_t& ph = ph_of<stage>(slot_of<T>(c));
  if (c.rank == 0) {
#ifdef DFN_FWD_STATS
    { const long long t0 = clock64(); mbar_wait_cluster(full, ph); c.w_full += clock64() - t0; }
#endif
    const uint32_t accfull = c.bar0 + 8u * PP::bar_accfull_i(T);
    const uint32_t b_lo = desc_lo(c.sbase + PP::kOffRing + stage * PP::kRingStageBytes, cb);
    const uint32_t a_lo = PP::template a_chunk_lo<J>(kc0, ab);
    const uint32_t a_last = PP::template a_chunk_lo<J>(kc0 + nchunk - 2, ab);
    constexpr uint32_t accf = kc0 > 0 ? 1u : 0u;
    stage_mma<nchunk / 2, B_STEP, last>(full, ph, tacc, a_lo, a_last, b_lo, hi, idesc, accf, empty, accfull);
  }
  ph ^= 1;
}
template <class PP, int J, int T, int... Ks>
__device__ __forceinline__ void issue_stages(IssueCtx& c, const typename PP::ABases& ab, uint32_t tacc,
                                             std::integer_sequence<int, Ks...>) {
  (issue_stage<PP, J, T, Ks>(c, ab, tacc), ...);
}

// Runs on the whole warp 1 (converged) of both CTAs: issuer in the leader, relay in the peer.
template <class PP, int J, int T>
__device__ __forceinline__ void issue_job(IssueCtx& c) {
  SlotCtx& sc = slot_of<T>(c);
  if (!sc.has) return;
  const uint32_t aready = c.bar0 + 8u * PP::bar_aready_i(T);
  if (c.rank == 0) {   // the A operand of this job has been published: 8 local epilogue warps + the peer's relay
#ifdef DFN_FWD_STATS
    const long long t0 = clock64();
#endif
    mbar_wait_cluster(aready, sc.pa);
#ifdef DFN_FWD_STATS
    c.w_aready += clock64() - t0;
#endif
    tc_fence_after();
  } else {
    // peer: relay "my 8 epilogue warps have published their tile" to the leader.  The epilogue warps arrive on their
    // OWN CTA's barrier (cheap CTA-scope release); the cluster-scope release is paid once, by a thread with no memory
    // traffic of its own - issued from the epilogue warps it made every publish wait for their outstanding global stores.
    if (c.lane == 0) {
      mbar_wait(aready, sc.pa);
      mbar_arrive_cluster(map_to_cta(aready, 0));
    }
    __syncwarp();
  }
  sc.pa ^= 1;
  const typename PP::ABases ab = PP::bases(c.sbase, T);
  const uint32_t tacc = c.tmem_base + T * 256;
  issue_stages<PP, J, T>(c, ab, tacc, std::make_integer_sequence<int, job_stages<PP>(J)>{});
}
template <class PP, int... Js>
__device__ __forceinline__ void issue_all(IssueCtx& c, std::integer_sequence<int, Js...>) {
  ((issue_job<PP, Js, 0>(c), issue_job<PP, Js, 1>(c)), ...);
}

// All pairs stream the same weight blocks in the same order: with ONE copy of the image every pair asks the same L2 slices
// for the same lines at about the same time.  The packed images are therefore kept in R identical replicas (different
// addresses hash to different slices) and pair p streams replica p % R.  DFN_WEIGHT_REPLICAS overrides R.
static inline int weight_replicas() {
  static const int r = [] { const char* e = getenv("DFN_WEIGHT_REPLICAS"); int v = e ? atoi(e) : 1; return v < 1 ? 1 : (v > 16 ? 16 : v); }();
  return r;
}

// tensor maps over the packed weight image (rebuilt per call: a few hundred ns of host time, no device work)
static inline int make_weight_maps(const void* image, unsigned long long image_bytes, WeightMaps* out) {
  static PFN_cuTensorMapEncodeTiled encode = nullptr;
  if (!encode) {
    cudaDriverEntryPointQueryResult q;
    void* fn = nullptr;
    DFN_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q), "mlp(driver entry point)");
    DFN_CHECK_ARG(fn && q == cudaDriverEntryPointSuccess, "mlp: cuTensorMapEncodeTiled not available");
    encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled>(fn);
  }
  for (int i = 0; i < 7; ++i) {
    const cuuint64_t gdim[2] = {128, image_bytes / 128};
    const cuuint64_t gstride[1] = {128};
    const cuuint32_t box[2] = {128, kMapRows[i]};
    const cuuint32_t estride[2] = {1, 1};
    const CUresult r = encode(&out->m[i], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(image), gdim, gstride, box, estride,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DFN_CHECK_ARG(r == CUDA_SUCCESS, "mlp: cuTensorMapEncodeTiled failed (%d)", static_cast<int>(r));
  }
  return DFN_OK;
}


}  // namespace mlp_pipe
