// Fused vanilla-NeRF MLP forward (reference scripts/model.py:139-187) for sm_100a.
//
// Persistent kernel, one CTA per SM, CTAs paired into clusters of two (tcgen05 cta_group::2): a pair works on
// 256-point super-tiles, CTA r owning rows [128r, 128r+128), and two super-tiles (slots X/Y) are in flight per pair.
//   warp 0      : weight producer (both CTAs) - streams ITS HALF (N/2 rows) of the packed bf16 weight blocks from L2
//                 into a shared memory ring with tensor-map TMA loads (cp.async.bulk.tensor ... .cta_group::2) whose
//                 complete_tx goes to the LEADER's full barrier, whichever CTA issued the load.  The pair
//                 shares the B operand, so each SM pulls only half of the 1.2 MB of weights per super-tile: the weight
//                 stream (L2 -> SM latency x ring depth) is what bounds this kernel, see DESIGN.md section 5.
//   warp 1      : leader CTA: MMA issuer - the warp runs converged and one elected lane issues
//                 tcgen05.mma.cta_group::2 (M=256, N=256/128, K=16, bf16 -> fp32 in TMEM); slot X accumulates in TMEM
//                 columns [0,256) of both CTAs, slot Y in [256,512).  One ring stage (wait, MMAs, commit) is a single
//                 asm block with the election inside.  Commits are multicast to the barriers of both CTAs.
//                 peer CTA: idle.
//   warps 2..9  : epilogue (both CTAs) - tcgen05.ld the accumulator, ReLU, round to bf16, write the next layer's A
//                 operand back to shared memory in the UMMA chunk layout (activations never leave the SM);
//                 positional encodings are computed here too.  Publishing a tile = one arrive per warp on the
//                 LEADER's aready barrier (remote arrive from the peer).
// While the tensor cores run layer L of slot X the epilogue warps drain layer L-1 of slot Y.
// Biases ride along as one extra K=16 MMA per layer against a constant "ones" tile (mlp_common.cuh).
// The sigma head (N=1) and color_layer.2 (N=3) are CUDA-core dot products inside the epilogues.
// TRAIN=true additionally stashes h1..h8, features and the colour hidden layer (bf16, chunk layout) for the
// backward kernels: each epilogue thread stores the 16-byte groups it has rounded straight to global memory
// (a warp covers 512 contiguous bytes per chunk), after the tile has been published to the issuer.
#include "mlp_pipe.cuh"
#include "rng.cuh"
#include <stdlib.h>
#include <string.h>
#include <utility>

namespace {
using namespace mlp;
using namespace umma;
using namespace mlp_pipe;

// Ring: 3 stages x 16 KB = 8 K-chunks (4 MMAs) of this CTA's half of B per stage.  6 x 8 KB (2 MMAs per stage, finer
// release) was measured 35 % slower: the issuer's per-stage cost (try_wait, fence, election, commits) is ~2 MMA times.
constexpr int kStages = 3;
constexpr int kFwdCPS = 8;
constexpr uint32_t kFwdStageBytes = 16384;
constexpr int kEpiWarps = 8;
constexpr int kEpiThreads = kEpiWarps * 32;
#ifndef DFN_FWD_STORE_WARP
#define DFN_FWD_STORE_WARP 1
#endif
#ifndef DFN_STORE_PIECES
#define DFN_STORE_PIECES 8
#endif
// > 0: pause that long between the pieces of a stash copy; < 0: wait until the previous piece has been read out of shared
// memory instead.  Round 3 sweep (profiles/r3_store_pacing_sweep.txt): a fixed 75-250 ns pause beats the read wait by 3-4 % on
// the fine forward (1.243 -> 1.188 ms at 200 ns); 400 ns and more make the store the bottleneck (1.64 ms at 400, 2.79 at 800).
#ifndef DFN_STORE_GAP_NS
#define DFN_STORE_GAP_NS 200
#endif
constexpr bool kUseStoreWarp = DFN_FWD_STORE_WARP != 0;
constexpr int kStoreWarp = 2 + kEpiWarps;     // TRAIN: ships published activation tiles to the stash (bulk shared -> global)
constexpr int kThreads = 64 + kEpiThreads + (kUseStoreWarp ? 32 : 0);   // producer, issuer / relay, 8 epilogue warps, stash store warp

// shared memory map (bytes), identical in both CTAs of a pair (the MMA addresses the peer's tiles by the same offsets)
constexpr uint32_t kOffAct = 0;                                  // 2 x 64 KB
constexpr uint32_t kOffAux = kOffAct + 2 * kActBytes;            // 2 x 16 KB
constexpr uint32_t kOffRing = kOffAux + 2 * kAuxBytes;           // kStages x 16 KB
constexpr uint32_t kOffOnes = kOffRing + kStages * kFwdStageBytes;  // 128 x 16 bf16, column 0 = 1 (bias operand)
constexpr uint32_t kOnesBytes = 2 * kChunkBytes;
constexpr uint32_t kOffConst = kOffOnes + kOnesBytes;
// partial head dot products of the upper column half, [2 slots][128 rows][3] f32 (sigma uses [..][0];
// reuse across j7 / j9 of a slot is ordered by the other slot's named barrier in between)
constexpr uint32_t kOffPart = kOffConst + kConstBytes;
constexpr uint32_t kOffBars = kOffPart + 2 * 128 * 3 * 4;
// barriers: full[slot][stage], empty[slot][stage] (one set per slot: a parity wait can only tell adjacent phases
// apart), accfull[slot], aready[slot] (leader only)
constexpr uint32_t kNumBars = 4 * kStages + 6;
__host__ __device__ constexpr uint32_t bar_full_i(int t, int s) { return t * kStages + s; }
__host__ __device__ constexpr uint32_t bar_empty_i(int t, int s) { return 2 * kStages + t * kStages + s; }
__host__ __device__ constexpr uint32_t bar_accfull_i(int t) { return 4 * kStages + t; }
__host__ __device__ constexpr uint32_t bar_aready_i(int t) { return 4 * kStages + 2 + t; }
__host__ __device__ constexpr uint32_t bar_stored_i(int t) { return 4 * kStages + 4 + t; }   // act tile of slot t may be overwritten
constexpr uint32_t kOffTmemPtr = kOffBars + kNumBars * 8;
constexpr uint32_t kSmemBytes = kOffTmemPtr + 16;
static_assert(kSmemBytes <= 232448, "shared memory budget exceeded");

// pipeline policy (mlp_pipe.cuh): shared-memory map, job table, A-operand addressing of the forward chain
struct FwdPipe {
  static constexpr int kNumJobs = mlp::kNumJobs;
  static constexpr int kStages = ::kStages;
  static constexpr int kCPS = kFwdCPS;
  static constexpr uint32_t kRingStageBytes = kFwdStageBytes;
  static constexpr uint32_t kOffRing = ::kOffRing;
  static __host__ __device__ constexpr uint32_t bar_full_i(int t, int s) { return ::bar_full_i(t, s); }
  static __host__ __device__ constexpr uint32_t bar_empty_i(int t, int s) { return ::bar_empty_i(t, s); }
  static __host__ __device__ constexpr uint32_t bar_accfull_i(int t) { return ::bar_accfull_i(t); }
  static __host__ __device__ constexpr uint32_t bar_aready_i(int t) { return ::bar_aready_i(t); }
  static __host__ __device__ constexpr int job_N(int j) { return mlp::job_N(j); }
  static __host__ __device__ constexpr int job_chunks(int j) { return mlp::job_kcb(j); }
  static __host__ __device__ constexpr uint32_t job_off(int j) { return mlp::job_off(j); }
  static __host__ __device__ constexpr uint32_t job_half_chunk(int j) { return mlp::job_half_chunk(j); }
  static __host__ __device__ constexpr uint32_t job_half_bytes(int j) { return mlp::job_half_bytes(j); }
  struct ABases { uint32_t act, aux, ones; };   // descriptor "lo" words of the slot's tiles
  static __device__ __forceinline__ ABases bases(uint32_t sbase, int T) {
    return ABases{desc_lo(sbase + kOffAct + T * kActBytes, kChunkBytes), desc_lo(sbase + kOffAux + T * kAuxBytes, kChunkBytes),
                  desc_lo(sbase + kOffOnes, kChunkBytes)};
  }
  // where chunk kc of job J's A operand lives: hidden activations | positional encoding | ones tile (bias slice)
  template <int J>
  static __device__ __forceinline__ uint32_t a_chunk_lo(int kc, const ABases& b) {
    constexpr int KC = job_kc(J), KC_ACT = job_kc_act(J);
    if (kc >= KC) return b.ones + (kc - KC) * (kChunkBytes >> 4);
    if (kc < KC_ACT) return b.act + kc * (kChunkBytes >> 4);
    return b.aux + (kc - KC_ACT) * (kChunkBytes >> 4);
  }
};

struct FwdArgs {
  PointSource src;
  long long P;
  int n_tiles;            // 128-point tiles, even (num_tiles(P))
  const uint8_t* wpack;   // packed forward weights
  const float* consts;    // fp32 consts block
  const float* noise;     // [P] sigma noise or null (model.py:160-162)
  const int* rng_step;    // non-null (and noise null): the noise is drawn in the kernel, N(0,1) = f(rng_seed, *rng_step, rng_stream, point)
  unsigned long long rng_seed;
  unsigned rng_stream;
  float* rgb;             // [P,3]
  float* sigma;           // [P]
  uint8_t* stash_h;       // TRAIN: [8][n_tiles][64 KB]
  uint8_t* stash_f;       // TRAIN: [n_tiles][64 KB]
  uint8_t* stash_c;       // TRAIN: [n_tiles][32 KB]
  uint8_t* stash_m;       // TRAIN: [8][n_tiles][4 KB] ReLU bit masks of h1..h8 (mlp_common.cuh)
  int skew_ns;            // start of pair p is delayed by (p % 8) * skew_ns (see dfn_mlp_forward)
  int replicas;           // copies of the packed weight image; pair p streams copy p % replicas
};

// ---- epilogue helpers ------------------------------------------------------------------------------
// the first NC packed 32-column groups of one row -> stash (global); a warp covers 512 contiguous bytes per chunk
template <int NC>
__device__ __forceinline__ void stash_row(uint8_t* stash, int col0, int row, const uint32_t (&keep)[64], uint64_t pol) {
#pragma unroll
  for (int g = 0; g < NC * 4; ++g) {
    const uint32_t off = static_cast<uint32_t>(col0 / 8 + g) * kChunkBytes + row * 16;
    st_global_v4_hint(stash + off, keep[g * 4], keep[g * 4 + 1], keep[g * 4 + 2], keep[g * 4 + 3], pol);
  }
}

// 32 accumulator columns -> optional ReLU, bf16 -> 4 x 16 B into the act tile
template <bool RELU>
__device__ __forceinline__ void emit32(const uint32_t (&v)[32], uint8_t* act, int col0, int row, uint32_t* packed /* 16 */) {
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    const float x0 = __uint_as_float(v[g * 8 + 0]), x1 = __uint_as_float(v[g * 8 + 1]);
    const float x2 = __uint_as_float(v[g * 8 + 2]), x3 = __uint_as_float(v[g * 8 + 3]);
    const float x4 = __uint_as_float(v[g * 8 + 4]), x5 = __uint_as_float(v[g * 8 + 5]);
    const float x6 = __uint_as_float(v[g * 8 + 6]), x7 = __uint_as_float(v[g * 8 + 7]);
    uint4 q;
    if (RELU) {
      q.x = pack_bf16x2_relu(x0, x1); q.y = pack_bf16x2_relu(x2, x3);
      q.z = pack_bf16x2_relu(x4, x5); q.w = pack_bf16x2_relu(x6, x7);
    } else {
      q.x = pack_bf16x2(x0, x1); q.y = pack_bf16x2(x2, x3);
      q.z = pack_bf16x2(x4, x5); q.w = pack_bf16x2(x6, x7);
    }
    packed[g * 4 + 0] = q.x; packed[g * 4 + 1] = q.y; packed[g * 4 + 2] = q.z; packed[g * 4 + 3] = q.w;
    const uint32_t off = static_cast<uint32_t>(col0 / 8 + g) * kChunkBytes + row * 16;
    if (act) *reinterpret_cast<uint4*>(act + off) = q;
  }
}

template <bool TRAIN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1) mlp_fwd_kernel(const FwdArgs a, const __grid_constant__ WeightMaps maps) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bar0 = sbase + kOffBars;
  auto bar_accfull = [&](int t) { return bar0 + 8u * bar_accfull_i(t); };
  auto bar_aready = [&](int t) { return bar0 + 8u * bar_aready_i(t); };
  auto bar_stored = [&](int t) { return bar0 + 8u * bar_stored_i(t); };
  float* consts = reinterpret_cast<float*>(smem + kOffConst);

  if (threadIdx.x == 0) {
    for (int t = 0; t < 2; ++t)
      for (int s = 0; s < kStages; ++s) {
        // full: used in the leader only (one arrive.expect_tx for the bytes of both CTAs' halves)
        mbar_init(bar0 + 8u * bar_full_i(t, s), 1);
        mbar_init(bar0 + 8u * bar_empty_i(t, s), 1);
      }
    for (int t = 0; t < 2; ++t) { mbar_init(bar_accfull(t), 1); mbar_init(bar_aready(t), kEpiWarps + (rank == 0 ? 1 : 0)); mbar_init(bar_stored(t), 1); }   // aready: + the peer's relay
    fence_mbar_init();
  }
  if (warp == 1) {   // one warp per CTA allocates all 512 TMEM columns of the pair
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sbase + kOffTmemPtr), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  for (int i = threadIdx.x; i < kConstFloats; i += kThreads) consts[i] = a.consts[i];
  // ones tile: [2 chunks][128 rows][8 bf16], column 0 of every row = 1.0
  for (int i = threadIdx.x; i < static_cast<int>(kOnesBytes / 16); i += kThreads)
    *reinterpret_cast<uint4*>(smem + kOffOnes + i * 16) = make_uint4(i < 128 ? 0x3F80u : 0u, 0u, 0u, 0u);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // barriers of both CTAs are initialised before any remote arrive / multicast commit
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem + kOffTmemPtr);

  const int n_pairs = gridDim.x >> 1;
  const int n_super = a.n_tiles >> 1;
  const int n_it = (n_super + 2 * n_pairs - 1) / (2 * n_pairs);

  if (warp == 0) {
    // ================= weight producer (this CTA's half of every block) =================
    if (lane == 0) {
      ProdState<FwdPipe> ps;
#pragma unroll
      for (int i = 0; i < kStages; ++i) { ps.use[0][i] = ps.use[1][i] = 0; ps.last[i] = -1; }
      ps.pol = make_evict_last_policy();
      ps.row0 = static_cast<uint32_t>((blockIdx.x >> 1) % a.replicas) * (kFwdPackBytes >> 7);
      // de-synchronise the pairs: all of them stream the same weight blocks in the same order, and in lockstep they all
      // hit the same L2 lines at the same moment
      for (int w = (blockIdx.x >> 1) & 7; a.skew_ns > 0 && w > 0; --w) __nanosleep(a.skew_ns);
      for (int it = 0; it < n_it; ++it) {
        const bool has[2] = {tile_of(it, 0, rank) < a.n_tiles, tile_of(it, 1, rank) < a.n_tiles};
        produce_all<FwdPipe>(maps, rank, sbase + kOffRing, bar0, ps, has, std::make_integer_sequence<int, kNumJobs>{});
      }
    }
  } else if (warp == 1) {
    // ================= leader: MMA issuer (whole warp, converged; one elected lane issues); peer: bookkeeping only =================
    IssueCtx c{};
    c.sbase = sbase; c.bar0 = bar0; c.tmem_base = tmem_base; c.lane = lane; c.rank = rank;
    c.s0 = SlotCtx{};
    c.s1 = c.s0;
#ifdef DFN_FWD_STATS
    c.w_full = c.w_aready = 0;
    const long long t_start = clock64();
#endif
    for (int it = 0; it < n_it; ++it) {
      c.s0.has = tile_of(it, 0, rank) < a.n_tiles;
      c.s1.has = tile_of(it, 1, rank) < a.n_tiles;
      issue_all<FwdPipe>(c, std::make_integer_sequence<int, kNumJobs>{});
    }
#ifdef DFN_FWD_STATS
    if (blockIdx.x == 0 && lane == 0)
      printf("fwd stats (leader of pair 0, cycles): issuer total %lld wait_aready %lld wait_weights %lld\n", clock64() - t_start,
             c.w_aready, c.w_full);
#endif
  } else if (kUseStoreWarp && warp == kStoreWarp) {
    // ================= TRAIN: stash store (one thread) =================
    // The activation tile a job's epilogue publishes (h1..h8, features) is exactly the stash tile (same chunk layout), so
    // it leaves the SM as bulk copies while the tensor cores read it; the epilogue warps issue no global stores for it.
    // The copy goes out in 8 KB pieces a fixed pause apart (DFN_STORE_GAP_NS): one 64 KB copy holds up the weight loads
    // queued behind it in the TMA unit (measured: fine pass 1.18 ms vs 1.07 ms in pieces).
    // `stored` tells the epilogues the tile may be overwritten; one arrive per (job, slot) keeps the aready phases
    // this thread has to observe in lockstep with them.
    if (TRAIN && lane == 0) {
      uint32_t pa[2] = {0, 0};
      const uint64_t pol = make_evict_first_policy();
      for (int it = 0; it < n_it; ++it)
        for (int j = 0; j < kNumJobs; ++j)
          for (int t = 0; t < 2; ++t) {
            const int tile = tile_of(it, t, rank);
            if (tile >= a.n_tiles) continue;
            mbar_wait(bar_aready(t), pa[t]);   // the A operand of job j: encodings (j = 0) or the output of job j-1
            pa[t] ^= 1;
            if (j >= 1) {
              uint8_t* dst = j <= 8 ? a.stash_h + ((static_cast<size_t>(j - 1) * a.n_tiles + tile) << 16)
                                    : a.stash_f + (static_cast<size_t>(tile) << 16);
              constexpr uint32_t kPiece = kActBytes / DFN_STORE_PIECES;
#pragma unroll 1
              for (int i = 0; i < DFN_STORE_PIECES; ++i) {
                bulk_s2g_hint(dst + i * kPiece, sbase + kOffAct + t * kActBytes + i * kPiece, kPiece, pol);
                bulk_commit();
                if (DFN_STORE_GAP_NS > 0 && i + 1 < DFN_STORE_PIECES) __nanosleep(DFN_STORE_GAP_NS);
                if (DFN_STORE_GAP_NS < 0 && i + 1 < DFN_STORE_PIECES) bulk_wait_read_all();
              }
              bulk_wait_read_all();
            }
            mbar_arrive(bar_stored(t));
          }
      bulk_wait_all();
    }
  } else {
    // ================= epilogue / PE warps =================
    const int ew = warp - 2;
    const int q = warp & 3;         // TMEM lane quadrant this warp may access
    const int hh = ew >> 2;         // column half
    const int row = q * 32 + lane;  // row of the tile == TMEM lane
    const uint32_t lane_sel = static_cast<uint32_t>(q * 32) << 16;
    float* part = reinterpret_cast<float*>(smem + kOffPart);
    uint32_t pf[2] = {0, 0}, pst[2] = {0, 0};
    const uint64_t pol = make_evict_first_policy();

    auto point_of = [&](int tile) {
      long long p = static_cast<long long>(tile) * kTile + row;
      return p < a.P ? p : a.P - 1;
    };
    // activations / encodings written -> visible to the async proxy (the leader's MMA reads both CTAs' tiles) -> one
    // arrive per warp on the leader's barrier
    auto publish = [&](int t) {
      tc_fence_before();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_aready(t));   // own CTA's barrier; the peer's warp 1 relays it to the leader
    };
    // initial PE for both slots
    for (int t = 0; t < 2; ++t) {
      const int tile = tile_of(0, t, rank);
      if (tile < a.n_tiles) {
        float x[3], d[3];
        load_point(a.src, point_of(tile), x, d);
        write_pe_row(smem + kOffAux + t * kAuxBytes, nullptr, row, hh, x, d);
      }
      publish(t);
    }

    for (int it = 0; it < n_it; ++it) {
      for (int j = 0; j < kNumJobs; ++j) {
        for (int t = 0; t < 2; ++t) {
          const int tile = tile_of(it, t, rank);
          if (tile >= a.n_tiles) continue;
          uint8_t* act = smem + kOffAct + t * kActBytes;
          uint8_t* aux = smem + kOffAux + t * kAuxBytes;
          const uint32_t tacc = tmem_base + t * 256 + lane_sel;
          const long long p = static_cast<long long>(tile) * kTile + row;
          // TRAIN: the bf16 tile this job produces (h1..h8, features, colour hidden) also goes to the stash
          uint8_t* st = nullptr;
          if (TRAIN) {
            if (j < 8) st = a.stash_h + ((static_cast<size_t>(j) * a.n_tiles + tile) << 16);
            else if (j == 8) st = a.stash_f + (static_cast<size_t>(tile) << 16);
            else st = a.stash_c + (static_cast<size_t>(tile) << 15);
          }
          // global inputs of the encodings this epilogue writes (view direction at job 4, the slot's next tile at job 9)
          // are fetched before waiting for the accumulator: their latency would otherwise sit between two jobs
          float xe[3] = {0.f, 0.f, 0.f}, de[3] = {0.f, 0.f, 0.f};
          const int ntile = tile_of(it + 1, t, rank);
          if (j == 4) load_point(a.src, p < a.P ? p : a.P - 1, xe, de);
          else if (j == kNumJobs - 1 && ntile < a.n_tiles) load_point(a.src, point_of(ntile), xe, de);
          mbar_wait(bar_accfull(t), pf[t]);
          pf[t] ^= 1;
          tc_fence_after();
          if (TRAIN && kUseStoreWarp) { mbar_wait(bar_stored(t), pst[t]); pst[t] ^= 1; }   // the store warp has read the previous tile out of act

          if (j < 9) {
            const int c0 = hh * 128;
            uint32_t v0[32], v1[32], keep[64];
            tmem_ld32(tacc + c0, v0);
            tmem_ld_wait();
#pragma unroll
            for (int cc = 0; cc < 4; ++cc) {
              uint32_t (&cur)[32] = (cc & 1) ? v1 : v0;
              uint32_t (&nxt)[32] = (cc & 1) ? v0 : v1;
              if (cc < 3) tmem_ld32(tacc + c0 + (cc + 1) * 32, nxt);
              const int col0 = c0 + cc * 32;
              uint32_t* pk = keep + cc * 16;
              if (j == 8) emit32<false>(cur, act, col0, row, pk);
              else emit32<true>(cur, act, col0, row, pk);
              if (cc < 3) tmem_ld_wait();
            }
            publish(t);
            // pos-PE is dead once job 4's MMAs finished: overlay the dir-PE (job 9 reads it; the publish of job 5's
            // epilogue orders it)
            if (j == 4) write_pe_row(nullptr, aux, row, hh, xe, de);
            if (TRAIN) {
              if (!kUseStoreWarp) stash_row<4>(st, c0, row, keep, pol);
              if (j < 8)   // ReLU bit mask for the data-gradient kernel: 16 B per thread instead of re-reading the tile
                st_global_v4_hint(a.stash_m + ((static_cast<size_t>(j) * a.n_tiles + tile) << 12) + (hh * 128 + row) * 16, relu_bits16(keep),
                                  relu_bits16(keep + 16), relu_bits16(keep + 32), relu_bits16(keep + 48), pol);
            }
            if (j == 7) {  // sigma head: relu(round_bf16(w.h8 + b) (+ noise)), after the publish: the 256 FMAs per thread
                           // would otherwise make this the longest epilogue of the chain.  Inference re-reads its own h8
                           // values from the act tile (nobody overwrites them before this thread's next epilogue).
              float sdot = 0.f;
#pragma unroll
              for (int cc = 0; cc < 4; ++cc) {
                const int col0 = c0 + cc * 32;
                uint32_t tmp[16];
                const uint32_t* pk = keep + cc * 16;
                if (!TRAIN) {
#pragma unroll
                  for (int g = 0; g < 4; ++g) {
                    const uint4 q = *reinterpret_cast<const uint4*>(act + static_cast<uint32_t>(col0 / 8 + g) * kChunkBytes + row * 16);
                    tmp[g * 4] = q.x; tmp[g * 4 + 1] = q.y; tmp[g * 4 + 2] = q.z; tmp[g * 4 + 3] = q.w;
                  }
                  pk = tmp;
                }
                const float* ws = consts + kConstWsig + col0;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                  sdot += bf16lo_to_f32(pk[i]) * ws[2 * i];
                  sdot += bf16hi_to_f32(pk[i]) * ws[2 * i + 1];
                }
              }
              if (hh == 1) part[(t * 128 + row) * 3] = sdot;
              named_bar_sync(1, kEpiThreads);
              if (hh == 0 && p < a.P) {
                float s = sdot + part[(t * 128 + row) * 3] + consts[kConstBsig];
                s = round_bf16(s);
                if (a.noise) s = round_bf16(s + round_bf16(a.noise[p]));
                else if (a.rng_step) s = round_bf16(s + round_bf16(rng::normal1(a.rng_seed, static_cast<uint32_t>(*a.rng_step), a.rng_stream, static_cast<unsigned long long>(p))));
                a.sigma[p] = fmaxf(s, 0.f);
              }
            }
          } else {
            // j == 9: colour hidden layer (N=128) + color_layer.2 + sigmoid
            float r0 = 0.f, r1 = 0.f, r2 = 0.f;
            const int c0 = hh * 64;
            uint32_t v0[32], v1[32], keep[64];
            tmem_ld32(tacc + c0, v0);
            tmem_ld32(tacc + c0 + 32, v1);
            tmem_ld_wait();
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
              uint32_t (&cur)[32] = cc ? v1 : v0;
              const int col0 = c0 + cc * 32;
              uint32_t* pk = keep + cc * 16;
              emit32<true>(cur, nullptr, col0, row, pk);
            }
            // next tile of this slot first: its pos-PE goes into aux (dir-PE was consumed by job 9's MMAs) and is published
            // before the colour head is evaluated, which nobody on the chip waits for
            if (ntile < a.n_tiles) {
              write_pe_row(aux, nullptr, row, hh, xe, de);
              publish(t);   // (no publish after the slot's last tile: nobody waits for it)
            }
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
              const uint32_t* pk = keep + cc * 16;
              const float* w0 = consts + kConstWc2 + c0 + cc * 32;
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const float lo = bf16lo_to_f32(pk[i]), hi = bf16hi_to_f32(pk[i]);
                r0 += lo * w0[2 * i] + hi * w0[2 * i + 1];
                r1 += lo * w0[128 + 2 * i] + hi * w0[128 + 2 * i + 1];
                r2 += lo * w0[256 + 2 * i] + hi * w0[256 + 2 * i + 1];
              }
            }
            if (TRAIN) stash_row<2>(st, c0, row, keep, pol);
            float* pr = part + (t * 128 + row) * 3;
            if (hh == 1) { pr[0] = r0; pr[1] = r1; pr[2] = r2; }
            named_bar_sync(1, kEpiThreads);
            if (hh == 0 && p < a.P) {
              const float own[3] = {r0, r1, r2};
#pragma unroll
              for (int o = 0; o < 3; ++o) {
                const float xo = round_bf16(own[o] + pr[o] + consts[kConstBc2 + o]);
                a.rgb[p * 3 + o] = 1.0f / (1.0f + __expf(-xo));
              }
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // the leader's MMAs read the peer's shared memory: nobody leaves before both are done
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
}

}  // namespace
// ---------------------------------------------------------------------------------------------------
// weight packing: flat fp32 parameters -> bf16 operand images + fp32 consts
// ---------------------------------------------------------------------------------------------------
namespace {

__device__ __forceinline__ float param_w(const float* params, int layer, int o, int i) {
  constexpr ParamOffsets po = param_offsets();
  return params[po.w[layer] + o * layer_in(layer) + i];
}

// Source of every element of the packed images: where in the flat fp32 parameter vector it comes from.  The layout only
// depends on the architecture, so the index arithmetic (a division chain per element: 33 us per network) runs ONCE into an
// int32 map; every later repack (after each optimiser step) is a plain gather.
//   map[idx] >= 0 : params[map[idx]]      map[idx] == -1 : zero
// idx runs over the forward image (bf16 elements), then the backward image; the fp32 consts block follows with the same coding.
__device__ __forceinline__ int pack_source(int idx) {
  constexpr ParamOffsets po = param_offsets();
  const int nf = kFwdPackBytes / 2, nb = kBwdPackBytes / 2;
  if (idx < nf) {
    int j = 0;
    while (j + 1 < kNumJobs && static_cast<uint32_t>(idx) * 2u >= job_off(j + 1)) ++j;
    const int e0 = idx - job_off(j) / 2;
    const int N = job_N(j), NH = N / 2;
    const int half_elems = job_half_bytes(j) / 2;
    const int r = e0 / half_elems, e = e0 % half_elems;   // r: which CTA of the pair holds this half
    const int kc = e / (NH * 8), n = r * NH + (e / 8) % NH, k = kc * 8 + (e & 7);
    const int layer = j < 8 ? j : (j == 8 ? 9 : 10);   // reference layer index of job j
    const int kdata = job_kc(j) * 8;
    if (k >= kdata) return k == kdata ? po.b[layer] + n : -1;   // bias slice: column 0 carries the bias, the other 15 are zero
    const int kin = layer_in(layer);
    return k < kin ? po.w[layer] + n * kin + k : -1;            // K padding (63 -> 64, 319 -> 320, 283 -> 288) is zero
  }
  if (idx < nf + nb) {
    const int bi = idx - nf;
    int b = 0;
    while (b + 1 < kNumBwdJobs && static_cast<uint32_t>(bi) * 2u >= bjob_off(b + 1)) ++b;
    const int e0 = bi - bjob_off(b) / 2;
    const int half_elems = bjob_bytes(b) / 4;
    const int r = e0 / half_elems, e = e0 % half_elems;   // r: which CTA of the pair holds this half of W^T's N rows
    const int kc = e / (128 * 8), i = r * 128 + (e / 8) % 128, o = kc * 8 + (e & 7);
    const int layer_of[kNumBwdJobs] = {10, 9, 7, 6, 5, 4, 3, 2, 1};
    const int layer = layer_of[b];
    return po.w[layer] + o * layer_in(layer) + i;
  }
  const int c = idx - nf - nb;
  if (c < kConstWc2) return po.w[8] + (c - kConstWsig);
  if (c < kConstBsig) return po.w[11] + (c - kConstWc2);
  if (c == kConstBsig) return po.b[8];
  if (c < kConstBc2 + 3) return po.b[11] + (c - kConstBc2);
  return -1;
}

constexpr int kPackElems = kFwdPackBytes / 2 + kBwdPackBytes / 2 + kConstFloats;

__global__ void pack_map_kernel(int* __restrict__ map) {
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < kPackElems; idx += gridDim.x * blockDim.x) map[idx] = pack_source(idx);
}

// map == nullptr: the sources are computed on the fly (dfn_mlp_pack_weights, the self-contained entry point)
__global__ void pack_weights_kernel(const float* __restrict__ params, const int* __restrict__ map, __nv_bfloat16* __restrict__ fwd,
                                    __nv_bfloat16* __restrict__ bwd, float* __restrict__ consts, int replicas) {
  const int nf = kFwdPackBytes / 2, nb = kBwdPackBytes / 2;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < kPackElems; idx += gridDim.x * blockDim.x) {
    const int src = map ? map[idx] : pack_source(idx);
    const float v = src >= 0 ? params[src] : 0.f;
    if (idx < nf) {
      const __nv_bfloat16 w = __float2bfloat16_rn(v);
      for (int r = 0; r < replicas; ++r) fwd[idx + static_cast<size_t>(r) * nf] = w;
    } else if (idx < nf + nb) {
      const __nv_bfloat16 w = __float2bfloat16_rn(v);
      for (int r = 0; r < replicas; ++r) bwd[idx - nf + static_cast<size_t>(r) * nb] = w;
    } else {
      consts[idx - nf - nb] = round_bf16(v);
    }
  }
}

}  // namespace

extern "C" long long dfn_mlp_fwd_pack_bytes(void) { return static_cast<long long>(kFwdPackBytes) * weight_replicas(); }
extern "C" long long dfn_mlp_bwd_pack_bytes(void) { return static_cast<long long>(kBwdPackBytes) * weight_replicas(); }
extern "C" long long dfn_mlp_consts_bytes(void) { return kConstBytes; }
extern "C" long long dfn_mlp_stash_bytes(long long P) { return stash_total_bytes(num_tiles(P)); }

extern "C" int dfn_mlp_pack_weights(const float* params, void* fwd_pack, void* bwd_pack, float* consts, void* stream) {
  DFN_CHECK_ARG(params && fwd_pack && bwd_pack && consts, "mlp_pack_weights: null pointer");
  pack_weights_kernel<<<dfn_num_sms() * 4, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      params, nullptr, static_cast<__nv_bfloat16*>(fwd_pack), static_cast<__nv_bfloat16*>(bwd_pack), consts, weight_replicas());
  DFN_CHECK_LAUNCH("mlp_pack_weights");
  return DFN_OK;
}

// The repack after every optimiser step as a gather through a precomputed source map (dfn_mlp_pack_map_bytes() bytes of int32,
// filled once by dfn_mlp_build_pack_map; the same map serves every network of this architecture on the device).
extern "C" long long dfn_mlp_pack_map_bytes(void) { return static_cast<long long>(kPackElems) * 4; }
extern "C" int dfn_mlp_build_pack_map(int* map, void* stream) {
  DFN_CHECK_ARG(map, "mlp_build_pack_map: null pointer");
  pack_map_kernel<<<dfn_num_sms() * 4, 256, 0, static_cast<cudaStream_t>(stream)>>>(map);
  DFN_CHECK_LAUNCH("mlp_build_pack_map");
  return DFN_OK;
}
extern "C" int dfn_mlp_pack_weights_mapped(const float* params, const int* map, void* fwd_pack, void* bwd_pack, float* consts,
                                           void* stream) {
  DFN_CHECK_ARG(params && map && fwd_pack && bwd_pack && consts, "mlp_pack_weights_mapped: null pointer");
  pack_weights_kernel<<<dfn_num_sms() * 4, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      params, map, static_cast<__nv_bfloat16*>(fwd_pack), static_cast<__nv_bfloat16*>(bwd_pack), consts, weight_replicas());
  DFN_CHECK_LAUNCH("mlp_pack_weights_mapped");
  return DFN_OK;
}

static int check_source(const dfn_point_source* s, long long P, const char* who) {
  DFN_CHECK_ARG(s, "%s: null point source", who);
  if (s->mode == 0) DFN_CHECK_ARG(s->pos && s->view, "%s: mode 0 needs pos and view", who);
  else if (s->mode == 1) DFN_CHECK_ARG(s->rays_o && s->rays_d && s->z && s->S > 0 && P % s->S == 0, "%s: mode 1 needs rays_o/rays_d/z and P %% S == 0", who);
  else if (s->mode == 2) DFN_CHECK_ARG(s->lin && s->view && s->R > 0 && s->p_offset >= 0 && s->p_offset + P <= 1LL * s->R * s->R * s->R, "%s: mode 2 needs lin/view and a range inside R^3", who);
  else DFN_CHECK_ARG(false, "%s: unknown point source mode %d", who, s->mode);
  return DFN_OK;
}

static int mlp_forward_impl(const dfn_point_source* src, long long P, const void* fwd_pack, const float* consts,
                            const float* noise, unsigned long long rng_seed, const int* rng_step, unsigned rng_stream, float* rgb,
                            float* sigma, void* stash, void* stream) {
  if (int rc = check_source(src, P, "mlp_forward")) return rc;
  DFN_CHECK_ARG(fwd_pack && consts && rgb && sigma, "mlp_forward: null pointer");
  DFN_CHECK_ARG(P >= 0 && P < (1LL << 31) * kTile, "mlp_forward: P out of range");
  if (P == 0) return DFN_OK;
  FwdArgs a{};
  static const int skew_ns = [] { const char* e = getenv("DFN_FWD_SKEW_NS"); return e ? atoi(e) : 0; }();   // measured: no gain in the forward (0 .. 8000 ns), 3 % in dgrad
  a.skew_ns = skew_ns;
  a.src = *src; a.P = P; a.n_tiles = static_cast<int>(num_tiles(P));
  a.wpack = static_cast<const uint8_t*>(fwd_pack); a.consts = consts; a.noise = noise; a.rgb = rgb; a.sigma = sigma;
  a.rng_seed = rng_seed; a.rng_step = rng_step; a.rng_stream = rng_stream;
  if (stash) {
    uint8_t* s = static_cast<uint8_t*>(stash);
    a.stash_h = s;
    a.stash_f = s + stash_off_feat(a.n_tiles);
    a.stash_c = s + stash_off_c(a.n_tiles);
    a.stash_m = s + stash_off_mask(a.n_tiles);
  }
  WeightMaps maps;
  a.replicas = weight_replicas();
  if (int rc = make_weight_maps(fwd_pack, static_cast<unsigned long long>(kFwdPackBytes) * a.replicas, &maps)) return rc;
  // one CTA per SM, paired: grid = 2 * min(#SM / 2, #super-tiles)
  const int pairs = a.n_tiles / 2 < dfn_mlp_sms() / 2 ? a.n_tiles / 2 : dfn_mlp_sms() / 2;
  const int grid = 2 * pairs;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (stash) {
    DFN_CUDA(cudaFuncSetAttribute(mlp_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes), "mlp_forward");
    mlp_fwd_kernel<true><<<grid, kThreads, kSmemBytes, st>>>(a, maps);
  } else {
    DFN_CUDA(cudaFuncSetAttribute(mlp_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes), "mlp_forward");
    mlp_fwd_kernel<false><<<grid, kThreads, kSmemBytes, st>>>(a, maps);
  }
  DFN_CHECK_LAUNCH("mlp_forward");
  return DFN_OK;
}

extern "C" int dfn_mlp_forward(const dfn_point_source* src, long long P, const void* fwd_pack, const float* consts,
                               const float* noise, float* rgb, float* sigma, void* stash, void* stream) {
  return mlp_forward_impl(src, P, fwd_pack, consts, noise, 0ull, nullptr, 0u, rgb, sigma, stash, stream);
}

// training-mode forward with the sigma noise of model.py:160-162 drawn in the kernel (Philox, csrc/rng.cuh):
// noise(point p) = N(0,1)(seed, *step_ptr, rng_stream, p), rounded to bf16 like randn_like on an autocast tensor
extern "C" int dfn_mlp_forward_rng(const dfn_point_source* src, long long P, const void* fwd_pack, const float* consts,
                                   unsigned long long seed, const int* step_ptr, unsigned rng_stream, float* rgb, float* sigma,
                                   void* stash, void* stream) {
  DFN_CHECK_ARG(step_ptr, "mlp_forward_rng: null step pointer");
  return mlp_forward_impl(src, P, fwd_pack, consts, nullptr, seed, step_ptr, rng_stream, rgb, sigma, stash, stream);
}
