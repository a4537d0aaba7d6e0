// Backward of the fused vanilla-NeRF MLP as ONE launch: data gradients and weight gradients run CONCURRENTLY on disjoint
// SMs and the dZ tiles go from one to the other through a small ring that stays resident in L2, instead of a 4 880 B/point
// round trip through HBM (dgrad wrote, wgrad re-read: 10.2 GB of the 20.9 GB a 4096-ray training step moved).
//
//   clusters [0, n_pairs)        : the data-gradient pipeline of mlp_bwd.cu (CTA pairs, tcgen05 cta_group::2, W^T streamed
//                                  by tensor-map TMA, 8 epilogue warps).  Every dZ tile an epilogue publishes in shared
//                                  memory is shipped by the store thread as ONE bulk copy into a ring slot and announced
//                                  with a release flag: per stream q (dZ of layer 0..7, dZ_feat, dZ_c0) a ticket counter
//                                  hands out slots in production order.
//   CTAs  [2 n_pairs, grid)      : the weight-gradient units of mlp_wgrad.cu (one (layer, part) per CTA, dW block resident
//                                  in all 512 TMEM columns, MN-major tcgen05.mma over the points, ONE atomic flush at the
//                                  end).  Part c of a unit with `split` parts consumes tickets c, c + split, ... of its
//                                  stream in order: acquire the flag, read the tile index, bulk-load the dZ tile from the
//                                  ring (an L2 hit) and the matching activation tile from the forward's stash (HBM), and
//                                  release the slot as soon as the tile sits in shared memory.
// All CTAs of the launch are co-resident (grid <= #SMs, one CTA per SM); producers block only on a full ring and consumers
// only on their own stream, so the wait-for graph is acyclic (induction over ticket numbers).  Every spin is bounded and
// traps instead of hanging.  Numerics are those of the two-launch path: the same bf16 dZ tiles, the same MMAs, the same fp32
// atomics - only the order of the fp32 atomic additions differs.
#include "mlp_pipe.cuh"
#include <stdlib.h>
#include <type_traits>

namespace {
using namespace mlp;
using namespace umma;
using namespace mlp_pipe;

// ---------------------------------------------------------------------------------------------------
// ring control block (global memory, zeroed by the host before every launch)
// ---------------------------------------------------------------------------------------------------
constexpr int kStreams = 10;          // q = 0..7: dZ of layer q (pos_layers.0 .. skip_layers.6); 8: dZ_feat; 9: dZ_c0 (32 KB)
constexpr int kMaxSlots = 256;
constexpr uint32_t kSlotBytes = 65536;
struct RingCtl {
  unsigned head[16];                  // tickets handed out per stream
  unsigned assigned[kStreams][kMaxSlots];   // ticket + 1 once the slot belongs to the ticket and tile[] is valid (its bytes are still on their way)
  unsigned ready[kStreams][kMaxSlots];   // ticket + 1 once the slot's bytes are in place
  unsigned done[kStreams][kMaxSlots];    // consumptions completed on the slot (monotonic)
  int tile[kStreams][kMaxSlots];         // which 128-point tile the slot holds
};
// dZ of skip_layers.0 feeds the 256-column block and the PE columns; a dZ_c0 ticket also tells the color_layer.2 unit that the
// tile's head gradients are in place
__host__ __device__ constexpr int stream_consumers(int q) { return (q == 4 || q == 9) ? 2 : 1; }

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned* p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_release_gpu_add(unsigned* p, unsigned v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

#ifdef DFN_FUSED_STATS
#define ST_DECL(...) long long __VA_ARGS__
#define ST_T0() const long long st_t0 = clock64()
#define ST_ADD(v) v += clock64() - st_t0
#else
#define ST_DECL(...)
#define ST_T0()
#define ST_ADD(v)
#endif

// bounded spin on a global counter: returns when *p >= want (acquire)
__device__ __forceinline__ void spin_until_ge(const unsigned* p, unsigned want, const char* what) {
  if (ld_acquire_gpu(p) >= want) return;
  const long long t0 = clock64();
  unsigned ns = 32;
  while (ld_acquire_gpu(p) < want) {
    __nanosleep(ns);
    if (ns < 256) ns <<= 1;
    if (clock64() - t0 > 4000000000LL) {
      printf("dfn: bwd ring timeout (%s, block %d thread %d, have %u want %u)\n", what, blockIdx.x, threadIdx.x, ld_acquire_gpu(p), want);
      __trap();
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// shared configuration of the merged kernel
// ---------------------------------------------------------------------------------------------------
constexpr int kEpiWarps = 8;                      // dgrad role: 4 TMEM lane quadrants x 2 column halves
constexpr int kColParts = kEpiWarps / 4;
constexpr int kGroups = 8 / kColParts;            // 32-column groups per epilogue thread
constexpr int kProChunks = 16 / kColParts;
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kStoreWarp = 2 + kEpiWarps;         // two store warps: kStoreWarp + t ships the tiles of slot t
constexpr int kThreads = 64 + kEpiThreads + 64;   // 384: producer, issuer, 8 epilogue / worker warps, 2 store warps

// ---- dgrad role: shared memory map ----
constexpr int kStages = 3;
constexpr uint32_t kBwdStageBytes = 32768;
constexpr uint32_t kOffAct = 0;
constexpr uint32_t kOffRing = kOffAct + 2 * kActBytes;
constexpr uint32_t kOffConst = kOffRing + kStages * kBwdStageBytes;
constexpr uint32_t kOffBars = kOffConst + kConstBytes;
constexpr uint32_t kNumBars = 4 * kStages + 8;
constexpr uint32_t kOffTmemPtr = kOffBars + kNumBars * 8;
constexpr uint32_t kDgradSmem = kOffTmemPtr + 16;
static_assert(kDgradSmem <= 232448, "shared memory budget exceeded");

struct BwdPipe {
  static constexpr int kNumJobs = kNumBwdJobs;
  static constexpr int kStages = ::kStages;
  static constexpr int kCPS = 16;
  static constexpr uint32_t kRingStageBytes = kBwdStageBytes;
  static constexpr uint32_t kOffRing = ::kOffRing;
  static __host__ __device__ constexpr uint32_t bar_full_i(int t, int s) { return t * kStages + s; }
  static __host__ __device__ constexpr uint32_t bar_empty_i(int t, int s) { return 2 * kStages + t * kStages + s; }
  static __host__ __device__ constexpr uint32_t bar_accfull_i(int t) { return 4 * kStages + t; }
  static __host__ __device__ constexpr uint32_t bar_aready_i(int t) { return 4 * kStages + 2 + t; }
  static __host__ __device__ constexpr int job_N(int) { return 256; }
  static __host__ __device__ constexpr int job_chunks(int b) { return bjob_kc(b); }
  static __host__ __device__ constexpr uint32_t job_off(int b) { return bjob_off(b); }
  static __host__ __device__ constexpr uint32_t job_half_chunk(int) { return 128u * 16u; }
  static __host__ __device__ constexpr uint32_t job_half_bytes(int b) { return bjob_bytes(b) / 2; }
  struct ABases { uint32_t act; };
  static __device__ __forceinline__ ABases bases(uint32_t sbase, int T) {
    return ABases{desc_lo(sbase + kOffAct + T * kActBytes, kChunkBytes)};
  }
  template <int J>
  static __device__ __forceinline__ uint32_t a_chunk_lo(int kc, const ABases& b) { return b.act + kc * (kChunkBytes >> 4); }
};

// ---- wgrad role: shared memory map ----
constexpr int kWorkers = 8;
constexpr int kWorkThreads = kWorkers * 32;
constexpr uint32_t kInBytes = 65536, kDzBytes = 32768, kPeBytes = 8192;
constexpr uint32_t kWNumBars = 32;
constexpr uint32_t kWOffBars = 0;                        // 1 KB header: barriers, recent tile indices, TMEM pointer
constexpr uint32_t kWOffTiles = kWOffBars + kWNumBars * 8;   // 8 x int: tile index of the producer's recent tickets
constexpr uint32_t kWOffTmemPtr = kWOffTiles + 64;
constexpr uint32_t kWOffDsig = 1024;                     // 2 x [128] f32: d sigma_pre of the tile (feature-layer unit)
constexpr uint32_t kWOffIn = 2048;                       // 2 x 64 KB
constexpr uint32_t kWOffDz = kWOffIn + 2 * kInBytes;     // 2 x 32 KB; big units: 3 x 32 KB (they use neither PE nor small tiles)
constexpr uint32_t kWOffPe = kWOffDz + 2 * kDzBytes;     // 2 x 8 KB (dir-PE tiles, colour unit)
constexpr uint32_t kWOffSmall = kWOffPe + 2 * kPeBytes;  // 2 x [128][4] f32
constexpr uint32_t kWgradSmem = kWOffDz + 3 * kDzBytes;
static_assert(kWOffSmall + 2 * 2048 <= kWgradSmem && kWgradSmem <= 232448, "shared memory budget exceeded");
constexpr uint32_t kSmemBytes = kDgradSmem > kWgradSmem ? kDgradSmem : kWgradSmem;

enum { TYPE_BIG = 0, TYPE_PE = 1, TYPE_COLOR = 2, TYPE_COLOR2 = 3 };   // COLOR: color_layer.0 (MMA); COLOR2: color_layer.2 (CUDA cores)

struct FusedArgs {
  // dgrad
  long long P;
  int n_tiles;           // even (num_tiles(P))
  const float* consts;
  const float* rgb;      // [P,3] forward output
  const float* sigma;    // [P]
  const float* d_rgb;    // [P,3]
  const float* d_sigma;  // [P]
  const uint8_t* stash;  // forward stash
  int skew_ns;
  int replicas;          // copies of the packed weight image; pair p streams copy p % replicas
  // hand-off
  RingCtl* ctl;
  uint8_t* ring;         // [kStreams][n_slots][64 KB]
  float* small;          // [n_tiles][128][4] f32: dZ_c2[0..2], d sigma_pre (full size, written once, 16 B/point)
  int n_slots;
  // wgrad
  dfn_point_source src;
  float* grads;
  int n_pairs;           // dgrad pairs = clusters [0, n_pairs)
  int big_split, big7_split, pe_split, col_split, col2_split;   // parts per unit: 8 big units, 2 PE units (layer 0 / the PE columns of layer 4), colour 0 / 2
};

__device__ __forceinline__ uint32_t relu_mask2(uint32_t h2) {
  return ((h2 & 0x0000FFFFu) ? 0x0000FFFFu : 0u) | ((h2 & 0xFFFF0000u) ? 0xFFFF0000u : 0u);
}

// tile of (iteration, slot) for a dgrad CTA: super-tile index * 2 + rank, over the dgrad pairs only
__device__ __forceinline__ int ftile_of(int it, int t, int rank, int n_pairs) {
  const int pair = blockIdx.x >> 1;
  return 2 * ((it * 2 + t) * n_pairs + pair) + rank;
}

// ===================================================================================================
// dgrad role
// ===================================================================================================
__device__ __forceinline__ void dgrad_role(const FusedArgs& a, const WeightMaps& maps, uint8_t* smem) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bar0 = sbase + kOffBars;
  auto bar_accfull = [&](int t) { return bar0 + 8u * BwdPipe::bar_accfull_i(t); };
  auto bar_aready = [&](int t) { return bar0 + 8u * BwdPipe::bar_aready_i(t); };
  auto bar_stored = [&](int t) { return bar0 + 8u * (4 * kStages + 4 + t); };   // the tile in act[t] has been copied out
  auto bar_last = [&](int t) { return bar0 + 8u * (4 * kStages + 6 + t); };     // the last job's dZ tile (layer 0) sits in act[t]
  float* consts = reinterpret_cast<float*>(smem + kOffConst);

  if (threadIdx.x == 0) {
    for (int t = 0; t < 2; ++t)
      for (int s = 0; s < kStages; ++s) {
        mbar_init(bar0 + 8u * BwdPipe::bar_full_i(t, s), 1);
        mbar_init(bar0 + 8u * BwdPipe::bar_empty_i(t, s), 1);
      }
    for (int t = 0; t < 2; ++t) {
      mbar_init(bar_accfull(t), 1);
      mbar_init(bar_aready(t), kEpiWarps + (rank == 0 ? 1 : 0));   // + the peer's relay
      mbar_init(bar_stored(t), 1);
      mbar_init(bar_last(t), kEpiWarps);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sbase + kOffTmemPtr), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  for (int i = threadIdx.x; i < kConstFloats; i += kThreads) consts[i] = a.consts[i];
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem + kOffTmemPtr);
  const int n_pairs = a.n_pairs;
  const int nt = a.n_tiles;
  const int n_it = ((nt >> 1) + 2 * n_pairs - 1) / (2 * n_pairs);
  auto tile_at = [&](int it, int t) { return ftile_of(it, t, rank, n_pairs); };

  if (warp == 0) {
    if (lane == 0) {  // ---- weight producer (this CTA's half of every W^T block)
      ProdState<BwdPipe> ps;
#pragma unroll
      for (int i = 0; i < kStages; ++i) { ps.use[0][i] = ps.use[1][i] = 0; ps.last[i] = -1; }
      ps.pol = make_evict_last_policy();
      ps.row0 = static_cast<uint32_t>((blockIdx.x >> 1) % a.replicas) * (kBwdPackBytes >> 7);
      for (int w = (blockIdx.x >> 1) & 7; a.skew_ns > 0 && w > 0; --w) __nanosleep(a.skew_ns);
      for (int it = 0; it < n_it; ++it) {
        const bool has[2] = {tile_at(it, 0) < nt, tile_at(it, 1) < nt};
        produce_all<BwdPipe>(maps, rank, sbase + kOffRing, bar0, ps, has, std::make_integer_sequence<int, kNumBwdJobs>{});
      }
    }
  } else if (warp == 1) {
    IssueCtx c{};
    c.sbase = sbase; c.bar0 = bar0; c.tmem_base = tmem_base; c.lane = lane; c.rank = rank;
    c.s0 = SlotCtx{};
    c.s1 = c.s0;
    for (int it = 0; it < n_it; ++it) {
      c.s0.has = tile_at(it, 0) < nt;
      c.s1.has = tile_at(it, 1) < nt;
      issue_all<BwdPipe>(c, std::make_integer_sequence<int, kNumBwdJobs>{});
    }
  } else if (warp >= kStoreWarp) {
    // ---- ring producers (one thread per slot): ticket -> free slot -> one bulk copy shared -> global (L2) -> release flag.
    // Per item the thread is busy ~4000 cycles (2200 until the tile has left shared memory, the rest until the write is
    // acknowledged and the flag is out); one thread for both slots was the bottleneck of the pair (20 items per ~84 000 cycles).
    if (lane == 0) {
      const int t = warp - kStoreWarp;
      uint32_t pa = 0, pl = 0;
      const unsigned ns = static_cast<unsigned>(a.n_slots);
      ST_DECL(c_ticket = 0, c_slot = 0, c_read = 0, c_write = 0, c_aready = 0, c_total = clock64());
      // One item = ticket -> free slot -> "assigned" flag (the consumer may fetch the matching stash tile) -> bulk copy ->
      // "ready" flag.  The first three steps do not need the tile's bytes: they run BEFORE the thread waits for the epilogues
      // (reserve), so that between "tile published" and "act[t] may be overwritten" there is only the copy itself.
      unsigned r_k = 0, r_slot = 0;
      int r_q = 0;
      auto reserve = [&](int q, int tile) {
        { ST_T0(); r_k = atomicAdd(&a.ctl->head[q], 1u); ST_ADD(c_ticket); }
        r_q = q;
        r_slot = r_k % ns;
        { ST_T0(); spin_until_ge(&a.ctl->done[q][r_slot], (r_k / ns) * stream_consumers(q), "slot free"); ST_ADD(c_slot); }
        a.ctl->tile[q][r_slot] = tile;
        st_release_gpu(&a.ctl->assigned[q][r_slot], r_k + 1u);
      };
      auto ship = [&](uint32_t bytes) {
        uint8_t* dst = a.ring + (static_cast<size_t>(r_q) * ns + r_slot) * kSlotBytes;
        { ST_T0();
        bulk_s2g(dst, sbase + kOffAct + t * kActBytes, bytes);
        bulk_commit();
        bulk_wait_read_all();
        ST_ADD(c_read); }
        mbar_arrive(bar_stored(t));      // act[t] may be overwritten
        { ST_T0();
        bulk_wait_all();                 // the bytes are in place (L2)
        fence_proxy_async_all();
        st_release_gpu(&a.ctl->ready[r_q][r_slot], r_k + 1u);
        ST_ADD(c_write); }
      };
      for (int it = 0; it < n_it; ++it) {
        const int tile = tile_at(it, t);
        if (tile >= nt) continue;
        for (int b = 0; b < kNumBwdJobs; ++b) {
          reserve(b == 0 ? 9 : (b == 1 ? 8 : 9 - b), tile);
          { ST_T0(); mbar_wait(bar_aready(t), pa); ST_ADD(c_aready); }   // A operand of job b: dZ_c0 (b = 0, written by the prologue) or job b-1's output
          pa ^= 1;
          ship(b == 0 ? kActBytes / 2 : kActBytes);
        }
        reserve(0, tile);
        { ST_T0(); mbar_wait(bar_last(t), pl); ST_ADD(c_aready); }        // the last job's output: dZ of layer 0 (nobody reads it on this SM)
        pl ^= 1;
        ship(kActBytes);
      }
#ifdef DFN_FUSED_STATS
      if (blockIdx.x == 0 || blockIdx.x == 17)
        printf("fused stats dgrad CTA %d store thread %d (cycles): total %lld wait_aready %lld ticket %lld slot_free %lld copy_read %lld write+release %lld\n",
               blockIdx.x, t, clock64() - c_total, c_aready, c_ticket, c_slot, c_read, c_write);
#endif
    }
  } else {
    // ---- epilogue warps
    const int ew = warp - 2, q = warp & 3, hh = ew >> 2;
    const int row = q * 32 + lane;
    const uint32_t lane_sel = static_cast<uint32_t>(q * 32) << 16;
    uint32_t pf[2] = {0, 0}, pst[2] = {0, 0};
    auto publish = [&](int t) {
      tc_fence_before();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_aready(t));
    };
    auto prologue = [&](int t, int tile) -> float {
      uint8_t* act = smem + kOffAct + t * kActBytes;
      const long long p = static_cast<long long>(tile) * kTile + row;
      const bool valid = p < a.P;
      float g[3] = {0.f, 0.f, 0.f};
      float ds = 0.f;
      if (valid) {
#pragma unroll
        for (int o = 0; o < 3; ++o) {
          const float r = a.rgb[p * 3 + o];
          g[o] = a.d_rgb[p * 3 + o] * r * (1.0f - r);
        }
        ds = a.sigma[p] > 0.f ? a.d_sigma[p] : 0.f;
      }
      if (hh == 0) reinterpret_cast<float4*>(a.small + static_cast<size_t>(tile) * 512)[row] = make_float4(g[0], g[1], g[2], ds);
      const uint8_t* cst = a.stash + (static_cast<size_t>(9) * nt << 16) + (static_cast<size_t>(tile) << 15);
      const float* w2 = consts + kConstWc2;
#pragma unroll
      for (int ch = 0; ch < kProChunks; ++ch) {
        const int chunk = hh * kProChunks + ch;
        const uint32_t off = chunk * kChunkBytes + row * 16;
        const uint4 cv = *reinterpret_cast<const uint4*>(cst + off);
        const uint32_t cw[4] = {cv.x, cv.y, cv.z, cv.w};
        uint32_t o4[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int k = chunk * 8 + 2 * i;
          const float d0 = g[0] * w2[k] + g[1] * w2[128 + k] + g[2] * w2[256 + k];
          const float d1 = g[0] * w2[k + 1] + g[1] * w2[128 + k + 1] + g[2] * w2[256 + k + 1];
          o4[i] = pack_bf16x2(d0, d1) & relu_mask2(cw[i]);
        }
        *reinterpret_cast<uint4*>(act + off) = make_uint4(o4[0], o4[1], o4[2], o4[3]);
      }
      return ds;
    };
    auto prefetch_prologue = [&](int tile) {
      const long long p = static_cast<long long>(tile) * kTile + row;
      const uint8_t* cst = a.stash + (static_cast<size_t>(9) * nt << 16) + (static_cast<size_t>(tile) << 15);
      if ((lane & 7) == 0) {
#pragma unroll
        for (int ch = 0; ch < kProChunks; ++ch) prefetch_l2(cst + (hh * kProChunks + ch) * kChunkBytes + row * 16);
      }
      if (hh == 0 && p < a.P && (lane & 3) == 0) {
        prefetch_l2(a.rgb + p * 3);
        prefetch_l2(a.d_rgb + p * 3);
        if ((lane & 15) == 0) { prefetch_l2(a.sigma + p); prefetch_l2(a.d_sigma + p); }
      }
    };
    ST_DECL(e_acc = 0, e_stored = 0, e_total = clock64());
    float ds0 = 0.f, ds1 = 0.f;
    uint32_t mk_next[kGroups];
#pragma unroll
    for (int i = 0; i < kGroups; ++i) mk_next[i] = 0;
    const int ec0 = hh * (256 / kColParts);
    auto load_mask = [&](const uint8_t* tile_masks) {
      const uint8_t* pm = tile_masks + ((ec0 >> 7) * 128 + row) * 16 + ((ec0 & 127) >> 5) * 4;
      const uint4 v = *reinterpret_cast<const uint4*>(pm);
      mk_next[0] = v.x; mk_next[1] = v.y; mk_next[2] = v.z; mk_next[3] = v.w;
    };
    static_assert(kGroups == 4, "mask load assumes four 32-column groups per thread");

    for (int t = 0; t < 2; ++t) {
      const int tile = tile_at(0, t);
      if (tile < nt) { const float ds = prologue(t, tile); if (t) ds1 = ds; else ds0 = ds; }
      publish(t);
    }

    for (int it = 0; it < n_it; ++it)
      for (int b = 0; b < kNumBwdJobs; ++b)
        for (int t = 0; t < 2; ++t) {
          const int tile = tile_at(it, t);
          if (tile >= nt) continue;
          uint8_t* act = smem + kOffAct + t * kActBytes;
          const uint32_t tacc = tmem_base + t * 256 + lane_sel;
          const int c0 = ec0;
          uint32_t mk[kGroups];
#pragma unroll
          for (int i = 0; i < kGroups; ++i) mk[i] = mk_next[i];
          {
            const int tile1 = tile_at(it, 1);
            const bool other = t == 0 && tile1 < nt;
            const int nb = other ? b : b + 1;
            const int ntl = other ? tile1 : tile_at(it, 0);
            if (nb >= 1 && nb < kNumBwdJobs)
              load_mask(a.stash + stash_off_mask(nt) + ((static_cast<size_t>(8 - nb) * nt + ntl) << 12));
          }
          const float dsig = t ? ds1 : ds0;
          if (b == 4) {
            const int ntile = tile_at(it + 1, t);
            if (ntile < nt) prefetch_prologue(ntile);
          }
          { ST_T0(); mbar_wait(bar_accfull(t), pf[t]); ST_ADD(e_acc); } pf[t] ^= 1;
          tc_fence_after();
          { ST_T0(); mbar_wait(bar_stored(t), pst[t]); ST_ADD(e_stored); } pst[t] ^= 1;   // the store thread has read the previous dZ tile out of act
          //   0: b = 0 (no mask)   1: b = 1 (+ d sigma_pre * density weights, mask)   2: b = 2..8 (mask)
          auto drain = [&](auto kind_c) {
            constexpr int KIND = decltype(kind_c)::value;
            uint8_t* out_row = act + static_cast<uint32_t>(c0 / 8) * kChunkBytes + row * 16;
            uint32_t v0[32], v1[32];
            tmem_ld32(tacc + c0, v0);
            tmem_ld_wait();
#pragma unroll
            for (int cc = 0; cc < kGroups; ++cc) {
              uint32_t (&cur)[32] = (cc & 1) ? v1 : v0;
              uint32_t (&nxt)[32] = (cc & 1) ? v0 : v1;
              const int col0 = c0 + cc * 32;
              if (cc < kGroups - 1) tmem_ld32(tacc + col0 + 32, nxt);
              const uint32_t mw = mk[cc];
#pragma unroll
              for (int g4 = 0; g4 < 4; ++g4) {
                float x[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) x[i] = __uint_as_float(cur[g4 * 8 + i]);
                if constexpr (KIND == 1) {
                  const float* ws = consts + kConstWsig + col0 + g4 * 8;
#pragma unroll
                  for (int i = 0; i < 8; ++i) x[i] += dsig * ws[i];
                }
                uint4 o;
                o.x = pack_bf16x2(x[0], x[1]); o.y = pack_bf16x2(x[2], x[3]);
                o.z = pack_bf16x2(x[4], x[5]); o.w = pack_bf16x2(x[6], x[7]);
                if constexpr (KIND >= 1) {
                  o.x &= relu_bits_expand(mw, g4 * 4); o.y &= relu_bits_expand(mw, g4 * 4 + 1);
                  o.z &= relu_bits_expand(mw, g4 * 4 + 2); o.w &= relu_bits_expand(mw, g4 * 4 + 3);
                }
                *reinterpret_cast<uint4*>(out_row + (cc * 4 + g4) * kChunkBytes) = o;
              }
              if (cc < kGroups - 1) tmem_ld_wait();
            }
          };
          if (b == 0) drain(std::integral_constant<int, 0>{});
          else if (b == 1) drain(std::integral_constant<int, 1>{});
          else drain(std::integral_constant<int, 2>{});
          if (b == kNumBwdJobs - 1) {
            // dZ of layer 0 sits in act[t]: hand it to the store thread, then (once it has left) start the slot's next tile
            tc_fence_before();
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_last(t));
            const int ntile = tile_at(it + 1, t);
            if (ntile < nt) {
              { ST_T0(); mbar_wait(bar_stored(t), pst[t]); ST_ADD(e_stored); } pst[t] ^= 1;
              const float ds = prologue(t, ntile);
              if (t) ds1 = ds; else ds0 = ds;
              publish(t);
            }
          } else {
            publish(t);
          }
        }
#ifdef DFN_FUSED_STATS
    if ((blockIdx.x == 0 || blockIdx.x == 17) && ew == 0 && lane == 0)
      printf("fused stats dgrad CTA %d epilogue warp 0 (cycles): total %lld wait_accfull %lld wait_stored %lld\n", blockIdx.x, clock64() - e_total, e_acc, e_stored);
#endif
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
}

// ===================================================================================================
// wgrad role
// ===================================================================================================
template <int NCH, int NW, int WS = 4>
__device__ __forceinline__ void wcolsum(const uint8_t* tile, const float* wrow, float (&acc)[NW][8], int wt) {
  constexpr int TPC = kWorkThreads / NCH;
  const int c = wt / TPC, r0 = wt % TPC;
#pragma unroll 4
  for (int r = r0; r < kTile; r += TPC) {
    const uint4 v = *reinterpret_cast<const uint4*>(tile + c * kChunkBytes + r * 16);
    const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
    float x[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) { x[2 * i] = bf16lo_to_f32(w4[i]); x[2 * i + 1] = bf16hi_to_f32(w4[i]); }
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      const float wv = wrow ? wrow[r * WS + w] : 1.0f;
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[w][i] += wv * x[i];
    }
  }
}
template <int NCH>
__device__ __forceinline__ void flush_colsum(float (&acc)[8], float* dst, int wt, int ncols_valid) {
  constexpr int TPC = kWorkThreads / NCH;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float v = acc[i];
#pragma unroll
    for (int o = TPC / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    acc[i] = v;
  }
  if (wt % TPC == 0) {
    const int c = wt / TPC;
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (c * 8 + i < ncols_valid) atomicAdd(dst + c * 8 + i, acc[i]);
  }
}

__device__ __forceinline__ void wgrad_role(const FusedArgs& a, uint8_t* smem) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bar0 = sbase + kWOffBars;
  auto bar_in_full = [&](int s) { return bar0 + 8u * s; };
  auto bar_in_empty = [&](int s) { return bar0 + 8u * (2 + s); };
  auto bar_dz_full = [&](int s) { return bar0 + 8u * (4 + s); };     // 4 stages
  auto bar_dz_empty = [&](int s) { return bar0 + 8u * (8 + s); };
  auto bar_pe_full = [&](int s) { return bar0 + 8u * (12 + s); };
  auto bar_pe_empty = [&](int s) { return bar0 + 8u * (14 + s); };
  const uint32_t bar_done = bar0 + 8u * 16;
  auto bar_tile = [&](int i) { return bar0 + 8u * (17 + i); };       // 8 entries: the tile index of a ticket is in s_tiles[i]
  volatile int* s_tiles = reinterpret_cast<volatile int*>(smem + kWOffTiles);
  constexpr ParamOffsets po = param_offsets();
  const int nt = a.n_tiles;
  const unsigned ns = static_cast<unsigned>(a.n_slots);

  // ---- which unit is this CTA? ----
  int type, job = 0, part, split, pe_layer = 0;
  {
    const int u = static_cast<int>(blockIdx.x) - 2 * a.n_pairs;
    const int nb7 = 7 * a.big_split, nb = nb7 + a.big7_split;   // the feature-layer unit (job 7) also reduces the density head: more parts
    if (u < nb7) { type = TYPE_BIG; job = u / a.big_split; part = u % a.big_split; split = a.big_split; }
    else if (u < nb) { type = TYPE_BIG; job = 7; part = u - nb7; split = a.big7_split; }
    else if (u < nb + 2 * a.pe_split) { type = TYPE_PE; pe_layer = (u - nb) < a.pe_split ? 0 : 4; part = (u - nb) % a.pe_split; split = a.pe_split; }
    else if (u < nb + 2 * a.pe_split + a.col_split) { type = TYPE_COLOR; part = u - nb - 2 * a.pe_split; split = a.col_split; }
    else { type = TYPE_COLOR2; part = u - nb - 2 * a.pe_split - a.col_split; split = a.col2_split; }
  }
  const int big_layer = job < 3 ? job + 1 : (job == 3 ? 4 : (job < 7 ? job + 1 : 9));
  const int big_in = job < 7 ? big_layer - 1 : 7;
  const int q = type == TYPE_BIG ? (job == 7 ? 8 : big_layer) : (type == TYPE_PE ? pe_layer : 9);   // the stream this unit consumes
  const bool active = part < split && part < nt;   // a unit with more parts than tickets has idle parts

  if (threadIdx.x == 0) {
    for (int s = 0; s < 4; ++s) { mbar_init(bar_dz_full(s), 1); mbar_init(bar_dz_empty(s), 1 + kWorkers); }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_in_full(s), type == TYPE_PE ? kWorkThreads : 1);
      mbar_init(bar_in_empty(s), type == TYPE_PE ? 1 : 1 + kWorkers);
      mbar_init(bar_pe_full(s), kWorkThreads);
      mbar_init(bar_pe_empty(s), 1);
    }
    mbar_init(bar_done, 1);
    for (int i = 0; i < 8; ++i) mbar_init(bar_tile(i), 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(sbase + kWOffTmemPtr, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem + kWOffTmemPtr);

  // dz ring in shared memory: two 32 KB stages; the PE unit's input tiles are computed, not loaded, and fill only 16 KB of
  // each 64 KB input slot, so it parks two more dz stages there
  // (the color_layer.2 unit has no input tiles either: four stages of c tiles)
  // big units: three stages, so that every 32 KB ring load (~1500 cycles from L2) has two MMA blocks (2048 cycles) to land
  const int n_dzs = (type == TYPE_PE || type == TYPE_COLOR2) ? 4 : (type == TYPE_BIG ? 3 : 2);
  auto dz_off = [&](int s) -> uint32_t {
    return (s < 2 || type == TYPE_BIG) ? kWOffDz + s * kDzBytes : kWOffIn + (s - 2) * kInBytes + 16384u;
  };
  const int n_load = (type == TYPE_COLOR || type == TYPE_COLOR2) ? 1 : 2;   // 32 KB dz-stage loads per ticket
  const bool from_ring = type != TYPE_COLOR2;                               // ... out of the ring slot (else: the c tile of the stash)

  if (active && warp == 0) {
    if (lane == 0) {  // ================= producer: wait for the ticket, then load =================
      int si = 0, sd = 0; uint32_t pi = 0, pd = 0;
      const uint64_t pol = make_evict_first_policy();   // stash tiles are read exactly once
      int kc = 0;
      ST_DECL(w_flag = 0, w_flag2 = 0, w_in = 0, w_dz = 0, w_total = clock64());
      // Order of the bulk loads in the SM's (in-order) TMA queue: the two 32 KB ring loads of ticket kc first - the MMAs wait
      // for them next -, then the 64 KB stash tile of ticket kc + 1, which has a whole ticket of MMAs to arrive.
      // phase 1 of a ticket: the slot has been assigned and the tile index is known (the bytes are still being copied; the
      // producer reserves the slot before its epilogues have even published the tile).  Units whose workers only recompute
      // encodings from the points learn the tile index right away (8-entry ring in shared memory, this thread runs at most 4
      // stages ahead); units whose workers read the head gradients of the tile (written by the dgrad prologue) learn it in
      // phase 2, which the producer releases after it has observed the tile's publication.
      const bool tile_early = type == TYPE_PE || type == TYPE_COLOR;
      auto wait_ticket = [&](int k, int kc_) -> int {
        const unsigned slot = static_cast<unsigned>(k) % ns;
        { ST_T0(); spin_until_ge(&a.ctl->assigned[q][slot], static_cast<unsigned>(k) + 1u, "ticket assigned"); ST_ADD(w_flag); }
        const int tile = __ldcg(&a.ctl->tile[q][slot]);   // written by another SM in this launch: read it from L2
        if (tile_early) { s_tiles[kc_ & 7] = tile; mbar_arrive(bar_tile(kc_ & 7)); }
        return tile;
      };
      // phase 2: the slot's bytes are in place
      auto wait_data = [&](int k, int kc_, int tile) {
        const unsigned slot = static_cast<unsigned>(k) % ns;
        { ST_T0(); spin_until_ge(&a.ctl->ready[q][slot], static_cast<unsigned>(k) + 1u, "ticket ready"); ST_ADD(w_flag2); }
        fence_proxy_async_all();   // the slot was written through another SM's async proxy
        if (!tile_early) { s_tiles[kc_ & 7] = tile; mbar_arrive(bar_tile(kc_ & 7)); }
        if (type == TYPE_COLOR2) red_release_gpu_add(&a.ctl->done[q][slot], 1u);   // only the ticket is consumed, not the slot's bytes
      };
      auto load_in = [&](int tile) {
        if (type == TYPE_BIG || type == TYPE_COLOR) {
          const uint8_t* in = type == TYPE_BIG ? a.stash + ((static_cast<size_t>(big_in) * nt + tile) << 16)
                                               : a.stash + ((static_cast<size_t>(8) * nt + tile) << 16);   // colour: features
          { ST_T0(); mbar_wait(bar_in_empty(si), pi ^ 1); ST_ADD(w_in); }
          mbar_arrive_expect_tx(bar_in_full(si), kInBytes);
          bulk_g2s_hint(sbase + kWOffIn + si * kInBytes, in, 32768, bar_in_full(si), pol);
          bulk_g2s_hint(sbase + kWOffIn + si * kInBytes + 32768, in + 32768, 32768, bar_in_full(si), pol);
          if (++si == 2) { si = 0; pi ^= 1; }
        }
      };
      int tile = wait_ticket(part, 0);
      load_in(tile);
      for (int k = part; k < nt; k += split, ++kc) {
        const unsigned slot = static_cast<unsigned>(k) % ns;
        const uint8_t* rslot = a.ring + (static_cast<size_t>(q) * ns + slot) * kSlotBytes;
        wait_data(k, kc, tile);
        for (int i = 0; i < n_load; ++i) {
          { ST_T0(); mbar_wait(bar_dz_empty(sd), pd ^ 1); ST_ADD(w_dz); }
          mbar_arrive_expect_tx(bar_dz_full(sd), kDzBytes);
          if (from_ring) bulk_g2s(sbase + dz_off(sd), rslot + static_cast<size_t>(i) * kDzBytes, kDzBytes, bar_dz_full(sd));
          else bulk_g2s_hint(sbase + dz_off(sd), a.stash + (static_cast<size_t>(9) * nt << 16) + (static_cast<size_t>(tile) << 15), kDzBytes,
                             bar_dz_full(sd), pol);   // color_layer.2 unit: the c tile
          if (++sd == n_dzs) { sd = 0; pd ^= 1; }
        }
        if (k + split < nt) {
          tile = wait_ticket(k + split, kc + 1);
          load_in(tile);
        }
      }
#ifdef DFN_FUSED_STATS
      if (part == 0)
        printf("fused stats wgrad CTA %d type %d job %d pe_layer %d producer (cycles): total %lld tickets %d wait_assigned %lld wait_ready %lld wait_in_empty %lld wait_dz_empty %lld\n",
               blockIdx.x, type, job, pe_layer, clock64() - w_total, kc, w_flag, w_flag2, w_in, w_dz);
#endif
    }
  } else if (active && warp == 1) {
    if (lane == 0) {  // ================= MMA issuer =================
      int si = 0, sd = 0, sp = 0; uint32_t pi = 0, pd = 0, pp = 0;
      auto mma_block = [&](uint32_t a_addr, uint32_t b_addr, uint32_t tcol, int N, bool first_tile) {
        const uint32_t idesc = instr_desc_bf16(128, N, 1, 1);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          mma_bf16(tmem_base + tcol, smem_desc(a_addr + k * 256, 128, kChunkBytes), smem_desc(b_addr + k * 256, 128, kChunkBytes),
                   idesc, (first_tile && k == 0) ? 0u : 1u);
      };
      ST_DECL(i_in = 0, i_dz = 0, i_total = clock64());
      for (int k = part; k < nt; k += split) {
        const bool first = k == part;
        const uint32_t in_addr = sbase + kWOffIn + si * kInBytes;
        if (type == TYPE_COLOR2) {              // no MMAs: account for the issuer's share of the stage's empty barrier
          mbar_wait(bar_dz_full(sd), pd);
          mbar_arrive(bar_dz_empty(sd));
          if (++sd == n_dzs) { sd = 0; pd ^= 1; }
          continue;
        }
        { ST_T0(); mbar_wait(bar_in_full(si), pi); ST_ADD(i_in); }
        tc_fence_after();
        if (type == TYPE_BIG) {
          for (int h = 0; h < 2; ++h) {
            { ST_T0(); mbar_wait(bar_dz_full(sd), pd); ST_ADD(i_dz); }

            tc_fence_after();
            mma_block(sbase + dz_off(sd), in_addr, h * 256, 256, first);
            mma_commit(bar_dz_empty(sd));
            if (++sd == n_dzs) { sd = 0; pd ^= 1; }
          }
        } else if (type == TYPE_PE) {
          for (int i = 0; i < 2; ++i) {
            mbar_wait(bar_dz_full(sd), pd);

            tc_fence_after();
            mma_block(sbase + dz_off(sd), in_addr, i * 64, 64, first);
            mma_commit(bar_dz_empty(sd));
            if (++sd == n_dzs) { sd = 0; pd ^= 1; }
          }
        } else {
          mbar_wait(bar_dz_full(sd), pd);
          mbar_wait(bar_pe_full(sp), pp);
          tc_fence_after();
          mma_block(sbase + dz_off(sd), in_addr, 0, 256, first);
          mma_block(sbase + dz_off(sd), sbase + kWOffPe + sp * kPeBytes, 256, 32, first);
          mma_commit(bar_dz_empty(sd));
          mma_commit(bar_pe_empty(sp));
          if (++sd == n_dzs) { sd = 0; pd ^= 1; }
          if (++sp == 2) { sp = 0; pp ^= 1; }
        }
        mma_commit(bar_in_empty(si));
        if (++si == 2) { si = 0; pi ^= 1; }
      }
      mma_commit(bar_done);
#ifdef DFN_FUSED_STATS
      if (part == 0 && type == TYPE_BIG)
        printf("fused stats wgrad CTA %d job %d issuer (cycles): total %lld wait_in_full %lld wait_dz_full %lld\n", blockIdx.x, job, clock64() - i_total, i_in, i_dz);
#endif
    }
  } else if (active && warp >= 2 && warp < 2 + kWorkers) {
    // ================= workers =================
    const int ww = warp - 2, wt = threadIdx.x - 64;
    int si = 0, sd = 0, sp = 0; uint32_t pi = 0, pd = 0, pp = 0;
    float bias_acc[2][8];
    float extra[3][8];
    float scal[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 8; ++i) { bias_acc[0][i] = bias_acc[1][i] = 0.f; extra[0][i] = extra[1][i] = extra[2][i] = 0.f; }
    float* s_small = reinterpret_cast<float*>(smem + kWOffSmall);
    const bool need_small = type == TYPE_COLOR2;          // staged in shared memory for the weighted column sums
    const bool need_point = type == TYPE_PE || type == TYPE_COLOR;
    const bool job7 = type == TYPE_BIG && job == 7;       // reads the tile's d sigma_pre straight from global (L2): no room for staging
    const bool need_tile = need_small || need_point || job7;
    static_assert(kWorkThreads == 256, "two small-grad floats per worker");
    float gn[2] = {0.f, 0.f}, xn[3] = {0.f, 0.f, 0.f}, dn[3] = {0.f, 0.f, 0.f};
    auto fetch_tile = [&](int tile) {
      if (need_small) {
        const float* g = a.small + static_cast<size_t>(tile) * 512;
        gn[0] = __ldcg(g + wt); gn[1] = __ldcg(g + wt + kWorkThreads);   // written by another SM in this launch: L2, not L1
      }
      if (job7 && wt < 128) gn[0] = __ldcg(a.small + static_cast<size_t>(tile) * 512 + wt * 4 + 3);   // d sigma_pre of row wt
      if (need_point) {
        long long p = static_cast<long long>(tile) * kTile + (wt & 127);
        if (p >= a.P) p = a.P - 1;
        load_point(a.src, p, xn, dn);
      }
    };
    // the tile index of local ticket kc: written by the producer thread once the ticket's flag has been observed
    auto tile_of_ticket = [&](int kc) -> int {
      mbar_wait(bar_tile(kc & 7), (kc >> 3) & 1);
      return s_tiles[kc & 7];
    };
    bool have = false;   // inputs of the current ticket were fetched during the previous one

    int kc = 0;
    for (int k = part; k < nt; k += split, ++kc) {
      float* small = s_small + (kc & 1) * 512;
      const int cur_tile = need_tile ? tile_of_ticket(kc) : 0;
      if (need_tile && !have) fetch_tile(cur_tile);
      const float g0 = gn[0], g1 = gn[1];
      const float x[3] = {xn[0], xn[1], xn[2]}, d[3] = {dn[0], dn[1], dn[2]};
      have = false;
      if (need_tile && k + split < nt && mbar_try_wait(bar_tile((kc + 1) & 7), ((kc + 1) >> 3) & 1)) {
        fetch_tile(s_tiles[(kc + 1) & 7]);     // the next ticket is already there: hide its global loads behind this tile
        have = true;
      }
      if (need_small) {
        small[wt] = g0; small[wt + kWorkThreads] = g1;
        named_bar_sync(2, kWorkThreads);
        if (wt < 128) {
          const float4 v = reinterpret_cast<const float4*>(small)[wt];
          scal[0] += v.x; scal[1] += v.y; scal[2] += v.z;
        }
      }
      float* dsig = reinterpret_cast<float*>(smem + kWOffDsig) + (kc & 1) * 128;
      if (job7) {
        if (wt < 128) {
          dsig[wt] = g0;
          scal[3] += g0;
        }
        named_bar_sync(2, kWorkThreads);
      }
      if (type == TYPE_PE || type == TYPE_COLOR) {
        const int row = wt & 127, hh = wt >> 7;
        if (type == TYPE_PE) {
          mbar_wait(bar_in_empty(si), pi ^ 1);
          write_pe_row(smem + kWOffIn + si * kInBytes, nullptr, row, hh, x, d);
          fence_proxy_async();
          mbar_arrive(bar_in_full(si));
        } else {
          mbar_wait(bar_pe_empty(sp), pp ^ 1);
          write_pe_row(nullptr, smem + kWOffPe + sp * kPeBytes, row, hh, x, d);
          fence_proxy_async();
          mbar_arrive(bar_pe_full(sp));
          if (++sp == 2) { sp = 0; pp ^= 1; }
        }
      }
      unsigned* done = &a.ctl->done[q][static_cast<unsigned>(k) % ns];   // released by ONE worker thread once the slot's bytes are in shared memory
      if (type == TYPE_BIG) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          mbar_wait(bar_dz_full(sd), pd);
          wcolsum<16, 1>(smem + dz_off(sd), nullptr, reinterpret_cast<float(&)[1][8]>(bias_acc[h]), wt);
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_dz_empty(sd));
          if (++sd == n_dzs) { sd = 0; pd ^= 1; }
        }
        if (wt == 0) red_release_gpu_add(done, 1u);   // (after the stage has been handed back: the fence stalls only this thread)
        mbar_wait(bar_in_full(si), pi);
        if (job == 7) {  // density_layer.weight grad: sum_p d sigma_pre[p] * h8[p,:]
          float tmp[1][8];
#pragma unroll
          for (int i = 0; i < 8; ++i) tmp[0][i] = 0.f;
          wcolsum<32, 1, 1>(smem + kWOffIn + si * kInBytes, dsig, tmp, wt);
#pragma unroll
          for (int i = 0; i < 8; ++i) extra[0][i] += tmp[0][i];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_in_empty(si));
      } else if (type == TYPE_PE) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          mbar_wait(bar_dz_full(sd), pd);
          if (pe_layer == 0) wcolsum<16, 1>(smem + dz_off(sd), nullptr, reinterpret_cast<float(&)[1][8]>(bias_acc[i]), wt);
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_dz_empty(sd));
          if (++sd == n_dzs) { sd = 0; pd ^= 1; }
        }
        if (wt == 0) red_release_gpu_add(done, 1u);
      } else if (type == TYPE_COLOR) {
        mbar_wait(bar_dz_full(sd), pd);  // dZ_c0 -> color_layer.0.bias grad
        wcolsum<16, 1>(smem + dz_off(sd), nullptr, reinterpret_cast<float(&)[1][8]>(bias_acc[0]), wt);
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_dz_empty(sd));
        if (++sd == n_dzs) { sd = 0; pd ^= 1; }
        if (wt == 0) red_release_gpu_add(done, 1u);
        mbar_wait(bar_in_full(si), pi);
        if (lane == 0) mbar_arrive(bar_in_empty(si));
      } else {
        mbar_wait(bar_dz_full(sd), pd);  // c tile -> color_layer.2.weight grad (3 weighted column sums)
        wcolsum<16, 3>(smem + dz_off(sd), small, extra, wt);
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_dz_empty(sd));
        if (++sd == n_dzs) { sd = 0; pd ^= 1; }
      }
      if (++si == 2) { si = 0; pi ^= 1; }
    }

    // ---- CUDA-core partial sums -> global
    if (type == TYPE_BIG) {
      float* gb = a.grads + po.b[big_layer];
      flush_colsum<16>(bias_acc[0], gb, wt, 128);
      flush_colsum<16>(bias_acc[1], gb + 128, wt, 128);
      if (job == 7) {
        flush_colsum<32>(extra[0], a.grads + po.w[8], wt, 256);
        float v = wt < 128 ? scal[3] : 0.f;
        v = warp_sum(v);
        if (lane == 0 && ww < 4) atomicAdd(a.grads + po.b[8], v);
      }
    } else if (type == TYPE_PE) {
      if (pe_layer == 0) {
        flush_colsum<16>(bias_acc[0], a.grads + po.b[0], wt, 128);
        flush_colsum<16>(bias_acc[1], a.grads + po.b[0] + 128, wt, 128);
      }
    } else if (type == TYPE_COLOR) {
      flush_colsum<16>(bias_acc[0], a.grads + po.b[10], wt, 128);
    } else {
#pragma unroll
      for (int o = 0; o < 3; ++o) {
        flush_colsum<16>(extra[o], a.grads + po.w[11] + o * 128, wt, 128);
        float v = wt < 128 ? scal[o] : 0.f;
        v = warp_sum(v);
        if (lane == 0 && ww < 4) atomicAdd(a.grads + po.b[11] + o, v);
      }
    }
    // ---- TMEM accumulators -> global (fp32 atomics)
    mbar_wait(bar_done, 0);
    tc_fence_after();
    const int qd = warp & 3, hh = ww >> 2;
    const uint32_t lane_sel = static_cast<uint32_t>(qd * 32) << 16;
    const int out_lo = qd * 32 + lane;
    const int ncols = type == TYPE_BIG ? 512 : (type == TYPE_PE ? 128 : (type == TYPE_COLOR ? 288 : 0));
    for (int c0 = hh * 32; c0 < ncols; c0 += 64) {
      uint32_t v[32];
      tmem_ld32(tmem_base + lane_sel + c0, v);
      tmem_ld_wait();
      float* dst; int in0, in_valid;
      if (type == TYPE_BIG) {
        const int half = c0 >> 8;
        const int stride = big_layer == 4 ? 319 : 256;
        in0 = c0 & 255; in_valid = 256;
        dst = a.grads + po.w[big_layer] + static_cast<size_t>(half * 128 + out_lo) * stride;
      } else if (type == TYPE_PE) {
        const int half = c0 >> 6;
        in0 = c0 & 63; in_valid = 63;
        if (pe_layer == 0) dst = a.grads + po.w[0] + static_cast<size_t>(half * 128 + out_lo) * 63;
        else dst = a.grads + po.w[4] + static_cast<size_t>(half * 128 + out_lo) * 319 + 256;
      } else {
        if (c0 < 256) { in0 = c0; in_valid = 256; dst = a.grads + po.w[10] + static_cast<size_t>(out_lo) * 283; }
        else { in0 = c0 - 256; in_valid = 27; dst = a.grads + po.w[10] + static_cast<size_t>(out_lo) * 283 + 256; }
      }
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (in0 + i < in_valid) atomicAdd(dst + in0 + i, __uint_as_float(v[i]));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
mlp_bwd_fused_kernel(const FusedArgs a, const __grid_constant__ WeightMaps maps) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if (static_cast<int>(blockIdx.x) < 2 * a.n_pairs) dgrad_role(a, maps, smem);
  else wgrad_role(a, smem);
}

}  // namespace

// ---------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------
static int fused_slots() {
  static const int n = [] { const char* e = getenv("DFN_BWD_RING_SLOTS"); int v = e ? atoi(e) : 48; return v < 4 ? 4 : (v > kMaxSlots ? kMaxSlots : v); }();
  return n;
}

extern "C" long long dfn_mlp_bwd_fused_ws_bytes(long long P) {
  const long long nt = num_tiles(P);
  const long long ctl = (static_cast<long long>(sizeof(RingCtl)) + 255) / 256 * 256;
  return ctl + nt * 2048 + static_cast<long long>(kStreams) * fused_slots() * kSlotBytes;
}

extern "C" int dfn_mlp_backward_fused(const dfn_point_source* src, long long P, const void* bwd_pack, const float* consts,
                                      const float* rgb, const float* sigma, const float* d_rgb, const float* d_sigma,
                                      const void* stash, void* workspace, float* grads, void* stream) {
  DFN_CHECK_ARG(src && bwd_pack && consts && rgb && sigma && d_rgb && d_sigma && stash && workspace && grads, "mlp_backward_fused: null pointer");
  DFN_CHECK_ARG(P >= 0 && P < (1LL << 31) * kTile, "mlp_backward_fused: P out of range");
  if (P == 0) return DFN_OK;
  FusedArgs a{};
  static const int skew_ns = [] { const char* e = getenv("DFN_DGRAD_SKEW_NS"); return e ? atoi(e) : 7000; }();
  a.skew_ns = skew_ns;
  a.P = P; a.n_tiles = static_cast<int>(num_tiles(P));
  a.consts = consts; a.rgb = rgb; a.sigma = sigma; a.d_rgb = d_rgb; a.d_sigma = d_sigma;
  a.stash = static_cast<const uint8_t*>(stash);
  a.src = *src; a.grads = grads;
  const long long ctl_bytes = (static_cast<long long>(sizeof(RingCtl)) + 255) / 256 * 256;
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  a.ctl = reinterpret_cast<RingCtl*>(ws);
  a.small = reinterpret_cast<float*>(ws + ctl_bytes);
  a.ring = ws + ctl_bytes + static_cast<long long>(a.n_tiles) * 2048;
  a.n_slots = fused_slots();
  // SM split: dgrad pairs vs. weight-gradient units.  DFN_BWD_SPLIT="pairs,big,big7,pe,col,col2" overrides.
  const int sms = dfn_num_sms();
  int pairs = 38, big = 5, big7 = 8, pe = 6, col = 6, col2 = 7;
  static const char* split_env = getenv("DFN_BWD_SPLIT");
  if (split_env) {
    int p_ = 0, b_ = 0, b7_ = 0, e_ = 0, c_ = 0, c2_ = 0;
    if (sscanf(split_env, "%d,%d,%d,%d,%d,%d", &p_, &b_, &b7_, &e_, &c_, &c2_) == 6 && p_ > 0 && b_ > 0 && b7_ > 0 && e_ > 0 && c_ > 0 && c2_ > 0) {
      pairs = p_; big = b_; big7 = b7_; pe = e_; col = c_; col2 = c2_;
    }
  }
  if (pairs > a.n_tiles / 2) pairs = a.n_tiles / 2;
  int units = 7 * big + big7 + 2 * pe + col + col2;
  units += units & 1;                                   // whole clusters of two
  DFN_CHECK_ARG(2 * pairs + units <= sms, "mlp_backward_fused: split %d pairs + %d units does not fit %d SMs", pairs, units, sms);
  a.n_pairs = pairs; a.big_split = big; a.big7_split = big7; a.pe_split = pe; a.col_split = col; a.col2_split = col2;
  WeightMaps maps;
  a.replicas = weight_replicas();
  if (int rc = make_weight_maps(bwd_pack, static_cast<unsigned long long>(kBwdPackBytes) * a.replicas, &maps)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DFN_CUDA(cudaMemsetAsync(a.ctl, 0, sizeof(RingCtl), st), "mlp_backward_fused(memset)");
  DFN_CUDA(cudaFuncSetAttribute(mlp_bwd_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes), "mlp_backward_fused");
  mlp_bwd_fused_kernel<<<2 * pairs + units, kThreads, kSmemBytes, st>>>(a, maps);
  DFN_CHECK_LAUNCH("mlp_backward_fused");
  return DFN_OK;
}
