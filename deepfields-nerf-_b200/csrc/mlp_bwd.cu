// Backward of the fused vanilla-NeRF MLP, part 1: the data-gradient chain (autograd of reference
// scripts/model.py:139-187 w.r.t. the hidden activations).  Same pipeline as the forward kernel (mlp_pipe.cuh):
// persistent CTA pairs (tcgen05 cta_group::2, M = 256), two 256-point super-tiles in flight, W^T streamed through
// a 6 x 16 KB ring by tensor-map TMA loads (each SM holds half of every block).  Per 128-point tile:
//   prologue : dZ_c2 = d_rgb * rgb(1-rgb)            (sigmoid', 3 columns, fp32 -> global)
//              dZ_c0 = (dZ_c2 * color_layer.2.weight) masked by c > 0   (CUDA cores, K = 3)
//   b0       : d feat = dZ_c0 * color_layer.0.weight[:, :256]           (tcgen05, K = 128)
//   b1       : d h8   = dZ_f * feature_layer.weight + d sigma_pre * density_layer.weight, masked by h8 > 0
//   b2..b8   : d h7 .. d h1 through skip_layers / pos_layers (skip_layers.0 uses its first 256 columns)
// Every dZ is rounded to bf16 (what bf16 autocast autograd does), written back to shared memory as the next
// job's A operand and stored to the dz scratch (chunk layout) for the weight-gradient kernel.  ReLU masks come
// from the bit tiles the forward wrote beside its activation stash (32 B per point and layer instead of 512 B).
#include "mlp_pipe.cuh"
#include <stdlib.h>
#include <type_traits>

namespace mlp {
// dz scratch layout (per n_tiles): [8][nt][64K] dZ_1..dZ_8 | [nt][64K] dZ_f | [nt][32K] dZ_c0 | [nt][128][4] f32
// (dZ_c2[0..2], d sigma_pre)
__host__ __device__ inline size_t dz_off_h(int l, int nt, int tile) { return (static_cast<size_t>(l) * nt + tile) << 16; }
__host__ __device__ inline size_t dz_off_f(int nt, int tile) { return (static_cast<size_t>(8) * nt + tile) << 16; }
__host__ __device__ inline size_t dz_off_c0(int nt, int tile) { return (static_cast<size_t>(9) * nt << 16) + (static_cast<size_t>(tile) << 15); }
__host__ __device__ inline size_t dz_off_small(int nt, int tile) {
  return (static_cast<size_t>(9) * nt << 16) + (static_cast<size_t>(nt) << 15) + static_cast<size_t>(tile) * 2048;
}
}  // namespace mlp

namespace {
using namespace mlp;
using namespace umma;
using namespace mlp_pipe;

constexpr int kStages = 3;
constexpr uint32_t kBwdStageBytes = 32768;   // 16 K-chunks = 8 MMAs of this CTA's half of W^T per stage
#ifndef DFN_DGRAD_EPI_WARPS
#define DFN_DGRAD_EPI_WARPS 16
#endif
constexpr int kEpiWarps = DFN_DGRAD_EPI_WARPS;   // 4 TMEM lane quadrants x kColParts column parts
constexpr int kColParts = kEpiWarps / 4;
constexpr int kGroups = 8 / kColParts;           // 32-column groups per epilogue thread
constexpr int kProChunks = 16 / kColParts;       // dZ_c0 chunks per thread in the prologue
static_assert(kEpiWarps == 8 || kEpiWarps == 16, "epilogue warps");
constexpr int kEpiThreads = kEpiWarps * 32;
#ifndef DFN_DZ_STORE_PIECES
#define DFN_DZ_STORE_PIECES 1
#endif
// > 0: pause between the pieces instead of waiting for the previous piece's shared-memory read (as the forward does)
#ifndef DFN_DZ_STORE_GAP_NS
#define DFN_DZ_STORE_GAP_NS -1
#endif
constexpr int kStoreWarp = 2 + kEpiWarps;     // ships published dZ tiles to the dz scratch (bulk shared -> global)
constexpr int kThreads = 64 + kEpiThreads + 32;   // producer, issuer, 8 epilogue warps, dz store warp

// shared memory map (bytes), identical in both CTAs of a pair
constexpr uint32_t kOffAct = 0;                                  // 2 x 64 KB (dZ tiles = A operands)
constexpr uint32_t kOffRing = kOffAct + 2 * kActBytes;           // kStages x 16 KB
constexpr uint32_t kOffConst = kOffRing + kStages * kBwdStageBytes;
constexpr uint32_t kOffBars = kOffConst + kConstBytes;
constexpr uint32_t kNumBars = 4 * kStages + 6;
constexpr uint32_t kOffTmemPtr = kOffBars + kNumBars * 8;
constexpr uint32_t kSmemBytes = kOffTmemPtr + 16;
static_assert(kSmemBytes <= 232448, "shared memory budget exceeded");

// pipeline policy: 9 jobs, all N = 256; K = 128 (b0) or 256; no bias slices; A operand = the slot's dZ tile
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

struct DgradArgs {
  long long P;
  int n_tiles;           // even (num_tiles(P))
  const float* consts;
  const float* rgb;      // [P,3] forward output
  const float* sigma;    // [P]
  const float* d_rgb;    // [P,3]
  const float* d_sigma;  // [P]
  const uint8_t* stash;  // forward stash
  uint8_t* dz;
  int skew_ns;           // start of pair p is delayed by (p % 8) * skew_ns (see mlp_fwd.cu)
  int replicas;          // copies of the packed weight image; pair p streams copy p % replicas
};

__device__ __forceinline__ uint32_t relu_mask2(uint32_t h2) {
  // 0xFFFF per bf16 half that is > 0 (post-ReLU activations are >= +0)
  return ((h2 & 0x0000FFFFu) ? 0x0000FFFFu : 0u) | ((h2 & 0xFFFF0000u) ? 0xFFFF0000u : 0u);
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1) mlp_dgrad_kernel(const DgradArgs a, const __grid_constant__ WeightMaps maps) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bar0 = sbase + kOffBars;
  auto bar_accfull = [&](int t) { return bar0 + 8u * BwdPipe::bar_accfull_i(t); };
  auto bar_aready = [&](int t) { return bar0 + 8u * BwdPipe::bar_aready_i(t); };
  auto bar_stored = [&](int t) { return bar0 + 8u * (4 * kStages + 4 + t); };   // dZ tile of slot t may be overwritten
  float* consts = reinterpret_cast<float*>(smem + kOffConst);

  if (threadIdx.x == 0) {
    for (int t = 0; t < 2; ++t)
      for (int s = 0; s < kStages; ++s) {
        mbar_init(bar0 + 8u * BwdPipe::bar_full_i(t, s), 1);    // leader only: expect_tx of both CTAs' halves
        mbar_init(bar0 + 8u * BwdPipe::bar_empty_i(t, s), 1);
      }
    for (int t = 0; t < 2; ++t) { mbar_init(bar_accfull(t), 1); mbar_init(bar_aready(t), kEpiWarps + (rank == 0 ? 1 : 0)); mbar_init(bar_stored(t), 1); }   // aready: + the peer's relay
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
  const int n_pairs = gridDim.x >> 1;
  const int nt = a.n_tiles;
  const int n_it = ((nt >> 1) + 2 * n_pairs - 1) / (2 * n_pairs);

  if (warp == 0) {
    if (lane == 0) {  // ---- weight producer (this CTA's half of every W^T block)
      ProdState<BwdPipe> ps;
#pragma unroll
      for (int i = 0; i < kStages; ++i) { ps.use[0][i] = ps.use[1][i] = 0; ps.last[i] = -1; }
      ps.pol = make_evict_last_policy();
      ps.row0 = static_cast<uint32_t>((blockIdx.x >> 1) % a.replicas) * (kBwdPackBytes >> 7);
      // de-synchronise the pairs: all of them stream the same weight blocks in the same order, and in lockstep they all
      // hit the same L2 lines at the same moment
      for (int w = (blockIdx.x >> 1) & 7; a.skew_ns > 0 && w > 0; --w) __nanosleep(a.skew_ns);
      for (int it = 0; it < n_it; ++it) {
        const bool has[2] = {tile_of(it, 0, rank) < nt, tile_of(it, 1, rank) < nt};
        produce_all<BwdPipe>(maps, rank, sbase + kOffRing, bar0, ps, has, std::make_integer_sequence<int, kNumBwdJobs>{});
      }
    }
  } else if (warp == 1) {
    // ---- leader: MMA issuer (whole warp converged, one elected lane issues); peer: bookkeeping only
    IssueCtx c{};
    c.sbase = sbase; c.bar0 = bar0; c.tmem_base = tmem_base; c.lane = lane; c.rank = rank;
    c.s0 = SlotCtx{};
    c.s1 = c.s0;
#ifdef DFN_FWD_STATS
    c.w_full = c.w_aready = 0;
    const long long t_start = clock64();
#endif
    for (int it = 0; it < n_it; ++it) {
      c.s0.has = tile_of(it, 0, rank) < nt;
      c.s1.has = tile_of(it, 1, rank) < nt;
      issue_all<BwdPipe>(c, std::make_integer_sequence<int, kNumBwdJobs>{});
    }
#ifdef DFN_FWD_STATS
    if (blockIdx.x == 0 && lane == 0)
      printf("dgrad stats (leader of pair 0, cycles): issuer total %lld wait_aready %lld wait_weights %lld\n", clock64() - t_start,
             c.w_aready, c.w_full);
#endif
  } else if (warp == kStoreWarp) {
    // ---- dz store (one thread): the dZ tile an epilogue publishes for the next job is exactly the tile the weight-gradient
    // kernel reads later (same chunk layout), so it leaves the SM as one 64 KB bulk copy and the epilogue warps never stall
    // on the store path.  `stored` tells them the tile may be overwritten; one arrive per (job, slot) keeps the aready
    // phases this thread has to observe in lockstep with the epilogues.
    if (lane == 0) {
      uint32_t pa[2] = {0, 0};
      const uint64_t pol = make_evict_first_policy();   // dz is written once and read once, much later
      for (int it = 0; it < n_it; ++it)
        for (int b = 0; b < kNumBwdJobs; ++b)
          for (int t = 0; t < 2; ++t) {
            const int tile = tile_of(it, t, rank);
            if (tile >= nt) continue;
            mbar_wait(bar_aready(t), pa[t]);   // A operand of job b: dZ_c0 (b = 0, stored by the prologue itself) or job b-1's output
            pa[t] ^= 1;
            if (b >= 1) {
              uint8_t* dst = b == 1 ? a.dz + dz_off_f(nt, tile) : a.dz + dz_off_h(9 - b, nt, tile);
              // (the forward ships its tiles in 8 KB pieces so as not to hold up the weight loads queued behind them; with
              // this kernel's 96 KB ring one 64 KB copy measured faster: 1.01 ms vs 1.07 ms in 8 pieces)
              constexpr uint32_t kPiece = kActBytes / DFN_DZ_STORE_PIECES;
#pragma unroll 1
              for (int i = 0; i < DFN_DZ_STORE_PIECES; ++i) {
                bulk_s2g_hint(dst + i * kPiece, sbase + kOffAct + t * kActBytes + i * kPiece, kPiece, pol);
                bulk_commit();
                if (DFN_DZ_STORE_GAP_NS > 0 && i + 1 < DFN_DZ_STORE_PIECES) __nanosleep(DFN_DZ_STORE_GAP_NS);
                else bulk_wait_read_all();
              }
            }
            mbar_arrive(bar_stored(t));
          }
      bulk_wait_all();
    }
  } else {
    // ---- epilogue warps
    const int ew = warp - 2, q = warp & 3, hh = ew >> 2;
    const int row = q * 32 + lane;
    const uint32_t lane_sel = static_cast<uint32_t>(q * 32) << 16;
    uint32_t pf[2] = {0, 0}, pst[2] = {0, 0};
    const uint64_t pol = make_evict_first_policy();   // dz is written once and read once, much later
    // dZ tile written -> visible to the async proxy (the leader's MMA reads both CTAs' tiles) -> one arrive per warp
    auto publish = [&](int t) {
      tc_fence_before();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_aready(t));   // own CTA's barrier; the peer's warp 1 relays it to the leader
    };

    // prologue of a tile: dZ_c0 -> act chunks 0..15 (+ dz scratch), small fp32 grads
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
      if (hh == 0) {
        float4* sm = reinterpret_cast<float4*>(a.dz + dz_off_small(nt, tile)) + row;
        *sm = make_float4(g[0], g[1], g[2], ds);
      }
      const uint8_t* cst = a.stash + (static_cast<size_t>(9) * nt << 16) + (static_cast<size_t>(tile) << 15);
      uint8_t* dzc0 = a.dz + dz_off_c0(nt, tile);
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
        st_global_v4_hint(dzc0 + off, o4[0], o4[1], o4[2], o4[3], pol);
      }
      return ds;
    };
    // L2 prefetch of everything the prologue of `tile` will read (issued a few jobs ahead: its loads sit on the
    // critical path between two tiles of a slot)
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
    float ds0 = 0.f, ds1 = 0.f;          // d sigma_pre of the slot's current tile (job 1 adds it into d h8)
    uint32_t mk_next[kGroups];   // ReLU bit masks (one word per 32-column group) for the NEXT epilogue in program order
#pragma unroll
    for (int i = 0; i < kGroups; ++i) mk_next[i] = 0;
    const int ec0 = hh * (256 / kColParts);   // first column this thread handles
    auto load_mask = [&](const uint8_t* tile_masks) {
      const uint8_t* pm = tile_masks + ((ec0 >> 7) * 128 + row) * 16 + ((ec0 & 127) >> 5) * 4;
      if constexpr (kGroups == 4) {
        const uint4 v = *reinterpret_cast<const uint4*>(pm);
        mk_next[0] = v.x; mk_next[1] = v.y; mk_next[2] = v.z; mk_next[3] = v.w;
      } else {
        const uint2 v = *reinterpret_cast<const uint2*>(pm);
        mk_next[0] = v.x; mk_next[1] = v.y;
      }
    };

    for (int t = 0; t < 2; ++t) {
      const int tile = tile_of(0, t, rank);
      if (tile < nt) { const float ds = prologue(t, tile); if (t) ds1 = ds; else ds0 = ds; }
      publish(t);
    }

#ifdef DFN_FWD_STATS
    long long e_wait = 0, e_stored = 0;
    const long long e_start = clock64();
#endif
    for (int it = 0; it < n_it; ++it)
      for (int b = 0; b < kNumBwdJobs; ++b)
        for (int t = 0; t < 2; ++t) {
          const int tile = tile_of(it, t, rank);
          if (tile >= nt) continue;
          uint8_t* act = smem + kOffAct + t * kActBytes;
          const uint32_t tacc = tmem_base + t * 256 + lane_sel;
          const int c0 = ec0;
          // ReLU mask of h_{l+1} (bit tile written by the forward, 16 B per thread): loaded during the previous epilogue;
          // now fetch the one the next epilogue in program order needs - (b, slot 1) or (b+1, slot 0)
          const int l = 8 - b;  // b >= 1
          uint32_t mk[kGroups];
#pragma unroll
          for (int i = 0; i < kGroups; ++i) mk[i] = mk_next[i];
          {
            const int tile1 = tile_of(it, 1, rank);
            const bool other = t == 0 && tile1 < nt;
            const int nb = other ? b : b + 1;
            const int ntl = other ? tile1 : tile_of(it, 0, rank);
            if (nb >= 1 && nb < kNumBwdJobs)
              load_mask(a.stash + stash_off_mask(nt) + ((static_cast<size_t>(8 - nb) * nt + ntl) << 12));
          }
          uint8_t* dst = b == 0 ? a.dz + dz_off_f(nt, tile) : a.dz + dz_off_h(l, nt, tile);
          const float dsig = t ? ds1 : ds0;
          if (b == 4) {
            const int ntile = tile_of(it + 1, t, rank);
            if (ntile < nt) prefetch_prologue(ntile);
          }
#ifdef DFN_FWD_STATS
          const long long tw0 = clock64();
#endif
          mbar_wait(bar_accfull(t), pf[t]); pf[t] ^= 1;
#ifdef DFN_FWD_STATS
          e_wait += clock64() - tw0;
#endif
          tc_fence_after();
#ifdef DFN_FWD_STATS
          const long long ts0 = clock64();
#endif
          mbar_wait(bar_stored(t), pst[t]); pst[t] ^= 1;   // the store warp has read the previous dZ tile out of act
#ifdef DFN_FWD_STATS
          e_stored += clock64() - ts0;
#endif
          // drain the accumulator: one instantiation per kind of job so that nothing job-dependent is tested per group
          //   0: b = 0 (no mask)   1: b = 1 (+ d sigma_pre * density weights, mask)   2: b = 2..7 (mask)   3: b = 8 (mask, to global)
          auto drain = [&](auto kind_c) {
            constexpr int KIND = decltype(kind_c)::value;
            uint8_t* out_row = (KIND == 3 ? dst : act) + static_cast<uint32_t>(c0 / 8) * kChunkBytes + row * 16;
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
                uint8_t* po = out_row + (cc * 4 + g4) * kChunkBytes;
                if constexpr (KIND < 3) *reinterpret_cast<uint4*>(po) = o;   // next job's A operand; shipped to dz by the store warp
                else st_global_v4_hint(po, o.x, o.y, o.z, o.w, pol);         // last job: nobody reads it on chip
              }
              if (cc < kGroups - 1) tmem_ld_wait();
            }
          };
          if (b == 0) drain(std::integral_constant<int, 0>{});
          else if (b == 1) drain(std::integral_constant<int, 1>{});
          else if (b < kNumBwdJobs - 1) drain(std::integral_constant<int, 2>{});
          else drain(std::integral_constant<int, 3>{});
          if (b == kNumBwdJobs - 1) {
            const int ntile = tile_of(it + 1, t, rank);
            if (ntile < nt) { const float ds = prologue(t, ntile); if (t) ds1 = ds; else ds0 = ds; publish(t); }
          } else {
            publish(t);
          }
        }
#ifdef DFN_FWD_STATS
    if (blockIdx.x < 2 && ew == 0 && lane == 0)
      printf("dgrad stats (CTA %d epilogue warp 0, cycles): total %lld wait_accfull %lld wait_stored %lld\n", blockIdx.x, clock64() - e_start, e_wait, e_stored);
#endif
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // the leader's MMAs read the peer's shared memory: nobody leaves before both are done
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
}

}  // namespace

int dfn_launch_wgrad(const dfn_point_source* src, long long P, int n_tiles, const uint8_t* stash, const uint8_t* dz,
                     float* grads, cudaStream_t st);

extern "C" long long dfn_mlp_dz_bytes(long long P) {
  const long long nt = num_tiles(P);
  return nt * (9LL * 65536 + 32768 + 2048);
}

extern "C" int dfn_mlp_backward_dgrad(long long P, const void* bwd_pack, const float* consts, const float* rgb,
                                     const float* sigma, const float* d_rgb, const float* d_sigma, const void* stash,
                                     void* dz, void* stream) {
  DFN_CHECK_ARG(bwd_pack && consts && rgb && sigma && d_rgb && d_sigma && stash && dz, "mlp_backward_dgrad: null pointer");
  DFN_CHECK_ARG(P >= 0 && P < (1LL << 31) * kTile, "mlp_backward_dgrad: P out of range");
  if (P == 0) return DFN_OK;
  DgradArgs a{};
  static const int skew_ns = [] { const char* e = getenv("DFN_DGRAD_SKEW_NS"); return e ? atoi(e) : 7000; }();   // measured: 0.984 -> 0.955 ms (fine), 0.358 -> 0.343 (coarse)
  a.skew_ns = skew_ns;
  a.P = P; a.n_tiles = static_cast<int>(num_tiles(P));
  a.consts = consts; a.rgb = rgb; a.sigma = sigma;
  a.d_rgb = d_rgb; a.d_sigma = d_sigma; a.stash = static_cast<const uint8_t*>(stash); a.dz = static_cast<uint8_t*>(dz);
  WeightMaps maps;
  a.replicas = weight_replicas();
  if (int rc = make_weight_maps(bwd_pack, static_cast<unsigned long long>(kBwdPackBytes) * a.replicas, &maps)) return rc;
  int pairs = a.n_tiles / 2 < dfn_mlp_sms() / 2 ? a.n_tiles / 2 : dfn_mlp_sms() / 2;
  static const int pairs_env = [] { const char* e = getenv("DFN_DGRAD_PAIRS"); return e ? atoi(e) : 0; }();   // bring-up: per-pair speed on fewer SMs
  if (pairs_env > 0 && pairs_env < pairs) pairs = pairs_env;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DFN_CUDA(cudaFuncSetAttribute(mlp_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes), "mlp_backward_dgrad");
  mlp_dgrad_kernel<<<2 * pairs, kThreads, kSmemBytes, st>>>(a, maps);
  DFN_CHECK_LAUNCH("mlp_backward_dgrad");
  return DFN_OK;
}

extern "C" int dfn_mlp_backward_wgrad(const dfn_point_source* src, long long P, const void* stash, const void* dz,
                                     float* grads, void* stream) {
  DFN_CHECK_ARG(src && stash && dz && grads, "mlp_backward_wgrad: null pointer");
  DFN_CHECK_ARG(P >= 0 && P < (1LL << 31) * kTile, "mlp_backward_wgrad: P out of range");
  if (P == 0) return DFN_OK;
  return dfn_launch_wgrad(src, P, static_cast<int>(num_tiles(P)), static_cast<const uint8_t*>(stash),
                          static_cast<const uint8_t*>(dz), grads, static_cast<cudaStream_t>(stream));
}

extern "C" int dfn_mlp_backward(const dfn_point_source* src, long long P, const void* bwd_pack, const float* consts,
                                const float* rgb, const float* sigma, const float* d_rgb, const float* d_sigma,
                                const void* stash, void* dz, float* grads, void* stream) {
  DFN_CHECK_ARG(src && grads, "mlp_backward: null pointer");
  if (int rc = dfn_mlp_backward_dgrad(P, bwd_pack, consts, rgb, sigma, d_rgb, d_sigma, stash, dz, stream)) return rc;
  return dfn_mlp_backward_wgrad(src, P, stash, dz, grads, stream);
}
