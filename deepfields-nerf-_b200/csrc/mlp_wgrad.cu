// Backward of the fused vanilla-NeRF MLP, part 2: weight / bias gradients.
//   dW_l[out,in] = sum_p dZ_l[p,out] * A_l[p,in]   (A_l = layer input: stashed activation or recomputed PE)
// Both operands are the [128 points x features] bf16 tiles the forward / dgrad kernels wrote in the chunk
// layout; read as MN-major tcgen05 operands the reduction runs over the points, so no transposition is
// ever materialised.  Each CTA owns one (job, tile-range) unit, accumulates its dW block in TMEM (fp32) over
// ALL its tiles and flushes once with atomics:
//   big jobs (8): a 256x256 block = two M=128 halves x N=256 = all 512 TMEM columns
//   PE job      : pos_layers.0.weight [256x63] and the PE columns of skip_layers.0.weight (N=64, PE recomputed)
//   colour job  : color_layer.0.weight [128 x (256 | 27)]
// Bias gradients (column sums of dZ), density_layer (N=1) and color_layer.2 (N=3) gradients are weighted
// column sums of tiles that are already in shared memory, done by the worker warps on the CUDA cores.
// Operand tiles stream through shared memory rings with 1-D bulk async copies (TMA engine).
#include "mlp_common.cuh"
#include <stdio.h>
#include <stdlib.h>

namespace mlp {
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

constexpr int kWorkers = 8;
constexpr int kWorkThreads = kWorkers * 32;
constexpr int kThreads = 64 + kWorkThreads;
constexpr uint32_t kInBytes = 65536, kDzBytes = 32768, kPeBytes = 8192;
constexpr uint32_t kOffIn = 0;                         // 2 x 64 KB
constexpr uint32_t kOffDz = kOffIn + 2 * kInBytes;     // 2 x 32 KB
constexpr uint32_t kOffPe = kOffDz + 2 * kDzBytes;     // 2 x 8 KB (dir-PE tiles, colour job)
constexpr uint32_t kOffSmall = kOffPe + 2 * kPeBytes;  // 2 x [128][4] f32
constexpr uint32_t kOffBars = kOffSmall + 2 * 2048;
constexpr uint32_t kOffTmemPtr = kOffBars + 20 * 8;
constexpr uint32_t kSmemBytes = kOffTmemPtr + 16;
static_assert(kSmemBytes <= 232448, "shared memory budget exceeded");

enum { TYPE_BIG = 0, TYPE_PE = 1, TYPE_COLOR = 2 };

struct WgradArgs {
  dfn_point_source src;
  long long P;
  int n_tiles;
  int big_split, pe_split, col_split;
  const uint8_t* stash;
  const uint8_t* dz;
  float* grads;
};

template <int NCH, int NW>
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
      const float wv = wrow ? wrow[r * 4 + w] : 1.0f;
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[w][i] += wv * x[i];
    }
  }
}

// reduce over the TPC consecutive lanes that share a chunk, then one lane adds 8 columns to global
template <int NCH>
__device__ __forceinline__ void flush_colsum(float (&acc)[8], float* dst /* column 0 of the chunk-major vector */, int wt,
                                             int ncols_valid) {
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

__global__ void __launch_bounds__(kThreads, 1) mlp_wgrad_kernel(const WgradArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bar0 = sbase + kOffBars;
  auto bar_in_full = [&](int s) { return bar0 + 8u * s; };
  auto bar_in_empty = [&](int s) { return bar0 + 8u * (2 + s); };
  auto bar_dz_full = [&](int s) { return bar0 + 8u * (s < 2 ? 4 + s : 11 + s); };    // stages 2, 3: barriers 13, 14
  auto bar_dz_empty = [&](int s) { return bar0 + 8u * (s < 2 ? 6 + s : 13 + s); };   // stages 2, 3: barriers 15, 16
  auto bar_pe_full = [&](int s) { return bar0 + 8u * (8 + s); };
  auto bar_pe_empty = [&](int s) { return bar0 + 8u * (10 + s); };
  const uint32_t bar_done = bar0 + 8u * 12;
  constexpr ParamOffsets po = param_offsets();
  const int nt = a.n_tiles;

  // ---- which unit is this CTA? -------------------------------------------------------------------
  int type, job = 0, part, split;
  {
    const int u = blockIdx.x;
    if (u < 8 * a.big_split) { type = TYPE_BIG; job = u / a.big_split; part = u % a.big_split; split = a.big_split; }
    else if (u < 8 * a.big_split + a.pe_split) { type = TYPE_PE; part = u - 8 * a.big_split; split = a.pe_split; }
    else { type = TYPE_COLOR; part = u - 8 * a.big_split - a.pe_split; split = a.col_split; }
  }
  // tiles part, part + split, ...: all units walk the tile list front to back at about the same pace, so that the dZ5
  // tiles the layer-4 unit and the PE unit both need are still in L2 for whichever of the two comes second
  const int tile_begin = part, tile_end = nt, tile_step = split;
  // big job table: layer, input stash index, dz source
  const int big_layer = job < 3 ? job + 1 : (job == 3 ? 4 : (job < 7 ? job + 1 : 9));
  const int big_in = job < 7 ? big_layer - 1 : 7;  // stash_h index of the input activation

  if (threadIdx.x == 0) {
    for (int s = 2; s < 4; ++s) { mbar_init(bar_dz_full(s), 1); mbar_init(bar_dz_empty(s), 1 + kWorkers); }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_in_full(s), type == TYPE_PE ? kWorkThreads : 1);
      mbar_init(bar_in_empty(s), type == TYPE_PE ? 1 : 1 + kWorkers);
      mbar_init(bar_dz_full(s), 1);
      mbar_init(bar_dz_empty(s), 1 + kWorkers);
      mbar_init(bar_pe_full(s), kWorkThreads);
      mbar_init(bar_pe_empty(s), 1);
    }
    mbar_init(bar_done, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(sbase + kOffTmemPtr, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem + kOffTmemPtr);

  auto in_src = [&](int tile) -> const uint8_t* {
    if (type == TYPE_BIG) return a.stash + ((static_cast<size_t>(big_in) * nt + tile) << 16);
    return a.stash + ((static_cast<size_t>(8) * nt + tile) << 16);  // colour job: features
  };
  // i-th dz-ring load of a tile
  auto dz_src = [&](int tile, int i) -> const uint8_t* {
    if (type == TYPE_BIG) {
      const size_t base = job == 7 ? dz_off_f(nt, tile) : dz_off_h(big_layer, nt, tile);
      return a.dz + base + static_cast<size_t>(i) * kDzBytes;
    }
    if (type == TYPE_PE) return a.dz + dz_off_h(i < 2 ? 0 : 4, nt, tile) + static_cast<size_t>(i & 1) * kDzBytes;
    return i == 0 ? a.dz + dz_off_c0(nt, tile)
                  : a.stash + (static_cast<size_t>(9) * nt << 16) + (static_cast<size_t>(tile) << 15);  // c tile
  };
  const int n_dz = type == TYPE_PE ? 4 : 2;
  // dz ring: two 32 KB stages; the PE unit's input tiles are computed, not loaded, and fill only 16 KB of each 64 KB
  // input slot, so it parks two more dz stages there and keeps a whole tile (128 KB) of loads in flight
  const int n_dzs = type == TYPE_PE ? 4 : 2;
  auto dz_off = [&](int s) -> uint32_t { return s < 2 ? kOffDz + s * kDzBytes : kOffIn + (s - 2) * kInBytes + 16384u; };

  if (warp == 0) {
    if (lane == 0) {  // ================= producer =================
      int si = 0, sd = 0; uint32_t pi = 0, pd = 0;
      const uint64_t pol = make_evict_first_policy();   // stash and dz tiles are read exactly once here
      for (int tile = tile_begin; tile < tile_end; tile += tile_step) {
        if (type != TYPE_PE) {
          mbar_wait(bar_in_empty(si), pi ^ 1);
          mbar_arrive_expect_tx(bar_in_full(si), kInBytes);
          bulk_g2s_hint(sbase + kOffIn + si * kInBytes, in_src(tile), 32768, bar_in_full(si), pol);
          bulk_g2s_hint(sbase + kOffIn + si * kInBytes + 32768, in_src(tile) + 32768, 32768, bar_in_full(si), pol);
          if (++si == 2) { si = 0; pi ^= 1; }
        }
        for (int i = 0; i < n_dz; ++i) {
          mbar_wait(bar_dz_empty(sd), pd ^ 1);
          mbar_arrive_expect_tx(bar_dz_full(sd), kDzBytes);
          const bool shared_dz = (type == TYPE_BIG && big_layer == 4) || (type == TYPE_PE && i >= 2);   // dZ5: read by two units
          if (shared_dz) bulk_g2s(sbase + dz_off(sd), dz_src(tile, i), kDzBytes, bar_dz_full(sd));
          else bulk_g2s_hint(sbase + dz_off(sd), dz_src(tile, i), kDzBytes, bar_dz_full(sd), pol);
          if (++sd == n_dzs) { sd = 0; pd ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {  // ================= MMA issuer =================
      int si = 0, sd = 0, sp = 0; uint32_t pi = 0, pd = 0, pp = 0;
      auto mma_block = [&](uint32_t a_addr, uint32_t b_addr, uint32_t tcol, int N, bool first_tile) {
        const uint32_t idesc = instr_desc_bf16(128, N, 1, 1);
#pragma unroll
        for (int k = 0; k < 8; ++k)  // 8 x 16 points
          mma_bf16(tmem_base + tcol, smem_desc(a_addr + k * 256, 128, kChunkBytes), smem_desc(b_addr + k * 256, 128, kChunkBytes),
                   idesc, (first_tile && k == 0) ? 0u : 1u);
      };
      for (int tile = tile_begin; tile < tile_end; tile += tile_step) {
        const bool first = tile == tile_begin;
        const uint32_t in_addr = sbase + kOffIn + si * kInBytes;
        mbar_wait(bar_in_full(si), pi);
        tc_fence_after();
        if (type == TYPE_BIG) {
          for (int h = 0; h < 2; ++h) {
            mbar_wait(bar_dz_full(sd), pd);
            tc_fence_after();
            mma_block(sbase + dz_off(sd), in_addr, h * 256, 256, first);
            mma_commit(bar_dz_empty(sd));
            if (++sd == n_dzs) { sd = 0; pd ^= 1; }
          }
        } else if (type == TYPE_PE) {
          for (int i = 0; i < 4; ++i) {
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
          mma_block(sbase + dz_off(sd), sbase + kOffPe + sp * kPeBytes, 256, 32, first);
          mma_commit(bar_dz_empty(sd));
          mma_commit(bar_pe_empty(sp));
          if (++sd == n_dzs) { sd = 0; pd ^= 1; }
          if (++sp == 2) { sp = 0; pp ^= 1; }
          // the c tile is consumed by the workers only: account for the MMA side of its empty barrier
          mbar_wait(bar_dz_full(sd), pd);
          mbar_arrive(bar_dz_empty(sd));
          if (++sd == n_dzs) { sd = 0; pd ^= 1; }
        }
        mma_commit(bar_in_empty(si));
        if (++si == 2) { si = 0; pi ^= 1; }
      }
      mma_commit(bar_done);
    }
  } else {
    // ================= workers =================
    const int ww = warp - 2, wt = threadIdx.x - 64;
    int si = 0, sd = 0, sp = 0; uint32_t pi = 0, pd = 0, pp = 0;
    float bias_acc[2][8];   // BIG: column sums of the two dZ halves; PE: layer-0 bias; COLOR: c0 bias in [0]
    float extra[3][8];      // BIG job 7: density weight grad (chunk-major over h8) in [0]; COLOR: color_layer.2 grads
    float scal[4] = {0.f, 0.f, 0.f, 0.f};  // sums of (dZ_c2[0..2], d sigma_pre)
#pragma unroll
    for (int i = 0; i < 8; ++i) { bias_acc[0][i] = bias_acc[1][i] = 0.f; extra[0][i] = extra[1][i] = extra[2][i] = 0.f; }
    float* s_small = reinterpret_cast<float*>(smem + kOffSmall);
    const bool need_small = (type == TYPE_BIG && job == 7) || type == TYPE_COLOR;

    // global inputs of a tile (the small fp32 grads, the points whose encodings are recomputed) are fetched one tile
    // ahead: their latency is otherwise paid once per tile, in series with the encoding and the column sums
    static_assert(kWorkThreads == 256, "two small-grad floats per worker");
    const bool need_point = type == TYPE_PE || type == TYPE_COLOR;
    float gn[2] = {0.f, 0.f}, xn[3] = {0.f, 0.f, 0.f}, dn[3] = {0.f, 0.f, 0.f};
    auto fetch_tile = [&](int tile) {
      if (need_small) {
        const float* g = reinterpret_cast<const float*>(a.dz + dz_off_small(nt, tile));
        gn[0] = g[wt]; gn[1] = g[wt + kWorkThreads];
      }
      if (need_point) {
        long long p = static_cast<long long>(tile) * kTile + (wt & 127);
        if (p >= a.P) p = a.P - 1;
        load_point(a.src, p, xn, dn);
      }
    };
    if (tile_begin < tile_end) fetch_tile(tile_begin);

    for (int tile = tile_begin, itn = 0; tile < tile_end; tile += tile_step, ++itn) {
      float* small = s_small + (itn & 1) * 512;
      const float g0 = gn[0], g1 = gn[1];
      const float x[3] = {xn[0], xn[1], xn[2]}, d[3] = {dn[0], dn[1], dn[2]};
      if (tile + tile_step < tile_end) fetch_tile(tile + tile_step);
      if (need_small) {
        small[wt] = g0; small[wt + kWorkThreads] = g1;
        named_bar_sync(2, kWorkThreads);
        if (wt < 128) {
          const float4 v = reinterpret_cast<const float4*>(small)[wt];
          if (type == TYPE_COLOR) { scal[0] += v.x; scal[1] += v.y; scal[2] += v.z; } else { scal[3] += v.w; }
        }
      }
      if (type == TYPE_PE || type == TYPE_COLOR) {
        // recompute the encodings of this tile's points (2 threads per row)
        const int row = wt & 127, hh = wt >> 7;
        if (type == TYPE_PE) {
          mbar_wait(bar_in_empty(si), pi ^ 1);
          write_pe_row(smem + kOffIn + si * kInBytes, nullptr, row, hh, x, d);
          fence_proxy_async();
          mbar_arrive(bar_in_full(si));
        } else {
          mbar_wait(bar_pe_empty(sp), pp ^ 1);
          write_pe_row(nullptr, smem + kOffPe + sp * kPeBytes, row, hh, x, d);
          fence_proxy_async();
          mbar_arrive(bar_pe_full(sp));
          if (++sp == 2) { sp = 0; pp ^= 1; }
        }
      }
      if (type == TYPE_BIG) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          mbar_wait(bar_dz_full(sd), pd);
          wcolsum<16, 1>(smem + dz_off(sd), nullptr, reinterpret_cast<float(&)[1][8]>(bias_acc[h]), wt);
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_dz_empty(sd));
          if (++sd == n_dzs) { sd = 0; pd ^= 1; }
        }
        mbar_wait(bar_in_full(si), pi);
        if (job == 7) {  // density_layer.weight grad: sum_p d sigma_pre[p] * h8[p,:]
          float tmp[1][8];
#pragma unroll
          for (int i = 0; i < 8; ++i) tmp[0][i] = 0.f;
          wcolsum<32, 1>(smem + kOffIn + si * kInBytes, small + 3, tmp, wt);
#pragma unroll
          for (int i = 0; i < 8; ++i) extra[0][i] += tmp[0][i];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_in_empty(si));
      } else if (type == TYPE_PE) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          mbar_wait(bar_dz_full(sd), pd);
          if (i < 2) wcolsum<16, 1>(smem + dz_off(sd), nullptr, reinterpret_cast<float(&)[1][8]>(bias_acc[i]), wt);
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_dz_empty(sd));
          if (++sd == n_dzs) { sd = 0; pd ^= 1; }
        }
      } else {
        mbar_wait(bar_dz_full(sd), pd);  // dZ_c0 -> color_layer.0.bias grad
        wcolsum<16, 1>(smem + dz_off(sd), nullptr, reinterpret_cast<float(&)[1][8]>(bias_acc[0]), wt);
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_dz_empty(sd));
        if (++sd == n_dzs) { sd = 0; pd ^= 1; }
        mbar_wait(bar_dz_full(sd), pd);  // c tile -> color_layer.2.weight grad (3 weighted column sums)
        wcolsum<16, 3>(smem + dz_off(sd), small, extra, wt);
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_dz_empty(sd));
        if (++sd == n_dzs) { sd = 0; pd ^= 1; }
        mbar_wait(bar_in_full(si), pi);
        if (lane == 0) mbar_arrive(bar_in_empty(si));
      }
      if (++si == 2) { si = 0; pi ^= 1; }
    }

    if (tile_end > tile_begin) {
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
        flush_colsum<16>(bias_acc[0], a.grads + po.b[0], wt, 128);
        flush_colsum<16>(bias_acc[1], a.grads + po.b[0] + 128, wt, 128);
      } else {
        flush_colsum<16>(bias_acc[0], a.grads + po.b[10], wt, 128);
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
      const int q = warp & 3, hh = ww >> 2;
      const uint32_t lane_sel = static_cast<uint32_t>(q * 32) << 16;
      const int out_lo = q * 32 + lane;
      const int ncols = type == TYPE_BIG ? 512 : (type == TYPE_PE ? 256 : 288);
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
          const int region = c0 >> 6, half = region & 1;
          in0 = c0 & 63; in_valid = 63;
          if (region < 2) dst = a.grads + po.w[0] + static_cast<size_t>(half * 128 + out_lo) * 63;
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
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

}  // namespace

int dfn_launch_wgrad(const dfn_point_source* src, long long P, int n_tiles, const uint8_t* stash, const uint8_t* dz,
                     float* grads, cudaStream_t st) {
  WgradArgs a{};
  a.src = *src; a.P = P; a.n_tiles = n_tiles; a.stash = stash; a.dz = dz; a.grads = grads;
  const int sms = dfn_mlp_sms();
  // 8 big jobs + PE + colour share the SMs in proportion to their (HBM-bound, roughly equal) per-tile cost
  int per = sms / 10;
  if (per < 1) per = 1;
  if (per > n_tiles) per = n_tiles;
  a.big_split = per;
  a.pe_split = per;
  a.col_split = per;
  int rest = sms - 10 * per;
  if (per < n_tiles && rest > 0) { a.big_split = per; a.pe_split = per + rest / 2; a.col_split = per + (rest - rest / 2); }
  if (a.pe_split > n_tiles) a.pe_split = n_tiles;
  if (a.col_split > n_tiles) a.col_split = n_tiles;
  // bring-up override of the split: DFN_WGRAD_SPLIT="big,pe,col"
  static const char* split_env = getenv("DFN_WGRAD_SPLIT");
  if (split_env) {
    int b = 0, p = 0, c = 0;
    if (sscanf(split_env, "%d,%d,%d", &b, &p, &c) == 3 && b > 0 && p > 0 && c > 0 && b <= n_tiles && p <= n_tiles && c <= n_tiles) {
      a.big_split = b; a.pe_split = p; a.col_split = c;
    }
  }
  const int grid = 8 * a.big_split + a.pe_split + a.col_split;
  DFN_CUDA(cudaFuncSetAttribute(mlp_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes), "mlp_backward");
  mlp_wgrad_kernel<<<grid, kThreads, kSmemBytes, st>>>(a);
  DFN_CHECK_LAUNCH("mlp_backward(wgrad)");
  return DFN_OK;
}
