// Fused vanilla-NeRF MLP forward (reference scripts/model.py:139-187) for sm_100a.
//
// Persistent kernel, one CTA per SM, two 128-point tiles (slots X/Y) in flight per CTA:
//   warp 0      : weight producer - streams packed bf16 weight blocks (16 KB) from L2 into a shared
//                 memory ring with 1-D bulk async copies (TMA engine, mbarrier complete_tx)
//   warp 1      : MMA issuer - one thread issues tcgen05.mma (M=128, N=256/128, K=16, bf16 -> fp32 in
//                 TMEM); slot X accumulates in TMEM columns [0,256), slot Y in [256,512)
//   warps 2..9  : epilogue - tcgen05.ld the accumulator, bias + ReLU, round to bf16, write the next
//                 layer's A operand back to shared memory in the UMMA chunk layout (activations never
//                 leave the SM); positional encodings are computed here too.
// While the tensor core runs layer L of slot X the epilogue warps drain layer L-1 of slot Y.
// The sigma head (N=1) and color_layer.2 (N=3) are CUDA-core dot products inside the epilogues.
// TRAIN=true additionally stashes h1..h8, features and the colour hidden layer (bf16, chunk layout)
// for the backward kernels.
#include "mlp_common.cuh"

namespace {
using namespace mlp;
using namespace umma;

constexpr int kStages = 3;
constexpr int kEpiWarps = 8;
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kThreads = 64 + kEpiThreads;

// shared memory map (bytes)
constexpr uint32_t kOffAct = 0;                                  // 2 x 64 KB
constexpr uint32_t kOffAux = kOffAct + 2 * kActBytes;            // 2 x 16 KB
constexpr uint32_t kOffRing = kOffAux + 2 * kAuxBytes;           // kStages x 16 KB
constexpr uint32_t kOffConst = kOffRing + kStages * kStageBytes;
// partial head dot products of the upper column half, [2 slots][128 rows][3] f32 (sigma uses [..][0];
// reuse across j7 / j9 of a slot is ordered by the other slot's named barrier in between)
constexpr uint32_t kOffPart = kOffConst + kConstBytes;
constexpr uint32_t kOffBars = kOffPart + 2 * 128 * 3 * 4;
constexpr uint32_t kNumBars = 2 * kStages + 4;
constexpr uint32_t kOffTmemPtr = kOffBars + kNumBars * 8;
constexpr uint32_t kSmemBytes = kOffTmemPtr + 16;
static_assert(kSmemBytes <= 232448, "shared memory budget exceeded");

struct FwdArgs {
  PointSource src;
  long long P;
  int n_tiles;
  const uint8_t* wpack;   // packed forward weights
  const float* consts;    // fp32 consts block
  const float* noise;     // [P] sigma noise or null (model.py:160-162)
  float* rgb;             // [P,3]
  float* sigma;           // [P]
  uint8_t* stash_h;       // TRAIN: [8][n_tiles][64 KB]
  uint8_t* stash_f;       // TRAIN: [n_tiles][64 KB]
  uint8_t* stash_c;       // TRAIN: [n_tiles][32 KB]
};

__device__ __forceinline__ int tile_of(int it, int t) { return (it * 2 + t) * gridDim.x + blockIdx.x; }

// ---- epilogue helpers ------------------------------------------------------------------------------
// 32 accumulator columns -> bias, optional ReLU, bf16 -> 4 x 16 B into the act tile (+ stash)
template <bool RELU>
__device__ __forceinline__ void emit32(const uint32_t (&v)[32], const float* bias, uint8_t* act, uint8_t* stash,
                                       int col0, int row, uint32_t (&packed)[16]) {
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    const float4 b0 = *reinterpret_cast<const float4*>(bias + col0 + g * 8);
    const float4 b1 = *reinterpret_cast<const float4*>(bias + col0 + g * 8 + 4);
    const float x0 = __uint_as_float(v[g * 8 + 0]) + b0.x, x1 = __uint_as_float(v[g * 8 + 1]) + b0.y;
    const float x2 = __uint_as_float(v[g * 8 + 2]) + b0.z, x3 = __uint_as_float(v[g * 8 + 3]) + b0.w;
    const float x4 = __uint_as_float(v[g * 8 + 4]) + b1.x, x5 = __uint_as_float(v[g * 8 + 5]) + b1.y;
    const float x6 = __uint_as_float(v[g * 8 + 6]) + b1.z, x7 = __uint_as_float(v[g * 8 + 7]) + b1.w;
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
    if (stash) *reinterpret_cast<uint4*>(stash + off) = q;
  }
}

template <bool TRAIN>
__global__ void __launch_bounds__(kThreads, 1) mlp_fwd_kernel(const FwdArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bar0 = sbase + kOffBars;
  auto bar_wfull = [&](int s) { return bar0 + 8u * s; };
  auto bar_wempty = [&](int s) { return bar0 + 8u * (kStages + s); };
  auto bar_accfull = [&](int t) { return bar0 + 8u * (2 * kStages + t); };
  auto bar_aready = [&](int t) { return bar0 + 8u * (2 * kStages + 2 + t); };
  float* consts = reinterpret_cast<float*>(smem + kOffConst);

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(bar_wfull(s), 1); mbar_init(bar_wempty(s), 1); }
    for (int t = 0; t < 2; ++t) { mbar_init(bar_accfull(t), 1); mbar_init(bar_aready(t), kEpiThreads); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(sbase + kOffTmemPtr, 512);
  for (int i = threadIdx.x; i < kConstFloats; i += kThreads) consts[i] = a.consts[i];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem + kOffTmemPtr);

  const int n_it = (a.n_tiles + 2 * gridDim.x - 1) / (2 * gridDim.x);

  if (warp == 0) {
    // ================= weight producer =================
    if (lane == 0) {
      int stage = 0; uint32_t ph = 0;
      for (int it = 0; it < n_it; ++it) {
        for (int j = 0; j < kNumJobs; ++j) {
          const uint32_t jbytes = job_bytes(j);
          const uint8_t* src = a.wpack + job_off(j);
          for (int t = 0; t < 2; ++t) {
            if (tile_of(it, t) >= a.n_tiles) continue;
            for (uint32_t off = 0; off < jbytes; off += kStageBytes) {
              const uint32_t nb = (jbytes - off) < kStageBytes ? (jbytes - off) : kStageBytes;
              mbar_wait(bar_wempty(stage), ph ^ 1);
              mbar_arrive_expect_tx(bar_wfull(stage), nb);
              bulk_g2s(sbase + kOffRing + stage * kStageBytes, src + off, nb, bar_wfull(stage));
              if (++stage == kStages) { stage = 0; ph ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      int stage = 0; uint32_t ph = 0;
      uint32_t pa[2] = {0, 0};
      for (int it = 0; it < n_it; ++it) {
        for (int j = 0; j < kNumJobs; ++j) {
          const int N = job_N(j), kc_total = job_kc(j), kc_act = job_kc_act(j);
          const uint32_t idesc = instr_desc_bf16(128, N, 0, 0);
          const uint32_t b_chunk = N * 16;                    // bytes per K chunk of the weight block
          const int chunks_per_stage = kStageBytes / b_chunk;  // 4 (N=256) or 8 (N=128)
          for (int t = 0; t < 2; ++t) {
            if (tile_of(it, t) >= a.n_tiles) continue;
            mbar_wait(bar_aready(t), pa[t]); pa[t] ^= 1;
            tc_fence_after();
            const uint32_t act = sbase + kOffAct + t * kActBytes;
            const uint32_t aux = sbase + kOffAux + t * kAuxBytes;
            const uint32_t tacc = tmem_base + t * 256;
            for (int kc0 = 0; kc0 < kc_total; kc0 += chunks_per_stage) {
              mbar_wait(bar_wfull(stage), ph);
              tc_fence_after();
              const uint32_t wst = sbase + kOffRing + stage * kStageBytes;
              const int nchunk = (kc_total - kc0) < chunks_per_stage ? (kc_total - kc0) : chunks_per_stage;
              for (int c = 0; c < nchunk; c += 2) {
                const int kc = kc0 + c;
                const uint32_t a_addr = kc < kc_act ? act + kc * kChunkBytes : aux + (kc - kc_act) * kChunkBytes;
                const uint64_t ad = smem_desc(a_addr, kChunkBytes, 128);
                const uint64_t bd = smem_desc(wst + c * b_chunk, b_chunk, 128);
                mma_bf16(tacc, ad, bd, idesc, kc > 0 ? 1u : 0u);
              }
              mma_commit(bar_wempty(stage));
              if (++stage == kStages) { stage = 0; ph ^= 1; }
            }
            mma_commit(bar_accfull(t));
          }
        }
      }
    }
  } else {
    // ================= epilogue / PE warps =================
    const int ew = warp - 2;
    const int q = warp & 3;         // TMEM lane quadrant this warp may access
    const int hh = ew >> 2;         // column half
    const int row = q * 32 + lane;  // row of the tile == TMEM lane
    const uint32_t lane_sel = static_cast<uint32_t>(q * 32) << 16;
    float* part = reinterpret_cast<float*>(smem + kOffPart);
    uint32_t pf[2] = {0, 0};

    auto point_of = [&](int tile) {
      long long p = static_cast<long long>(tile) * kTile + row;
      return p < a.P ? p : a.P - 1;
    };
    // initial PE for both slots
    for (int t = 0; t < 2; ++t) {
      const int tile = tile_of(0, t);
      if (tile < a.n_tiles) {
        float x[3], d[3];
        load_point(a.src, point_of(tile), x, d);
        write_pe_row(smem + kOffAux + t * kAuxBytes, nullptr, row, hh, x, d);
      }
      fence_proxy_async();
      mbar_arrive(bar_aready(t));
    }

    for (int it = 0; it < n_it; ++it) {
      for (int j = 0; j < kNumJobs; ++j) {
        for (int t = 0; t < 2; ++t) {
          const int tile = tile_of(it, t);
          if (tile >= a.n_tiles) continue;
          uint8_t* act = smem + kOffAct + t * kActBytes;
          uint8_t* aux = smem + kOffAux + t * kAuxBytes;
          const uint32_t tacc = tmem_base + t * 256 + lane_sel;
          const long long p = static_cast<long long>(tile) * kTile + row;
          mbar_wait(bar_accfull(t), pf[t]); pf[t] ^= 1;
          tc_fence_after();

          if (j < 9) {
            const float* bias = consts + kConstBias + j * 256;
            uint8_t* stash = nullptr;
            if (TRAIN) {
              stash = (j < 8) ? a.stash_h + ((static_cast<size_t>(j) * a.n_tiles + tile) << 16)
                              : a.stash_f + (static_cast<size_t>(tile) << 16);
            }
            float sdot = 0.f;
            const int c0 = hh * 128;
            uint32_t v0[32], v1[32], pk[16];
            tmem_ld32(tacc + c0, v0);
            tmem_ld_wait();
#pragma unroll
            for (int cc = 0; cc < 4; ++cc) {
              uint32_t (&cur)[32] = (cc & 1) ? v1 : v0;
              uint32_t (&nxt)[32] = (cc & 1) ? v0 : v1;
              if (cc < 3) tmem_ld32(tacc + c0 + (cc + 1) * 32, nxt);
              const int col0 = c0 + cc * 32;
              if (j == 8) emit32<false>(cur, bias, act, stash, col0, row, pk);
              else emit32<true>(cur, bias, act, stash, col0, row, pk);
              if (j == 7) {
                const float* ws = consts + kConstWsig + col0;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                  sdot += bf16lo_to_f32(pk[i]) * ws[2 * i];
                  sdot += bf16hi_to_f32(pk[i]) * ws[2 * i + 1];
                }
              }
              if (cc < 3) tmem_ld_wait();
            }
            if (j == 4) {  // pos-PE is dead once job 4's MMAs finished: overlay the dir-PE
              float x[3], d[3];
              load_point(a.src, p < a.P ? p : a.P - 1, x, d);
              write_pe_row(nullptr, aux, row, hh, x, d);
            }
            tc_fence_before();
            fence_proxy_async();
            mbar_arrive(bar_aready(t));
            if (j == 7) {  // sigma head: relu(round_bf16(w.h8 + b) (+ noise))
              if (hh == 1) part[(t * 128 + row) * 3] = sdot;
              named_bar_sync(1, kEpiThreads);
              if (hh == 0 && p < a.P) {
                float s = sdot + part[(t * 128 + row) * 3] + consts[kConstBsig];
                s = round_bf16(s);
                if (a.noise) s = round_bf16(s + round_bf16(a.noise[p]));
                a.sigma[p] = fmaxf(s, 0.f);
              }
            }
          } else {
            // j == 9: colour hidden layer (N=128) + color_layer.2 + sigmoid
            const float* bias = consts + kConstBias9;
            uint8_t* stash = TRAIN ? a.stash_c + (static_cast<size_t>(tile) << 15) : nullptr;
            float r0 = 0.f, r1 = 0.f, r2 = 0.f;
            const int c0 = hh * 64;
            uint32_t v0[32], v1[32], pk[16];
            tmem_ld32(tacc + c0, v0);
            tmem_ld32(tacc + c0 + 32, v1);
            tmem_ld_wait();
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
              uint32_t (&cur)[32] = cc ? v1 : v0;
              const int col0 = c0 + cc * 32;
              emit32<true>(cur, bias, nullptr, stash, col0, row, pk);
              const float* w0 = consts + kConstWc2 + col0;
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const float lo = bf16lo_to_f32(pk[i]), hi = bf16hi_to_f32(pk[i]);
                r0 += lo * w0[2 * i] + hi * w0[2 * i + 1];
                r1 += lo * w0[128 + 2 * i] + hi * w0[128 + 2 * i + 1];
                r2 += lo * w0[256 + 2 * i] + hi * w0[256 + 2 * i + 1];
              }
            }
            // next tile of this slot: its pos-PE goes into aux (dir-PE was consumed by job 9's MMAs)
            const int ntile = tile_of(it + 1, t);
            if (ntile < a.n_tiles) {
              float x[3], d[3];
              load_point(a.src, point_of(ntile), x, d);
              write_pe_row(aux, nullptr, row, hh, x, d);
            }
            tc_fence_before();
            fence_proxy_async();
            mbar_arrive(bar_aready(t));
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
  if (warp == 1) tmem_dealloc(tmem_base, 512);
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

__global__ void pack_weights_kernel(const float* __restrict__ params, __nv_bfloat16* __restrict__ fwd,
                                    __nv_bfloat16* __restrict__ bwd, float* __restrict__ consts) {
  constexpr ParamOffsets po = param_offsets();
  const int nf = kFwdPackBytes / 2, nb = kBwdPackBytes / 2;
  const int total = nf + nb + kConstFloats;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    if (idx < nf) {
      int j = 0;
      while (j + 1 < kNumJobs && static_cast<uint32_t>(idx) * 2u >= job_off(j + 1)) ++j;
      const int e = idx - job_off(j) / 2;
      const int N = job_N(j);
      const int kc = e / (N * 8), n = (e / 8) % N, k = kc * 8 + (e & 7);
      float v = 0.f;
      if (j == 0) { if (k < 63) v = param_w(params, 0, n, k); }
      else if (j <= 3) v = param_w(params, j, n, k);
      else if (j == 4) { if (k < 256) v = param_w(params, 4, n, k); else if (k - 256 < 63) v = param_w(params, 4, n, k); }
      else if (j <= 7) v = param_w(params, j, n, k);
      else if (j == 8) v = param_w(params, 9, n, k);
      else { if (k < 256) v = param_w(params, 10, n, k); else if (k - 256 < 27) v = param_w(params, 10, n, k); }
      fwd[idx] = __float2bfloat16_rn(v);
    } else if (idx < nf + nb) {
      const int bi = idx - nf;
      int b = 0;
      while (b + 1 < kNumBwdJobs && static_cast<uint32_t>(bi) * 2u >= bjob_off(b + 1)) ++b;
      const int e = bi - bjob_off(b) / 2;
      const int kc = e / (256 * 8), i = (e / 8) % 256, o = kc * 8 + (e & 7);
      const int layer_of[kNumBwdJobs] = {10, 9, 7, 6, 5, 4, 3, 2, 1};
      bwd[bi] = __float2bfloat16_rn(param_w(params, layer_of[b], o, i));
    } else {
      const int c = idx - nf - nb;
      float v = 0.f;
      if (c < kConstBias9) {
        const int j = c / 256, n = c % 256;
        const int layer = j < 8 ? j : 9;
        v = params[po.b[layer] + n];
      } else if (c < kConstWsig) v = params[po.b[10] + (c - kConstBias9)];
      else if (c < kConstWc2) v = round_bf16(params[po.w[8] + (c - kConstWsig)]);
      else if (c < kConstBsig) v = round_bf16(params[po.w[11] + (c - kConstWc2)]);
      else if (c == kConstBsig) v = params[po.b[8]];
      else if (c < kConstBc2 + 3) v = params[po.b[11] + (c - kConstBc2)];
      consts[c] = v;
    }
  }
}

}  // namespace

extern "C" long long dfn_mlp_fwd_pack_bytes(void) { return kFwdPackBytes; }
extern "C" long long dfn_mlp_bwd_pack_bytes(void) { return kBwdPackBytes; }
extern "C" long long dfn_mlp_consts_bytes(void) { return kConstBytes; }
extern "C" long long dfn_mlp_stash_bytes(long long P) {
  const long long nt = (P + kTile - 1) / kTile;
  return nt * (8LL * 65536 + 65536 + 32768);
}

extern "C" int dfn_mlp_pack_weights(const float* params, void* fwd_pack, void* bwd_pack, float* consts, void* stream) {
  DFN_CHECK_ARG(params && fwd_pack && bwd_pack && consts, "mlp_pack_weights: null pointer");
  pack_weights_kernel<<<dfn_num_sms() * 4, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      params, static_cast<__nv_bfloat16*>(fwd_pack), static_cast<__nv_bfloat16*>(bwd_pack), consts);
  DFN_CHECK_LAUNCH("mlp_pack_weights");
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

extern "C" int dfn_mlp_forward(const dfn_point_source* src, long long P, const void* fwd_pack, const float* consts,
                               const float* noise, float* rgb, float* sigma, void* stash, void* stream) {
  if (int rc = check_source(src, P, "mlp_forward")) return rc;
  DFN_CHECK_ARG(fwd_pack && consts && rgb && sigma, "mlp_forward: null pointer");
  DFN_CHECK_ARG(P >= 0 && P < (1LL << 31) * kTile, "mlp_forward: P out of range");
  if (P == 0) return DFN_OK;
  FwdArgs a{};
  a.src = *src; a.P = P; a.n_tiles = static_cast<int>((P + kTile - 1) / kTile);
  a.wpack = static_cast<const uint8_t*>(fwd_pack); a.consts = consts; a.noise = noise; a.rgb = rgb; a.sigma = sigma;
  if (stash) {
    uint8_t* s = static_cast<uint8_t*>(stash);
    a.stash_h = s;
    a.stash_f = s + 8LL * a.n_tiles * 65536;
    a.stash_c = a.stash_f + 1LL * a.n_tiles * 65536;
  }
  const int grid = a.n_tiles < dfn_num_sms() ? a.n_tiles : dfn_num_sms();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (stash) {
    DFN_CUDA(cudaFuncSetAttribute(mlp_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes), "mlp_forward");
    mlp_fwd_kernel<true><<<grid, kThreads, kSmemBytes, st>>>(a);
  } else {
    DFN_CUDA(cudaFuncSetAttribute(mlp_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes), "mlp_forward");
    mlp_fwd_kernel<false><<<grid, kThreads, kSmemBytes, st>>>(a);
  }
  DFN_CHECK_LAUNCH("mlp_forward");
  return DFN_OK;
}
