// Instant-NGP variant, training backward in ONE launch (scripts/instantngp.py:41-104 under autograd; arithmetic restated
// from tiny-cuda-nn, parity UNPINNED - see ngp.cu).  No activation is stashed and no [P,32] encoding or encoding gradient ever
// reaches HBM: a warp re-encodes its 32 points straight into mma.sync A fragments (the gathers are L2 hits: the table is
// 26 MB), recomputes both tiny MLPs in registers exactly as ngp_forward_kernel does, then walks the chain backwards:
//   d rgb -> sigmoid' -> colour MLP (64 <- 64 <- 32) -> features (+ d sigma on feature 0) -> density MLP (64 <- 32) -> encoding
// Data gradients  dA = dZ W    : mma.m16n8k16 against TRANSPOSED weight images in shared memory, ReLU masks taken from the
//                                recomputed activations in the same register layout;
// weight gradients dW = dZ^T A : mma.m16n8k16 with both operands transposed in registers (movmatrix.m8n8.trans), reduction over
//                                the tile's 16 points, accumulated per CTA in shared memory (fp32, red.shared) and flushed once;
// table gradient               : the d-encoding accumulators ARE laid out as (point, level) pairs - the same pairs the lane
//                                encoded - so the trilinear scatter-add (x-pair vector reductions, as hashgrid_backward_kernel)
//                                runs straight from registers, in fp32.
// Gradients are computed on loss_scale * d out (fp16 range, tcnn's default 128) and divided by it in fp32 on the way out.
#include "ngp_common.cuh"

namespace {
using namespace ngp;

constexpr int kThreads = 256;

__device__ __forceinline__ uint32_t trans8x8(uint32_t x) {
  uint32_t r;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(r) : "r"(x));
  return r;
}
// fp32 accumulators of 2*KB n-tiles -> gradient w.r.t. the pre-activation (ReLU mask from the packed fp16 activations, which
// sit in the same element order) -> fp16 A fragments of the next backward layer
template <int KB>
__device__ __forceinline__ void mask_repack(const float (&c)[2 * KB][4], const uint32_t (&act)[KB][4], uint32_t (&out)[KB][4]) {
#pragma unroll
  for (int kb = 0; kb < KB; ++kb) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int j = 2 * kb + (r >> 1), e = 2 * (r & 1);
      const uint32_t a = act[kb][r];
      const float lo = (a & 0x0000FFFFu) ? c[j][e] : 0.f;      // post-ReLU activations are >= +0: non-zero bits == positive
      const float hi = (a & 0xFFFF0000u) ? c[j][e + 1] : 0.f;
      out[kb][r] = pack_h2(lo, hi);
    }
  }
}
// dW[16 ob + .., 8 nj + ..] += dZ^T (16 outs x 16 points) * A (16 points x 8 ins) for every (ob, nj); operands arrive transposed
// (TD: out blocks, TX: in blocks); dWs: shared-memory accumulator [OUT][IN + 8] fp32
template <int KBO, int KBI>
__device__ __forceinline__ void wgrad_tile(const uint32_t (&TD)[KBO][4], const uint32_t (&TX)[KBI][4], float* dWs, int g, int t) {
  constexpr int LD = 16 * KBI + 8;
#pragma unroll
  for (int ob = 0; ob < KBO; ++ob) {
    const uint32_t a[4] = {TD[ob][0], TD[ob][2], TD[ob][1], TD[ob][3]};   // rows = outs, cols = points
#pragma unroll
    for (int nj = 0; nj < 2 * KBI; ++nj) {
      float c[4] = {0.f, 0.f, 0.f, 0.f};
      mma16816(c, a, TX[nj >> 1][2 * (nj & 1)], TX[nj >> 1][2 * (nj & 1) + 1]);
      float* d = dWs + (16 * ob + g) * LD + 8 * nj + 2 * t;
      atomicAdd(d, c[0]);
      atomicAdd(d + 1, c[1]);
      atomicAdd(d + 8 * LD, c[2]);
      atomicAdd(d + 8 * LD + 1, c[3]);
    }
  }
}
template <int KB>
__device__ __forceinline__ void trans_frag(const uint32_t (&x)[KB][4], uint32_t (&tx)[KB][4]) {
#pragma unroll
  for (int kb = 0; kb < KB; ++kb)
#pragma unroll
    for (int r = 0; r < 4; ++r) tx[kb][r] = trans8x8(x[kb][r]);
}

// transposed weight images (halves): image of a [OUT][IN] layer = [IN rows][OUT + 8]; only the first 16 input columns of the
// colour MLP's first layer carry a gradient that is needed (the SH half of its input has none to pass on)
constexpr int kTD0 = 0;                      // Wd0^T [32][72]
constexpr int kTD1 = kTD0 + 32 * 72;         // Wd1^T [64][24]
constexpr int kTC0 = kTD1 + 64 * 24;         // Wc0[:, :16]^T [16][72]
constexpr int kTC1 = kTC0 + 16 * 72;         // Wc1^T [64][72]
constexpr int kTC2 = kTC1 + 64 * 72;         // Wc2^T [64][24]
constexpr int kTHalves = kTC2 + 64 * 24;
// fp32 weight-gradient accumulators [OUT][IN + 8]
constexpr int kGD0 = 0, kGD1 = kGD0 + 64 * 40, kGC0 = kGD1 + 16 * 72, kGC1 = kGC0 + 64 * 40, kGC2 = kGC1 + 64 * 72,
              kGFloats = kGC2 + 16 * 72;

__device__ __forceinline__ void stage_transposed(__half* dst, const __half* __restrict__ src, int out, int in_total, int in_used) {
  const int ld = out + 8;
  for (int i = threadIdx.x; i < out * in_used; i += blockDim.x) {
    const int o = i / in_used, k = i % in_used;
    dst[k * ld + o] = src[o * in_total + k];
  }
}

struct NgpBwdArgs {
  const float* pos;        // [P,3]
  const float* dirs;       // [P,3] or one direction (dirs_stride 0)
  int dirs_stride, dirs_group;
  long long P;
  const __half2* table;
  const __half* dmlp_w;
  const __half* cmlp_w;
  const float* d_rgb;      // [P,3] fp32: d loss / d rgb
  const float* d_sigma;    // [P]   fp32: d loss / d sigma (= feature 0)
  float loss_scale;
  float* d_table;          // [n_entries,2] fp32, accumulated
  float* d_dmlp;           // [3072] fp32, accumulated
  float* d_cmlp;           // [7168] fp32, accumulated
};

__global__ void __launch_bounds__(kThreads, 1) ngp_backward_kernel(const NgpBwdArgs a, const __grid_constant__ GridDev grid) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __half* W = reinterpret_cast<__half*>(smem_raw);
  __half* WT = W + kWHalves;
  float* dWs = reinterpret_cast<float*>(smem_raw + (kWHalves + kTHalves) * 2);
  LevelDev* lv = reinterpret_cast<LevelDev*>(dWs + kGFloats);

  stage_weights(W + kOffD0, a.dmlp_w, 64, 32);
  stage_weights(W + kOffD1, a.dmlp_w + 64 * 32, 16, 64);
  stage_weights(W + kOffC0, a.cmlp_w, 64, 32);
  stage_weights(W + kOffC1, a.cmlp_w + 64 * 32, 64, 64);
  stage_weights(W + kOffC2, a.cmlp_w + 64 * 32 + 64 * 64, 16, 64);
  stage_transposed(WT + kTD0, a.dmlp_w, 64, 32, 32);
  stage_transposed(WT + kTD1, a.dmlp_w + 64 * 32, 16, 64, 64);
  stage_transposed(WT + kTC0, a.cmlp_w, 64, 32, 16);
  stage_transposed(WT + kTC1, a.cmlp_w + 64 * 32, 64, 64, 64);
  stage_transposed(WT + kTC2, a.cmlp_w + 64 * 32 + 64 * 64, 16, 64, 64);
  for (int i = threadIdx.x; i < kGFloats; i += blockDim.x) dWs[i] = 0.f;
  if (threadIdx.x < kLevels) lv[threadIdx.x] = grid.lv[threadIdx.x];
  __syncthreads();

  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const long long P = a.P;
  const long long n_tiles = (P + 15) / 16;                       // 16-point tiles: one per warp iteration
  const long long warp_global = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long warp_stride = (long long)gridDim.x * (blockDim.x >> 5);
  const float inv_scale = 1.0f / a.loss_scale;
  float2* d_table = reinterpret_cast<float2*>(a.d_table);

  for (long long tile = warp_global; tile < n_tiles; tile += warp_stride) {
    const long long p0 = tile * 16;
    // ---- re-encode the 16 points into A fragments (as ngp_forward_kernel) ----
    uint32_t enc[2][4];
    float px[2], py[2], pz[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      long long p = p0 + 8 * h + g;
      p = p < P ? p : P - 1;
      px[h] = __ldg(a.pos + 3 * p), py[h] = __ldg(a.pos + 3 * p + 1), pz[h] = __ldg(a.pos + 3 * p + 2);
      Cell cell[4];
      uint32_t raw[4][8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const LevelDev L = lv[8 * (j >> 1) + 4 * (j & 1) + t];
        cell[j] = locate(px[h], py[h], pz[h], L.scale);
        gather_level_uniform<false>(a.table + L.offset, L, cell[j], raw[j]);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) enc[j >> 1][2 * (j & 1) + h] = blend_level(cell[j], raw[j]);
    }
    const long long pr[2] = {p0 + g, p0 + 8 + g};

    // ---- forward recompute ----
    uint32_t hd1[4][4], feat[1][4], cin[2][4], hc1[4][4], hc2[4][4];
    float c3[1][4];
    {
      float c1[8][4];
      layer<32, 8>(enc, W + kOffD0, g, t, c1);
      relu_repack<4>(c1, hd1);
      float c2[2][4];
      layer<64, 2>(hd1, W + kOffD1, g, t, c2);
      relu_repack<1>(c2, feat);
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) cin[0][r] = feat[0][r];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      long long p = pr[h] < P ? pr[h] : P - 1;
      const float* d = a.dirs + (long long)a.dirs_stride * dir_index(p, a.dirs_group);
      float o[16];
      sh4(__ldg(d), __ldg(d + 1), __ldg(d + 2), o);
      float s0 = o[0], s1 = o[1], s2 = o[8], s3 = o[9];
      if (t == 1) s0 = o[2], s1 = o[3], s2 = o[10], s3 = o[11];
      if (t == 2) s0 = o[4], s1 = o[5], s2 = o[12], s3 = o[13];
      if (t == 3) s0 = o[6], s1 = o[7], s2 = o[14], s3 = o[15];
      cin[1][h] = pack_h2(s0, s1);
      cin[1][2 + h] = pack_h2(s2, s3);
    }
    {
      float c1[8][4];
      layer<32, 8>(cin, W + kOffC0, g, t, c1);
      relu_repack<4>(c1, hc1);
      float c2[8][4];
      layer<64, 8>(hc1, W + kOffC1, g, t, c2);
      relu_repack<4>(c2, hc2);
    }
    layer<64, 1>(hc2, W + kOffC2, g, t, c3);

    // ---- d x3 = loss_scale * d rgb * rgb (1 - rgb): columns 0..2 of the padded 16-wide output ----
    uint32_t dx3[1][4] = {{0u, 0u, 0u, 0u}};
    float dsig[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const bool valid = pr[h] < P;
      dsig[h] = valid ? a.loss_scale * __ldg(a.d_sigma + pr[h]) : 0.f;
      float g0 = 0.f, g1 = 0.f;
      if (valid && t < 2) {
        const float v0 = 1.0f / (1.0f + __expf(-c3[0][2 * h])), v1 = 1.0f / (1.0f + __expf(-c3[0][2 * h + 1]));
        const float* dr = a.d_rgb + 3 * pr[h];
        if (t == 0) { g0 = a.loss_scale * __ldg(dr) * v0 * (1.0f - v0); g1 = a.loss_scale * __ldg(dr + 1) * v1 * (1.0f - v1); }
        else { g0 = a.loss_scale * __ldg(dr + 2) * v0 * (1.0f - v0); }
      }
      dx3[0][h] = pack_h2(g0, g1);
    }

    // ---- colour MLP backward ----
    uint32_t dz2[4][4], dz1[4][4], dfe[1][4];
    {
      float acc[8][4];
      layer<16, 8>(dx3, WT + kTC2, g, t, acc);          // d hc2 = d x3 * Wc2
      mask_repack<4>(acc, hc2, dz2);
    }
    {
      uint32_t TD[1][4], TX[4][4];
      trans_frag<1>(dx3, TD);
      trans_frag<4>(hc2, TX);
      wgrad_tile<1, 4>(TD, TX, dWs + kGC2, g, t);        // d Wc2 += d x3^T hc2
    }
    {
      float acc[8][4];
      layer<64, 8>(dz2, WT + kTC1, g, t, acc);           // d hc1 = d z2 * Wc1
      mask_repack<4>(acc, hc1, dz1);
    }
    {
      uint32_t TD[4][4], TX[4][4];
      trans_frag<4>(dz2, TD);
      trans_frag<4>(hc1, TX);
      wgrad_tile<4, 4>(TD, TX, dWs + kGC1, g, t);        // d Wc1 += d z2^T hc1
    }
    {
      uint32_t TD[4][4], TX[2][4];
      trans_frag<4>(dz1, TD);
      trans_frag<2>(cin, TX);
      wgrad_tile<4, 2>(TD, TX, dWs + kGC0, g, t);        // d Wc0 += d z1^T [features | SH]
    }
    {
      float acc[2][4];
      layer<64, 2>(dz1, WT + kTC0, g, t, acc);           // d features = d z1 * Wc0[:, :16]
      if (t == 0) { acc[0][0] += dsig[0]; acc[0][2] += dsig[1]; }   // sigma IS feature 0 (instantngp.py:79)
      mask_repack<1>(acc, feat, dfe);
    }
    // ---- density MLP backward ----
    uint32_t dzd[4][4];
    {
      float acc[8][4];
      layer<16, 8>(dfe, WT + kTD1, g, t, acc);           // d hd1 = d feat * Wd1
      mask_repack<4>(acc, hd1, dzd);
    }
    {
      uint32_t TD[1][4], TX[4][4];
      trans_frag<1>(dfe, TD);
      trans_frag<4>(hd1, TX);
      wgrad_tile<1, 4>(TD, TX, dWs + kGD1, g, t);        // d Wd1 += d feat^T hd1
    }
    {
      uint32_t TD[4][4], TX[2][4];
      trans_frag<4>(dzd, TD);
      trans_frag<2>(enc, TX);
      wgrad_tile<4, 2>(TD, TX, dWs + kGD0, g, t);        // d Wd0 += d zd^T enc
    }
    float denc[4][4];
    layer<64, 4>(dzd, WT + kTD0, g, t, denc);            // d enc = d zd * Wd0: (row g / g+8, level 4j + t, features 0 / 1)

    // ---- trilinear scatter-add into the table gradient, straight from the accumulators ----
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (pr[h] >= P) continue;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float d0 = denc[j][2 * h] * inv_scale, d1 = denc[j][2 * h + 1] * inv_scale;
        if (d0 == 0.f && d1 == 0.f) continue;
        const LevelDev L = lv[4 * j + t];
        const Cell c = locate(px[h], py[h], pz[h], L.scale);
        float2* base = d_table + L.offset;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint32_t y = c.gy + (k & 1), z = c.gz + (k >> 1);
          const uint32_t i0 = grid_index(c.gx, y, z, L), i1 = grid_index(c.gx + 1u, y, z, L);
          const float w0 = corner_weight(c, 2 * k), w1 = corner_weight(c, 2 * k + 1);
          if (i1 == (i0 ^ 1u)) {
            const bool odd = i0 & 1u;
            const float wa = odd ? w1 : w0, wb = odd ? w0 : w1;
            atomicAdd(reinterpret_cast<float4*>(base + (i0 & ~1u)), make_float4(wa * d0, wa * d1, wb * d0, wb * d1));
          } else {
            atomicAdd(base + i0, make_float2(w0 * d0, w0 * d1));
            atomicAdd(base + i1, make_float2(w1 * d0, w1 * d1));
          }
        }
      }
    }
  }

  // ---- flush the CTA's weight gradients (unscaled) ----
  __syncthreads();
  auto flush = [&](const float* src, float* dst, int out, int in) {
    const int ld = in + 8;
    for (int i = threadIdx.x; i < out * in; i += blockDim.x) {
      const float v = src[(i / in) * ld + (i % in)] * inv_scale;
      if (v != 0.f) atomicAdd(dst + i, v);
    }
  };
  flush(dWs + kGD0, a.d_dmlp, 64, 32);
  flush(dWs + kGD1, a.d_dmlp + 64 * 32, 16, 64);
  flush(dWs + kGC0, a.d_cmlp, 64, 32);
  flush(dWs + kGC1, a.d_cmlp + 64 * 32, 64, 64);
  flush(dWs + kGC2, a.d_cmlp + 64 * 32 + 64 * 64, 16, 64);
}

}  // namespace

// d loss / d (hash table, density MLP, colour MLP) of InstantNGP.forward on P points, accumulated into fp32 buffers, from
// d loss / d rgb [P,3] and d loss / d sigma [P] (fp32).  The forward is recomputed (nothing is stashed).
extern "C" int dfn_ngp_backward(const float* pos, const float* dirs, int dirs_stride, int dirs_group, long long P, const void* table_f16,
                                const dfn_hashgrid* g, const void* dmlp_w_f16, const void* cmlp_w_f16, const float* d_rgb,
                                const float* d_sigma, float loss_scale, float* d_table, float* d_dmlp, float* d_cmlp, void* stream) {
  DFN_CHECK_ARG(g != nullptr && g->n_entries > 0, "dfn_ngp_backward: grid descriptor not initialised (dfn_hashgrid_init)");
  DFN_CHECK_ARG(P >= 0, "dfn_ngp_backward: P < 0");
  if (P == 0) return DFN_OK;
  DFN_CHECK_ARG(pos && dirs && table_f16 && dmlp_w_f16 && cmlp_w_f16 && d_rgb && d_sigma && d_table && d_dmlp && d_cmlp,
                "dfn_ngp_backward: null pointer");
  DFN_CHECK_ARG(((uintptr_t)table_f16 & 7) == 0 && ((uintptr_t)d_table & 15) == 0, "dfn_ngp_backward: table / d_table alignment");
  DFN_CHECK_ARG(dirs_stride == 0 || dirs_stride == 3, "dfn_ngp_backward: dirs_stride must be 3 (per point) or 0 (one view)");
  DFN_CHECK_ARG(loss_scale > 0.f, "dfn_ngp_backward: loss_scale must be positive");
  NgpBwdArgs a{};
  a.pos = pos; a.dirs = dirs; a.dirs_stride = dirs_stride; a.dirs_group = dirs_group < 1 ? 1 : dirs_group; a.P = P; a.table = (const __half2*)table_f16;
  a.dmlp_w = (const __half*)dmlp_w_f16; a.cmlp_w = (const __half*)cmlp_w_f16; a.d_rgb = d_rgb; a.d_sigma = d_sigma;
  a.loss_scale = loss_scale; a.d_table = d_table; a.d_dmlp = d_dmlp; a.d_cmlp = d_cmlp;
  const size_t smem = (kWHalves + kTHalves) * 2 + kGFloats * 4 + kLevels * sizeof(LevelDev);
  DFN_CUDA(cudaFuncSetAttribute(ngp_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "dfn_ngp_backward");
  const long long tiles = (P + 15) / 16;
  long long blocks = (tiles + kThreads / 32 - 1) / (kThreads / 32);
  int occ = 1;
  DFN_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ngp_backward_kernel, kThreads, smem), "dfn_ngp_backward");
  const long long cap = static_cast<long long>(occ < 1 ? 1 : occ) * dfn_num_sms();
  if (blocks > cap) blocks = cap;
  ngp_backward_kernel<<<(int)blocks, kThreads, smem, (cudaStream_t)stream>>>(a, to_dev(g));
  DFN_CHECK_LAUNCH("dfn_ngp_backward");
  return DFN_OK;
}
