// Instant-NGP variant, training backward in ONE launch (scripts/instantngp.py:41-104 under autograd; arithmetic restated
// from tiny-cuda-nn, parity UNPINNED - see ngp.cu).  No activation is stashed and no [P,32] encoding or encoding gradient ever
// reaches HBM: a warp re-encodes its 32 points straight into mma.sync A fragments (the gathers are L2 hits: the table is
// 26 MB), recomputes both tiny MLPs in registers exactly as ngp_forward_kernel does, then walks the chain backwards:
//   d rgb -> sigmoid' -> colour MLP (64 <- 64 <- 32) -> features (+ d sigma on feature 0) -> density MLP (64 <- 32) -> encoding
// Data gradients  dA = dZ W    : mma.m16n8k16 against TRANSPOSED weight images in shared memory, ReLU masks taken from the
//                                recomputed activations in the same register layout;
// weight gradients dW = dZ^T A : every warp parks its tile's activations and pre-activation gradients (fp16) in shared memory;
//                                after a CTA barrier warp w owns 10 of the 80 (16 out x 8 in) output tiles and reduces them
//                                over ALL 128 points of the round with mma.m16n8k16 (operands through ldmatrix.trans),
//                                accumulating in REGISTERS across rounds; one global flush at the end.  (A first version
//                                reduced per 16-point tile into shared memory with red.shared.f32: 640 shared atomics per
//                                point, 2 cycles per lane - ten times the cost of everything else in the kernel.)
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
// ---- staging of a warp's fragments for the CTA-wide weight-gradient reduction ----
// one A-layout fragment set (KB k-blocks of 16 columns, 16 rows) -> shared memory [128 points][LD halves], rows 16 w + ..
template <int KB, int LD>
__device__ __forceinline__ void park(const uint32_t (&x)[KB][4], __half* dst, int row0, int g, int t) {
#pragma unroll
  for (int kb = 0; kb < KB; ++kb)
#pragma unroll
    for (int r = 0; r < 4; ++r)
      *reinterpret_cast<uint32_t*>(dst + (row0 + g + 8 * (r & 1)) * LD + 16 * kb + 8 * (r >> 1) + 2 * t) = x[kb][r];
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], const __half* p) {
  const uint32_t addr = static_cast<uint32_t>(__cvta_generic_to_shared(p));
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x2_trans(uint32_t (&r)[2], const __half* p) {
  const uint32_t addr = static_cast<uint32_t>(__cvta_generic_to_shared(p));
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(addr));
}
// A' = dZ^T block (16 outs from column out0, 16 points from row pt0) of Ds [points][LD]: matrices (pts 0-7, outs 0-7),
// (pts 0-7, outs 8-15), (pts 8-15, outs 0-7), (pts 8-15, outs 8-15), each transposed on load = the A fragment registers
template <int LD>
__device__ __forceinline__ void load_at(uint32_t (&a)[4], const __half* Ds, int pt0, int out0, int lane) {
  const int mi = lane >> 3, r = lane & 7;
  ldsm_x4_trans(a, Ds + (pt0 + r + 8 * (mi >> 1)) * LD + out0 + 8 * (mi & 1));
}
// B for two 8-wide input tiles (columns in0 .. in0 + 15) of Xs [points][LD]: {b0, b1} of tile 0, {b0, b1} of tile 1
template <int LD>
__device__ __forceinline__ void load_b2(uint32_t (&b)[4], const __half* Xs, int pt0, int in0, int lane) {
  const int mi = lane >> 3, r = lane & 7;
  ldsm_x4_trans(b, Xs + (pt0 + r + 8 * (mi & 1)) * LD + in0 + 8 * (mi >> 1));
}
template <int LD>
__device__ __forceinline__ void load_b1(uint32_t (&b)[2], const __half* Xs, int pt0, int in0, int lane) {
  const int mi = (lane >> 3) & 1, r = lane & 7;
  ldsm_x2_trans(b, Xs + (pt0 + r + 8 * mi) * LD + in0);
}

// transposed weight images (halves): image of a [OUT][IN] layer = [IN rows][OUT + 8]; only the first 16 input columns of the
// colour MLP's first layer carry a gradient that is needed (the SH half of its input has none to pass on)
constexpr int kTD0 = 0;                      // Wd0^T [32][72]
constexpr int kTD1 = kTD0 + 32 * 72;         // Wd1^T [64][24]
constexpr int kTC0 = kTD1 + 64 * 24;         // Wc0[:, :16]^T [16][72]
constexpr int kTC1 = kTC0 + 16 * 72;         // Wc1^T [64][72]
constexpr int kTC2 = kTC1 + 64 * 72;         // Wc2^T [64][24]
constexpr int kTHalves = kTC2 + 64 * 24;
// staging of one round (8 warps x 16 points) for the weight gradients: activations X and pre-activation gradients D, fp16,
// [128 points][columns + 8] (row strides of 48 / 80 / 144 bytes: ldmatrix reads 8 rows without bank conflicts)
constexpr int kPts = 128;
constexpr int kSEnc = 0;                        // enc  [128][40]
constexpr int kSHd1 = kSEnc + kPts * 40;        // hd1  [128][72]
constexpr int kSCin = kSHd1 + kPts * 72;        // cin  [128][40]  (features | SH)
constexpr int kSHc1 = kSCin + kPts * 40;        // hc1  [128][72]
constexpr int kSHc2 = kSHc1 + kPts * 72;        // hc2  [128][72]
constexpr int kSDzd = kSHc2 + kPts * 72;        // d pre hd1 [128][72]
constexpr int kSDfe = kSDzd + kPts * 72;        // d pre feat [128][24]
constexpr int kSDz1 = kSDfe + kPts * 24;        // d pre hc1 [128][72]
constexpr int kSDz2 = kSDz1 + kPts * 72;        // d pre hc2 [128][72]
constexpr int kSDx3 = kSDz2 + kPts * 72;        // d x3 [128][24]
constexpr int kSHalves = kSDx3 + kPts * 24;

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
  __half* ST = WT + kTHalves;
  LevelDev* lv = reinterpret_cast<LevelDev*>(ST + kSHalves);

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
  if (threadIdx.x < kLevels) lv[threadIdx.x] = grid.lv[threadIdx.x];
  __syncthreads();

  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const long long P = a.P;
  const long long n_tiles = (P + 15) / 16;                       // 16-point tiles: one per warp and round
  const int warp = threadIdx.x >> 5;
  const long long n_rounds = (n_tiles + 8ll * gridDim.x - 1) / (8ll * gridDim.x);
  const float inv_scale = 1.0f / a.loss_scale;
  float2* d_table = reinterpret_cast<float2*>(a.d_table);
  // this warp's share of the weight gradients: Wc1 (ob = w/2, 4 input tiles), Wd0 and Wc0 (ob = w/2, 2 input tiles each),
  // Wd1 and Wc2 (input tile w)
  float acc_c1[4][4], acc_d0[2][4], acc_c0[2][4], acc_d1[4], acc_c2[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    acc_d1[i] = acc_c2[i] = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) acc_c1[j][i] = 0.f;
    acc_d0[0][i] = acc_d0[1][i] = acc_c0[0][i] = acc_c0[1][i] = 0.f;
  }

  for (long long round = 0; round < n_rounds; ++round) {
    const long long tile = (round * gridDim.x + blockIdx.x) * 8 + warp;
    // (a warp without a tile in the last round runs on clamped points with zero output gradients: it parks zeros)
    const long long p0 = (tile < n_tiles ? tile : n_tiles - 1) * 16;
    const bool live = tile < n_tiles;
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
      const bool valid = live && pr[h] < P;
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
    park<1, 24>(dx3, ST + kSDx3, 16 * warp, g, t);
    park<4, 72>(hc2, ST + kSHc2, 16 * warp, g, t);
    {
      float acc[8][4];
      layer<64, 8>(dz2, WT + kTC1, g, t, acc);           // d hc1 = d z2 * Wc1
      mask_repack<4>(acc, hc1, dz1);
    }
    park<4, 72>(dz2, ST + kSDz2, 16 * warp, g, t);
    park<4, 72>(hc1, ST + kSHc1, 16 * warp, g, t);
    park<4, 72>(dz1, ST + kSDz1, 16 * warp, g, t);
    park<2, 40>(cin, ST + kSCin, 16 * warp, g, t);
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
    park<1, 24>(dfe, ST + kSDfe, 16 * warp, g, t);
    park<4, 72>(hd1, ST + kSHd1, 16 * warp, g, t);
    park<4, 72>(dzd, ST + kSDzd, 16 * warp, g, t);
    park<2, 40>(enc, ST + kSEnc, 16 * warp, g, t);
    float denc[4][4];
    layer<64, 4>(dzd, WT + kTD0, g, t, denc);            // d enc = d zd * Wd0: (row g / g+8, level 4j + t, features 0 / 1)

    // ---- trilinear scatter-add into the table gradient, straight from the accumulators ----
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (!live || pr[h] >= P) continue;
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

    // ---- weight gradients of the round: dW += D^T X over the 128 parked points, this warp's 10 output tiles ----
    __syncthreads();
    {
      const int ob = warp >> 1, hw = warp & 1;
#pragma unroll 2
      for (int ks = 0; ks < 8; ++ks) {
        const int pt0 = 16 * ks;
        uint32_t at[4], b4[4], b2[2];
        load_at<72>(at, ST + kSDz2, pt0, 16 * ob, lane);                      // d Wc1 [64 x 64]
#pragma unroll
        for (int pr2 = 0; pr2 < 2; ++pr2) {
          load_b2<72>(b4, ST + kSHc1, pt0, 32 * hw + 16 * pr2, lane);
          mma16816(acc_c1[2 * pr2], at, b4[0], b4[1]);
          mma16816(acc_c1[2 * pr2 + 1], at, b4[2], b4[3]);
        }
        load_at<72>(at, ST + kSDzd, pt0, 16 * ob, lane);                      // d Wd0 [64 x 32]
        load_b2<40>(b4, ST + kSEnc, pt0, 16 * hw, lane);
        mma16816(acc_d0[0], at, b4[0], b4[1]);
        mma16816(acc_d0[1], at, b4[2], b4[3]);
        load_at<72>(at, ST + kSDz1, pt0, 16 * ob, lane);                      // d Wc0 [64 x 32]
        load_b2<40>(b4, ST + kSCin, pt0, 16 * hw, lane);
        mma16816(acc_c0[0], at, b4[0], b4[1]);
        mma16816(acc_c0[1], at, b4[2], b4[3]);
        load_at<24>(at, ST + kSDfe, pt0, 0, lane);                            // d Wd1 [16 x 64]
        load_b1<72>(b2, ST + kSHd1, pt0, 8 * warp, lane);
        mma16816(acc_d1, at, b2[0], b2[1]);
        load_at<24>(at, ST + kSDx3, pt0, 0, lane);                            // d Wc2 [16 x 64]
        load_b1<72>(b2, ST + kSHc2, pt0, 8 * warp, lane);
        mma16816(acc_c2, at, b2[0], b2[1]);
      }
    }
    __syncthreads();   // the staging arrays are rewritten by the next round
  }

  // ---- flush this warp's weight-gradient tiles (unscaled): c[0], c[1] = (out g, ins 2t, 2t+1), c[2], c[3] = out g + 8 ----
  {
    const int ob = warp >> 1, hw = warp & 1;
    auto put = [&](float* W0, int in_dim, int out0, int in0, const float (&c)[4]) {
      float* d = W0 + (out0 + g) * in_dim + in0 + 2 * t;
      if (c[0] != 0.f) atomicAdd(d, c[0] * inv_scale);
      if (c[1] != 0.f) atomicAdd(d + 1, c[1] * inv_scale);
      if (c[2] != 0.f) atomicAdd(d + 8 * in_dim, c[2] * inv_scale);
      if (c[3] != 0.f) atomicAdd(d + 8 * in_dim + 1, c[3] * inv_scale);
    };
#pragma unroll
    for (int j = 0; j < 4; ++j) put(a.d_cmlp + 64 * 32, 64, 16 * ob, 32 * hw + 8 * j, acc_c1[j]);
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      put(a.d_dmlp, 32, 16 * ob, 16 * hw + 8 * j, acc_d0[j]);
      put(a.d_cmlp, 32, 16 * ob, 16 * hw + 8 * j, acc_c0[j]);
    }
    put(a.d_dmlp + 64 * 32, 64, 0, 8 * warp, acc_d1);
    put(a.d_cmlp + 64 * 32 + 64 * 64, 64, 0, 8 * warp, acc_c2);
  }
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
  const size_t smem = (kWHalves + kTHalves + kSHalves) * 2 + kLevels * sizeof(LevelDev);
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
