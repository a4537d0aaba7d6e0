// Shared definitions for the vanilla-NeRF MLP kernels (reference scripts/model.py:56-187,
// hidden_dim = 256, pos_L = 10, dir_L = 4).
//
// Tile = 128 points.  Every operand tile lives in the chunk layout of umma.cuh:
//   [cols/8 chunks][128 rows][8 bf16]  (2 KB per chunk for a 128-row tile).
//
// Packed forward weights (bf16), per model, job j (one tcgen05 accumulation chain):
//   j0 : pos_layers.0      N=256  K= 64 (63 PE columns + 1 zero)
//   j1-3: pos_layers.2/4/6 N=256  K=256
//   j4 : skip_layers.0     N=256  K=320 (256 hidden | 63 PE + 1 zero)
//   j5-7: skip_layers.2/4/6
//   j8 : feature_layer     N=256  K=256
//   j9 : color_layer.0     N=128  K=288 (256 features | 27 dir-PE + 5 zero)
// each stored as K/8 chunks of [N rows][8] bf16 (K-major B operand, LBO = N*16, SBO = 128).
// The sigma head (density_layer, N=1) and color_layer.2 (N=3) are evaluated on the CUDA cores inside
// the epilogues of j7 / j9 from the fp32 "consts" block.
#pragma once
#include "dfn_common.cuh"
#include "umma.cuh"
#include "../../include/deepfields_b200.h"

namespace mlp {

constexpr int kTile = 128;
constexpr int kHidden = 256;
constexpr int kPosL = 10, kDirL = 4;
constexpr int kPosDim = 63, kDirDim = 27;
constexpr int kNumJobs = 10;
constexpr uint32_t kChunkBytes = kTile * 16;        // one 8-column chunk of a 128-row tile
constexpr uint32_t kActBytes = 32 * kChunkBytes;    // 128 x 256 bf16 = 64 KB
constexpr uint32_t kAuxBytes = 8 * kChunkBytes;     // 128 x 64 bf16  = 16 KB
constexpr uint32_t kStageBytes = 16384;

// forward job table.  Every job's weight block ends with one extra K=16 slice that carries the layer bias
// (column k=0 = bf16(bias), the rest zero); its A operand is a constant "ones" tile (column 0 = 1), so the bias
// is added by the tensor core exactly as bf16 autocast does it (F.linear casts the bias to bf16) and the
// epilogue has no per-column loads.
__host__ __device__ constexpr int job_N(int j) { return j == 9 ? 128 : 256; }
__host__ __device__ constexpr int job_kc(int j) { return j == 0 ? 8 : (j == 4 ? 40 : (j == 9 ? 36 : 32)); }  // data chunks
__host__ __device__ constexpr int job_kcb(int j) { return job_kc(j) + 2; }                                   // + bias slice
__host__ __device__ constexpr int job_kc_act(int j) { return j == 0 ? 0 : 32; }  // chunks read from the act tile
__host__ __device__ constexpr uint32_t job_bytes(int j) { return static_cast<uint32_t>(job_kcb(j)) * job_N(j) * 16u; }
__host__ __device__ constexpr uint32_t job_off(int j) {
  uint32_t o = 0;
  for (int i = 0; i < j; ++i) o += job_bytes(i);
  return o;
}
constexpr uint32_t kFwdPackBytes = job_off(kNumJobs);
// The forward kernel runs as CTA pairs (tcgen05 cta_group::2, M = 256): CTA r of a pair holds rows [r*N/2, (r+1)*N/2)
// of every weight block, so a job's image is stored as two halves, each [K/8 chunks][N/2 rows][8] bf16:
//   byte offset = job_off(j) + r * job_bytes(j)/2 + kc * job_half_chunk(j) + n_local * 16 + (k % 8) * 2
__host__ __device__ constexpr uint32_t job_half_chunk(int j) { return static_cast<uint32_t>(job_N(j) / 2) * 16u; }  // bytes per K chunk per CTA
__host__ __device__ constexpr uint32_t job_half_bytes(int j) { return job_bytes(j) / 2; }

// Number of 128-point tiles the MLP kernels work on for P points: whole 256-point pairs (the forward processes
// two tiles per CTA pair), so the stash / dz layouts of all three kernels agree.
__host__ __device__ constexpr long long num_tiles(long long P) { return 2 * ((P + 2 * kTile - 1) / (2 * kTile)); }

// dgrad job table (backward chain; B operand = W^T, reduction over the layer's outputs)
//   b0: d feat = dZ_c0 [128x128] * color_layer.0[:, :256]      K=128 N=256
//   b1: d h8   = dZ_f  * feature_layer                         K=256 N=256
//   b2..b4: d h7,h6,h5 = dZ_8,7,6 * skip_layers.6/4/2
//   b5: d h4   = dZ_5 * skip_layers.0[:, :256]
//   b6..b8: d h3,h2,h1 = dZ_4,3,2 * pos_layers.6/4/2
constexpr int kNumBwdJobs = 9;
__host__ __device__ constexpr int bjob_kc(int b) { return b == 0 ? 16 : 32; }
__host__ __device__ constexpr uint32_t bjob_bytes(int b) { return static_cast<uint32_t>(bjob_kc(b)) * 256u * 16u; }
__host__ __device__ constexpr uint32_t bjob_off(int b) {
  uint32_t o = 0;
  for (int i = 0; i < b; ++i) o += bjob_bytes(i);
  return o;
}
constexpr uint32_t kBwdPackBytes = bjob_off(kNumBwdJobs);

// fp32 consts block (floats): the two small heads that run on the CUDA cores
constexpr int kConstWsig = 0;                       // 256, bf16-rounded density_layer.weight
constexpr int kConstWc2 = kConstWsig + 256;         // 3 x 128, bf16-rounded color_layer.2.weight
constexpr int kConstBsig = kConstWc2 + 384;         // 1, bf16-rounded density_layer.bias
constexpr int kConstBc2 = kConstBsig + 1;           // 3, bf16-rounded color_layer.2.bias
constexpr int kConstFloats = ((kConstBc2 + 3 + 3) / 4) * 4;
constexpr uint32_t kConstBytes = kConstFloats * 4;

// flat fp32 parameter order == reference state_dict / named_parameters order (model.py:99-137)
//   pos_layers.{0,2,4,6}.{weight,bias}, skip_layers.{0,2,4,6}.{weight,bias}, density_layer.{w,b},
//   feature_layer.{w,b}, color_layer.{0,2}.{w,b}
struct ParamOffsets {
  int w[12];
  int b[12];
  int total;
};
__host__ __device__ constexpr ParamOffsets param_offsets() {
  ParamOffsets p{};
  const int outs[12] = {256, 256, 256, 256, 256, 256, 256, 256, 1, 256, 128, 3};
  const int ins[12] = {63, 256, 256, 256, 319, 256, 256, 256, 256, 256, 283, 128};
  int o = 0;
  for (int i = 0; i < 12; ++i) {
    p.w[i] = o; o += outs[i] * ins[i];
    p.b[i] = o; o += outs[i];
  }
  p.total = o;
  return p;
}
constexpr int kNumParams = param_offsets().total;  // 595,844
__host__ __device__ constexpr int layer_in(int i) {
  const int ins[12] = {63, 256, 256, 256, 319, 256, 256, 256, 256, 256, 283, 128};
  return ins[i];
}
__host__ __device__ constexpr int layer_out(int i) {
  const int outs[12] = {256, 256, 256, 256, 256, 256, 256, 256, 1, 256, 128, 3};
  return outs[i];
}

// Activation stash written by a training forward, per nt = num_tiles(P) tiles (bytes):
//   [8][nt][64 KB] h1..h8 | [nt][64 KB] features | [nt][32 KB] colour hidden | [8][nt][4 KB] ReLU bit masks of h1..h8
// A mask tile is [2 column halves][128 rows][4 words]; word g of (half hh, row r) covers columns hh*128 + g*32 .. +31 in
// the order of the packed bf16 pairs: bit i = column 2i, bit 16+i = column 2i+1 of the group (i < 16).
__host__ __device__ constexpr long long stash_off_feat(long long nt) { return 8 * nt * 65536; }
__host__ __device__ constexpr long long stash_off_c(long long nt) { return 9 * nt * 65536; }
__host__ __device__ constexpr long long stash_off_mask(long long nt) { return 9 * nt * 65536 + nt * 32768; }
__host__ __device__ constexpr long long stash_total_bytes(long long nt) { return stash_off_mask(nt) + 8 * nt * 4096; }
// 16 packed bf16 pairs (post-ReLU, i.e. >= +0) -> 32 "is positive" bits in the order above
__device__ __forceinline__ uint32_t relu_bits16(const uint32_t* pk) {
  uint32_t m = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) m = (m >> 1) | ((pk[i] + 0x7FFF7FFFu) & 0x80008000u);   // 0x0001..0x7F80 + 0x7FFF sets bit 15
  return m;
}
// the 0xFFFF-per-half AND mask of packed pair i from such a word
// (bit i -> bit 7, bit 16+i -> bit 23, then a byte permute in sign-replicate mode: 3 instructions with the AND)
__device__ __forceinline__ uint32_t relu_bits_expand(uint32_t m, int i) {
  const uint32_t t = i <= 7 ? m << (7 - i) : m >> (i - 7);
  uint32_t r;
  asm("prmt.b32 %0, %1, %1, 0xAA88;" : "=r"(r) : "r"(t));
  return r;
}

// how a tile's points are produced: the public dfn_point_source (include/deepfields_b200.h)
using PointSource = dfn_point_source;

__device__ __forceinline__ void load_point(const PointSource& s, long long p, float (&x)[3], float (&d)[3]) {
  if (s.mode == 0) {
#pragma unroll
    for (int c = 0; c < 3; ++c) { x[c] = s.pos[p * 3 + c]; d[c] = s.view[p * 3 + c]; }
  } else if (s.mode == 1) {
    // (a 64-bit division costs ~100 instructions on the epilogue's critical path; indices almost always fit 32 bits)
    const long long ray = (p >> 32) == 0 ? static_cast<long long>(static_cast<uint32_t>(p) / static_cast<uint32_t>(s.S)) : p / s.S;
    const float zz = s.z[p];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      d[c] = s.rays_d[ray * 3 + c];
      x[c] = __fadd_rn(s.rays_o[ray * 3 + c], __fmul_rn(d[c], zz));
    }
  } else {
    const long long g = s.p_offset + p;
    const long long RR = static_cast<long long>(s.R) * s.R;
    if (((g | RR) >> 32) == 0) {
      const uint32_t g32 = static_cast<uint32_t>(g), R32 = static_cast<uint32_t>(s.R);
      const uint32_t i0 = g32 / static_cast<uint32_t>(RR), rem = g32 - i0 * static_cast<uint32_t>(RR);
      const uint32_t i1 = rem / R32;
      x[0] = s.lin[i0]; x[1] = s.lin[i1]; x[2] = s.lin[rem - i1 * R32];
    } else {
      x[0] = s.lin[g / RR]; x[1] = s.lin[(g / s.R) % s.R]; x[2] = s.lin[g % s.R];
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) d[c] = s.view[c];
  }
}

// Positional encoding of one point into `vals` in the reference column order
// [x | sin(x_d 2^k pi) d-major,k-minor | cos(...)] (model.py:42-53).  The reference angle at level k is
// exactly 2^k * fl(x * fl(pi)) (power-of-two scaling is exact), so sin/cos at k>0 follow from the
// level-0 pair by the double-angle recurrence.
template <int L>
__device__ __forceinline__ void posenc_point(const float (&x)[3], float* vals /* 3 + 6L */) {
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    vals[d] = x[d];
    float s, c;
    sincosf(__fmul_rn(x[d], 3.14159274101257324f), &s, &c);
#pragma unroll
    for (int k = 0; k < L; ++k) {
      vals[3 + d * L + k] = s;
      vals[3 + 3 * L + d * L + k] = c;
      const float s2 = 2.0f * s * c;
      c = 1.0f - 2.0f * s * s;
      s = s2;
    }
  }
}

// Write this thread's half of the PE tiles of row `r`: thread half hh owns columns [32hh,32hh+32) of
// the 64-wide pos-PE and columns [16hh,16hh+16) of the 32-wide dir-PE.
__device__ __forceinline__ void write_pe_row(uint8_t* tile_pos /* 8 chunks */, uint8_t* tile_dir /* 4 chunks or null */,
                                             int r, int hh, const float (&x)[3], const float (&d)[3]) {
  if (tile_pos) {
    float v[64];
    posenc_point<kPosL>(x, v);
    v[63] = 0.f;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      uint32_t w[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float a = hh ? v[32 + g * 8 + 2 * i] : v[g * 8 + 2 * i];
        const float b = hh ? v[32 + g * 8 + 2 * i + 1] : v[g * 8 + 2 * i + 1];
        w[i] = pack_bf16x2(a, b);
      }
      *reinterpret_cast<uint4*>(tile_pos + (hh * 4 + g) * kChunkBytes + r * 16) = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
  if (tile_dir) {
    float v[32];
    posenc_point<kDirL>(d, v);
#pragma unroll
    for (int i = kDirDim; i < 32; ++i) v[i] = 0.f;
#pragma unroll
    for (int g = 0; g < 2; ++g) {
      uint32_t w[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float a = hh ? v[16 + g * 8 + 2 * i] : v[g * 8 + 2 * i];
        const float b = hh ? v[16 + g * 8 + 2 * i + 1] : v[g * 8 + 2 * i + 1];
        w[i] = pack_bf16x2(a, b);
      }
      *reinterpret_cast<uint4*>(tile_dir + (hh * 2 + g) * kChunkBytes + r * 16) = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
}

}  // namespace mlp
