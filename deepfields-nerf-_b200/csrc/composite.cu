// Alpha compositing (reference scripts/train.py:78-106 raw2outputs), forward, backward and a fused
// forward + MSE + backward ("render loss") kernel.  One warp per ray; S <= 256 samples, each lane owns
// a contiguous run of ceil(S/32) samples so the exclusive transmittance product is a local product
// followed by one warp shuffle scan.  HBM-bound: every input element is read once (coalesced, staged
// through shared memory), every output written once.
#include "dfn_common.cuh"

namespace {

constexpr int kWarpsPerBlock = 4;
constexpr int kMaxS = 256;
constexpr int kMaxPerLane = kMaxS / 32;

struct CompositeArgs {
  const void* rgb;      // [B,S,3] fp32 or bf16
  const void* sigma;    // [B,S]   fp32 or bf16
  const float* z;       // [B,S]
  int B, S;
  int in_bf16;
  int white_bg;
  float dist_last;
  // forward outputs (nullable)
  float* rgb_map;   // [B,3]
  float* weights;   // [B,S]
  float* acc;       // [B]
  float* depth;     // [B]
  // backward inputs
  const float* d_rgb_map;  // [B,3]   (mode 1)
  const float* d_weights;  // [B,S]   nullable
  const float* target;     // [B,3]   (mode 2: fused loss)
  float loss_scale;        // dL/d(rgb_map) = loss_scale * 2 (rgb_map - target) / (3B)
  float* loss_sum;         // [1] += sum over rays of squared error (mode 2)
  // backward outputs
  float* d_rgb;    // [B,S,3] fp32
  float* d_sigma;  // [B,S]   fp32
};

__device__ __forceinline__ float ld_in(const void* p, size_t i, int bf16) {
  return bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]) : reinterpret_cast<const float*>(p)[i];
}

// MODE 0: forward only; 1: backward from d_rgb_map (+d_weights); 2: forward + MSE + backward
template <int MODE>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) composite_kernel(CompositeArgs a) {
  extern __shared__ float smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = a.S;
  float* sz = smem + warp * (5 * S);
  float* ss = sz + S;
  float* sc = ss + S;  // 3S
  const int n = (S + 31) >> 5;

  for (int ray = blockIdx.x * kWarpsPerBlock + warp; ray < a.B; ray += gridDim.x * kWarpsPerBlock) {
    const size_t base = static_cast<size_t>(ray) * S;
    for (int i = lane; i < S; i += 32) {
      sz[i] = a.z[base + i];
      ss[i] = ld_in(a.sigma, base + i, a.in_bf16);
    }
    for (int i = lane; i < 3 * S; i += 32) sc[i] = ld_in(a.rgb, base * 3 + i, a.in_bf16);
    __syncwarp();

    float alpha[kMaxPerLane], ex[kMaxPerLane], delta[kMaxPerLane], keep[kMaxPerLane], T[kMaxPerLane];
    float local = 1.0f;
    const int i0 = lane * n;
#pragma unroll
    for (int k = 0; k < kMaxPerLane; ++k) {
      const int i = i0 + k;
      if (k < n && i < S) {
        const float d = (i + 1 < S) ? __fsub_rn(sz[i + 1], sz[i]) : a.dist_last;
        const float e = expf(-__fmul_rn(ss[i], d));
        delta[k] = d;
        ex[k] = e;
        alpha[k] = 1.0f - e;
        keep[k] = __fadd_rn(1.0f - alpha[k], 1e-10f);
        local *= keep[k];
      } else {
        delta[k] = 0.f; ex[k] = 1.f; alpha[k] = 0.f; keep[k] = 1.f;
      }
    }
    // exclusive product scan over lanes
    float incl = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl *= t;
    }
    float run = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane == 0) run = 1.0f;

    float w[kMaxPerLane];
    float m0 = 0.f, m1 = 0.f, m2 = 0.f, accv = 0.f, dep = 0.f;
#pragma unroll
    for (int k = 0; k < kMaxPerLane; ++k) {
      const int i = i0 + k;
      T[k] = run;
      w[k] = alpha[k] * run;
      run *= keep[k];
      if (k < n && i < S) {
        m0 += w[k] * sc[3 * i];
        m1 += w[k] * sc[3 * i + 1];
        m2 += w[k] * sc[3 * i + 2];
        accv += w[k];
        dep += w[k] * sz[i];
      } else {
        w[k] = 0.f;
      }
    }
    m0 = warp_sum(m0); m1 = warp_sum(m1); m2 = warp_sum(m2);
    accv = warp_sum(accv); dep = warp_sum(dep);
    // Per-sample outputs leave through shared memory: a lane owns a CONTIGUOUS run of samples (that is what makes the
    // transmittance scan cheap), so direct stores would hit 32 different sectors per instruction.  sz is dead after this
    // point (depth is summed), ss after the alphas: they stage weights / d_sigma; the colours stay until d_rgb is out.
    if (MODE != 1 && a.weights) {
      __syncwarp();
#pragma unroll
      for (int k = 0; k < kMaxPerLane; ++k)
        if (k < n && i0 + k < S) sz[i0 + k] = w[k];
      __syncwarp();
      for (int i = lane; i < S; i += 32) a.weights[base + i] = sz[i];
    }
    if (a.white_bg) { m0 += 1.0f - accv; m1 += 1.0f - accv; m2 += 1.0f - accv; }
    if (MODE != 1 && lane == 0) {
      if (a.rgb_map) { a.rgb_map[ray * 3] = m0; a.rgb_map[ray * 3 + 1] = m1; a.rgb_map[ray * 3 + 2] = m2; }
      if (a.acc) a.acc[ray] = accv;
      if (a.depth) a.depth[ray] = dep;
    }
    if (MODE == 0) { __syncwarp(); continue; }

    float g0, g1, g2;
    if (MODE == 1) {
      g0 = a.d_rgb_map[ray * 3]; g1 = a.d_rgb_map[ray * 3 + 1]; g2 = a.d_rgb_map[ray * 3 + 2];
    } else {
      const float e0 = m0 - a.target[ray * 3], e1 = m1 - a.target[ray * 3 + 1], e2 = m2 - a.target[ray * 3 + 2];
      const float c = a.loss_scale * 2.0f / (3.0f * static_cast<float>(a.B));
      g0 = c * e0; g1 = c * e1; g2 = c * e2;
      if (lane == 0 && a.loss_sum) atomicAdd(a.loss_sum, e0 * e0 + e1 * e1 + e2 * e2);
    }
    const float gwb = a.white_bg ? (g0 + g1 + g2) : 0.f;
    float gw[kMaxPerLane];
    float lsum = 0.f;
#pragma unroll
    for (int k = 0; k < kMaxPerLane; ++k) {
      const int i = i0 + k;
      gw[k] = 0.f;
      if (k < n && i < S) {
        float g = g0 * sc[3 * i] + g1 * sc[3 * i + 1] + g2 * sc[3 * i + 2] - gwb;
        if (a.d_weights) g += a.d_weights[base + i];
        gw[k] = g;
        lsum += g * w[k];
      }
    }
    // d_rgb[i][c] = w[i] * g_c: weights -> shared memory (sz; any staged forward weights have been written out), then one
    // coalesced pass over the 3S floats
    __syncwarp();
#pragma unroll
    for (int k = 0; k < kMaxPerLane; ++k)
      if (k < n && i0 + k < S) sz[i0 + k] = w[k];
    __syncwarp();
    for (int j = lane; j < 3 * S; j += 32) {
      const int i = j / 3, c = j - 3 * i;
      a.d_rgb[base * 3 + j] = sz[i] * (c == 0 ? g0 : (c == 1 ? g1 : g2));
    }
    // exclusive suffix sum over lanes: suffix(lane) = sum over lanes > lane of lsum
    float incl_s = lsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float t = __shfl_down_sync(0xffffffffu, incl_s, o);
      if (lane + o < 32) incl_s += t;
    }
    float suf = __shfl_down_sync(0xffffffffu, incl_s, 1);
    if (lane == 31) suf = 0.f;
#pragma unroll
    for (int k = kMaxPerLane - 1; k >= 0; --k) {
      const int i = i0 + k;
      if (k < n && i < S) {
        const float dalpha = gw[k] * T[k] - suf / keep[k];
        ss[i] = dalpha * delta[k] * ex[k];
        suf += gw[k] * w[k];
      }
    }
    __syncwarp();
    for (int i = lane; i < S; i += 32) a.d_sigma[base + i] = ss[i];
    __syncwarp();
  }
}

int launch(const CompositeArgs& a, int mode, cudaStream_t st) {
  DFN_CHECK_ARG(a.S >= 1 && a.S <= kMaxS, "composite: S=%d outside [1,%d]", a.S, kMaxS);
  DFN_CHECK_ARG(a.B >= 0, "composite: negative B");
  if (a.B == 0) return DFN_OK;
  const int blocks_needed = (a.B + kWarpsPerBlock - 1) / kWarpsPerBlock;
  const int grid = blocks_needed < dfn_num_sms() * 8 ? blocks_needed : dfn_num_sms() * 8;
  const size_t smem = static_cast<size_t>(kWarpsPerBlock) * 5 * a.S * sizeof(float);
  if (mode == 0) composite_kernel<0><<<grid, kWarpsPerBlock * 32, smem, st>>>(a);
  else if (mode == 1) composite_kernel<1><<<grid, kWarpsPerBlock * 32, smem, st>>>(a);
  else composite_kernel<2><<<grid, kWarpsPerBlock * 32, smem, st>>>(a);
  DFN_CHECK_LAUNCH("composite");
  return DFN_OK;
}

}  // namespace

extern "C" int dfn_composite_forward(const void* rgb, const void* sigma, const float* z, int B, int S, int in_bf16,
                                     int white_bg, float dist_last, float* rgb_map, float* weights, float* acc,
                                     float* depth, void* stream) {
  DFN_CHECK_ARG(rgb && sigma && z, "composite_forward: null input");
  CompositeArgs a{};
  a.rgb = rgb; a.sigma = sigma; a.z = z; a.B = B; a.S = S; a.in_bf16 = in_bf16; a.white_bg = white_bg;
  a.dist_last = dist_last; a.rgb_map = rgb_map; a.weights = weights; a.acc = acc; a.depth = depth;
  return launch(a, 0, static_cast<cudaStream_t>(stream));
}

extern "C" int dfn_composite_backward(const void* rgb, const void* sigma, const float* z, int B, int S, int in_bf16,
                                      int white_bg, float dist_last, const float* d_rgb_map, const float* d_weights,
                                      float* d_rgb, float* d_sigma, void* stream) {
  DFN_CHECK_ARG(rgb && sigma && z && d_rgb_map && d_rgb && d_sigma, "composite_backward: null pointer");
  CompositeArgs a{};
  a.rgb = rgb; a.sigma = sigma; a.z = z; a.B = B; a.S = S; a.in_bf16 = in_bf16; a.white_bg = white_bg;
  a.dist_last = dist_last; a.d_rgb_map = d_rgb_map; a.d_weights = d_weights; a.d_rgb = d_rgb; a.d_sigma = d_sigma;
  return launch(a, 1, static_cast<cudaStream_t>(stream));
}

extern "C" int dfn_render_loss(const void* rgb, const void* sigma, const float* z, const float* target, int B, int S,
                               int in_bf16, int white_bg, float dist_last, float loss_scale, float* rgb_map,
                               float* weights, float* loss_sum, float* d_rgb, float* d_sigma, void* stream) {
  DFN_CHECK_ARG(rgb && sigma && z && target && d_rgb && d_sigma, "render_loss: null pointer");
  CompositeArgs a{};
  a.rgb = rgb; a.sigma = sigma; a.z = z; a.B = B; a.S = S; a.in_bf16 = in_bf16; a.white_bg = white_bg;
  a.dist_last = dist_last; a.target = target; a.loss_scale = loss_scale; a.rgb_map = rgb_map; a.weights = weights;
  a.loss_sum = loss_sum; a.d_rgb = d_rgb; a.d_sigma = d_sigma;
  return launch(a, 2, static_cast<cudaStream_t>(stream));
}
