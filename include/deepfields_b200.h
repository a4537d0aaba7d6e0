/* libdeepfields_b200.so -- C ABI of the B200-native NeRF render/train hot path.
 *
 * The reference (mohamedhaiham94/DeepFields-NeRF-) has no FFI of its own: its hot path is Python calling
 * torch ops.  Each entry point below replaces the reference Python function cited beside it; the
 * reference-side binding is the ctypes stub in INTEGRATION.md.  Conventions:
 *   - plain pointers + sizes only, all pointers are DEVICE pointers unless stated otherwise;
 *   - every call enqueues on `stream` (a cudaStream_t passed as void*), never synchronises, never
 *     allocates, never retains a pointer after return;
 *   - return 0 on success, non-zero on bad arguments / launch failure; dfn_last_error() gives the
 *     thread-local message;
 *   - there is no CPU fallback.
 */
#ifndef DEEPFIELDS_B200_H
#define DEEPFIELDS_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* dfn_last_error(void);
int dfn_abi_version(void);
/* SM budget of the MLP launches (dfn_mlp_forward*, dfn_mlp_backward*) of the CALLING THREAD: their persistent grids are sized to
 * n_sms instead of the whole device (0 = whole device), so that two of them can be resident at once on two streams. */
int dfn_set_sm_limit(int n_sms);

/* ---- sampling ---------------------------------------------------------------------------------- */
/* StratifiedSampler.__call__  scripts/sampling.py:10-21.  near/far [B], t_vals [S] (the sampler's linspace),
 * t_rand [B,S] (the torch.rand draw of :20) -> z [B,S].  Bit-exact with the reference's torch ops. */
int dfn_stratified_sample(const float* near, const float* far, const float* t_vals, const float* t_rand, int B, int S,
                          float* z_out, void* stream);
/* pts = o + d*z, dirs = d broadcast  scripts/sampling.py:24-31 / :43-48.  pts/dirs [B*S,3], either may be NULL. */
int dfn_points_on_rays(const float* rays_o, const float* rays_d, const float* z, int B, int S, float* pts, float* dirs,
                       void* stream);
/* sample_pdf  scripts/train.py:31-75 with the draw u [B,NS] passed in.  bins [B,NB] (or, bins_mid=1, z [B,NB+1]
 * whose mid-points are the bins, train.py:157); weights row r starts at weights + r*w_stride + w_offset and has
 * NB-1 entries (train.py:158 passes weights[...,1:-1]).  inds_out (int64 searchsorted result) and cdf_out
 * [B,NB] are optional.  The cdf prefixes equal a sequential double-precision cumsum bit for bit when the weights are
 * compositing weights (row sum <= 1); for arbitrary non-negative weights they are within one fp32 ulp of it. */
int dfn_sample_pdf(const float* bins, int bins_mid, const float* weights, long long w_stride, int w_offset,
                   const float* u, int B, int NB, int NS, float* samples, long long* inds_out, float* cdf_out,
                   void* stream);
/* scripts/train.py:56-73 on a caller-supplied cdf [B,NB] (parity tier A: bit-exact indices). */
int dfn_sample_pdf_with_cdf(const float* bins, const float* cdf, const float* u, int B, int NB, int NS, float* samples,
                            long long* inds_out, void* stream);
/* torch.sort(cat([z_vals, z_samples]))  scripts/sampling.py:40.  NA+NB <= 256. */
int dfn_merge_sort_z(const float* z_coarse, int NA, const float* z_fine, int NB, int B, float* z_out, void* stream);
/* sample_pdf on (mid-points of z, weights[:, 1:-1]) + sort(cat(z, samples)) in one launch (scripts/train.py:157-158 and
 * scripts/sampling.py:40): z [B,NA], weights [B,NA] = the full coarse weights, u [B,NS]; z_all [B,NA+NS]; z_new [B,NS] or NULL.
 * Bit-identical to dfn_sample_pdf followed by dfn_merge_sort_z. */
int dfn_sample_merge(const float* z, const float* weights, const float* u, int B, int NA, int NS, float* z_all, float* z_new,
                     void* stream);

/* ---- in-kernel random draws (csrc/rng.cuh: Philox4x32-10, draw = f(seed, *step_ptr, stream id, item)) -----------------
 * The reference draws inside sampling.py:20 (torch.rand), train.py:53 (torch.rand) and model.py:162 (randn_like); the entry
 * points above take those draws as tensors (parity tests inject the reference's), the ones below make them in the kernel:
 * no draw ever reaches HBM, CUDA-graph replays advance through the step counter in device memory (step_ptr = the counter
 * dfn_adamw_step increments), ranks use different seeds. */
int dfn_stratified_sample_rng(const float* near, const float* far, const float* t_vals, unsigned long long seed,
                              const int* step_ptr, int B, int S, float* z_out, void* stream);
int dfn_sample_merge_rng(const float* z, const float* weights, unsigned long long seed, const int* step_ptr, int B, int NA,
                         int NS, float* z_all, float* z_new, void* stream);
/* what the kernels draw, written out: uniforms [n_items,4], normals [n_items] (either may be NULL); rng_stream: 1 jitter,
 * 2 inverse-CDF draws, 3 / 4 sigma noise of the coarse / fine network */
int dfn_philox_dump(unsigned long long seed, const int* step_ptr, unsigned rng_stream, long long n_items, float* uniforms,
                    float* normals, void* stream);

/* ---- encoding ---------------------------------------------------------------------------------- */
/* OptimizedPositionalEncoding.forward  scripts/model.py:32-53.  x [P,3], freqs [L] -> out [P, 3*inc + 6L]. */
int dfn_posenc(const float* x, const float* freqs, long long P, int L, int include_input, float* out, void* stream);

/* ---- compositing ------------------------------------------------------------------------------- */
/* raw2outputs  scripts/train.py:78-106 (dist_last 1e10; train_no_amp.py:84 uses 1e6).  rgb [B,S,3], sigma [B,S]
 * fp32 (in_bf16=0) or bf16 (in_bf16=1); z [B,S].  Outputs (each optional): rgb_map [B,3], weights [B,S],
 * acc [B], depth [B]. */
int dfn_composite_forward(const void* rgb, const void* sigma, const float* z, int B, int S, int in_bf16, int white_bg,
                          float dist_last, float* rgb_map, float* weights, float* acc, float* depth, void* stream);
/* autograd of the above: d_rgb_map [B,3] (+ optional d_weights [B,S]) -> d_rgb [B,S,3], d_sigma [B,S] fp32. */
int dfn_composite_backward(const void* rgb, const void* sigma, const float* z, int B, int S, int in_bf16, int white_bg,
                           float dist_last, const float* d_rgb_map, const float* d_weights, float* d_rgb,
                           float* d_sigma, void* stream);
/* compositing + MSELoss(mean) + their backward in one pass (train.py:154,179,182-184):
 * loss_sum[0] += sum of squared errors of this batch; d_rgb/d_sigma = d(loss_scale * mean sq err). */
int dfn_render_loss(const void* rgb, const void* sigma, const float* z, const float* target, int B, int S, int in_bf16,
                    int white_bg, float dist_last, float loss_scale, float* rgb_map, float* weights, float* loss_sum,
                    float* d_rgb, float* d_sigma, void* stream);

/* ---- vanilla NeRF MLP (OptimizedVanillaNeRF, scripts/model.py:56-187; hidden 256, pos_L 10, dir_L 4) ---- */
#define DFN_MLP_NUM_PARAMS 595844 /* flat fp32, reference named_parameters() order */
#define DFN_MLP_TILE 128

/* where the P points of a query come from */
typedef struct dfn_point_source {
  int mode;            /* 0: pos/view arrays; 1: rays (pos = o + d*z, view = d); 2: dense grid          */
  const float* pos;    /* mode 0: [P,3]                                                                   */
  const float* view;   /* mode 0: [P,3]; mode 2: [3] canonical view (extract_vol.py:291-292)             */
  const float* rays_o; /* mode 1: [B,3]                                                                   */
  const float* rays_d; /* mode 1: [B,3]                                                                   */
  const float* z;      /* mode 1: [B,S], P = B*S                                                          */
  int S;               /* mode 1                                                                          */
  const float* lin;    /* mode 2: [R] = torch.linspace(min,max,R) (extract_vol.py:71); point index g ->   */
  int R;               /*         (lin[g/R^2], lin[(g/R)%R], lin[g%R])  (ij meshgrid, :74-77)             */
  long long p_offset;  /* mode 2: flat grid index of point 0 (slab start)                                 */
} dfn_point_source;

/* sizes (bytes) of the buffers the caller allocates */
long long dfn_mlp_fwd_pack_bytes(void);
long long dfn_mlp_bwd_pack_bytes(void);
long long dfn_mlp_consts_bytes(void);
long long dfn_mlp_stash_bytes(long long P); /* activations kept by a training forward for the backward */
long long dfn_mlp_dz_bytes(long long P);    /* pre-activation gradients kept by dgrad for wgrad        */

/* fp32 parameters (flat, DFN_MLP_NUM_PARAMS) -> bf16 operand images for the kernels.  Called after
 * load_state_dict and after every optimiser step. */
int dfn_mlp_pack_weights(const float* params, void* fwd_pack, void* bwd_pack, float* consts, void* stream);
/* the same repack as a gather through a precomputed source map (the layout depends on the architecture only): build the map
 * once per device (dfn_mlp_pack_map_bytes() bytes of int32), then repack with it after every optimiser step */
long long dfn_mlp_pack_map_bytes(void);
int dfn_mlp_build_pack_map(int* map, void* stream);
int dfn_mlp_pack_weights_mapped(const float* params, const int* map, void* fwd_pack, void* bwd_pack, float* consts, void* stream);

/* OptimizedVanillaNeRF.forward  scripts/model.py:173-187 on P points: rgb [P,3], sigma [P] (fp32 values of the
 * bf16 computation).  noise [P] or NULL = the randn_like draw of model.py:162 (training mode).
 * stash != NULL keeps the activations the backward needs (dfn_mlp_stash_bytes). */
int dfn_mlp_forward(const dfn_point_source* src, long long P, const void* fwd_pack, const float* consts,
                    const float* noise, float* rgb, float* sigma, void* stash, void* stream);
/* the same in training mode with the sigma noise drawn in the kernel: noise(point p) = N(0,1)(seed, *step_ptr, rng_stream, p) */
int dfn_mlp_forward_rng(const dfn_point_source* src, long long P, const void* fwd_pack, const float* consts,
                        unsigned long long seed, const int* step_ptr, unsigned rng_stream, float* rgb, float* sigma, void* stash,
                        void* stream);
/* backward of the above w.r.t. the parameters: d_rgb [P,3], d_sigma [P] (fp32), rgb/sigma = forward outputs,
 * stash from the forward, dz = scratch of dfn_mlp_dz_bytes(P).  grads [DFN_MLP_NUM_PARAMS] fp32 is ACCUMULATED
 * into (zero it first). */
int dfn_mlp_backward(const dfn_point_source* src, long long P, const void* bwd_pack, const float* consts,
                     const float* rgb, const float* sigma, const float* d_rgb, const float* d_sigma, const void* stash,
                     void* dz, float* grads, void* stream);

/* the two halves of dfn_mlp_backward as separate launches (stage timing; same buffers): the data-gradient chain
 * (writes dz) and the weight/bias-gradient reduction (reads stash + dz, accumulates into grads). */
int dfn_mlp_backward_dgrad(long long P, const void* bwd_pack, const float* consts, const float* rgb, const float* sigma,
                           const float* d_rgb, const float* d_sigma, const void* stash, void* dz, void* stream);
int dfn_mlp_backward_wgrad(const dfn_point_source* src, long long P, const void* stash, const void* dz, float* grads,
                           void* stream);

/* dfn_mlp_backward as ONE launch: the data-gradient pipeline and the weight-gradient units run concurrently on disjoint
 * SMs and the dZ tiles go from one to the other through a ring that stays resident in L2 (csrc/mlp_bwd_fused.cu), so they
 * never make the 2 x 4 880 B/point round trip through HBM.  workspace = dfn_mlp_bwd_fused_ws_bytes(P) bytes (ring control
 * block, ring, 16 B/point of head gradients); same inputs, same accumulation into grads as dfn_mlp_backward. */
long long dfn_mlp_bwd_fused_ws_bytes(long long P);
int dfn_mlp_backward_fused(const dfn_point_source* src, long long P, const void* bwd_pack, const float* consts,
                           const float* rgb, const float* sigma, const float* d_rgb, const float* d_sigma, const void* stash,
                           void* workspace, float* grads, void* stream);

/* single 128 x N x K tcgen05 GEMM on operand tiles in the chunk layout (bring-up / descriptor self test) */
int dfn_umma_selftest(const void* a_tile, const void* b_tile, int N, int K, int a_mn_major, int b_mn_major,
                      unsigned a_lbo, unsigned a_sbo, unsigned b_lbo, unsigned b_sbo, unsigned a_kstep, unsigned b_kstep,
                      float* d_out, void* stream);


/* ---- dense grid / volume (scripts/extract_vol.py:57-206) ---------------------------------------- */
int dfn_grid_points(const float* lin, int R, long long start, long long count, float* out, void* stream);
int dfn_histogram(const float* v, long long n, float lo, float hi, int nbins, unsigned long long* hist, void* stream);
int dfn_pack_volume(const float* sigma, const float* rgb, long long n, float thr, unsigned char* occ, unsigned char* col,
                    void* stream);

/* ---- ray generation in front of the path (scripts/precompute_rays.py) ---------------------------- */
/* get_rays (precompute_rays.py:67-119): every pixel ray of M cameras in the flattened [M,H,W] order.  kinv = HOST pointer
 * to the 9 doubles of inv(K) (row-major); poses = DEVICE [M][4][4] doubles (camera-to-world, row-major); colmap applies
 * F = diag(1,-1,-1); fp64 arithmetic like the reference, fp32 results ([M*H*W,3] each). */
int dfn_generate_rays(const double* kinv, const double* poses, int M, int H, int W, int colmap, int normalize,
                      float* rays_o, float* rays_d, void* stream);
/* ray_cube_intersection_numpy (precompute_rays.py:10-64) on fp32 rays: intersects [N] u8 (taken before t_near is
 * clamped to 0), t_near, t_far [N]; numpy's NaN-propagating minimum / maximum. */
int dfn_ray_cube_intersect(const float* rays_o, const float* rays_d, long long n_rays, float cube_min, float cube_max,
                           unsigned char* intersects, float* t_near, float* t_far, void* stream);
/* precompute_save_ray (precompute_rays.py:206-228) without the .npz round trip: generate, intersect, keep the hits in
 * order.  images = DEVICE [M,H,W,3] u8 or null (then rgbs must be null); outputs sized for M*H*W rays; *count
 * (device int64) receives the number kept; workspace of dfn_precompute_rays_workspace_bytes(M*H*W) bytes. */
long long dfn_precompute_rays_workspace_bytes(long long n_rays);
int dfn_precompute_rays(const double* kinv, const double* poses, int M, int H, int W, int colmap, int normalize,
                        const unsigned char* images, float cube_min, float cube_max, float* rays_o, float* rays_d,
                        float* rgbs, float* t_near, float* t_far, long long* count, void* workspace, void* stream);

/* ---- occupied voxels as a point set, the step after extract_vol (scripts/post_process_vol.py:37-50, :77-107) ---- */
/* argwhere(occupancy) in C order -> points [N,3] fp64 = (index / (R-1)) * 2 - 1, rgbs [N,3] fp32 = rgb_volume / max(rgb_volume)
 * gathered at those voxels; aabb_min / aabb_max = HOST pointers to 3 doubles (closed box, slice_aabb) or both null.
 * occupancy [R^3] u8, rgb_volume [R^3,3] u8 on the device; outputs sized for R^3 points; *count (device int64) = N;
 * workspace of dfn_volume_points_workspace_bytes(R^3) bytes. */
long long dfn_volume_points_workspace_bytes(long long n_voxels);
/* the two halves of dfn_volume_points, for callers that size their outputs by the count: _count fills the workspace and
 * *count, _emit writes exactly that many points from the same workspace */
int dfn_volume_points_count(const unsigned char* occupancy, const unsigned char* rgb_volume, int R, const double* aabb_min,
                            const double* aabb_max, long long* count, void* workspace, void* stream);
int dfn_volume_points_emit(const unsigned char* occupancy, const unsigned char* rgb_volume, int R, double* points, float* rgbs,
                           const void* workspace, void* stream);
int dfn_volume_points(const unsigned char* occupancy, const unsigned char* rgb_volume, int R, const double* aabb_min,
                      const double* aabb_max, double* points, float* rgbs, long long* count, void* workspace, void* stream);

/* ---- early ray termination for the inference render (BASELINE config 5; compositing = scripts/train.py:78-106) ---- */
/* After the fine network has been queried on samples [s0, s0+C) of the n active rays (active [n] int32 ray ids or null =
 * identity; seg_rgb [n,C,3], seg_sigma [n,C] in list order): scatter them into rgb_full [B,S,3] / sigma_full [B,S],
 * T_run[ray] *= prod(1 - alpha + 1e-10) over the segment (train.py:89-95, last interval dist_last), alive[k] = T > eps. */
int dfn_ert_update(const int* active, int n, const float* seg_rgb, const float* seg_sigma, const float* z, int S, int s0, int C,
                   float dist_last, float eps, float* rgb_full, float* sigma_full, float* T_run, unsigned char* alive,
                   void* stream);
/* ordered compaction: idx_out = idx_in[flags != 0] (idx_in null = the positions themselves); *count device int64 */
long long dfn_compact_indices_workspace_bytes(long long n);
int dfn_compact_indices(const unsigned char* flags, const int* idx_in, long long n, int* idx_out, long long* count,
                        void* workspace, void* stream);
/* dfn_compact_indices without a workspace argument.  No counterpart in the reference (which has no compaction and no C
 * interface); it exists for hosts that cannot give this library device memory beyond the input and output buffers: a render
 * thread inside an engine whose arena already owns the whole GPU (cudaMalloc fails there), a frame loop that cannot afford the
 * device-wide stall of cudaMalloc / cudaFree, a plug-in that is handed buffers and may not create any.  The scratch
 * (573 696 bytes) is a fixed array inside the library image: every GPU on which the library gets loaded carries it until the
 * process ends, whether this function is called or not.  Limits and rules:
 *   - n <= DFN_COMPACT_STATIC_MAX_N (2^22 flags, enough for the 1920 x 1080 rays of a frame); a longer list is refused with an
 *     error message, nothing is launched;
 *   - the call itself allocates nothing (no cudaMalloc / cudaFree, no allocating library routine); flags, idx_in, idx_out and
 *     count are still the caller's;
 *   - the first call on each GPU looks up the array's device address (cudaGetSymbolAddress, with that GPU current) and keeps
 *     it; make that first call during start-up, before any stream capture begins;
 *   - all calls on one GPU share the one array: two calls whose kernels may overlap on the GPU (different streams without an
 *     ordering between them, or two host threads) corrupt each other's counts -- order them on one stream or with events.
 *     The array keeps whatever the last call wrote; the function does not rely on it (every word it reads it has written
 *     earlier in the same call), so no clearing is needed between calls.
 * Output is identical to dfn_compact_indices. */
#define DFN_COMPACT_STATIC_MAX_N (1LL << 22)
long long dfn_compact_indices_static_max_n(void); /* DFN_COMPACT_STATIC_MAX_N as the loaded library was built */
int dfn_compact_indices_static(const unsigned char* flags, const int* idx_in, long long n, int* idx_out, long long* count,
                               void* stream);

/* ---- optimiser (scripts/train.py:244-259, :321-322) ---------------------------------------------- */
/* AdamW on a flat buffer; lr_table/beta1_table [table_len] hold the OneCycleLR schedule, *step_ptr (device int,
 * 0-based) selects the entry and is incremented. */
int dfn_adamw_step(float* p, const float* g, float* m, float* v, long long n, const float* lr_table,
                   const float* beta1_table, int* step_ptr, int table_len, float beta2, float eps, float weight_decay,
                   float grad_scale, void* stream);
/* the same update that also refreshes an fp16 copy of the parameters (the operands of the Instant-NGP kernels) and is skipped,
 * step counter included, when *skip_flag != 0 (nullable); dfn_nonfinite_flag sets flag[0] = 1 iff a gradient is inf / NaN -
 * the guard torch.amp.GradScaler.step applies in scripts/train.py:262-323 */
int dfn_adamw_step_shadow(float* p, const float* g, float* m, float* v, long long n, const float* lr_table,
                          const float* beta1_table, int* step_ptr, int table_len, float beta2, float eps, float weight_decay,
                          float grad_scale, void* shadow_f16, const int* skip_flag, void* stream);
int dfn_nonfinite_flag(const float* g, long long n, int* flag, void* stream);

/* ---- Instant-NGP variant (scripts/instantngp.py:7-104; arithmetic of tiny-cuda-nn 74ea173d restated, parity unpinned) ---- */
#define DFN_NGP_LEVELS 16
/* Level table of the reference's HashGrid config (instantngp.py:23-33: 16 levels x 2 features, base resolution 16,
 * per-level scale 1.5, log2_hashmap_size 15/17/19 for small/medium/large, :13-19).  size/offset count 2-feature entries. */
typedef struct dfn_hashgrid {
  float scale[DFN_NGP_LEVELS];
  unsigned res[DFN_NGP_LEVELS];
  unsigned size[DFN_NGP_LEVELS];
  unsigned offset[DFN_NGP_LEVELS];
  long long n_entries;
  int log2_hashmap_size;
} dfn_hashgrid;
/* host only (no GPU needed): fills *g, returns the total number of entries (params = 2 x entries) or -1 */
long long dfn_hashgrid_init(dfn_hashgrid* g, int log2_hashmap_size);
/* self.pos_encoding(pos), instantngp.py:77: pos [P,3] fp32 in [0,1]^3, table fp16 [n_entries,2] -> out fp16 [P,32]
 * (feature = 2*level + f) */
int dfn_hashgrid_forward(const float* pos, long long P, const void* table_f16, const dfn_hashgrid* g, void* out_f16, void* stream);
/* its backward: d_table [n_entries,2] fp32 += grad_scale * scatter(d_out fp16 [P,32]); the caller zeroes d_table */
int dfn_hashgrid_backward(const float* pos, long long P, const void* d_out_f16, const dfn_hashgrid* g, float grad_scale,
                          float* d_table, void* stream);
/* self.dir_encoding(viewdirs), instantngp.py:98: SphericalHarmonics degree 4, dirs [P,3] fp32 in [0,1] -> fp16 [P,16] */
int dfn_sh4_encode(const float* dirs, long long P, void* out_f16, void* stream);
/* InstantNGP.forward / query_density in one launch (instantngp.py:66-104), inference: dmlp_w fp16 [64*32 + 16*64],
 * cmlp_w fp16 [64*32 + 64*64 + 16*64] (tcnn FullyFusedMLP layout: row-major [out,in] per layer, no biases, last layer
 * padded to 16 rows); dirs_stride 3 = per-point directions, 0 = one canonical view (extract_vol.py:100);
 * features fp16 [P,16] (sigma = column 0) and/or rgb fp16 [P,3]; either output may be null. */
int dfn_ngp_forward(const float* pos, const float* dirs, int dirs_stride, long long P, const void* table_f16,
                    const dfn_hashgrid* g, const void* dmlp_w_f16, const void* cmlp_w_f16, void* features_f16, void* rgb_f16,
                    void* stream);
/* the same query with fp32 outputs for the compositing kernels of the train step: rgb [P,3], sigma [P] (= feature 0) hold the
 * fp16 values.  dirs_group: that many consecutive points share one direction (the samples of a ray; 1 = per point). */
int dfn_ngp_forward_f32(const float* pos, const float* dirs, int dirs_stride, int dirs_group, long long P, const void* table_f16,
                        const dfn_hashgrid* g, const void* dmlp_w_f16, const void* cmlp_w_f16, float* rgb, float* sigma, void* stream);
/* training backward of InstantNGP.forward in ONE launch (csrc/ngp_train.cu): the forward is recomputed in registers (nothing is
 * stashed), d loss / d rgb [P,3] and d loss / d sigma [P] (fp32) are pushed through both tiny MLPs with mma.sync and scattered
 * into the fp32 gradients of the hash table [n_entries,2], the density MLP [3072] and the colour MLP [7168] (all ACCUMULATED).
 * loss_scale (tcnn's default 128) keeps the fp16 gradient fragments in range; the results are unscaled. */
int dfn_ngp_backward(const float* pos, const float* dirs, int dirs_stride, int dirs_group, long long P, const void* table_f16,
                     const dfn_hashgrid* g, const void* dmlp_w_f16, const void* cmlp_w_f16, const float* d_rgb, const float* d_sigma,
                     float loss_scale, float* d_table, float* d_dmlp, float* d_cmlp, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DEEPFIELDS_B200_H */
