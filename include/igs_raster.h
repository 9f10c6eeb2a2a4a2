/*
 * igs_raster.h — C-ABI of the B200-native differentiable Gaussian-splat rasterizer.
 *
 * This is the drop-in boundary for the one hot path of IGS's per-key-frame refine loop
 * (reference: /root/reference/infer_batch.py:279-324).  Every entry point below replaces one
 * reference interface; `DGR/` abbreviates
 * submodules/RaDe-GS/submodules/diff-gaussian-rasterization/ in the reference tree.
 *
 *   igs_rasterize_forward   <- CudaRasterizer::Rasterizer::forward   (DGR/cuda_rasterizer/rasterizer.h:31-63,
 *                                                                     DGR/cuda_rasterizer/rasterizer_impl.cu:254-425)
 *                              as called by RasterizeGaussiansCUDA    (DGR/rasterize_points.cu:35-133)
 *   igs_rasterize_backward  <- CudaRasterizer::Rasterizer::backward  (DGR/cuda_rasterizer/rasterizer.h:65-110,
 *                                                                     DGR/cuda_rasterizer/rasterizer_impl.cu:429-571)
 *                              as called by RasterizeGaussiansBackwardCUDA (DGR/rasterize_points.cu:135-246)
 *   igs_mark_visible        <- CudaRasterizer::Rasterizer::markVisible (DGR/cuda_rasterizer/rasterizer.h:24-29,
 *                                                                     DGR/cuda_rasterizer/rasterizer_impl.cu:176-188)
 *   igs_decode_binning      <- (test/inspection only) exposes what the reference keeps in BinningState /
 *                              ImageState (DGR/cuda_rasterizer/rasterizer_impl.h:64-86): sorted 64-bit keys,
 *                              point_list, tile ranges, n_contrib.
 *   igs_adam_step(_groups)  <- torch.optim.Adam(eps=1e-15) over the refine parameters
 *                              (igs/models/gaussian_model.py:348; infer_batch.py:323)
 *   igs_refine_loss         <- 0.8*L1 + 0.2*(1-SSIM) (infer_batch.py:301-305, igs/utils/loss_utils.py:16-62)
 *   igs_depth_normal_loss   <- depth-normal consistency loss of frame-0 training (RADE/train.py:143-157,
 *                              RADE/utils/graphics_utils.py:97-126)
 *   igs_activate_forward/backward <- exp / normalize / sigmoid activations (igs/models/gaussian_model.py:89-127)
 *   igs_densify_stats / igs_densify_gather <- densification statistics and the parameter / optimizer-state rebuild
 *                              of densify_and_prune (igs/models/gaussian_model.py:466-666,865-868)
 *
 * Conventions
 *   - All pointers are DEVICE pointers unless stated otherwise; float = IEEE fp32; layouts are the
 *     reference's: means3D [P][3], shs [P][M][3], scales [P][3], rotations [P][4] (r,x,y,z; NOT
 *     normalised inside, reference forward.cu:279), opacities [P], cov3D_precomp [P][6],
 *     colors_precomp [P][3], image maps [V][C][H][W].
 *   - A NULL pointer means "absent", exactly as the reference branches on nullptr
 *     (rasterizer_impl.cu:394,540; forward.cu:363,407; backward.cu:622,626).
 *   - V (num_views) generalises the reference's one-view call: viewmatrix/projmatrix/cam_pos hold
 *     V cameras ([V][16], [V][16], [V][3], reference row-vector convention) that share W, H and
 *     tan_fov; outputs and radii get a leading [V] dimension.  V == 1 is bit-for-bit the reference call.
 *     In backward, per-Gaussian gradients are SUMMED over the V views (one refine step on V views).
 *   - Scratch memory is owned by the caller and obtained through igs_alloc_fn callbacks, mirroring
 *     the reference's std::function<char*(size_t)> resize functionals (rasterize_points.cu:27-33).
 *     The three forward buffers are opaque, self-describing (header + (P,V,R,W*H)) and must be handed
 *     back unchanged to igs_rasterize_backward (reference: ctx.save_for_backward,
 *     diff_gaussian_rasterization_rade/__init__.py:98-100).
 *   - All work is enqueued on `stream` (a cudaStream_t passed as void*).  In exact mode the only host
 *     synchronisation in forward is one 4-byte D2H read of the instance count (the reference has the
 *     same one, rasterizer_impl.cu:354); in capacity mode (igs_forward_args.instance_capacity) there is none.
 *   - Return value: 0 on success, negative igs_status otherwise.  igs_last_error() gives a message.
 *     There is NO CPU fallback: without a CUDA device every compute entry point returns
 *     IGS_ERR_CUDA.
 */
#ifndef IGS_RASTER_H_INCLUDED
#define IGS_RASTER_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IGS_RASTER_ABI_VERSION 6

typedef enum igs_status {
  IGS_OK = 0,
  IGS_ERR_INVALID = -1,     /* bad argument (reference: AT_ERROR / std::runtime_error)            */
  IGS_ERR_CUDA = -2,        /* CUDA runtime error (reference: thrown only when debug=true)        */
  IGS_ERR_UNSUPPORTED = -3, /* e.g. prefiltered=true or M > 16                       */
  IGS_ERR_ALLOC = -4        /* allocator callback returned NULL                                    */
} igs_status;

/* Scratch allocator: must return a device pointer to at least `bytes` bytes, 256-B aligned,
 * valid until the matching backward has run.  `user` is passed through unchanged. */
typedef char* (*igs_alloc_fn)(void* user, size_t bytes);

typedef struct igs_forward_args {
  int P;              /* number of Gaussians                                        */
  int D;              /* active SH degree (0..3)                                    */
  int M;              /* SH coefficients per Gaussian stored in `shs` (<= 16)       */
  int V;              /* number of camera views (>= 1)                              */
  int width, height;  /* image size shared by all V views                           */
  const float* background;     /* [3]                                               */
  const float* means3D;        /* [P][3]                                            */
  const float* shs;            /* [P][M][3] or NULL                                 */
  const float* colors_precomp; /* [P][3] or NULL (V==1 only)                        */
  const float* opacities;      /* [P]                                               */
  const float* scales;         /* [P][3] or NULL                                    */
  float scale_modifier;
  const float* rotations;      /* [P][4] or NULL                                    */
  const float* cov3D_precomp;  /* [P][6] or NULL                                    */
  const float* viewmatrix;     /* [V][16]                                           */
  const float* projmatrix;     /* [V][16]                                           */
  const float* cam_pos;        /* [V][3]                                            */
  float tan_fovx, tan_fovy;
  float kernel_size;
  int prefiltered;             /* must be 0                            */
  int require_coord, require_depth;
  int debug;                   /* 1: synchronise + check after every stage          */
  /* outputs; every pixel of every requested map is written, maps that are NOT requested (require_coord /
   * require_depth = 0) are left untouched and must be pre-zeroed by the caller to read like the reference's */
  float* out_color;   /* [V][3][H][W] */
  float* out_coord;   /* [V][3][H][W] */
  float* out_mcoord;  /* [V][3][H][W] */
  float* out_depth;   /* [V][1][H][W] */
  float* out_mdepth;  /* [V][1][H][W] */
  float* out_alpha;   /* [V][1][H][W] */
  float* out_normal;  /* [V][3][H][W] */
  int* radii;         /* [V][P]       */
  /* Optional (NULL = whole image): DEVICE array [V][2] of tile-row windows {row_begin, row_end} in units of 16-pixel
   * tile rows.  View v is binned and blended only in tile rows [row_begin, row_end); pixels outside receive the
   * background, radii report visibility inside the window, and backward returns the gradient of exactly the rendered
   * pixels.  Per-pixel results inside the window are identical to the full render.  Used to shard ONE view over
   * several GPUs (igs_b200.refine.shard_units); the reference has no equivalent. */
  const int* tile_rows;
  /* Capacity mode (0 = exact mode, the reference's behaviour: one host read of the instance count R, reference
   * rasterizer_impl.cu:354).  With instance_capacity > 0 the call makes NO host round-trip: the binning buffer is sized
   * for `instance_capacity` instances, *num_rendered returns that capacity (hand it to igs_rasterize_backward as R), and
   * every launch reads the true count on the device.  If the true count exceeds the capacity the surplus instances (the
   * farthest Gaussians' tiles) are dropped and status[1] is set: the caller must discard the result and call again with
   * a larger capacity (igs_adam_step_groups can be given the flag so that the parameter update becomes a no-op on the
   * device).  status: optional DEVICE array of 2 words, written by every call: {true instance count, overflow flag}.
   * A forward + backward in capacity mode can be captured in a CUDA graph. */
  int instance_capacity;
  unsigned int* status;
} igs_forward_args;

/* Forward.  *num_rendered receives R = total number of (Gaussian, tile) instances over all views
 * (the reference's return value, rasterizer_impl.cu:424). */
int igs_rasterize_forward(const igs_forward_args* args,
                          igs_alloc_fn geom_alloc, void* geom_user,
                          igs_alloc_fn binning_alloc, void* binning_user,
                          igs_alloc_fn image_alloc, void* image_user,
                          void* stream, int* num_rendered);

typedef struct igs_backward_args {
  int P, D, M, V, R;
  int width, height;
  const float* background;
  const float* means3D;
  const float* shs;
  const float* colors_precomp;
  const float* scales;
  float scale_modifier;
  const float* rotations;
  const float* cov3D_precomp;
  const float* viewmatrix;
  const float* projmatrix;
  const float* cam_pos;
  float tan_fovx, tan_fovy;
  float kernel_size;
  const int* radii;            /* [V][P] from forward                                */
  const float* alphas;         /* out_alpha of forward [V][1][H][W]                   */
  const float* normalmap;      /* out_normal of forward [V][3][H][W]                  */
  char* geom_buffer;           /* the three forward scratch buffers                   */
  char* binning_buffer;
  char* image_buffer;
  /* cotangents; a NULL pointer means "identically zero" (autograd gave no gradient) */
  const float* dL_dout_color;  /* [V][3][H][W] */
  const float* dL_dout_coord;
  const float* dL_dout_mcoord;
  const float* dL_dout_depth;
  const float* dL_dout_mdepth;
  const float* dL_dout_alpha;
  const float* dL_dout_normal;
  /* gradients (every element is written; no pre-zeroing required) */
  float* dL_dmeans2D;   /* [V][P][3]  (.z = accumulated |grad|)            */
  float* dL_dcolors;    /* [P][3]     only meaningful with colors_precomp (V==1)       */
  float* dL_dopacity;   /* [P]                                                         */
  float* dL_dmeans3D;   /* [P][3]                                                      */
  float* dL_dcov3D;     /* [P][6]                                                      */
  float* dL_dsh;        /* [P][M][3] or NULL                                           */
  float* dL_dscales;    /* [P][3] or NULL                                              */
  float* dL_drotations; /* [P][4] or NULL                                              */
  int require_coord, require_depth;
  int debug;
} igs_backward_args;

int igs_rasterize_backward(const igs_backward_args* args,
                           igs_alloc_fn temp_alloc, void* temp_user,
                           void* stream);

/* present[i] = (view-space z of means3D[i] > 0.2)  — reference in_frustum(), auxiliary.h:155-180 */
int igs_mark_visible(int P, const float* means3D, const float* viewmatrix,
                     const float* projmatrix, uint8_t* present, void* stream);

/* Inspection of the opaque forward buffers (parity tests).  Any output pointer may be NULL.
 *   keys       [R]  u64  (tile << 32) | depth bits, in sorted order      (reference point_list_keys)
 *   point_list [R]  u32  Gaussian index per sorted instance               (reference point_list)
 *   ranges     [V*tiles] uint2 {start,end} into the sorted list           (reference ranges)
 *   n_contrib  [V][2][H*W] u32 last / median contributor                  (reference n_contrib)
 *   tiles_touched [V][P] u32, depths [V][P] f32, means2D [V][P][2], conic_opacity [V][P][4], rgb [V][P][3]
 */
typedef struct igs_decode_out {
  uint64_t* keys;
  uint32_t* point_list;
  uint32_t* ranges;
  uint32_t* n_contrib;
  uint32_t* tiles_touched;
  float* depths;
  float* means2D;
  float* conic_opacity;
  float* rgb;
  float* cov3D;          /* [P][6] */
  float* view_points;    /* [V][P][3] */
  float* camera_planes;  /* [V][P][6] */
  float* ray_planes;     /* [V][P][2] */
  float* ts;             /* [V][P]    */
  float* normals;        /* [V][P][3] */
} igs_decode_out;

int igs_decode_binning(int P, int V, int R, int width, int height,
                       const char* geom_buffer, const char* binning_buffer,
                       const char* image_buffer, const igs_decode_out* out, void* stream);

/* Fused Adam over one parameter group (torch.optim.Adam semantics, no weight decay, no amsgrad):
 *   m = b1*m + (1-b1)*g ; v = b2*v + (1-b2)*g*g ;
 *   p -= lr/(1-b1^t) * m / (sqrt(v)/sqrt(1-b2^t) + eps)                                        */
int igs_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq,
                  size_t n, float lr, float beta1, float beta2, float eps, int step, void* stream);

/* Adam over `ngroups` (<= 8) CONTIGUOUS parameter groups of one flat buffer in ONE launch: group k covers
 * sizes[k] floats after the previous group and uses learning rate lrs[k] (HOST arrays).  Same arithmetic as
 * igs_adam_step; replaces the per-group loop of torch.optim.Adam over GaussianModel.training_setup's groups
 * (igs/models/gaussian_model.py:265-348). */
int igs_adam_step_groups(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int ngroups,
                         const size_t* sizes, const float* lrs, float beta1, float beta2, float eps, int step,
                         void* stream);
/* The same step guarded on the DEVICE: if skip_if_nonzero != NULL and *skip_if_nonzero != 0 when the kernel runs, nothing
 * is updated (capacity mode of igs_rasterize_forward: pass &status[1]).  step_dev (optional DEVICE int): when given, the
 * bias corrections use *step_dev + 1 and the kernel increments it if it updates — the step count then lives on the device
 * and a skipped step is not counted, which is what a CUDA-graph replay of the iteration needs; `step` is ignored then. */
int igs_adam_step_groups_guarded(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int ngroups,
                                 const size_t* sizes, const float* lrs, float beta1, float beta2, float eps, int step,
                                 const unsigned int* skip_if_nonzero, int* step_dev, void* stream);

/* Fused refine loss over V images [V][3][H][W] (replaces infer_batch.py:301-305 with igs/utils/loss_utils.py:16-62):
 *   *loss_sum = sum_v [ lambda * mean|pred_v - gt_v| + (1 - lambda) * (1 - mean SSIM(pred_v, gt_v)) ]
 *   dL_dpred  = grad_scale * d(*loss_sum) / d(pred)     (11x11 Gaussian window, sigma 1.5, zero padding)
 * grad_scale is the upstream gradient of loss_sum (1/total_views when the step's loss is the mean over views);
 * loss_sum itself is NOT scaled.  loss_sum is a DEVICE float; temp_alloc provides 36*V*H*W bytes (+ a few KB).
 * loss_rows: optional HOST array [V][2] (NULL = every row; V <= 64 otherwise): only the loss terms (L1 and SSIM-map
 * pixels) of rows [loss_rows[2v], loss_rows[2v+1]) of view v are summed — the share of a rank that renders a band of
 * the view (igs_forward_args.tile_rows must then cover these rows +- 5, the SSIM window radius). */
int igs_refine_loss(const float* pred, const float* gt, int V, int height, int width, float lambda_l1,
                    float grad_scale, const int* loss_rows, float* loss_sum, float* dL_dpred, igs_alloc_fn temp_alloc,
                    void* temp_user, void* stream);

/* Depth-normal consistency loss of RaDe-GS frame-0 training, value and gradients in one launch (replaces
 * RADE/train.py:143-157 with RADE/utils/graphics_utils.py:97-126 depths_double_to_points / point_double_to_normal):
 *   *loss_sum = sum_v [ (1 - depth_ratio) * mean(1 - <normal_v, n(map0_v)>) + depth_ratio * mean(1 - <normal_v, n(map1_v)>) ]
 * where n(.) is the normalised cross product of the central differences of the back-projected points (zero on the
 * one-pixel border).  use_coord == 0: map0 / map1 are the expected / median DEPTH maps [V][H][W] and points are
 * depth * K^-1 (x + 0.5, y + 0.5, 1) with fx = W / (2 tan_fovx); use_coord != 0: they are the expected / median COORD
 * maps [V][3][H][W] (RADE/train.py:150-153).  dL_dmap0 / dL_dmap1 / dL_dnormal (same shapes as the inputs) receive
 * grad_scale * d(*loss_sum)/d(.); *loss_sum (DEVICE float) is not scaled.  temp_alloc provides a few KB. */
int igs_depth_normal_loss(int use_coord, const float* map0, const float* map1, const float* normal, int V, int height,
                          int width, float tan_fovx, float tan_fovy, float depth_ratio, float grad_scale,
                          float* loss_sum, float* dL_dmap0, float* dL_dmap1, float* dL_dnormal, igs_alloc_fn temp_alloc,
                          void* temp_user, void* stream);

/* The exchange step of the view-sharded refine iteration as ONE kernel over NVLink / NVSwitch multicast (no counterpart
 * in the reference, which has no multi-GPU refine path; it replaces ncclAllReduce + the replicated Adam step of
 * BASELINE's schedule).  Rank r owns floats [begin, end) of the packed record of n floats; for them the kernel
 * reduces the gradient over all ranks with multimem.ld_reduce (grad_mc: MULTICAST address of the symmetric gradient
 * buffers), applies the Adam step of igs_adam_step_groups (same arithmetic; exp_avg / exp_avg_sq are touched in [begin,
 * end) only) to `param` (this rank's own parameter buffer) and broadcasts the result with multimem.st (param_mc:
 * MULTICAST address of the symmetric parameter buffers).  The 4 floats at grad[n .. n+4) of every rank hold {partial
 * loss, 1.0f if this rank's forward overflowed its instance capacity else 0, 0, 0}: *loss_out receives the sum of the
 * partial losses, *skipped_out is 1 if any rank overflowed — then NO rank updates anything.  n, begin, end: multiples
 * of 4; all buffers 16-byte aligned.  The caller orders the ranks: every rank's gradients (and tail slot) must be
 * complete before any rank launches, and all launches complete before parameters are read or gradients rewritten
 * (a cross-rank stream barrier before and after; igs_b200/refine.py uses the symmetric-memory signal pads). */
int igs_fused_reduce_adam(const float* grad_mc, float* param_mc, const float* param, float* exp_avg, float* exp_avg_sq,
                          size_t n, size_t begin, size_t end, int ngroups, const size_t* sizes, const float* lrs,
                          float beta1, float beta2, float eps, int step, float* loss_out, unsigned int* skipped_out,
                          void* stream);

/* igs_depth_normal_loss for callers that cannot allocate (CUDA-graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments,
 * arithmetic and results as igs_depth_normal_loss, minus the allocator callback: the scratch (one double per 32x8-pixel
 * tile for the deterministic loss reduction) lives in a file-scope __device__ array of IGS_NOALLOC_MAX_TILES doubles
 * (512 KiB, held by every GPU that loads the library until the process exits), and the call makes no cudaMalloc /
 * cudaFree / allocating library call.  Sizes with V * ceil(W/32) * ceil(H/8) > IGS_NOALLOC_MAX_TILES are rejected with
 * IGS_ERR_UNSUPPORTED (e.g. 12 views of 1352x1014 or 2 views of 3840x2160 fit).  All calls on one device share the
 * array: calls that overlap on the GPU (unordered streams) race on it and must be serialised by the caller.  The first
 * call on a device resolves the array's address (cudaGetSymbolAddress) and should therefore happen outside a capture. */
#define IGS_NOALLOC_MAX_TILES 65536
int igs_depth_normal_loss_noalloc(int use_coord, const float* map0, const float* map1, const float* normal, int V,
                                  int height, int width, float tan_fovx, float tan_fovy, float depth_ratio,
                                  float grad_scale, float* loss_sum, float* dL_dmap0, float* dL_dmap1, float* dL_dnormal,
                                  void* stream);

/* Parameter activations of the refine loop and their chain rule (igs/models/gaussian_model.py:89-127, i.e.
 * RaDe-GS scene/gaussian_model.py setup_functions: scaling = exp, rotation = F.normalize (eps 1e-12),
 * opacity = sigmoid), one launch each instead of one torch kernel + autograd node per activation.
 *   forward : scales[P][3] = exp(log_scales), rotations[P][4] = raw_rot / max(|raw_rot|, 1e-12),
 *             opacities[P] = sigmoid(opacity_logit)
 *   backward: dL_dlog_scales = dL_dscales * scales;  dL_draw_rot = (dL_drot - y (y . dL_drot)) / |raw_rot|;
 *             dL_dlogit = dL_dopac * o (1 - o).  Output pointers may alias the matching dL_d* input (in place). */
int igs_activate_forward(int P, const float* log_scales, const float* raw_rot, const float* opacity_logit,
                         float* scales, float* rotations, float* opacities, void* stream);
int igs_activate_backward(int P, const float* scales, const float* raw_rot, const float* opacities,
                          const float* dL_dscales, const float* dL_drot, const float* dL_dopac,
                          float* dL_dlog_scales, float* dL_draw_rot, float* dL_dlogit, void* stream);

/* Densification statistics of one step, all V views at once (igs/models/gaussian_model.py:865-868
 * `add_densification_stats` and the max_radii2D update of infer_batch.py:312-313, which the reference runs once per
 * single-view iteration):  for every view v with radii[v][g] > 0:  grad_accum[g] += |dL_dmeans2D[v][g].xy|,
 * denom[g] += 1, max_radii2D[g] = max(max_radii2D[g], radii[v][g]).  dL_dmeans2D is [V][P][3], radii [V][P].
 * grad_scale multiplies the norm: a step whose loss is the MEAN over its V views hands in gradients that are 1/V of the
 * reference's per-view gradients (infer_batch.py:302-313 does not divide), so the caller passes V to keep
 * densify_grad_threshold on the reference's scale. */
int igs_densify_stats(int P, int V, const float* dL_dmeans2D, const int* radii, float grad_scale, float* grad_accum,
                      float* denom, float* max_radii2D, void* stream);

/* Re-layout of the packed parameter record and both Adam moments after densify / prune, in one pass (replaces the
 * torch.cat / boolean-index rebuild of every nn.Parameter and optimizer state in igs/models/gaussian_model.py:466-557,
 * three times per densify_and_prune call).  New Gaussian i takes the record of old Gaussian src[i]:
 *   kind[i] == 0  keep:  parameters and moments carried over
 *   kind[i] == 1  clone: parameters copied, moments zero                       (densify_and_clone, :619-638)
 *   kind[i] == 2  split child: parameters copied, moments zero,
 *                 xyz = build_rotation(rotation) * samples[i] + xyz,  log_scale = log(exp(log_scale) / split_div)
 *                                                                               (densify_and_split, :586-615)
 * off_old / off_new (HOST arrays of 5) are the element offsets of (xyz, rotation, shs[48], opacity, scaling) inside
 * the old / new records; the old and new arrays must not overlap. */
int igs_densify_gather(int P_old, int P_new, const int* src, const unsigned char* kind, const float* samples,
                       float split_div, const size_t* off_old, const size_t* off_new, const float* param_old,
                       const float* exp_avg_old, const float* exp_avg_sq_old, float* param_new, float* exp_avg_new,
                       float* exp_avg_sq_new, void* stream);

/* Instrumentation for bench.py's roofline (no reference equivalent): replays the forward traversal of a finished
 * igs_rasterize_forward call without producing any image and ADDS to counts (DEVICE array of 4 u64):
 *   [0] (8x4 pixel block, instance) pairs that passed the block-level cull and were evaluated per pixel ("hits")
 *   [1] (pixel, instance) pairs whose exponent was evaluated on a pixel that was still open   (E': MUFU work)
 *   [2] (pixel, instance) pairs that were blended                                             (C: contributing pairs)
 *   [3] instances staged (tile-list entries visited by their tile's CTA)                                         */
int igs_count_blend_pairs(int P, int V, int R, int width, int height, const char* geom_buffer,
                          const char* binning_buffer, const char* image_buffer, unsigned long long* counts,
                          void* stream);

const char* igs_last_error(void);
int igs_abi_version(void);
/* Cumulative number of CUDA kernels this library has launched in this process (bench.py's
 * `gpu_launches` evidence; the reference has no equivalent). */
unsigned long long igs_kernel_launches(void);

/* Per-stage device timing for bench.py's roofline (no reference equivalent).  When enabled, every
 * stage is bracketed by a CUDA event pair on the launching stream; igs_stage_timing_read() resolves
 * them (it synchronises on the recorded events), fills ms[i] / counts[i] for stage i < n with the
 * accumulated milliseconds and number of launches, optionally resets the accumulators, and returns
 * the number of stages.  igs_stage_name(i) names stage i ("blend_fwd", "tile_sort", ...). */
void igs_stage_timing_enable(int on);
int igs_stage_timing_read(double* ms, unsigned long long* counts, int n, int reset);
const char* igs_stage_name(int i);

#ifdef __cplusplus
}
#endif
#endif /* IGS_RASTER_H_INCLUDED */
