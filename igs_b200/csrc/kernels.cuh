// kernels.cuh — parameter blocks and host launchers shared between api.cu and the kernel files.
#pragma once
#include "common.cuh"

namespace igs {

struct ProjectParams {
  int P, V, D, M;
  int W, H;
  int tiles_x, tiles_y;
  float tan_fovx, tan_fovy, focal_x, focal_y, kernel_size;
  const float* means3D;
  const float* opacities;
  const float* shs;
  const float* colors_precomp;
};

struct BlendFwdParams {
  int W, H, tiles_x, tiles_y;
  float focal_x, focal_y;
  const uint2* ranges;
  const uint32_t* point_list;
  const SplatHead* head;
  const SplatTail* tail;
  const float* bg;
  float* out_color;
  float* out_coord;
  float* out_mcoord;
  float* out_depth;
  float* out_mdepth;
  float* out_alpha;
  float* out_normal;
  uint32_t* n_contrib;
  float* accum_coord;
  float* accum_depth;
  float* normal_length;
  unsigned long long* counters;  // instrumentation build only (igs_count_blend_pairs)
  uint8_t* hitmask;              // [R] by list position: bit w = the instance can reach block w of its tile (nullptr: not written)
};

struct BlendBwdParams {
  int W, H, tiles_x, tiles_y;
  float focal_x, focal_y;
  float half_w, half_h;  // 0.5*W, 0.5*H (d ndc->pixel), kept in the constant bank
  const uint2* ranges;
  const uint32_t* point_list;
  const uint32_t* point_e;
  const SplatHead* head;
  const SplatTail* tail;
  const float* bg;
  const float* alphas;
  const float* normalmap;
  const uint32_t* n_contrib;
  const float* accum_coord;
  const float* accum_depth;
  const float* normal_length;
  const float* dL_color;
  const float* dL_coord;
  const float* dL_mcoord;
  const float* dL_depth;
  const float* dL_mdepth;
  const float* dL_alpha;
  const float* dL_normal;
  float* grec;  // [R][NS]
  const uint8_t* hitmask;  // [R] block-hit bytes written by the forward pass (nullptr: the backward culls by itself)
};

struct GatherBwdParams {
  int P, V, W, H;
  int ns;  // record layout: 16 = ten colour-only sums, 32 = with the geometry sums
  int rec_floats;  // stride of grec in floats (12 / 16 for ns == 16, 32 for ns == 32): blend_bwd_record_floats()
  uint32_t capacity;  // emission slots the record buffer holds (R)
  float focal_x, focal_y, tan_fovx, tan_fovy, kernel_size;
  int skip_geo;       // set by the launcher: 1 = colour-only records, the normal / plane block is skipped (preprocess_bwd.cu)
  const float* means3D;
  const Camera* cams;
  const float* cov3D;
  const GaussEig* eig;
  const SplatHead* head;
  const int* radii;
  const uint32_t* tiles;
  const uint32_t* rank;
  const uint32_t* sorted_gid;  // [V*P] depth-sorted position -> global Gaussian id
  const uint32_t* inst_off;
  const uint8_t* clamped;
  const float* grec;
  float* sums;         // [V*P][12 | 28] per-(view, Gaussian) sums of the records, by depth-sorted position
  float* dL_dmeans2D;  // [V*P][3]
  float* partial;      // [V*P][16]: dmean3, dcov6, dRGB3(clamp applied), dopacity, rawRGB3
};

struct FinishBwdParams {
  int P, V, D, M;
  float scale_modifier;
  const float* means3D;
  const float* shs;
  const float* scales;
  const float* rotations;
  const Camera* cams;
  const int* radii;
  const float* partial;
  float* dL_dmeans3D;
  float* dL_dcov3D;
  float* dL_dopacity;
  float* dL_dcolors;
  float* dL_dsh;
  float* dL_dscales;
  float* dL_drotations;
};

void launch_gauss_setup(int P, const float* scales, float mod, const float* rotations, const float* cov3D_precomp,
                        float* cov3D_out, GaussEig* eig, cudaStream_t s);
void launch_project(const ProjectParams& p, const GeomBuf& g, int* radii_out, cudaStream_t s);
void launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s);

void launch_exclusive_scan(const uint32_t* in, uint32_t* out, size_t n, uint32_t* block_sums, uint32_t* total,
                           cudaStream_t s);
template <typename K>
int radix_sort_pairs(K* keys[2], uint32_t* vals[2], size_t n, int end_bit, uint32_t* hist, const uint32_t* n_dev,
                     cudaStream_t s);
void launch_gather_rank(size_t VP, const uint32_t* sorted_gid, const uint32_t* tiles, uint32_t* tiles_sorted,
                        uint32_t* rank, cudaStream_t s);
void launch_emit_instances(size_t VP, uint32_t P, int tiles_x, int tiles_y, const Camera* cams,
                           const uint32_t* sorted_gid,
                           const uint32_t* inst_off, const SplatHead* head, const int* radii, uint32_t* tkey,
                           uint32_t* tval, uint32_t* inst_gid, uint32_t capacity, cudaStream_t s);
void launch_finalize_bins(uint32_t R, const uint32_t* n_dev, const uint32_t* tkey, const uint32_t* point_e,
                          const uint32_t* inst_gid, uint32_t* point_list, uint2* ranges, cudaStream_t s);
void launch_decode_keys(uint32_t R, uint32_t tiles_per_view, uint32_t P, const uint32_t* tkey,
                        const uint32_t* point_list, const float* depth, uint64_t* keys, uint32_t* plist_local,
                        cudaStream_t s);
bool blend_fwd_writes_hitmask();   // whether the forward kernel in use fills BlendFwdParams::hitmask (IGS_HITMASK=0: off)
void launch_blend_fwd(const BlendFwdParams& p, int V, bool coord, bool depth, cudaStream_t s);
int blend_bwd_record_floats(bool geograd);   // stride (floats) of the gradient records the selected backward kernel writes
void launch_blend_bwd(const BlendBwdParams& p, int V, bool coord, bool depth, bool geograd, cudaStream_t s);
void launch_gather_bwd(const GatherBwdParams& p, cudaStream_t s);
void launch_finish_bwd(const FinishBwdParams& p, cudaStream_t s);
size_t refine_loss_scratch_bytes(int V, int H, int W);
void launch_refine_loss(const float* pred, const float* gt, int V, int H, int W, float lambda, float grad_scale, const int* loss_rows_host, float* loss_sum,
                        float* dpred, char* scratch, cudaStream_t s);
size_t geo_loss_scratch_bytes(int V, int H, int W);
void launch_geo_loss(bool coord, const float* m0, const float* m1, const float* normal, int V, int H, int W, float tan_fovx,
                     float tan_fovy, float depth_ratio, float grad_scale, float* loss_sum, float* d_m0, float* d_m1,
                     float* d_normal, char* scratch, cudaStream_t s);
void launch_activate_fwd(int P, const float* log_scales, const float* raw_rot, const float* logit, float* scales,
                         float* rot, float* opac, cudaStream_t s);
void launch_activate_bwd(int P, const float* scales, const float* raw_rot, const float* opac, const float* dL_dscales,
                         const float* dL_drot, const float* dL_dopac, float* dL_dlog_scales, float* dL_draw_rot,
                         float* dL_dlogit, cudaStream_t s);
void launch_densify_stats(int P, int V, const float* g2d, const int* radii, float grad_scale, float* grad_accum,
                          float* denom, float* max_radii, cudaStream_t s);
void launch_densify_gather(int P_old, int P_new, const int* src, const uint8_t* kind, const float* samples,
                           float split_div, const size_t* off_old, const size_t* off_new, const float* param_old,
                           const float* m_old, const float* v_old, float* param_new, float* m_new, float* v_new,
                           cudaStream_t s);
void launch_adam_groups(float* p, const float* g, float* m, float* v, int ngroups, const size_t* sizes,
                        const float* step_sizes, float b1, float b2, float eps, float inv_sqrt_bc2, cudaStream_t s);
// guarded / device-step variant: lrs instead of step sizes, bias corrections computed on the device from *step_dev
void launch_adam_groups_guarded(float* p, const float* g, float* m, float* v, int ngroups, const size_t* sizes,
                                const float* lrs, float b1, float b2, float eps, int step, const unsigned int* skip,
                                int* step_dev, cudaStream_t s);
void launch_blend_count(const BlendFwdParams& p, int V, unsigned long long* counts, cudaStream_t s);

}  // namespace igs
