// api.cu — C-ABI entry points (include/igs_raster.h): workspace carving and stage orchestration.
// Mirrors the role of CudaRasterizer::Rasterizer::{forward,backward,markVisible}
// (DGR/cuda_rasterizer/rasterizer_impl.cu:176-188,254-425,429-571) for V >= 1 views.
#include "../../include/igs_raster.h"
#include "common.cuh"
#include "kernels.cuh"
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <vector>

namespace igs {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// ---- per-stage device timing (bench.py's roofline evidence; off by default) ----
// When enabled, an event pair is recorded on the launching stream around every stage; durations are
// resolved lazily in igs_stage_timing_read() so that no synchronisation is added to the timed path.
static const char* const kStageNames[IGS_NUM_STAGES] = {
    "gauss_setup", "project", "depth_sort", "scan_rank", "emit", "tile_sort", "finalize_bins", "blend_fwd",
    "blend_bwd", "gather_bwd", "finish_bwd", "adam", "loss", "activate"};
struct StageTimer {
  std::mutex mu;
  bool on = false;
  std::vector<cudaEvent_t> pool;
  struct Pending { int stage; cudaEvent_t a, b; };
  std::vector<Pending> pending;
  double ms[IGS_NUM_STAGES] = {0};
  unsigned long long count[IGS_NUM_STAGES] = {0};
  cudaEvent_t get() {
    if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
  }
  void drain() {
    for (auto& q : pending) {
      float t = 0.f;
      if (cudaEventSynchronize(q.b) == cudaSuccess && cudaEventElapsedTime(&t, q.a, q.b) == cudaSuccess) {
        ms[q.stage] += t; count[q.stage]++;
      }
      pool.push_back(q.a); pool.push_back(q.b);
    }
    pending.clear();
  }
};
static StageTimer g_timer;
StageScope::StageScope(int stage, cudaStream_t s) : stage_(stage), s_(s), a_(nullptr) {
  if (!g_timer.on) return;
  std::lock_guard<std::mutex> lk(g_timer.mu);
  a_ = g_timer.get();
  cudaEventRecord((cudaEvent_t)a_, s_);
}
StageScope::~StageScope() {
  if (!a_) return;
  std::lock_guard<std::mutex> lk(g_timer.mu);
  cudaEvent_t b = g_timer.get();
  cudaEventRecord(b, s_);
  g_timer.pending.push_back({stage_, (cudaEvent_t)a_, b});
  if (g_timer.pending.size() > 4096) g_timer.drain();
}

// ---- small kernels local to this file ----
__global__ void pack_cameras_kernel(int V, const float* __restrict__ view, const float* __restrict__ proj,
                                    const float* __restrict__ campos, const int* __restrict__ tile_rows, int tiles_y,
                                    Camera* __restrict__ cams) {
  int v = blockIdx.x;
  int t = threadIdx.x;
  if (v >= V) return;
  if (t < 16) cams[v].view[t] = view[v * 16 + t];
  else if (t < 32) cams[v].proj[t - 16] = proj[v * 16 + (t - 16)];
  else if (t < 35) cams[v].campos[t - 32] = campos[v * 3 + (t - 32)];
  else if (t == 35) {
    int r0 = 0, r1 = tiles_y;
    if (tile_rows != nullptr) {
      r0 = min(tiles_y, max(0, tile_rows[2 * v]));
      r1 = min(tiles_y, max(r0, tile_rows[2 * v + 1]));
    }
    cams[v].row0 = r0;
    cams[v].row1 = r1;
  }
}

// capacity == 0: R = the instance total.  capacity != 0 (capacity mode): the buffers were carved for `capacity`
// instances, R = capacity; the true total and the overflow flag go to pad[0] / pad[1] (and to `status`, if given).
__global__ void write_header_kernel(Header* h, Header v, const uint32_t* total, uint32_t capacity, uint32_t* status) {
  const uint32_t t = total ? *total : 0u;
  v.R = capacity ? capacity : t;
  v.pad[0] = t;
  v.pad[1] = (capacity != 0u && t > capacity) ? 1u : 0u;
  *h = v;
  if (status) { status[0] = t; status[1] = v.pad[1]; }
}
__global__ void write_tag_kernel(BufTag* t, BufTag v) { *t = v; }

__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, size_t n, float b1, float b2, float eps, float step_size,
                            float inv_sqrt_bc2) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    float gi = g[i];
    float mi = m[i] + (1.0f - b1) * (gi - m[i]);          // torch: exp_avg.lerp_(grad, 1-beta1)
    float vi = b2 * v[i] + (1.0f - b2) * gi * gi;          // torch: exp_avg_sq.mul_(b2).addcmul_(g,g,1-b2)
    m[i] = mi;
    v[i] = vi;
    float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
    p[i] = p[i] - step_size * (mi / denom);
  }
}

// Host-side shadow of the headers written by forward, keyed by the geometry-buffer address: lets backward validate
// its three buffers (and learn which ping-pong half holds the sorted list) WITHOUT a device-to-host read and the
// stream synchronisation that goes with it.  A miss (e.g. backward in another process) falls back to the D2H path.
struct ShadowEntry { Header hv; const void* bin; const void* img; };
static std::mutex g_shadow_mu;
static std::unordered_map<const void*, ShadowEntry> g_shadow;

static int debug_sync(bool debug, cudaStream_t s, const char* stage) {
  if (!debug) return IGS_OK;
  cudaError_t e = cudaStreamSynchronize(s);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("CUDA error after stage '%s': %s", stage, cudaGetErrorString(e));
    return IGS_ERR_CUDA;
  }
  return IGS_OK;
}

}  // namespace igs

using namespace igs;

extern "C" {

const char* igs_last_error(void) { return igs::g_err; }
int igs_abi_version(void) { return IGS_RASTER_ABI_VERSION; }
int igs_refine_loss(const float* pred, const float* gt, int V, int height, int width, float lambda_l1, float grad_scale,
                    const int* loss_rows, float* loss_sum, float* dL_dpred, igs_alloc_fn temp_alloc, void* temp_user,
                    void* stream) {
  if (V < 0 || height <= 0 || width <= 0 || !temp_alloc) { set_error("bad argument"); return IGS_ERR_INVALID; }
  if (loss_rows && V > 64) { set_error("loss_rows supports at most 64 views per call"); return IGS_ERR_UNSUPPORTED; }
  if ((unsigned long long)height * (unsigned long long)width >= (1ull << 32)) { set_error("H*W exceeds the 32-bit in-plane offset range"); return IGS_ERR_UNSUPPORTED; }
  if (V == 0) return IGS_OK;
  if (!pred || !gt || !loss_sum || !dL_dpred) { set_error("null pointer"); return IGS_ERR_INVALID; }
  size_t bytes = refine_loss_scratch_bytes(V, height, width);
  char* tmp = temp_alloc(temp_user, bytes + 256);
  if (!tmp) { set_error("temp allocator returned NULL"); return IGS_ERR_ALLOC; }
  tmp = (char*)(((uintptr_t)tmp + 255) & ~uintptr_t(255));
  {
    StageScope st(IGS_STAGE_LOSS, (cudaStream_t)stream);
    launch_refine_loss(pred, gt, V, height, width, lambda_l1, grad_scale, loss_rows, loss_sum, dL_dpred, tmp,
                       (cudaStream_t)stream);
  }
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}

int igs_depth_normal_loss(int use_coord, const float* map0, const float* map1, const float* normal, int V, int height,
                          int width, float tan_fovx, float tan_fovy, float depth_ratio, float grad_scale,
                          float* loss_sum, float* dL_dmap0, float* dL_dmap1, float* dL_dnormal, igs_alloc_fn temp_alloc,
                          void* temp_user, void* stream) {
  if (V < 0 || height <= 0 || width <= 0 || !temp_alloc || !(tan_fovx > 0.0f) || !(tan_fovy > 0.0f)) {
    set_error("bad argument");
    return IGS_ERR_INVALID;
  }
  if (V > 65535) { set_error("at most 65535 views per call"); return IGS_ERR_UNSUPPORTED; }
  if (V == 0) return IGS_OK;
  if (!map0 || !map1 || !normal || !loss_sum || !dL_dmap0 || !dL_dmap1 || !dL_dnormal) { set_error("null pointer"); return IGS_ERR_INVALID; }
  char* tmp = temp_alloc(temp_user, geo_loss_scratch_bytes(V, height, width) + 256);
  if (!tmp) { set_error("temp allocator returned NULL"); return IGS_ERR_ALLOC; }
  tmp = (char*)(((uintptr_t)tmp + 255) & ~uintptr_t(255));
  {
    StageScope st(IGS_STAGE_LOSS, (cudaStream_t)stream);
    launch_geo_loss(use_coord != 0, map0, map1, normal, V, height, width, tan_fovx, tan_fovy, depth_ratio, grad_scale,
                    loss_sum, dL_dmap0, dL_dmap1, dL_dnormal, tmp, (cudaStream_t)stream);
  }
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}

void igs_stage_timing_enable(int on) {
  std::lock_guard<std::mutex> lk(igs::g_timer.mu);
  igs::g_timer.on = on != 0;
}
int igs_stage_timing_read(double* ms, unsigned long long* counts, int n, int reset) {
  std::lock_guard<std::mutex> lk(igs::g_timer.mu);
  igs::g_timer.drain();
  for (int i = 0; i < n && i < IGS_NUM_STAGES; i++) {
    if (ms) ms[i] = igs::g_timer.ms[i];
    if (counts) counts[i] = igs::g_timer.count[i];
  }
  if (reset) for (int i = 0; i < IGS_NUM_STAGES; i++) { igs::g_timer.ms[i] = 0; igs::g_timer.count[i] = 0; }
  return IGS_NUM_STAGES;
}
const char* igs_stage_name(int i) { return (i >= 0 && i < IGS_NUM_STAGES) ? igs::kStageNames[i] : ""; }
unsigned long long igs_kernel_launches(void) { return igs::g_launches.load(std::memory_order_relaxed); }

int igs_rasterize_forward(const igs_forward_args* a, igs_alloc_fn geom_alloc, void* geom_user,
                          igs_alloc_fn bin_alloc, void* bin_user, igs_alloc_fn img_alloc, void* img_user,
                          void* stream, int* num_rendered) {
  cudaStream_t s = (cudaStream_t)stream;
  if (num_rendered) *num_rendered = 0;
  if (!a || !geom_alloc || !bin_alloc || !img_alloc) { set_error("null argument"); return IGS_ERR_INVALID; }
  if (a->P < 0 || a->V < 1 || a->width <= 0 || a->height <= 0) { set_error("bad sizes P=%d V=%d W=%d H=%d", a->P, a->V, a->width, a->height); return IGS_ERR_INVALID; }
  if (a->prefiltered) { set_error("prefiltered=True is not supported (the reference mis-routes it, rasterizer_impl.cu:344-345)"); return IGS_ERR_UNSUPPORTED; }
  if (a->shs == nullptr && a->colors_precomp == nullptr) { set_error("provide SHs or precomputed colors"); return IGS_ERR_INVALID; }
  if (a->shs != nullptr && a->colors_precomp == nullptr && (a->M < 1 || a->M > 16)) { set_error("M=%d out of range", a->M); return IGS_ERR_UNSUPPORTED; }
  if (a->cov3D_precomp == nullptr && (a->scales == nullptr || a->rotations == nullptr)) { set_error("provide scales+rotations or cov3D_precomp"); return IGS_ERR_INVALID; }
  if (a->P == 0) return IGS_OK;  // reference: skip everything, outputs stay zero (rasterize_points.cu:89-91)
  {
    int M = a->M;
    int need = (a->D + 1) * (a->D + 1);
    if (a->colors_precomp == nullptr && (a->D < 0 || a->D > 3 || need > M)) { set_error("SH degree %d needs %d coefficients, got %d", a->D, need, M); return IGS_ERR_INVALID; }
  }
  const int P = a->P, V = a->V, W = a->width, H = a->height;
  const size_t VP = (size_t)V * P;
  if (VP >= 0xFFFFFFFFull) { set_error("V*P too large"); return IGS_ERR_UNSUPPORTED; }
  const int tiles_x = (W + IGS_TILE_X - 1) / IGS_TILE_X, tiles_y = (H + IGS_TILE_Y - 1) / IGS_TILE_Y;
  const size_t tiles = (size_t)tiles_x * tiles_y, N = (size_t)W * H;
  const float focal_y = H / (2.0f * a->tan_fovy), focal_x = W / (2.0f * a->tan_fovx);
  const bool dbg = a->debug != 0;

  GeomBuf gsz = carve_geom(nullptr, P, V);
  char* gptr = geom_alloc(geom_user, gsz.bytes);
  if (!gptr) { set_error("geometry allocator returned NULL for %zu bytes", gsz.bytes); return IGS_ERR_ALLOC; }
  GeomBuf G = carve_geom(gptr, P, V);
  ImgBuf isz = carve_img(nullptr, V, N, tiles);
  char* iptr = img_alloc(img_user, isz.bytes);
  if (!iptr) { set_error("image allocator returned NULL for %zu bytes", isz.bytes); return IGS_ERR_ALLOC; }
  ImgBuf I = carve_img(iptr, V, N, tiles);
  count_launch();
  write_tag_kernel<<<1, 1, 0, s>>>(I.tag, BufTag{IGS_TAG_IMG, 0u, (uint32_t)V, (uint32_t)N});

  count_launch();
  pack_cameras_kernel<<<V, 64, 0, s>>>(V, a->viewmatrix, a->projmatrix, a->cam_pos, a->tile_rows, tiles_y, G.cams);
  { StageScope st(IGS_STAGE_GAUSS_SETUP, s);
  launch_gauss_setup(P, a->scales, a->scale_modifier, a->rotations, a->cov3D_precomp, G.cov3D, G.eig, s); }
  if (int e = debug_sync(dbg, s, "gauss_setup")) return e;

  ProjectParams pp;
  pp.P = P; pp.V = V; pp.D = a->D; pp.M = a->M; pp.W = W; pp.H = H; pp.tiles_x = tiles_x; pp.tiles_y = tiles_y;
  pp.tan_fovx = a->tan_fovx; pp.tan_fovy = a->tan_fovy; pp.focal_x = focal_x; pp.focal_y = focal_y;
  pp.kernel_size = a->kernel_size;
  pp.means3D = a->means3D; pp.opacities = a->opacities; pp.shs = a->shs; pp.colors_precomp = a->colors_precomp;
  { StageScope st(IGS_STAGE_PROJECT, s); launch_project(pp, G, a->radii, s); }
  if (int e = debug_sync(dbg, s, "project")) return e;

  // ---- sort A: (view, depth) ----
  int cur;
  { StageScope st(IGS_STAGE_DEPTH_SORT, s);
    cur = radix_sort_pairs<uint64_t>(G.skey, G.sval, VP, 31 + (V > 1 ? bits_for((uint64_t)V) : 0), G.hist, nullptr, s); }
  if (int e = debug_sync(dbg, s, "depth sort")) return e;
  const uint32_t* sorted_gid = G.sval[cur];
  uint32_t* d_total = G.scan_tmp + scan_blocks(VP + 1);
  { StageScope st(IGS_STAGE_SCAN_RANK, s);
    launch_gather_rank(VP, sorted_gid, G.tiles, G.inst_off, G.rank, s);
    launch_exclusive_scan(G.inst_off, G.inst_off, VP + 1, G.scan_tmp, d_total, s); }

  Header hv;
  memset(&hv, 0, sizeof(hv));
  hv.magic = IGS_MAGIC; hv.P = P; hv.V = V; hv.N = (uint32_t)N; hv.tiles_x = tiles_x; hv.tiles_y = tiles_y;
  hv.flags = (a->require_coord ? 1u : 0u) | (a->require_depth ? 2u : 0u) | (a->scales ? 8u : 0u) | (a->shs ? 16u : 0u) |
             ((uint32_t)cur << 8);
  // ---- the instance count ----
  // Exact mode (instance_capacity == 0, the reference's behaviour, rasterizer_impl.cu:354): one host read of R through a
  // pinned word, the binning buffers are sized for exactly R.  Capacity mode (instance_capacity > 0): NO host round-trip —
  // the buffers are sized for the caller's capacity, every downstream launch is sized for it and reads the true count
  // on the device; instances past the capacity are dropped and flagged (header pad[1], args->status[1]).
  const bool capmode = a->instance_capacity > 0;
  const uint32_t* n_dev = nullptr;
  uint32_t R = 0;
  if (!capmode) {
    static thread_local uint32_t* pinned_R = nullptr;
    if (pinned_R == nullptr && cudaHostAlloc(reinterpret_cast<void**>(&pinned_R), 64, cudaHostAllocPortable) != cudaSuccess) {
      pinned_R = nullptr;
      (void)cudaGetLastError();
    }
    uint32_t R_pageable = 0;
    uint32_t* R_dst = pinned_R ? pinned_R : &R_pageable;
    IGS_CUDA_CHECK(cudaMemcpyAsync(R_dst, d_total, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    IGS_CUDA_CHECK(cudaStreamSynchronize(s));
    R = *R_dst;
    if (R > 0x7FFFFFFFu) { set_error("too many instances: %u", R); return IGS_ERR_UNSUPPORTED; }
  } else {
    R = (uint32_t)a->instance_capacity;
    n_dev = d_total;
  }

  BinBuf bsz = carve_bin(nullptr, R);
  char* bptr = bin_alloc(bin_user, bsz.bytes);
  if (!bptr) { set_error("binning allocator returned NULL for %zu bytes", bsz.bytes); return IGS_ERR_ALLOC; }
  BinBuf B = carve_bin(bptr, R);
  count_launch();
  write_tag_kernel<<<1, 1, 0, s>>>(B.tag, BufTag{IGS_TAG_BIN, 0u, R, (uint32_t)V});

  IGS_CUDA_CHECK(cudaMemsetAsync(I.ranges, 0, sizeof(uint2) * V * tiles, s));
  int bcur = 0;
  if (R > 0) {
    { StageScope st(IGS_STAGE_EMIT, s);
      launch_emit_instances(VP, (uint32_t)P, tiles_x, tiles_y, G.cams, sorted_gid, G.inst_off, G.head, G.radii, B.tkey[0],
                            B.tval[0], B.inst_gid, R, s); }
    if (int e = debug_sync(dbg, s, "emit")) return e;
    { StageScope st(IGS_STAGE_TILE_SORT, s);
      bcur = radix_sort_pairs<uint32_t>(B.tkey, B.tval, R, bits_for((uint64_t)V * tiles), B.hist, n_dev, s); }
    if (int e = debug_sync(dbg, s, "tile sort")) return e;
    { StageScope st(IGS_STAGE_FINALIZE, s);
      launch_finalize_bins(R, n_dev, B.tkey[bcur], B.tval[bcur], B.inst_gid, B.point_list, I.ranges, s); }
    if (int e = debug_sync(dbg, s, "finalize bins")) return e;
  }
  // remember which ping-pong buffer holds the sorted instances (bit 16 of flags)
  {
    hv.flags |= ((uint32_t)bcur << 16);
    if (R > 0 && blend_fwd_writes_hitmask()) hv.flags |= (1u << 17);   // block-hit bytes in the other tile-key buffer
    count_launch();
    write_header_kernel<<<1, 1, 0, s>>>(G.hdr, hv, d_total, capmode ? R : 0u, a->status);
    hv.R = R;
    std::lock_guard<std::mutex> lk(g_shadow_mu);
    if (g_shadow.size() > 4096) g_shadow.clear();
    g_shadow[(const void*)gptr] = ShadowEntry{hv, (const void*)bptr, (const void*)iptr};
  }

  BlendFwdParams bp;
  bp.counters = nullptr;
  // Block-hit bytes for the backward pass live in the tile-key buffer the sort did not finish in (dead after the sort);
  // instances the forward never stages (behind the point where a tile's pixels are all opaque) keep the zero.
  bp.hitmask = reinterpret_cast<uint8_t*>(B.tkey[bcur ^ 1]);   // (R == 0: no tile has a list, nothing is written)
  if (R > 0) IGS_CUDA_CHECK(cudaMemsetAsync(bp.hitmask, 0, R, s));
  bp.W = W; bp.H = H; bp.tiles_x = tiles_x; bp.tiles_y = tiles_y; bp.focal_x = focal_x; bp.focal_y = focal_y;
  bp.ranges = I.ranges; bp.point_list = B.point_list; bp.head = G.head; bp.tail = G.tail; bp.bg = a->background;
  bp.out_color = a->out_color; bp.out_coord = a->out_coord; bp.out_mcoord = a->out_mcoord; bp.out_depth = a->out_depth;
  bp.out_mdepth = a->out_mdepth; bp.out_alpha = a->out_alpha; bp.out_normal = a->out_normal;
  bp.n_contrib = I.n_contrib; bp.accum_coord = I.accum_coord; bp.accum_depth = I.accum_depth;
  bp.normal_length = I.normal_length;
  { StageScope st(IGS_STAGE_BLEND_FWD, s); launch_blend_fwd(bp, V, a->require_coord != 0, a->require_depth != 0, s); }
  if (int e = debug_sync(dbg, s, "blend forward")) return e;
  IGS_CUDA_CHECK(cudaGetLastError());
  if (num_rendered) *num_rendered = (int)R;
  return IGS_OK;
}

int igs_rasterize_backward(const igs_backward_args* a, igs_alloc_fn temp_alloc, void* temp_user, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!a || !temp_alloc) { set_error("null argument"); return IGS_ERR_INVALID; }
  if (a->P < 0 || a->V < 1) { set_error("bad sizes"); return IGS_ERR_INVALID; }
  const int P = a->P, V = a->V, W = a->width, H = a->height;
  if (P == 0) return IGS_OK;
  if (!a->geom_buffer || !a->binning_buffer || !a->image_buffer || !a->dL_dout_color) { set_error("missing buffer / colour cotangent"); return IGS_ERR_INVALID; }
  const size_t VP = (size_t)V * P;
  const int tiles_x = (W + IGS_TILE_X - 1) / IGS_TILE_X, tiles_y = (H + IGS_TILE_Y - 1) / IGS_TILE_Y;
  const size_t tiles = (size_t)tiles_x * tiles_y, N = (size_t)W * H;
  const float focal_y = H / (2.0f * a->tan_fovy), focal_x = W / (2.0f * a->tan_fovx);
  const bool dbg = a->debug != 0;
  const uint32_t R = (uint32_t)a->R;

  GeomBuf G = carve_geom(a->geom_buffer, P, V);
  BinBuf B = carve_bin(a->binning_buffer, R);
  ImgBuf I = carve_img(a->image_buffer, V, N, tiles);
  Header hv;
  BufTag tb, ti;
  bool shadow_hit = false;
  {
    std::lock_guard<std::mutex> lk(g_shadow_mu);
    auto it = g_shadow.find((const void*)a->geom_buffer);
    if (it != g_shadow.end() && it->second.bin == (const void*)a->binning_buffer && it->second.img == (const void*)a->image_buffer) {
      hv = it->second.hv;
      tb = BufTag{IGS_TAG_BIN, 0u, hv.R, hv.V};
      ti = BufTag{IGS_TAG_IMG, 0u, hv.V, hv.N};
      shadow_hit = true;
    }
  }
  if (!shadow_hit) {
    IGS_CUDA_CHECK(cudaMemcpyAsync(&hv, G.hdr, sizeof(Header), cudaMemcpyDeviceToHost, s));
    IGS_CUDA_CHECK(cudaMemcpyAsync(&tb, B.tag, sizeof(BufTag), cudaMemcpyDeviceToHost, s));
    IGS_CUDA_CHECK(cudaMemcpyAsync(&ti, I.tag, sizeof(BufTag), cudaMemcpyDeviceToHost, s));
    IGS_CUDA_CHECK(cudaStreamSynchronize(s));
  }
  if (tb.magic != IGS_TAG_BIN || tb.a != R || ti.magic != IGS_TAG_IMG || ti.a != (uint32_t)V || ti.b != (uint32_t)N) {
    set_error("binning / image buffer does not belong to this call (tags %x/%u %x/%u/%u, expected R %u V %d N %zu)",
              tb.magic, tb.a, ti.magic, ti.a, ti.b, R, V, N);
    return IGS_ERR_INVALID;
  }
  if (hv.magic != IGS_MAGIC || hv.P != (uint32_t)P || hv.V != (uint32_t)V || hv.N != (uint32_t)N || hv.R != R) {
    set_error("geometry buffer does not match this call (magic %x P %u V %u N %u R %u vs P %d V %d N %zu R %u)",
              hv.magic, hv.P, hv.V, hv.N, hv.R, P, V, N, R);
    return IGS_ERR_INVALID;
  }
  const int bcur = (hv.flags >> 16) & 1;
  const bool coord = a->require_coord != 0, depth = a->require_depth != 0;
  const bool any_geo_cot = a->dL_dout_coord || a->dL_dout_mcoord || a->dL_dout_depth || a->dL_dout_mdepth || a->dL_dout_normal;
  const bool geograd = any_geo_cot && (coord || depth);
  const int ns = geograd ? 32 : 16;

  // temp: gradient records + per-(view,Gaussian) partials
  const int rec_floats = blend_bwd_record_floats(geograd);
  size_t grec_bytes = ((size_t)(R ? R : 1) * rec_floats * sizeof(float) + 255) & ~size_t(255);
  size_t part_bytes = VP * 16 * sizeof(float);
  size_t sums_bytes = VP * (ns == 16 ? 12 : 28) * sizeof(float);
  char* tmp = temp_alloc(temp_user, grec_bytes + part_bytes + sums_bytes + 512);
  if (!tmp) { set_error("temp allocator returned NULL"); return IGS_ERR_ALLOC; }
  tmp = (char*)(((uintptr_t)tmp + 255) & ~uintptr_t(255));
  float* grec = (float*)tmp;
  float* partial = (float*)(tmp + grec_bytes);
  float* sums = (float*)(tmp + grec_bytes + part_bytes);

  if (R > 0) {
    BlendBwdParams bp;
    bp.W = W; bp.H = H; bp.tiles_x = tiles_x; bp.tiles_y = tiles_y; bp.focal_x = focal_x; bp.focal_y = focal_y;
    bp.half_w = 0.5f * W; bp.half_h = 0.5f * H;
    bp.ranges = I.ranges; bp.point_list = B.point_list; bp.point_e = B.tval[bcur]; bp.head = G.head; bp.tail = G.tail;
    bp.bg = a->background; bp.alphas = a->alphas; bp.normalmap = a->normalmap; bp.n_contrib = I.n_contrib;
    bp.accum_coord = I.accum_coord; bp.accum_depth = I.accum_depth; bp.normal_length = I.normal_length;
    bp.dL_color = a->dL_dout_color; bp.dL_coord = a->dL_dout_coord; bp.dL_mcoord = a->dL_dout_mcoord;
    bp.dL_depth = a->dL_dout_depth; bp.dL_mdepth = a->dL_dout_mdepth; bp.dL_alpha = a->dL_dout_alpha;
    bp.dL_normal = a->dL_dout_normal; bp.grec = grec;
    bp.hitmask = (hv.flags & (1u << 17)) ? reinterpret_cast<const uint8_t*>(B.tkey[bcur ^ 1]) : nullptr;
    { StageScope st(IGS_STAGE_BLEND_BWD, s); launch_blend_bwd(bp, V, coord, depth, geograd, s); }
    if (int e = debug_sync(dbg, s, "blend backward")) return e;
  }

  GatherBwdParams gp;
  gp.P = P; gp.V = V; gp.W = W; gp.H = H; gp.ns = ns; gp.rec_floats = rec_floats; gp.capacity = R; gp.skip_geo = 0; gp.focal_x = focal_x; gp.focal_y = focal_y;
  gp.tan_fovx = a->tan_fovx; gp.tan_fovy = a->tan_fovy; gp.kernel_size = a->kernel_size;
  gp.means3D = a->means3D; gp.cams = G.cams; gp.cov3D = G.cov3D; gp.eig = G.eig; gp.head = G.head; gp.radii = G.radii;
  gp.tiles = G.tiles; gp.rank = G.rank; gp.sorted_gid = G.sval[(hv.flags >> 8) & 1]; gp.inst_off = G.inst_off; gp.clamped = G.clamped; gp.grec = grec; gp.sums = sums;
  gp.dL_dmeans2D = a->dL_dmeans2D; gp.partial = partial;
  { StageScope st(IGS_STAGE_GATHER_BWD, s); launch_gather_bwd(gp, s); }
  if (int e = debug_sync(dbg, s, "gather backward")) return e;

  FinishBwdParams fp;
  fp.P = P; fp.V = V; fp.D = a->D; fp.M = a->M; fp.scale_modifier = a->scale_modifier; fp.means3D = a->means3D;
  fp.shs = a->shs; fp.scales = a->scales; fp.rotations = a->rotations; fp.cams = G.cams; fp.radii = G.radii;
  fp.partial = partial; fp.dL_dmeans3D = a->dL_dmeans3D; fp.dL_dcov3D = a->dL_dcov3D; fp.dL_dopacity = a->dL_dopacity;
  fp.dL_dcolors = a->dL_dcolors; fp.dL_dsh = a->shs ? a->dL_dsh : nullptr; fp.dL_dscales = a->dL_dscales;
  fp.dL_drotations = a->dL_drotations;
  { StageScope st(IGS_STAGE_FINISH_BWD, s); launch_finish_bwd(fp, s); }
  if (int e = debug_sync(dbg, s, "finish backward")) return e;
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}

int igs_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present,
                     void* stream) {
  (void)projmatrix;
  if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) { set_error("bad argument"); return IGS_ERR_INVALID; }
  launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream);
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}

int igs_decode_binning(int P, int V, int R, int width, int height, const char* geom_buffer, const char* binning_buffer,
                       const char* image_buffer, const igs_decode_out* out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!out || !geom_buffer || !image_buffer || P <= 0 || V < 1) { set_error("bad argument"); return IGS_ERR_INVALID; }
  const int tiles_x = (width + IGS_TILE_X - 1) / IGS_TILE_X, tiles_y = (height + IGS_TILE_Y - 1) / IGS_TILE_Y;
  const size_t tiles = (size_t)tiles_x * tiles_y, N = (size_t)width * height, VP = (size_t)V * P;
  GeomBuf G = carve_geom(const_cast<char*>(geom_buffer), P, V);
  ImgBuf I = carve_img(const_cast<char*>(image_buffer), V, N, tiles);
  Header hv;
  IGS_CUDA_CHECK(cudaMemcpyAsync(&hv, G.hdr, sizeof(Header), cudaMemcpyDeviceToHost, s));
  IGS_CUDA_CHECK(cudaStreamSynchronize(s));
  if (hv.magic != IGS_MAGIC || hv.R != (uint32_t)R) { set_error("buffer mismatch in decode"); return IGS_ERR_INVALID; }
  const int bcur = (hv.flags >> 16) & 1;
  if (R > 0 && binning_buffer) {
    BinBuf B = carve_bin(const_cast<char*>(binning_buffer), R);
    const uint32_t n_true = hv.pad[0] < (uint32_t)R ? hv.pad[0] : (uint32_t)R;   // capacity mode: R is the capacity
    launch_decode_keys(n_true, (uint32_t)tiles, (uint32_t)P, B.tkey[bcur], B.point_list, G.depth, out->keys,
                       out->point_list, s);
  }
  if (out->ranges) IGS_CUDA_CHECK(cudaMemcpyAsync(out->ranges, I.ranges, sizeof(uint2) * V * tiles, cudaMemcpyDeviceToDevice, s));
  if (out->n_contrib) IGS_CUDA_CHECK(cudaMemcpyAsync(out->n_contrib, I.n_contrib, sizeof(uint32_t) * V * 2 * N, cudaMemcpyDeviceToDevice, s));
  if (out->tiles_touched) IGS_CUDA_CHECK(cudaMemcpyAsync(out->tiles_touched, G.tiles, sizeof(uint32_t) * VP, cudaMemcpyDeviceToDevice, s));
  if (out->depths) IGS_CUDA_CHECK(cudaMemcpyAsync(out->depths, G.depth, sizeof(float) * VP, cudaMemcpyDeviceToDevice, s));
  if (out->cov3D) IGS_CUDA_CHECK(cudaMemcpyAsync(out->cov3D, G.cov3D, sizeof(float) * 6 * P, cudaMemcpyDeviceToDevice, s));
  // strided extractions from the packed records
  if (out->means2D) IGS_CUDA_CHECK(cudaMemcpy2DAsync(out->means2D, 8, G.head, sizeof(SplatHead), 8, VP, cudaMemcpyDeviceToDevice, s));
  if (out->conic_opacity) IGS_CUDA_CHECK(cudaMemcpy2DAsync(out->conic_opacity, 16, (const char*)G.head + 8, sizeof(SplatHead), 16, VP, cudaMemcpyDeviceToDevice, s));
  if (out->rgb) {
    IGS_CUDA_CHECK(cudaMemcpy2DAsync(out->rgb, 12, (const char*)G.head + 24, sizeof(SplatHead), 8, VP, cudaMemcpyDeviceToDevice, s));
    IGS_CUDA_CHECK(cudaMemcpy2DAsync((char*)out->rgb + 8, 12, G.tail, sizeof(SplatTail), 4, VP, cudaMemcpyDeviceToDevice, s));
  }
  if (out->ts) IGS_CUDA_CHECK(cudaMemcpy2DAsync(out->ts, 4, (const char*)G.tail + 4, sizeof(SplatTail), 4, VP, cudaMemcpyDeviceToDevice, s));
  if (out->ray_planes) IGS_CUDA_CHECK(cudaMemcpy2DAsync(out->ray_planes, 8, (const char*)G.tail + 8, sizeof(SplatTail), 8, VP, cudaMemcpyDeviceToDevice, s));
  if (out->normals) IGS_CUDA_CHECK(cudaMemcpy2DAsync(out->normals, 12, (const char*)G.tail + 16, sizeof(SplatTail), 12, VP, cudaMemcpyDeviceToDevice, s));
  if (out->view_points) IGS_CUDA_CHECK(cudaMemcpy2DAsync(out->view_points, 12, (const char*)G.tail + 28, sizeof(SplatTail), 12, VP, cudaMemcpyDeviceToDevice, s));
  if (out->camera_planes) IGS_CUDA_CHECK(cudaMemcpy2DAsync(out->camera_planes, 24, (const char*)G.tail + 40, sizeof(SplatTail), 24, VP, cudaMemcpyDeviceToDevice, s));
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}

int igs_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, size_t n, float lr, float beta1,
                  float beta2, float eps, int step, void* stream) {
  if (n == 0) return IGS_OK;
  if (!param || !grad || !exp_avg || !exp_avg_sq || step < 1) { set_error("bad argument"); return IGS_ERR_INVALID; }
  double bc1 = 1.0 - pow((double)beta1, (double)step);
  double bc2 = 1.0 - pow((double)beta2, (double)step);
  float step_size = (float)((double)lr / bc1);
  float inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
  size_t nb = (n + 255) / 256;
  if (nb > 148 * 16) nb = 148 * 16;
  count_launch();
  {
    StageScope st(IGS_STAGE_ADAM, (cudaStream_t)stream);
    adam_kernel<<<(unsigned)nb, 256, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, n, beta1, beta2, eps,
                                                                  step_size, inv_sqrt_bc2);
  }
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}

int igs_adam_step_groups(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int ngroups,
                         const size_t* sizes, const float* lrs, float beta1, float beta2, float eps, int step,
                         void* stream) {
  if (ngroups == 0) return IGS_OK;
  if (ngroups < 0 || ngroups > 8 || !sizes || !lrs || !param || !grad || !exp_avg || !exp_avg_sq || step < 1) {
    set_error("bad argument");
    return IGS_ERR_INVALID;
  }
  const double bc1 = 1.0 - pow((double)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)beta2, (double)step);
  float step_sizes[8];
  for (int k = 0; k < ngroups; k++) step_sizes[k] = (float)((double)lrs[k] / bc1);
  {
    StageScope st(IGS_STAGE_ADAM, (cudaStream_t)stream);
    launch_adam_groups(param, grad, exp_avg, exp_avg_sq, ngroups, sizes, step_sizes, beta1, beta2, eps,
                       (float)(1.0 / sqrt(bc2)), (cudaStream_t)stream);
  }
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}

int igs_adam_step_groups_guarded(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int ngroups,
                                 const size_t* sizes, const float* lrs, float beta1, float beta2, float eps, int step,
                                 const unsigned int* skip_if_nonzero, int* step_dev, void* stream) {
  if (ngroups == 0) return IGS_OK;
  if (ngroups < 0 || ngroups > 8 || !sizes || !lrs || !param || !grad || !exp_avg || !exp_avg_sq ||
      (step_dev == nullptr && step < 1)) {
    set_error("bad argument");
    return IGS_ERR_INVALID;
  }
  {
    StageScope st(IGS_STAGE_ADAM, (cudaStream_t)stream);
    launch_adam_groups_guarded(param, grad, exp_avg, exp_avg_sq, ngroups, sizes, lrs, beta1, beta2, eps, step,
                               skip_if_nonzero, step_dev, (cudaStream_t)stream);
  }
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}

int igs_count_blend_pairs(int P, int V, int R, int width, int height, const char* geom_buffer, const char* binning_buffer,
                          const char* image_buffer, unsigned long long* counts, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (!geom_buffer || !image_buffer || !counts || P <= 0 || V < 1 || R < 0) { set_error("bad argument"); return IGS_ERR_INVALID; }
  if (R == 0) return IGS_OK;
  if (!binning_buffer) { set_error("bad argument"); return IGS_ERR_INVALID; }
  const int tiles_x = (width + IGS_TILE_X - 1) / IGS_TILE_X, tiles_y = (height + IGS_TILE_Y - 1) / IGS_TILE_Y;
  const size_t tiles = (size_t)tiles_x * tiles_y, N = (size_t)width * height;
  GeomBuf G = carve_geom(const_cast<char*>(geom_buffer), P, V);
  BinBuf B = carve_bin(const_cast<char*>(binning_buffer), (size_t)R);
  ImgBuf I = carve_img(const_cast<char*>(image_buffer), V, N, tiles);
  Header hv;
  IGS_CUDA_CHECK(cudaMemcpyAsync(&hv, G.hdr, sizeof(Header), cudaMemcpyDeviceToHost, s));
  IGS_CUDA_CHECK(cudaStreamSynchronize(s));
  if (hv.magic != IGS_MAGIC || hv.R != (uint32_t)R || hv.P != (uint32_t)P || hv.V != (uint32_t)V) { set_error("buffer mismatch in count"); return IGS_ERR_INVALID; }
  BlendFwdParams bp;
  memset(&bp, 0, sizeof(bp));
  bp.W = width; bp.H = height; bp.tiles_x = tiles_x; bp.tiles_y = tiles_y; bp.focal_x = 1.0f; bp.focal_y = 1.0f;
  bp.ranges = I.ranges; bp.point_list = B.point_list; bp.head = G.head; bp.tail = G.tail;
  launch_blend_count(bp, V, counts, s);
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}

int igs_activate_forward(int P, const float* log_scales, const float* raw_rot, const float* opacity_logit, float* scales,
                         float* rotations, float* opacities, void* stream) {
  if (P < 0) { set_error("bad argument"); return IGS_ERR_INVALID; }
  if (P == 0) return IGS_OK;
  if (!log_scales || !raw_rot || !opacity_logit || !scales || !rotations || !opacities) {
    set_error("null pointer");
    return IGS_ERR_INVALID;
  }
  {
    StageScope st(IGS_STAGE_ACTIVATE, (cudaStream_t)stream);
    launch_activate_fwd(P, log_scales, raw_rot, opacity_logit, scales, rotations, opacities, (cudaStream_t)stream);
  }
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}

int igs_activate_backward(int P, const float* scales, const float* raw_rot, const float* opacities,
                          const float* dL_dscales, const float* dL_drot, const float* dL_dopac, float* dL_dlog_scales,
                          float* dL_draw_rot, float* dL_dlogit, void* stream) {
  if (P < 0) { set_error("bad argument"); return IGS_ERR_INVALID; }
  if (P == 0) return IGS_OK;
  if (!scales || !raw_rot || !opacities || !dL_dscales || !dL_drot || !dL_dopac || !dL_dlog_scales || !dL_draw_rot ||
      !dL_dlogit) {
    set_error("null pointer");
    return IGS_ERR_INVALID;
  }
  {
    StageScope st(IGS_STAGE_ACTIVATE, (cudaStream_t)stream);
    launch_activate_bwd(P, scales, raw_rot, opacities, dL_dscales, dL_drot, dL_dopac, dL_dlog_scales, dL_draw_rot,
                        dL_dlogit, (cudaStream_t)stream);
  }
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}

int igs_densify_stats(int P, int V, const float* dL_dmeans2D, const int* radii, float grad_scale, float* grad_accum,
                      float* denom, float* max_radii2D, void* stream) {
  if (P < 0 || V < 0) { set_error("bad argument"); return IGS_ERR_INVALID; }
  if (P == 0 || V == 0) return IGS_OK;
  if (!dL_dmeans2D || !radii || !grad_accum || !denom || !max_radii2D) { set_error("null pointer"); return IGS_ERR_INVALID; }
  launch_densify_stats(P, V, dL_dmeans2D, radii, grad_scale, grad_accum, denom, max_radii2D, (cudaStream_t)stream);
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}

int igs_densify_gather(int P_old, int P_new, const int* src, const unsigned char* kind, const float* samples,
                       float split_div, const size_t* off_old, const size_t* off_new, const float* param_old,
                       const float* exp_avg_old, const float* exp_avg_sq_old, float* param_new, float* exp_avg_new,
                       float* exp_avg_sq_new, void* stream) {
  if (P_old < 0 || P_new < 0) { set_error("bad argument"); return IGS_ERR_INVALID; }
  if (P_new == 0) return IGS_OK;
  if (!src || !kind || !off_old || !off_new || !param_old || !exp_avg_old || !exp_avg_sq_old || !param_new ||
      !exp_avg_new || !exp_avg_sq_new) {
    set_error("null pointer");
    return IGS_ERR_INVALID;
  }
  if (!(split_div > 0.0f)) { set_error("split_div must be positive"); return IGS_ERR_INVALID; }
  launch_densify_gather(P_old, P_new, src, kind, samples, split_div, off_old, off_new, param_old, exp_avg_old,
                        exp_avg_sq_old, param_new, exp_avg_new, exp_avg_sq_new, (cudaStream_t)stream);
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}

}  // extern "C"
