// no_alloc.cu — igs_depth_normal_loss_noalloc: the depth-normal consistency loss (geo_loss.cu) for callers that
// cannot allocate (CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference.  The existing entry point
// igs_depth_normal_loss obtains its scratch (one double per 32x8-pixel CTA for the deterministic loss reduction)
// through the caller's allocator callback; this variant keeps that scratch in a file-scope __device__ array of
// IGS_NOALLOC_MAX_TILES doubles (512 KiB, held by every GPU that loads the library until the process exits) and
// calls no cudaMalloc / cudaFree / allocating library routine.
//
// In host code `g_tile_partials` is only a host-side stand-in for the device array, NOT a device address: the address
// is resolved once per device with cudaGetSymbolAddress (with that device current) and that pointer is what the kernels
// receive.  Every call on a device shares the one array and it keeps the previous call's contents: the prefix a call
// uses is cleared on its stream first (the kernel overwrites every entry it later sums, so this is belt and braces), and
// calls that OVERLAP on the GPU (different streams without ordering) race on it — serialise them.
#include "../../include/igs_raster.h"
#include "common.cuh"
#include "kernels.cuh"
#include <mutex>

namespace igs {

static __device__ __align__(256) double g_tile_partials[IGS_NOALLOC_MAX_TILES];

#define IGS_NOALLOC_MAX_DEVICES 64
static std::mutex g_noalloc_mu;
static double* g_partials_addr[IGS_NOALLOC_MAX_DEVICES] = {nullptr};

// Device address of g_tile_partials on the CURRENT device (resolved once per device; nullptr + error message on failure).
static double* tile_partials_on_current_device() {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { set_error("cudaGetDevice failed: %s", cudaGetErrorString(e)); return nullptr; }
  if (dev < 0 || dev >= IGS_NOALLOC_MAX_DEVICES) { set_error("device ordinal %d out of range", dev); return nullptr; }
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (g_partials_addr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_tile_partials);
    if (e != cudaSuccess || p == nullptr) {
      set_error("cudaGetSymbolAddress(g_tile_partials) failed on device %d: %s", dev, cudaGetErrorString(e));
      (void)cudaGetLastError();
      return nullptr;
    }
    g_partials_addr[dev] = static_cast<double*>(p);
  }
  return g_partials_addr[dev];
}

}  // namespace igs

using namespace igs;

extern "C" int igs_depth_normal_loss_noalloc(int use_coord, const float* map0, const float* map1, const float* normal,
                                             int V, int height, int width, float tan_fovx, float tan_fovy,
                                             float depth_ratio, float grad_scale, float* loss_sum, float* dL_dmap0,
                                             float* dL_dmap1, float* dL_dnormal, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (V < 0 || height <= 0 || width <= 0 || !(tan_fovx > 0.0f) || !(tan_fovy > 0.0f)) {
    set_error("bad argument");
    return IGS_ERR_INVALID;
  }
  if (V > 65535) { set_error("at most 65535 views per call"); return IGS_ERR_UNSUPPORTED; }
  if (V == 0) return IGS_OK;
  if (!map0 || !map1 || !normal || !loss_sum || !dL_dmap0 || !dL_dmap1 || !dL_dnormal) { set_error("null pointer"); return IGS_ERR_INVALID; }
  // one partial per CTA of the loss kernel: 32 x 8 pixel tiles (geo_loss.cu GL_TX / GL_TY)
  const unsigned long long tiles = (unsigned long long)V * (unsigned long long)((width + 31) / 32) *
                                   (unsigned long long)((height + 7) / 8);
  if (tiles > (unsigned long long)IGS_NOALLOC_MAX_TILES) {
    set_error("igs_depth_normal_loss_noalloc: V * ceil(W/32) * ceil(H/8) = %llu tiles exceeds the static scratch of %d "
              "(IGS_NOALLOC_MAX_TILES); use igs_depth_normal_loss with an allocator", tiles, IGS_NOALLOC_MAX_TILES);
    return IGS_ERR_UNSUPPORTED;
  }
  if (geo_loss_scratch_bytes(V, height, width) > sizeof(double) * IGS_NOALLOC_MAX_TILES + 256) {  // same bound, via the launcher's own formula
    set_error("igs_depth_normal_loss_noalloc: scratch requirement exceeds the static array");
    return IGS_ERR_UNSUPPORTED;
  }
  double* scratch = tile_partials_on_current_device();
  if (scratch == nullptr) return IGS_ERR_CUDA;
  IGS_CUDA_CHECK(cudaMemsetAsync(scratch, 0, (size_t)tiles * sizeof(double), s));   // no allocation; capturable
  // (no StageScope here: the stage timers create CUDA events on demand)
  launch_geo_loss(use_coord != 0, map0, map1, normal, V, height, width, tan_fovx, tan_fovy, depth_ratio, grad_scale,
                  loss_sum, dL_dmap0, dL_dmap1, dL_dnormal, reinterpret_cast<char*>(scratch), s);
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}
