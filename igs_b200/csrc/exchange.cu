// exchange.cu — the exchange step of the view-sharded refine iteration as ONE kernel over NVLink / NVSwitch multicast.
//
// The reference has no multi-GPU refine path (no collective on the hot path); BASELINE's view-sharded
// schedule adds one exchange: per-Gaussian gradients are summed over the ranks, then every rank takes the same Adam
// step (igs/models/gaussian_model.py:348, infer_batch.py:323).  Round 1 did that as ncclAllReduce(70.8 MB) + a
// replicated Adam launch: at 8 GPUs 0.33 ms + 0.12 ms of a 4.2 ms step, fully exposed and not shrinking with N.
//
// Here rank r owns the r-th 1/N slice of the packed [P*59] record and ONE kernel per rank does, for its slice,
//     g  = multimem.ld_reduce.add(grad)        the switch sums the N ranks' gradients on the way in (NVLS)
//     p' = Adam(p, g, m, v)                    moments live only on the owner: 1/N of the Adam work per rank
//     multimem.st(param) = p'                  the switch broadcasts the updated slice to all N parameter buffers
// i.e. reduce-scatter + optimizer + all-gather fused, the transfers overlapping the arithmetic element by element.
// Both buffers are symmetric allocations bound to a multicast object (torch.distributed._symmetric_memory on the host
// side: plumbing); the kernel takes their multicast addresses.  Every element is reduced exactly once (by its owner)
// and broadcast, so the parameters stay bit-identical across ranks by construction.
//
// A 16-byte tail slot behind the gradient record carries {partial loss, overflow flag}: the same ld_reduce gives every
// rank the step's loss and tells it — identically on all ranks — whether some rank's capacity-mode forward overflowed
// (then nobody updates), which replaces two more NCCL calls (4-byte all-reduces) per step.
//
// Ordering between ranks (all gradients written before the reduce; all broadcasts landed before the next forward) is
// the caller's: a cross-rank barrier on the stream before and after the launch (symmetric-memory signal pads).
#include "../../include/igs_raster.h"
#include "common.cuh"
#include "kernels.cuh"
#include <cmath>

namespace igs {

struct ExchangeGroups {
  unsigned long long end[8];   // end offset (in floats) of each parameter group in the packed record
  float step_size[8];          // lr / (1 - beta1^t)
  int n;
};

__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__device__ __forceinline__ float adam_one(float p, float g, float& m, float& v, float step_size, float b1, float b2, float eps,
                                          float inv_sqrt_bc2) {
  // identical arithmetic to adam_groups_kernel (activations.cu): torch's lerp / addcmul forms
  const float mi = m + (1.0f - b1) * (g - m);
  const float vi = b2 * v + (1.0f - b2) * g * g;
  m = mi; v = vi;
  const float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
  return p - step_size * (mi / denom);
}

__global__ void __launch_bounds__(256)
fused_reduce_adam_kernel(const float* __restrict__ grad_mc, float* __restrict__ param_mc, const float* __restrict__ param,
                         float* __restrict__ m, float* __restrict__ v, unsigned long long begin4, unsigned long long end4,
                         ExchangeGroups gr, float b1, float b2, float eps, float inv_sqrt_bc2, unsigned long long tail_off,
                         float* __restrict__ loss_out, unsigned int* __restrict__ skipped_out) {
  __shared__ float s_tail[2];
  if (threadIdx.x == 0) {
    const float4 t = multimem_ld_reduce_add(grad_mc + tail_off);   // {sum of partial losses, number of overflowed ranks, -, -}
    s_tail[0] = t.x; s_tail[1] = t.y;
    if (blockIdx.x == 0) { *loss_out = t.x; *skipped_out = t.y > 0.0f ? 1u : 0u; }
  }
  __syncthreads();
  if (s_tail[1] > 0.0f) return;   // some rank's forward overflowed its capacity: nobody steps (same decision everywhere)
  // Four independent 16-byte items per thread and trip: the reduce is a round trip through the switch (microseconds),
  // so the loads of all four are issued before the first is used.
  constexpr int U = 4;
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  for (unsigned long long base4 = begin4 + (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; base4 < end4; base4 += U * stride) {
    float4 g[U], p[U], mm[U], vv[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const unsigned long long i4 = base4 + u * stride;
      if (i4 < end4) g[u] = multimem_ld_reduce_add(grad_mc + 4ull * i4);
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const unsigned long long i4 = base4 + u * stride;
      if (i4 < end4) {
        p[u] = *reinterpret_cast<const float4*>(param + 4ull * i4);
        mm[u] = *reinterpret_cast<const float4*>(m + 4ull * i4);
        vv[u] = *reinterpret_cast<const float4*>(v + 4ull * i4);
      }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const unsigned long long i4 = base4 + u * stride;
      if (i4 >= end4) continue;
      const unsigned long long i = 4ull * i4;
      float ss[4];
#pragma unroll
      for (int c = 0; c < 4; c++) {   // group boundaries need not be multiples of four
        float sz = gr.step_size[0];
#pragma unroll
        for (int k = 1; k < 8; k++)
          if (k < gr.n && i + c >= gr.end[k - 1]) sz = gr.step_size[k];
        ss[c] = sz;
      }
      p[u].x = adam_one(p[u].x, g[u].x, mm[u].x, vv[u].x, ss[0], b1, b2, eps, inv_sqrt_bc2);
      p[u].y = adam_one(p[u].y, g[u].y, mm[u].y, vv[u].y, ss[1], b1, b2, eps, inv_sqrt_bc2);
      p[u].z = adam_one(p[u].z, g[u].z, mm[u].z, vv[u].z, ss[2], b1, b2, eps, inv_sqrt_bc2);
      p[u].w = adam_one(p[u].w, g[u].w, mm[u].w, vv[u].w, ss[3], b1, b2, eps, inv_sqrt_bc2);
      *reinterpret_cast<float4*>(m + i) = mm[u];
      *reinterpret_cast<float4*>(v + i) = vv[u];
      multimem_st(param_mc + i, p[u]);
    }
  }
}

}  // namespace igs

using namespace igs;

extern "C" int igs_fused_reduce_adam(const float* grad_mc, float* param_mc, const float* param, float* exp_avg,
                                     float* exp_avg_sq, size_t n, size_t begin, size_t end, int ngroups, const size_t* sizes,
                                     const float* lrs, float beta1, float beta2, float eps, int step, float* loss_out,
                                     unsigned int* skipped_out, void* stream) {
  if (!grad_mc || !param_mc || !param || !exp_avg || !exp_avg_sq || !sizes || !lrs || !loss_out || !skipped_out ||
      ngroups < 1 || ngroups > 8 || step < 1) {
    set_error("bad argument");
    return IGS_ERR_INVALID;
  }
  if ((n & 3u) || (begin & 3u) || (end & 3u) || begin > end || end > n) {
    set_error("n, begin and end must be multiples of 4 floats with begin <= end <= n (got %zu, %zu, %zu)", n, begin, end);
    return IGS_ERR_INVALID;
  }
  if ((reinterpret_cast<uintptr_t>(grad_mc) | reinterpret_cast<uintptr_t>(param_mc) | reinterpret_cast<uintptr_t>(param) |
       reinterpret_cast<uintptr_t>(exp_avg) | reinterpret_cast<uintptr_t>(exp_avg_sq)) & 15u) {
    set_error("buffers must be 16-byte aligned");
    return IGS_ERR_INVALID;
  }
  ExchangeGroups gr;
  gr.n = ngroups;
  const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
  size_t off = 0;
  for (int k = 0; k < 8; k++) {
    if (k < ngroups) { off += sizes[k]; gr.step_size[k] = (float)((double)lrs[k] / bc1); } else gr.step_size[k] = 0.0f;
    gr.end[k] = off;
  }
  if (off != n) { set_error("group sizes must add up to n"); return IGS_ERR_INVALID; }
  const size_t n4 = (end - begin) / 4;
  size_t nb = (n4 + 4 * 256 - 1) / (4 * 256);   // four items per thread
  if (nb > 148 * 8) nb = 148 * 8;
  if (nb == 0) nb = 1;   // the tail slot is still reduced (loss / overflow flag)
  count_launch();
  {
    StageScope st(IGS_STAGE_ADAM, (cudaStream_t)stream);
    fused_reduce_adam_kernel<<<(unsigned)nb, 256, 0, (cudaStream_t)stream>>>(
        grad_mc, param_mc, param, exp_avg, exp_avg_sq, (unsigned long long)(begin / 4), (unsigned long long)(end / 4), gr, beta1,
        beta2, eps, (float)(1.0 / sqrt(bc2)), (unsigned long long)n, loss_out, skipped_out);
  }
  IGS_CUDA_CHECK(cudaGetLastError());
  return IGS_OK;
}
