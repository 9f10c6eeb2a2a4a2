// activations.cu — the refine loop's parameter activations and their chain rule, one launch each.
//
// Reference (RaDe-GS scene/gaussian_model.py, mirrored by igs/models/gaussian_model.py:89-127):
//   get_scaling  = exp(_scaling)                 scaling_activation   = torch.exp
//   get_rotation = normalize(_rotation)          rotation_activation  = torch.nn.functional.normalize (eps 1e-12)
//   get_opacity  = sigmoid(_opacity)             opacity_activation   = torch.sigmoid
// The reference runs them as separate torch kernels (plus their autograd nodes) every iteration; here the
// forward is one launch over P Gaussians and the backward one launch that writes straight into the packed
// [P*59] gradient record, so the refine step needs no autograd graph at all.
#include "common.cuh"
#include <cmath>
#include "kernels.cuh"

namespace igs {

__global__ void __launch_bounds__(256)
activate_fwd_kernel(int P, const float* __restrict__ log_scales, const float* __restrict__ raw_rot,
                    const float* __restrict__ logit, float* __restrict__ scales, float* __restrict__ rot,
                    float* __restrict__ opac) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  // torch.exp / torch.sigmoid on CUDA are expf-based fp32
  scales[3 * i + 0] = expf(log_scales[3 * i + 0]);
  scales[3 * i + 1] = expf(log_scales[3 * i + 1]);
  scales[3 * i + 2] = expf(log_scales[3 * i + 2]);
  // raw_rot may be a slice of a packed record at any 4-byte offset: scalar loads
  const float4 q = make_float4(raw_rot[4 * i], raw_rot[4 * i + 1], raw_rot[4 * i + 2], raw_rot[4 * i + 3]);
  const float n = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
  rot[4 * i] = q.x / n; rot[4 * i + 1] = q.y / n; rot[4 * i + 2] = q.z / n; rot[4 * i + 3] = q.w / n;
  opac[i] = 1.0f / (1.0f + expf(-logit[i]));
}

// In-place safe: every output element depends only on the same Gaussian's inputs, read before written.
__global__ void __launch_bounds__(256)
activate_bwd_kernel(int P, const float* __restrict__ scales, const float* __restrict__ raw_rot,
                    const float* __restrict__ opac, const float* dL_dscales, const float* dL_drot,
                    const float* dL_dopac, float* dL_dlog_scales, float* dL_draw_rot, float* dL_dlogit) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  // d exp(x) = exp(x)
  const float g0 = dL_dscales[3 * i + 0], g1 = dL_dscales[3 * i + 1], g2 = dL_dscales[3 * i + 2];
  dL_dlog_scales[3 * i + 0] = g0 * scales[3 * i + 0];
  dL_dlog_scales[3 * i + 1] = g1 * scales[3 * i + 1];
  dL_dlog_scales[3 * i + 2] = g2 * scales[3 * i + 2];
  // y = x / max(|x|, eps):  dx = (dy - y (y . dy)) / |x|   (|x| >= eps); below eps the divisor is the constant eps
  const float4 q = make_float4(raw_rot[4 * i], raw_rot[4 * i + 1], raw_rot[4 * i + 2], raw_rot[4 * i + 3]);
  const float4 d = make_float4(dL_drot[4 * i], dL_drot[4 * i + 1], dL_drot[4 * i + 2], dL_drot[4 * i + 3]);
  const float len = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
  float4 o;
  if (len > 1e-12f) {
    const float inv = 1.0f / len;
    const float yx = q.x * inv, yy = q.y * inv, yz = q.z * inv, yw = q.w * inv;
    const float dt = yx * d.x + yy * d.y + yz * d.z + yw * d.w;
    o = make_float4((d.x - yx * dt) * inv, (d.y - yy * dt) * inv, (d.z - yz * dt) * inv, (d.w - yw * dt) * inv);
  } else {
    o = make_float4(d.x * 1e12f, d.y * 1e12f, d.z * 1e12f, d.w * 1e12f);
  }
  dL_draw_rot[4 * i] = o.x; dL_draw_rot[4 * i + 1] = o.y; dL_draw_rot[4 * i + 2] = o.z; dL_draw_rot[4 * i + 3] = o.w;
  // d sigmoid = s (1 - s)
  const float s = opac[i];
  dL_dlogit[i] = dL_dopac[i] * s * (1.0f - s);
}

void launch_activate_fwd(int P, const float* log_scales, const float* raw_rot, const float* logit, float* scales,
                         float* rot, float* opac, cudaStream_t s) {
  if (P <= 0) return;
  count_launch();
  activate_fwd_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, log_scales, raw_rot, logit, scales, rot, opac);
}

void launch_activate_bwd(int P, const float* scales, const float* raw_rot, const float* opac, const float* dL_dscales,
                         const float* dL_drot, const float* dL_dopac, float* dL_dlog_scales, float* dL_draw_rot,
                         float* dL_dlogit, cudaStream_t s) {
  if (P <= 0) return;
  count_launch();
  activate_bwd_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, scales, raw_rot, opac, dL_dscales, dL_drot, dL_dopac,
                                                     dL_dlog_scales, dL_draw_rot, dL_dlogit);
}

// Adam over up to IGS_ADAM_MAX_GROUPS contiguous parameter groups of one flat buffer, one launch.
struct AdamGroups {
  int n;
  unsigned long long end[8];  // exclusive end offset of each group (groups are contiguous from 0)
  float step_size[8];
};

__global__ void __launch_bounds__(256)
adam_groups_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                   size_t n, AdamGroups gr, float b1, float b2, float eps, float inv_sqrt_bc2) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    float step_size = gr.step_size[0];
#pragma unroll
    for (int k = 1; k < 8; k++)
      if (k < gr.n && i >= gr.end[k - 1]) step_size = gr.step_size[k];
    const float gi = g[i];
    const float mi = m[i] + (1.0f - b1) * (gi - m[i]);
    const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
    p[i] = p[i] - step_size * (mi / denom);
  }
}

void launch_adam_groups(float* p, const float* g, float* m, float* v, int ngroups, const size_t* sizes,
                        const float* step_sizes, float b1, float b2, float eps, float inv_sqrt_bc2, cudaStream_t s) {
  AdamGroups gr;
  gr.n = ngroups;
  size_t off = 0;
  for (int k = 0; k < 8; k++) {
    if (k < ngroups) { off += sizes[k]; gr.step_size[k] = step_sizes[k]; } else gr.step_size[k] = 0.0f;
    gr.end[k] = off;
  }
  if (off == 0) return;
  size_t nb = (off + 255) / 256;
  if (nb > 148 * 16) nb = 148 * 16;
  count_launch();
  adam_groups_kernel<<<(unsigned)nb, 256, 0, s>>>(p, g, m, v, off, gr, b1, b2, eps, inv_sqrt_bc2);
}

// Guarded Adam: the same update, but (a) a no-op when *skip != 0 (a capacity-mode forward that overflowed) and (b) with the
// step count optionally read from device memory, so that a CUDA-graph replay advances it (bias corrections in double
// like torch's Python-side `1 - beta ** step`).
__global__ void __launch_bounds__(256)
adam_groups_guarded_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                           size_t n, AdamGroups gr, float b1, float b2, float eps, float inv_sqrt_bc2_host,
                           const unsigned int* __restrict__ skip, const int* __restrict__ step_dev) {
  // step_dev == nullptr: gr.step_size[] / inv_sqrt_bc2_host are the host-computed step sizes (exactly the unguarded
  // kernel's arithmetic); otherwise gr.step_size[] holds the learning rates and the corrections are computed here.
  if (skip != nullptr && *skip != 0u) return;
  __shared__ float s_step[8];
  __shared__ float s_inv_sqrt_bc2;
  if (threadIdx.x == 0) {
    if (step_dev != nullptr) {
      const int t = *step_dev + 1;
      const double bc1 = 1.0 - pow((double)b1, (double)t);
      const double bc2 = 1.0 - pow((double)b2, (double)t);
      for (int k = 0; k < 8; k++) s_step[k] = (float)((double)gr.step_size[k] / bc1);
      s_inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
    } else {
      for (int k = 0; k < 8; k++) s_step[k] = gr.step_size[k];
      s_inv_sqrt_bc2 = inv_sqrt_bc2_host;
    }
  }
  __syncthreads();
  const float inv_sqrt_bc2 = s_inv_sqrt_bc2;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    float step_size = s_step[0];
#pragma unroll
    for (int k = 1; k < 8; k++)
      if (k < gr.n && i >= gr.end[k - 1]) step_size = s_step[k];
    const float gi = g[i];
    const float mi = m[i] + (1.0f - b1) * (gi - m[i]);
    const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
    p[i] = p[i] - step_size * (mi / denom);
  }
}

__global__ void adam_step_advance_kernel(int* step_dev, const unsigned int* skip) {
  if (skip != nullptr && *skip != 0u) return;
  *step_dev += 1;
}

void launch_adam_groups_guarded(float* p, const float* g, float* m, float* v, int ngroups, const size_t* sizes,
                                const float* lrs, float b1, float b2, float eps, int step, const unsigned int* skip,
                                int* step_dev, cudaStream_t s) {
  AdamGroups gr;
  gr.n = ngroups;
  size_t off = 0;
  const double bc1 = 1.0 - pow((double)b1, (double)step), bc2 = 1.0 - pow((double)b2, (double)step);
  for (int k = 0; k < 8; k++) {
    if (k < ngroups) {
      off += sizes[k];
      gr.step_size[k] = step_dev ? lrs[k] : (float)((double)lrs[k] / bc1);
    } else gr.step_size[k] = 0.0f;
    gr.end[k] = off;
  }
  if (off == 0) return;
  size_t nb = (off + 255) / 256;
  if (nb > 148 * 16) nb = 148 * 16;
  count_launch();
  adam_groups_guarded_kernel<<<(unsigned)nb, 256, 0, s>>>(p, g, m, v, off, gr, b1, b2, eps,
                                                          step_dev ? 0.0f : (float)(1.0 / sqrt(bc2)), skip, step_dev);
  if (step_dev != nullptr) {   // after every block has read the step (stream order)
    count_launch();
    adam_step_advance_kernel<<<1, 1, 0, s>>>(step_dev, skip);
  }
}

}  // namespace igs
