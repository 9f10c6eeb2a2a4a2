// densify.cu — device side of the refine loop's densify / prune.
//
// Reference: igs/models/gaussian_model.py:466-666 (`_prune_optimizer`, `cat_tensors_to_optimizer`,
// `densify_and_clone`, `densify_and_split`, `densify_and_prune`) and :865-868 (`add_densification_stats`), driven by
// infer_batch.py:308-321.  The reference rebuilds every nn.Parameter and both Adam moments with torch.cat / boolean
// indexing three times per call (clone, split, prune) and then calls empty_cache().  Here the host computes ONE
// source map (which old Gaussian each new slot comes from, and how) from [P]-sized vectors, and one kernel re-lays
// out the packed parameter record and both Adam moments in a single pass:
//   kind 0  keep        parameters and moments carried over
//   kind 1  clone       parameters copied, moments zero          (densify_and_clone + cat_tensors_to_optimizer)
//   kind 2  split child parameters copied, moments zero, xyz = R(rotation) * sample + xyz,
//                       log_scale = log(exp(log_scale) / (0.8 N))                      (densify_and_split)
#include "common.cuh"
#include "kernels.cuh"

namespace igs {

// One step's densification statistics for all V views (the reference updates them once per single-view iteration):
//   grad_accum[g] += grad_scale * |dL_dmeans2D[v][g].xy| and denom[g] += 1 for every view v that saw g (radii > 0),
//   max_radii2D[g] = max(max_radii2D[g], radii[v][g]).
__global__ void __launch_bounds__(256)
densify_stats_kernel(int P, int V, const float* __restrict__ g2d, const int* __restrict__ radii, float grad_scale,
                     float* __restrict__ grad_accum, float* __restrict__ denom, float* __restrict__ max_radii) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= P) return;
  float acc = grad_accum[g], den = denom[g], mr = max_radii[g];
  for (int v = 0; v < V; v++) {
    const size_t i = (size_t)v * P + g;
    const int r = radii[i];
    if (r > 0) {
      const float gx = g2d[3 * i], gy = g2d[3 * i + 1];
      acc += grad_scale * sqrtf(gx * gx + gy * gy);   // torch.norm(grad[:, :2], dim=-1), on the per-view scale
      den += 1.0f;
      mr = fmaxf(mr, (float)r);
    }
  }
  grad_accum[g] = acc; denom[g] = den; max_radii[g] = mr;
}

struct DensifyGatherParams {
  int P_old, P_new;
  const int* src;            // [P_new] source Gaussian in the old arrays
  const uint8_t* kind;       // [P_new] 0 keep, 1 clone, 2 split child
  const float* samples;      // [P_new][3] N(0, scale) draws for split children (ignored otherwise; may be null)
  float split_div;           // 0.8 N
  // element offsets of (xyz, rotation, shs, opacity, scaling) in the old / new packed records
  unsigned long long off_old[5], off_new[5];
  const float* param_old; const float* m_old; const float* v_old;
  float* param_new; float* m_new; float* v_new;
};

__device__ __forceinline__ void role_of(int j, int& role, int& c, int& w) {
  // float j of the 59 per Gaussian, in the order xyz(3) rotation(4) shs(48) opacity(1) scaling(3)
  if (j < 3) { role = 0; c = j; w = 3; }
  else if (j < 7) { role = 1; c = j - 3; w = 4; }
  else if (j < 55) { role = 2; c = j - 7; w = 48; }
  else if (j < 56) { role = 3; c = 0; w = 1; }
  else { role = 4; c = j - 56; w = 3; }
}

__global__ void __launch_bounds__(256)
densify_gather_kernel(DensifyGatherParams p) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (size_t)p.P_new * 59) return;
  const int i = (int)(t / 59), j = (int)(t - (size_t)i * 59);
  int role, c, w;
  role_of(j, role, c, w);
  const int s = p.src[i];
  const int kind = p.kind[i];
  const size_t io = p.off_old[role] + (size_t)s * w + c;
  const size_t in = p.off_new[role] + (size_t)i * w + c;
  float val = p.param_old[io];
  if (kind == 2 && p.samples != nullptr) {
    if (role == 0) {
      // build_rotation(_rotation) (igs/utils/general_utils.py:79-100) row c, times the sample, plus the parent's xyz
      const float* q = p.param_old + p.off_old[1] + (size_t)s * 4;
      const float n = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
      const float r = q[0] / n, x = q[1] / n, y = q[2] / n, z = q[3] / n;
      float r0, r1, r2;
      if (c == 0) { r0 = 1.0f - 2.0f * (y * y + z * z); r1 = 2.0f * (x * y - r * z); r2 = 2.0f * (x * z + r * y); }
      else if (c == 1) { r0 = 2.0f * (x * y + r * z); r1 = 1.0f - 2.0f * (x * x + z * z); r2 = 2.0f * (y * z - r * x); }
      else { r0 = 2.0f * (x * z - r * y); r1 = 2.0f * (y * z + r * x); r2 = 1.0f - 2.0f * (x * x + y * y); }
      const float* sm = p.samples + (size_t)i * 3;
      val = (r0 * sm[0] + r1 * sm[1] + r2 * sm[2]) + val;
    } else if (role == 4) {
      val = logf(expf(val) / p.split_div);   // scaling_inverse_activation(get_scaling / (0.8 N))
    }
  }
  p.param_new[in] = val;
  p.m_new[in] = kind == 0 ? p.m_old[io] : 0.0f;
  p.v_new[in] = kind == 0 ? p.v_old[io] : 0.0f;
}

void launch_densify_stats(int P, int V, const float* g2d, const int* radii, float grad_scale, float* grad_accum,
                          float* denom, float* max_radii, cudaStream_t s) {
  if (P <= 0 || V <= 0) return;
  count_launch();
  densify_stats_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, V, g2d, radii, grad_scale, grad_accum, denom, max_radii);
}

void launch_densify_gather(int P_old, int P_new, const int* src, const uint8_t* kind, const float* samples,
                           float split_div, const size_t* off_old, const size_t* off_new, const float* param_old,
                           const float* m_old, const float* v_old, float* param_new, float* m_new, float* v_new,
                           cudaStream_t s) {
  if (P_new <= 0) return;
  DensifyGatherParams p;
  p.P_old = P_old; p.P_new = P_new; p.src = src; p.kind = kind; p.samples = samples; p.split_div = split_div;
  for (int k = 0; k < 5; k++) { p.off_old[k] = off_old[k]; p.off_new[k] = off_new[k]; }
  p.param_old = param_old; p.m_old = m_old; p.v_old = v_old;
  p.param_new = param_new; p.m_new = m_new; p.v_new = v_new;
  const size_t n = (size_t)P_new * 59;
  count_launch();
  densify_gather_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(p);
}

}  // namespace igs
