// geo_loss.cu — depth-normal consistency loss of RaDe-GS frame-0 training, value and gradients in one launch.
//
// Reference (torch, ~20 kernels + their autograd nodes per iteration): RADE/train.py:143-157 with
// RADE/utils/graphics_utils.py:97-126 (`depths_double_to_points`, `point_double_to_normal`,
// `depth_double_to_normal`):
//   points_k = depth_k * K^-1 (x + 0.5, y + 0.5, 1)                  k = 0 expected, 1 median   (or the coord maps)
//   a = P(y+1, x) - P(y-1, x),  b = P(y, x+1) - P(y, x-1)             interior pixels only (the reference's "dx", "dy")
//   n_k = normalize(a x b) (eps 1e-12), 0 on the one-pixel border
//   loss = (1 - r) * mean_HW(1 - <normal, n_0>) + r * mean_HW(1 - <normal, n_1>),  r = depth_ratio = 0.6
// Outputs dL/d(depth_0), dL/d(depth_1) (or dL/d(coord_k) [3][H][W]) and dL/d(normal), all scaled by `grad_scale`.
//
// One CTA per 32x8 pixel tile.  Points of the tile + 2-pixel halo are staged in shared memory; the per-pixel
// cross-product adjoints (da, db) are computed once for the tile + 1-pixel halo and staged too; each pixel then
// GATHERS its point gradient from its four neighbours (no atomics):
//   dP(y,x) = da(y-1,x) - da(y+1,x) + db(y,x-1) - db(y,x+1).
// HBM-bound stencil: 5 floats in (2 depths + normal), 5 floats out per pixel (depth variant).
#include "common.cuh"
#include "kernels.cuh"

namespace igs {

#define GL_TX 32
#define GL_TY 8
#define GL_PW (GL_TX + 4)   // points: halo 2
#define GL_PH (GL_TY + 4)
#define GL_AW (GL_TX + 2)   // adjoints: halo 1
#define GL_AH (GL_TY + 2)

struct GeoLossParams {
  const float* m0;      // depth variant: expected depth [V][H][W]; coord variant: expected coord [V][3][H][W]
  const float* m1;      // median depth / coord
  const float* normal;  // [V][3][H][W]
  float* d_m0;
  float* d_m1;
  float* d_normal;
  double* partial;      // one per CTA
  int V, H, W;
  float inv_fx, inv_fy, cx, cy;   // ray(x, y) = ((x + 0.5) * inv_fx - cx, (y + 0.5) * inv_fy - cy, 1)
  float w0, w1;                   // (1 - ratio) / (H W), ratio / (H W)
  float grad_scale;
};

__device__ __forceinline__ void cross3(const float a[3], const float b[3], float c[3]) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}

template <bool COORD>
__global__ void __launch_bounds__(256)
geo_loss_kernel(GeoLossParams p) {
  __shared__ float s_pt[2][3][GL_PH][GL_PW];       // points of both maps
  __shared__ float s_ad[2][6][GL_AH][GL_AW + 1];   // (da, db) of both maps
  __shared__ double s_red[8];
  const int v = blockIdx.z;
  const int x0 = blockIdx.x * GL_TX, y0 = blockIdx.y * GL_TY;
  const size_t N = (size_t)p.H * p.W;
  const float* M[2] = {p.m0 + (size_t)v * (COORD ? 3 : 1) * N, p.m1 + (size_t)v * (COORD ? 3 : 1) * N};
  // ---- 1. points of the tile + 2-pixel halo (zeros outside the image; never used there) ----
  for (int f = threadIdx.x; f < GL_PH * GL_PW; f += 256) {
    const int ly = f / GL_PW, lx = f - ly * GL_PW;
    const int gy = y0 - 2 + ly, gx = x0 - 2 + lx;
    const bool in = gy >= 0 && gy < p.H && gx >= 0 && gx < p.W;
    const size_t o = in ? (size_t)gy * p.W + gx : 0;
#pragma unroll
    for (int k = 0; k < 2; k++) {
      float q0 = 0, q1 = 0, q2 = 0;
      if (in) {
        if constexpr (COORD) {
          q0 = M[k][o]; q1 = M[k][N + o]; q2 = M[k][2 * N + o];
        } else {
          const float d = M[k][o];
          q0 = d * (((float)gx + 0.5f) * p.inv_fx - p.cx);
          q1 = d * (((float)gy + 0.5f) * p.inv_fy - p.cy);
          q2 = d;
        }
      }
      s_pt[k][0][ly][lx] = q0; s_pt[k][1][ly][lx] = q1; s_pt[k][2][ly][lx] = q2;
    }
  }
  __syncthreads();
  // ---- 2. normals, loss terms, dL/dnormal and cross-product adjoints on the tile + 1-pixel halo ----
  double acc = 0.0;
  const float* NM = p.normal + (size_t)v * 3 * N;
  for (int f = threadIdx.x; f < GL_AH * GL_AW; f += 256) {
    const int ly = f / GL_AW, lx = f - ly * GL_AW;
    const int gy = y0 - 1 + ly, gx = x0 - 1 + lx;
    const bool in = gy >= 0 && gy < p.H && gx >= 0 && gx < p.W;
    const bool interior = gy >= 1 && gy < p.H - 1 && gx >= 1 && gx < p.W - 1;
    const bool own = in && ly >= 1 && ly <= GL_TY && lx >= 1 && lx <= GL_TX;   // inside this CTA's tile proper
    float nm[3] = {0, 0, 0};
    const size_t o = in ? (size_t)gy * p.W + gx : 0;
    if (in) { nm[0] = NM[o]; nm[1] = NM[N + o]; nm[2] = NM[2 * N + o]; }
    float dn[3] = {0, 0, 0};
#pragma unroll
    for (int k = 0; k < 2; k++) {
      const float wk = k == 0 ? p.w0 : p.w1;
      float da[3] = {0, 0, 0}, db[3] = {0, 0, 0};
      float dotv = 0.0f;
      if (interior) {
        const int py = ly + 1, px = lx + 1;   // position in the point tile
        float a[3], b[3], c[3];
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
          a[ch] = s_pt[k][ch][py + 1][px] - s_pt[k][ch][py - 1][px];
          b[ch] = s_pt[k][ch][py][px + 1] - s_pt[k][ch][py][px - 1];
        }
        cross3(a, b, c);
        const float len = sqrtf(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
        const float den = fmaxf(len, 1e-12f);
        const float n[3] = {c[0] / den, c[1] / den, c[2] / den};
        dotv = nm[0] * n[0] + nm[1] * n[1] + nm[2] * n[2];
        // dL/dn = -wk * normal;  through v / max(|v|, eps)
        float g[3] = {-wk * nm[0], -wk * nm[1], -wk * nm[2]};
        float dc[3];
        if (len > 1e-12f) {
          const float ng = n[0] * g[0] + n[1] * g[1] + n[2] * g[2];
#pragma unroll
          for (int ch = 0; ch < 3; ch++) dc[ch] = (g[ch] - n[ch] * ng) / den;
        } else {
#pragma unroll
          for (int ch = 0; ch < 3; ch++) dc[ch] = g[ch] / den;
        }
        cross3(b, dc, da);   // c = a x b:  dL/da = b x dc,  dL/db = dc x a
        cross3(dc, a, db);
#pragma unroll
        for (int ch = 0; ch < 3; ch++) dn[ch] -= wk * n[ch];
      }
      if (own) acc += (double)(wk * (1.0f - dotv));
#pragma unroll
      for (int ch = 0; ch < 3; ch++) {
        s_ad[k][ch][ly][lx] = da[ch];
        s_ad[k][3 + ch][ly][lx] = db[ch];
      }
    }
    if (own) {
      float* DN = p.d_normal + (size_t)v * 3 * N;
      DN[o] = dn[0] * p.grad_scale; DN[N + o] = dn[1] * p.grad_scale; DN[2 * N + o] = dn[2] * p.grad_scale;
    }
  }
  __syncthreads();
  // ---- 3. gather the point gradients ----
  {
    const int lx = threadIdx.x & 31, ly = threadIdx.x >> 5;
    const int gy = y0 + ly, gx = x0 + lx;
    if (gy < p.H && gx < p.W) {
      const size_t o = (size_t)gy * p.W + gx;
      const int ay = ly + 1, ax = lx + 1;
#pragma unroll
      for (int k = 0; k < 2; k++) {
        float dP[3];
#pragma unroll
        for (int ch = 0; ch < 3; ch++)
          dP[ch] = s_ad[k][ch][ay - 1][ax] - s_ad[k][ch][ay + 1][ax] + s_ad[k][3 + ch][ay][ax - 1] - s_ad[k][3 + ch][ay][ax + 1];
        float* D = (k == 0 ? p.d_m0 : p.d_m1) + (size_t)v * (COORD ? 3 : 1) * N;
        if constexpr (COORD) {
          D[o] = dP[0] * p.grad_scale; D[N + o] = dP[1] * p.grad_scale; D[2 * N + o] = dP[2] * p.grad_scale;
        } else {
          const float rx = ((float)gx + 0.5f) * p.inv_fx - p.cx, ry = ((float)gy + 0.5f) * p.inv_fy - p.cy;
          D[o] = (dP[0] * rx + dP[1] * ry + dP[2]) * p.grad_scale;
        }
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0;
    for (int w = 0; w < 8; w++) t += s_red[w];
    p.partial[((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(256)
geo_loss_final_kernel(const double* __restrict__ partial, size_t n, float* __restrict__ out) {
  __shared__ double red[8];
  double acc = 0.0;
  for (size_t i = threadIdx.x; i < n; i += 256) acc += partial[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0;
    for (int w = 0; w < 8; w++) t += red[w];
    *out = (float)t;
  }
}

size_t geo_loss_scratch_bytes(int V, int H, int W) {
  size_t nblk = (size_t)V * ((W + GL_TX - 1) / GL_TX) * ((H + GL_TY - 1) / GL_TY);
  return nblk * sizeof(double) + 256;
}

void launch_geo_loss(bool coord, const float* m0, const float* m1, const float* normal, int V, int H, int W, float tan_fovx,
                     float tan_fovy, float depth_ratio, float grad_scale, float* loss_sum, float* d_m0, float* d_m1,
                     float* d_normal, char* scratch, cudaStream_t s) {
  GeoLossParams p;
  p.m0 = m0; p.m1 = m1; p.normal = normal; p.d_m0 = d_m0; p.d_m1 = d_m1; p.d_normal = d_normal;
  p.partial = reinterpret_cast<double*>(scratch);
  p.V = V; p.H = H; p.W = W;
  // fx = W / (2 tan(FoVx/2)); intrins_inv = [[1/fx, 0, -W/(2 fx)], [0, 1/fy, -H/(2 fy)], [0, 0, 1]]
  const float fx = (float)W / (2.0f * tan_fovx), fy = (float)H / (2.0f * tan_fovy);
  p.inv_fx = 1.0f / fx; p.inv_fy = 1.0f / fy;
  p.cx = (float)W / (2.0f * fx); p.cy = (float)H / (2.0f * fy);
  const float inv_n = 1.0f / ((float)H * (float)W);
  p.w0 = (1.0f - depth_ratio) * inv_n; p.w1 = depth_ratio * inv_n;
  p.grad_scale = grad_scale;
  dim3 grid((W + GL_TX - 1) / GL_TX, (H + GL_TY - 1) / GL_TY, V);
  count_launch(2);
  if (coord) geo_loss_kernel<true><<<grid, 256, 0, s>>>(p);
  else geo_loss_kernel<false><<<grid, 256, 0, s>>>(p);
  geo_loss_final_kernel<<<1, 256, 0, s>>>(p.partial, (size_t)grid.x * grid.y * grid.z, loss_sum);
}

}  // namespace igs
