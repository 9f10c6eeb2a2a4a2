// loss.cu — fused refine loss  sum_v [ lambda * L1_v + (1 - lambda) * (1 - SSIM_v) ]  and its gradient w.r.t.
// the rendered image, replacing the torch ops of the reference refine loop (infer_batch.py:301-305,
// igs/utils/loss_utils.py:16-62: l1_loss, ssim with an 11x11 Gaussian window sigma 1.5, zero padding,
// C1 = 0.01^2, C2 = 0.03^2, mean over 3*H*W).
//
// Two tiled passes per (view, channel) plane, both separable (11 horizontal + 11 vertical taps through shared
// memory) instead of the reference's five dense 121-tap grouped convolutions forward and five more backward:
//   pass 1: mu1, mu2, E[x^2], E[y^2], E[xy] -> SSIM value, |x-y|, and the three per-pixel partials
//           dS/dmu1, dS/dE11, dS/dE12 (12 B/pixel of scratch);
//   pass 2: dS/dx(p) = (G * dmu1)(p) + 2 x(p) (G * dE11)(p) + y(p) (G * dE12)(p)   (the window is symmetric and the
//           padding is zero, so the adjoint of the convolution is the same convolution), combined with the L1 sign.
// The scalar loss is reduced deterministically: per-block partial sums in double, then one final block.
#include "common.cuh"
#include "kernels.cuh"

namespace igs {

#define LOSS_BX 32
#define LOSS_BY 16
#define LOSS_R 5
#define LOSS_TW (LOSS_BX + 2 * LOSS_R)  // 42
#define LOSS_TH (LOSS_BY + 2 * LOSS_R)  // 26

struct LossParams {
  int planes, H, W;        // planes = V*3
  int plane_per_view;      // 3
  float win[11];           // gauss(x) = exp(-(x-5)^2 / (2*1.5^2)) / sum, computed like loss_utils.py:23-25
  float lambda;
  float inv_n;             // grad_scale / (3*H*W)
  const float* pred;
  const float* gt;
  float* dmu1;             // scratch [planes][H][W]
  float* dE11;
  float* dE12;
  float* dpred;
  double* partial;         // [gridDim.x*gridDim.y*gridDim.z]
};

__device__ __forceinline__ float ld_pad(const float* __restrict__ p, int x, int y, int W, int H) {
  return (x >= 0 && x < W && y >= 0 && y < H) ? p[(size_t)y * W + x] : 0.0f;
}

// Register-tiled separable filter helpers.  Shared input tiles have a row stride of LOSS_TS floats (16-B aligned
// rows); the horizontal pass gives each thread 4 adjacent output columns of one row (14 inputs, loaded as three
// 16-B and one 8-B shared loads), the vertical pass gives each thread 4 adjacent output rows of one column
// (14 inputs reused across the 4 outputs) — ~4x fewer shared-memory instructions than one output per thread.
#define LOSS_TS 44

__device__ __forceinline__ void load14(const float* row, float (&x)[14]) {
  const float4 a = *reinterpret_cast<const float4*>(row), b = *reinterpret_cast<const float4*>(row + 4),
               c = *reinterpret_cast<const float4*>(row + 8);
  const float2 d = *reinterpret_cast<const float2*>(row + 12);
  x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w; x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
  x[8] = c.x; x[9] = c.y; x[10] = c.z; x[11] = c.w; x[12] = d.x; x[13] = d.y;
}

__global__ void __launch_bounds__(256)
loss_pass1_kernel(LossParams p) {
  __shared__ __align__(16) float sx[LOSS_TH][LOSS_TS];
  __shared__ __align__(16) float sy[LOSS_TH][LOSS_TS];
  __shared__ __align__(16) float hb[5][LOSS_TH][LOSS_BX];
  __shared__ double red[8];
  const int plane = blockIdx.z;
  const float* __restrict__ X = p.pred + (size_t)plane * p.H * p.W;
  const float* __restrict__ Y = p.gt + (size_t)plane * p.H * p.W;
  const int x0 = blockIdx.x * LOSS_BX, y0 = blockIdx.y * LOSS_BY;
  const int tid = threadIdx.x;
  // tile + halo, one row per warp iteration (no index division; zero padding outside the image)
  for (int ty = tid >> 5; ty < LOSS_TH; ty += 8) {
    const int gy = y0 + ty - LOSS_R;
    const bool rowok = gy >= 0 && gy < p.H;
    const size_t ro = (size_t)(rowok ? gy : 0) * p.W;
    for (int tx = tid & 31; tx < LOSS_TW; tx += 32) {
      const int gx = x0 + tx - LOSS_R;
      const bool ok = rowok && gx >= 0 && gx < p.W;
      sx[ty][tx] = ok ? X[ro + gx] : 0.0f;
      sy[ty][tx] = ok ? Y[ro + gx] : 0.0f;
    }
  }
  __syncthreads();
  // horizontal: task = (row r, column group cg of 4 outputs)
  if (tid < LOSS_TH * (LOSS_BX / 4)) {
    const int r = tid / (LOSS_BX / 4), cg = tid - r * (LOSS_BX / 4);
    float xv[14], yv[14];
    load14(&sx[r][cg * 4], xv);
    load14(&sy[r][cg * 4], yv);
    float o[5][4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      float a0 = 0, a1 = 0, a2 = 0, a3 = 0, a4 = 0;
#pragma unroll
      for (int k = 0; k < 11; k++) {
        const float w = p.win[k], xx = xv[j + k], yy = yv[j + k];
        a0 = fmaf(w, xx, a0); a1 = fmaf(w, yy, a1);
        a2 = fmaf(w, xx * xx, a2); a3 = fmaf(w, yy * yy, a3); a4 = fmaf(w, xx * yy, a4);
      }
      o[0][j] = a0; o[1][j] = a1; o[2][j] = a2; o[3][j] = a3; o[4][j] = a4;
    }
#pragma unroll
    for (int q = 0; q < 5; q++)
      *reinterpret_cast<float4*>(&hb[q][r][cg * 4]) = make_float4(o[q][0], o[q][1], o[q][2], o[q][3]);
  }
  __syncthreads();
  // vertical + SSIM + partials: task = (column c, row group rg of 4 outputs)
  double acc = 0.0;
  const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
  if (tid < LOSS_BX * (LOSS_BY / 4)) {
    const int c = tid % LOSS_BX, rg = tid / LOSS_BX;
    float m[5][4];
#pragma unroll
    for (int q = 0; q < 5; q++) {
      float in[14];
#pragma unroll
      for (int k = 0; k < 14; k++) in[k] = hb[q][rg * 4 + k][c];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        float a = 0;
#pragma unroll
        for (int k = 0; k < 11; k++) a = fmaf(p.win[k], in[j + k], a);
        m[q][j] = a;
      }
    }
    const int gx = x0 + c;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int gy = y0 + rg * 4 + j;
      if (gx >= p.W || gy >= p.H) continue;
      const float mu1 = m[0][j], mu2 = m[1][j], e11 = m[2][j], e22 = m[3][j], e12 = m[4][j];
      const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
      const float s1 = e11 - mu1_sq, s2 = e22 - mu2_sq, s12 = e12 - mu12;
      const float A = 2.0f * mu12 + C1, B = 2.0f * s12 + C2, Cc = mu1_sq + mu2_sq + C1, D = s1 + s2 + C2;
      const float inv = 1.0f / (Cc * D);
      const float ssim = A * B * inv;
      // partials of ssim w.r.t. mu1 (E's held fixed), E11, E12
      const float dmu1 = (2.0f * mu2 * B - 2.0f * mu2 * A) * inv - ssim * (2.0f * mu1 * D - 2.0f * mu1 * Cc) * inv;
      const float d11 = -ssim / D;
      const float d12 = 2.0f * A * inv;
      const size_t o = (size_t)plane * p.H * p.W + (size_t)gy * p.W + gx;
      p.dmu1[o] = dmu1; p.dE11[o] = d11; p.dE12[o] = d12;
      const float xv = sx[rg * 4 + j + LOSS_R][c + LOSS_R], yv = sy[rg * 4 + j + LOSS_R][c + LOSS_R];
      acc += (double)(p.lambda * fabsf(xv - yv)) - (double)((1.0f - p.lambda) * ssim);
    }
  }
  // block reduce (fixed order)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((tid & 31) == 0) red[tid >> 5] = acc;
  __syncthreads();
  if (tid == 0) {
    double t = 0;
    for (int w = 0; w < 8; w++) t += red[w];
    p.partial[((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(256)
loss_pass2_kernel(LossParams p) {
  __shared__ __align__(16) float s0[LOSS_TH][LOSS_TS];
  __shared__ __align__(16) float s1[LOSS_TH][LOSS_TS];
  __shared__ __align__(16) float s2[LOSS_TH][LOSS_TS];
  __shared__ __align__(16) float hb[3][LOSS_TH][LOSS_BX];
  const int plane = blockIdx.z;
  const size_t po = (size_t)plane * p.H * p.W;
  const int x0 = blockIdx.x * LOSS_BX, y0 = blockIdx.y * LOSS_BY;
  const int tid = threadIdx.x;
  for (int ty = tid >> 5; ty < LOSS_TH; ty += 8) {
    const int gy = y0 + ty - LOSS_R;
    const bool rowok = gy >= 0 && gy < p.H;
    const size_t ro = po + (size_t)(rowok ? gy : 0) * p.W;
    for (int tx = tid & 31; tx < LOSS_TW; tx += 32) {
      const int gx = x0 + tx - LOSS_R;
      const bool ok = rowok && gx >= 0 && gx < p.W;
      s0[ty][tx] = ok ? p.dmu1[ro + gx] : 0.0f;
      s1[ty][tx] = ok ? p.dE11[ro + gx] : 0.0f;
      s2[ty][tx] = ok ? p.dE12[ro + gx] : 0.0f;
    }
  }
  __syncthreads();
  if (tid < LOSS_TH * (LOSS_BX / 4)) {
    const int r = tid / (LOSS_BX / 4), cg = tid - r * (LOSS_BX / 4);
    float v0[14], v1[14], v2[14];
    load14(&s0[r][cg * 4], v0);
    load14(&s1[r][cg * 4], v1);
    load14(&s2[r][cg * 4], v2);
    float o[3][4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      float a0 = 0, a1 = 0, a2 = 0;
#pragma unroll
      for (int k = 0; k < 11; k++) {
        const float w = p.win[k];
        a0 = fmaf(w, v0[j + k], a0); a1 = fmaf(w, v1[j + k], a1); a2 = fmaf(w, v2[j + k], a2);
      }
      o[0][j] = a0; o[1][j] = a1; o[2][j] = a2;
    }
#pragma unroll
    for (int q = 0; q < 3; q++)
      *reinterpret_cast<float4*>(&hb[q][r][cg * 4]) = make_float4(o[q][0], o[q][1], o[q][2], o[q][3]);
  }
  __syncthreads();
  if (tid < LOSS_BX * (LOSS_BY / 4)) {
    const int c = tid % LOSS_BX, rg = tid / LOSS_BX;
    float m[3][4];
#pragma unroll
    for (int q = 0; q < 3; q++) {
      float in[14];
#pragma unroll
      for (int k = 0; k < 14; k++) in[k] = hb[q][rg * 4 + k][c];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        float a = 0;
#pragma unroll
        for (int k = 0; k < 11; k++) a = fmaf(p.win[k], in[j + k], a);
        m[q][j] = a;
      }
    }
    const int gx = x0 + c;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int gy = y0 + rg * 4 + j;
      if (gx >= p.W || gy >= p.H) continue;
      const size_t o = po + (size_t)gy * p.W + gx;
      const float xv = p.pred[o], yv = p.gt[o];
      const float dssim = m[0][j] + 2.0f * xv * m[1][j] + yv * m[2][j];
      const float d = xv - yv;
      const float sgn = (d > 0.0f) ? 1.0f : ((d < 0.0f) ? -1.0f : 0.0f);   // torch: sign(0) = 0
      p.dpred[o] = (p.lambda * sgn - (1.0f - p.lambda) * dssim) * p.inv_n;
    }
  }
}

__global__ void __launch_bounds__(256)
loss_final_kernel(const double* __restrict__ partial, size_t n, double inv_n, double constant, float* __restrict__ out) {
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
    *out = (float)(t * inv_n + constant);
  }
}

size_t refine_loss_scratch_bytes(int V, int H, int W) {
  size_t planes = (size_t)V * 3, N = (size_t)H * W;
  size_t nblk = planes * ((W + LOSS_BX - 1) / LOSS_BX) * ((H + LOSS_BY - 1) / LOSS_BY);
  return 3 * planes * N * sizeof(float) + 256 + nblk * sizeof(double) + 256;
}

void launch_refine_loss(const float* pred, const float* gt, int V, int H, int W, float lambda, float grad_scale,
                        float* loss_sum,
                        float* dpred, char* scratch, cudaStream_t s) {
  LossParams p;
  {
    float gf[11], fsum = 0;
    for (int x = 0; x < 11; x++) gf[x] = (float)exp(-(double)((x - 5) * (x - 5)) / (2.0 * 1.5 * 1.5));
    for (int x = 0; x < 11; x++) fsum += gf[x];  // torch: float32 tensor divided by its float32 sum
    for (int x = 0; x < 11; x++) p.win[x] = gf[x] / fsum;
  }
  p.planes = V * 3; p.H = H; p.W = W; p.plane_per_view = 3; p.lambda = lambda;
  p.inv_n = grad_scale / (3.0f * (float)H * (float)W);  // only dL_dpred uses it; the loss value is finished in double
  p.pred = pred; p.gt = gt; p.dpred = dpred;
  size_t N = (size_t)H * W, planes = (size_t)V * 3;
  p.dmu1 = reinterpret_cast<float*>(scratch);
  p.dE11 = p.dmu1 + planes * N;
  p.dE12 = p.dE11 + planes * N;
  size_t off = (3 * planes * N * sizeof(float) + 255) & ~size_t(255);
  p.partial = reinterpret_cast<double*>(scratch + off);
  dim3 grid((W + LOSS_BX - 1) / LOSS_BX, (H + LOSS_BY - 1) / LOSS_BY, V * 3);
  count_launch(3);
  loss_pass1_kernel<<<grid, 256, 0, s>>>(p);
  loss_pass2_kernel<<<grid, 256, 0, s>>>(p);
  size_t nblk = (size_t)grid.x * grid.y * grid.z;
  loss_final_kernel<<<1, 256, 0, s>>>(p.partial, nblk, 1.0 / (3.0 * (double)H * (double)W), (double)V * (1.0 - (double)lambda),
                                      loss_sum);
}

}  // namespace igs
