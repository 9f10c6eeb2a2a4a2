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
#include "blend_common.cuh"
#include "kernels.cuh"
#include <cstdlib>

namespace igs {

#define IGS_LOSS_MAX_ROW_VIEWS 64
#define LOSS_BX 32
#define LOSS_BY 32
#define LOSS_R 5
#define LOSS_TH (LOSS_BY + 2 * LOSS_R)  // 42 rows: tile + halo
#define LOSS_TC 48                      // staged columns: x0-8 .. x0+39 (16-B aligned superset of the 42-column halo)
#define LOSS_X0 8                       // staged column of image column x0
// Row strides in shared memory are chosen for the 16-byte accesses of the horizontal pass.  A quarter-warp (the unit a
// 16-byte shared access is served in) covers column groups {0..3} (or {4..7}) of TWO adjacent rows (loss_htask); its
// eight 16-byte chunks fall into eight different 16-byte bank groups when the row stride is an ODD number of chunks for
// the pair tiles (adjacent column groups are 2 chunks apart) and 4 mod 8 chunks for the scalar tiles (1 chunk apart).
// ncu r02: the unpadded 48 / 32-column rows made every such access a 2-way conflict (35-39 % of all wavefronts).
#define LOSS_TCP 50                     // staged pair tiles: 48 columns + 2 pad = 25 chunks
#define LOSS_HP 34                      // horizontally filtered pair tiles: 32 columns + 2 pad = 17 chunks

struct LossParams {
  int planes, H, W;        // planes = V*3
  int plane_per_view;      // 3
  float win[11];           // gauss(x) = exp(-(x-5)^2 / (2*1.5^2)) / sum, computed like loss_utils.py:23-25
  float lambda;
  float inv_n;             // grad_scale / (3*H*W)
  const float* pred;
  const float* gt;
  float2* d01;             // scratch [planes][H][W] pairs (dS/dmu1, dS/dE11): pass 2 copies them straight into shared memory
  float* dE12;             // scratch [planes][H][W]
  float* dpred;
  double* partial;         // [gridDim.x*gridDim.y*gridDim.z]
  // Optional per-view pixel-row ranges: the loss of view v is summed over rows [rows[2v], rows[2v+1]) only (a rank
  // that renders a band of a view owns the loss terms of that band's rows).  has_rows == 0: all rows.
  int has_rows;
  int rows[2 * IGS_LOSS_MAX_ROW_VIEWS];
};

__device__ __forceinline__ void loss_row_range(const LossParams& p, int plane, int& r0, int& r1) {
  r0 = 0; r1 = p.H;
  if (p.has_rows) { const int v = plane / p.plane_per_view; r0 = p.rows[2 * v]; r1 = p.rows[2 * v + 1]; }
}

// Register-tiled separable filters on fp32x2 pairs.  The kernels are instruction-issue bound (ncu: DRAM 15-18 % of
// peak), so
//  * the five (pass 1) / three (pass 2) filtered quantities are carried as packed pairs — (mu1, mu2), (E11, E22) and
//    E12;  (dmu1, dE11) and dE12 — and one FFMA2 advances two of them; every element is still one fused multiply-add
//    per tap, rounded exactly like the scalar instruction;
//  * the tile + halo is staged with 16-byte global loads of an aligned 48-column superset (VEC: W % 4 == 0 and
//    16-byte aligned planes; each float4 is then entirely inside or entirely outside the image), which cut the
//    staging loop from 30-48 % of all executed instructions to a few percent;
//  * the horizontal pass gives each thread 4 adjacent output columns of one row (16 input pairs from eight 16-B
//    shared loads), the vertical pass 4 adjacent output rows of one column (14 inputs reused across the 4 outputs).
__device__ __forceinline__ void load16_pairs(const float2* row, f32x2 (&x)[16]) {
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const float4 a = *reinterpret_cast<const float4*>(row + 2 * i);
    x[2 * i] = pack2(a.x, a.y);
    x[2 * i + 1] = pack2(a.z, a.w);
  }
}
__device__ __forceinline__ void load20(const float* row, float (&x)[20]) {
#pragma unroll
  for (int i = 0; i < 5; i++) {
    const float4 a = *reinterpret_cast<const float4*>(row + 4 * i);
    x[4 * i] = a.x; x[4 * i + 1] = a.y; x[4 * i + 2] = a.z; x[4 * i + 3] = a.w;
  }
}
__device__ __forceinline__ void store4_pairs(float2* dst, const f32x2 (&o)[4]) {
  float a, b, c, d;
  unpack2(o[0], a, b); unpack2(o[1], c, d);
  *reinterpret_cast<float4*>(dst) = make_float4(a, b, c, d);
  unpack2(o[2], a, b); unpack2(o[3], c, d);
  *reinterpret_cast<float4*>(dst + 2) = make_float4(a, b, c, d);
}

// Horizontal-pass task t -> (row r, column group cg of 4 outputs): lanes 0-3 / 4-7 of a quarter-warp take column groups
// {0..3} (or {4..7}) of rows 2k / 2k+1 (see the stride note above).  t < LOSS_TH * 8.
__device__ __forceinline__ void loss_htask(int t, int& r, int& cg) {
  cg = (t & 3) | ((t >> 1) & 4);
  r = ((t >> 4) << 1) | ((t >> 2) & 1);
}

// Stage rows [y0-5, y0+37) x columns [x0-8, x0+40) of two planes A, B as interleaved pairs (zero outside the image).
template <bool VEC>
__device__ __forceinline__ void stage_pairs(float2 (*dst)[LOSS_TCP], const float* __restrict__ A,
                                            const float* __restrict__ B, int x0, int y0, int W, int H, int tid) {
  if constexpr (VEC) {
    // both iterations' loads are issued before the first store: one DRAM round trip per CTA instead of two
    constexpr int NT = LOSS_TH * (LOSS_TC / 4);
    float4 a[2], b[2];
#pragma unroll
    for (int it = 0; it < 2; it++) {
      const int t = tid + it * 256;
      const int ty = t / (LOSS_TC / 4), g = t - ty * (LOSS_TC / 4);
      const int gy = y0 + ty - LOSS_R, gx = x0 - LOSS_X0 + 4 * g;
      a[it] = make_float4(0, 0, 0, 0); b[it] = a[it];
      if (t < NT && gy >= 0 && gy < H && gx >= 0 && gx < W) {
        const uint32_t o = (uint32_t)gy * (uint32_t)W + (uint32_t)gx;
        a[it] = *reinterpret_cast<const float4*>(A + o);
        b[it] = *reinterpret_cast<const float4*>(B + o);
      }
    }
#pragma unroll
    for (int it = 0; it < 2; it++) {
      const int t = tid + it * 256;
      if (t >= NT) break;
      const int ty = t / (LOSS_TC / 4), g = t - ty * (LOSS_TC / 4);
      float4* d = reinterpret_cast<float4*>(&dst[ty][4 * g]);
      const float4 lo = make_float4(a[it].x, b[it].x, a[it].y, b[it].y), hi = make_float4(a[it].z, b[it].z, a[it].w, b[it].w);
      // a quarter-warp's eight threads are 2 chunks apart: the upper four store their second chunk first, so the eight
      // stores of one instruction hit eight different bank groups
      const int sw = (g >> 2) & 1;
      d[sw] = sw ? hi : lo;
      d[sw ^ 1] = sw ? lo : hi;
    }
  } else {
    for (int t = tid; t < LOSS_TH * LOSS_TC; t += 256) {
      const int ty = t / LOSS_TC, tx = t - ty * LOSS_TC;
      const int gy = y0 + ty - LOSS_R, gx = x0 - LOSS_X0 + tx;
      const bool ok = gy >= 0 && gy < H && gx >= 0 && gx < W;
      const uint32_t o = ok ? (uint32_t)gy * (uint32_t)W + (uint32_t)gx : 0u;
      dst[ty][tx] = make_float2(ok ? A[o] : 0.0f, ok ? B[o] : 0.0f);
    }
  }
}
// Pass 2: stage the same window of the partials — pairs D = (dS/dmu1, dS/dE11) and the scalar plane E = dS/dE12.
// VEC: asynchronous 16-byte global -> shared copies (cp.async, zero-filled outside the image): nothing passes through
// registers and all of a thread's six copies are in flight at once (ncu r02: the register-staged loop made this kernel
// latency bound — long-scoreboard 14.5 stalled warps per issue, 40 % issue utilisation).
// Partials exist for the rows [r0, r1) this call owns only (pass 1 writes nothing else): every other row is staged as
// zeros without being read, so the scratch needs no zero-fill in band mode.
template <bool VEC>
__device__ __forceinline__ void stage_partials(float2 (*dp)[LOSS_TCP], float (*ds)[LOSS_TC], const float2* __restrict__ D,
                                               const float* __restrict__ E, int x0, int y0, int W, int r0, int r1, int tid) {
  if constexpr (VEC) {
    for (int t = tid; t < LOSS_TH * (LOSS_TC / 2); t += 256) {   // 16 bytes = 2 pairs
      const int ty = t / (LOSS_TC / 2), g = t - ty * (LOSS_TC / 2);
      const int gy = y0 + ty - LOSS_R, gx = x0 - LOSS_X0 + 2 * g;
      const bool ok = gy >= r0 && gy < r1 && gx >= 0 && gx < W;
      cp_async16_zfill(smem_addr(&dp[ty][2 * g]), D + (ok ? (uint32_t)gy * (uint32_t)W + (uint32_t)gx : 0u), ok);
    }
    for (int t = tid; t < LOSS_TH * (LOSS_TC / 4); t += 256) {   // 16 bytes = 4 scalars
      const int ty = t / (LOSS_TC / 4), g = t - ty * (LOSS_TC / 4);
      const int gy = y0 + ty - LOSS_R, gx = x0 - LOSS_X0 + 4 * g;
      const bool ok = gy >= r0 && gy < r1 && gx >= 0 && gx < W;
      cp_async16_zfill(smem_addr(&ds[ty][4 * g]), E + (ok ? (uint32_t)gy * (uint32_t)W + (uint32_t)gx : 0u), ok);
    }
    cp_async_commit();
  } else {
    for (int t = tid; t < LOSS_TH * LOSS_TC; t += 256) {
      const int ty = t / LOSS_TC, tx = t - ty * LOSS_TC;
      const int gy = y0 + ty - LOSS_R, gx = x0 - LOSS_X0 + tx;
      const bool ok = gy >= r0 && gy < r1 && gx >= 0 && gx < W;
      const uint32_t o = ok ? (uint32_t)gy * (uint32_t)W + (uint32_t)gx : 0u;
      dp[ty][tx] = ok ? D[o] : make_float2(0.0f, 0.0f);
      ds[ty][tx] = ok ? E[o] : 0.0f;
    }
  }
}

template <bool VEC>
__global__ void __launch_bounds__(256)
loss_pass1_kernel(LossParams p) {
  __shared__ __align__(16) float2 sxy[LOSS_TH][LOSS_TCP];          // (pred, gt)
  __shared__ __align__(16) float2 hbA[LOSS_TH][LOSS_HP];           // horizontally filtered (x, y)
  __shared__ __align__(16) float2 hbB[LOSS_TH][LOSS_HP];           // (x^2, y^2)
  __shared__ __align__(16) float hbC[LOSS_TH][LOSS_BX];            // x y; odd rows with the two 16-column halves swapped
  __shared__ double red[8];
  const int plane = blockIdx.z;
  const size_t po = (size_t)plane * p.H * p.W;
  const int x0 = blockIdx.x * LOSS_BX, y0 = blockIdx.y * LOSS_BY;
  const int tid = threadIdx.x;
  int r0, r1;
  loss_row_range(p, plane, r0, r1);
  if (y0 >= r1 || y0 + LOSS_BY <= r0) {  // no owned row in this tile: nothing to write (pass 2 reads owned rows only)
    if (tid == 0) p.partial[((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = 0.0;
    return;
  }
  stage_pairs<VEC>(sxy, p.pred + po, p.gt + po, x0, y0, p.W, p.H, tid);
  __syncthreads();
  f32x2 w2[11];
#pragma unroll
  for (int k = 0; k < 11; k++) w2[k] = pack2(p.win[k], p.win[k]);
  // horizontal: task = (row r, column group cg of 4 outputs); taps of output column 4cg+j are staged columns
  // 4cg+3+j .. 4cg+13+j = v[1+j .. 11+j] of the 16 pairs loaded from the (16-B aligned) column 4cg+2
  for (int t = tid; t < LOSS_TH * (LOSS_BX / 4); t += 256) {
    int r, cg;
    loss_htask(t, r, cg);
    f32x2 v[16], sq[16];
    float pr[16];
    load16_pairs(&sxy[r][cg * 4 + 2], v);
#pragma unroll
    for (int k = 1; k < 15; k++) {
      float xx, yy;
      unpack2(v[k], xx, yy);
      sq[k] = mul2(v[k], v[k]);
      pr[k] = xx * yy;
    }
    f32x2 oA[4], oB[4];
    float oC[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      f32x2 a = pack2(0.0f, 0.0f), b = pack2(0.0f, 0.0f);
      float c = 0.0f;
#pragma unroll
      for (int k = 0; k < 11; k++) {
        a = fma2(w2[k], v[1 + j + k], a);
        b = fma2(w2[k], sq[1 + j + k], b);
        c = fmaf(p.win[k], pr[1 + j + k], c);
      }
      oA[j] = a; oB[j] = b; oC[j] = c;
    }
    store4_pairs(&hbA[r][cg * 4], oA);
    store4_pairs(&hbB[r][cg * 4], oB);
    *reinterpret_cast<float4*>(&hbC[r][(cg ^ ((r & 1) << 2)) * 4]) = make_float4(oC[0], oC[1], oC[2], oC[3]);
  }
  __syncthreads();
  // vertical + SSIM + partials: task = (column c, row group rg of 4 outputs), all 256 threads
  double acc = 0.0;
  const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
  {
    const int c = tid % LOSS_BX, rg = tid / LOSS_BX;
    f32x2 mA[4], mB[4];
    float mC[4];
    {
      f32x2 in[14];
#pragma unroll
      for (int k = 0; k < 14; k++) { const float2 t = hbA[rg * 4 + k][c]; in[k] = pack2(t.x, t.y); }
#pragma unroll
      for (int j = 0; j < 4; j++) {
        f32x2 a = pack2(0.0f, 0.0f);
#pragma unroll
        for (int k = 0; k < 11; k++) a = fma2(w2[k], in[j + k], a);
        mA[j] = a;
      }
#pragma unroll
      for (int k = 0; k < 14; k++) { const float2 t = hbB[rg * 4 + k][c]; in[k] = pack2(t.x, t.y); }
#pragma unroll
      for (int j = 0; j < 4; j++) {
        f32x2 a = pack2(0.0f, 0.0f);
#pragma unroll
        for (int k = 0; k < 11; k++) a = fma2(w2[k], in[j + k], a);
        mB[j] = a;
      }
      float ins[14];
#pragma unroll
      for (int k = 0; k < 14; k++) ins[k] = hbC[rg * 4 + k][c ^ ((k & 1) << 4)];   // (row parity = k parity)
#pragma unroll
      for (int j = 0; j < 4; j++) {
        float a = 0.0f;
#pragma unroll
        for (int k = 0; k < 11; k++) a = fmaf(p.win[k], ins[j + k], a);
        mC[j] = a;
      }
    }
    const int gx = x0 + c;
    float accf = 0.0f;
    float2* __restrict__ o_01 = p.d01 + po;
    float* __restrict__ o_12 = p.dE12 + po;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int gy = y0 + rg * 4 + j;
      if (gx >= p.W || gy >= r1 || gy < r0) continue;
      float mu1, mu2, e11, e22;
      unpack2(mA[j], mu1, mu2);
      unpack2(mB[j], e11, e22);
      const float e12 = mC[j];
      const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
      const float s1 = e11 - mu1_sq, s2 = e22 - mu2_sq, s12 = e12 - mu12;
      const float A = 2.0f * mu12 + C1, B = 2.0f * s12 + C2, Cc = mu1_sq + mu2_sq + C1, D = s1 + s2 + C2;
      const float inv = 1.0f / (Cc * D);
      const float ssim = A * B * inv;
      // partials of ssim w.r.t. mu1 (E's held fixed), E11, E12;  1 / D = Cc * inv (one division per pixel)
      const float dmu1 = 2.0f * inv * (mu2 * (B - A) - ssim * mu1 * (D - Cc));
      const float d11 = -ssim * Cc * inv;
      const float d12 = 2.0f * A * inv;
      const uint32_t o = (uint32_t)gy * (uint32_t)p.W + (uint32_t)gx;
      o_01[o] = make_float2(dmu1, d11); o_12[o] = d12;
      const float2 xy = sxy[rg * 4 + j + LOSS_R][c + LOSS_X0];
      accf += p.lambda * fabsf(xy.x - xy.y) - (1.0f - p.lambda) * ssim;   // four pixels in fp32, then double
    }
    acc = (double)accf;
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

template <bool VEC>
__global__ void __launch_bounds__(256)
loss_pass2_kernel(LossParams p) {
  __shared__ __align__(16) float2 s01[LOSS_TH][LOSS_TCP];          // (dS/dmu1, dS/dE11)
  __shared__ __align__(16) float s2[LOSS_TH][LOSS_TC];             // dS/dE12 (48 columns = 12 chunks: 4 mod 8)
  __shared__ __align__(16) float2 hb01[LOSS_TH][LOSS_HP];
  __shared__ __align__(16) float hb2[LOSS_TH][LOSS_BX];            // odd rows with the two 16-column halves swapped
  const int plane = blockIdx.z;
  const size_t po = (size_t)plane * p.H * p.W;
  const int x0 = blockIdx.x * LOSS_BX, y0 = blockIdx.y * LOSS_BY;
  const int tid = threadIdx.x;
  int r0, r1;
  loss_row_range(p, plane, r0, r1);
  if (y0 >= r1 + LOSS_R || y0 + LOSS_BY <= r0 - LOSS_R) {  // out of reach of every owned row: dL/dpred is zero here
    const int c = tid % LOSS_BX, rg = tid / LOSS_BX;
    float* __restrict__ O = p.dpred + po;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int gx = x0 + c, gy = y0 + rg * 4 + j;
      if (gx < p.W && gy < p.H) O[(uint32_t)gy * (uint32_t)p.W + (uint32_t)gx] = 0.0f;
    }
    return;
  }
  stage_partials<VEC>(s01, s2, p.d01 + po, p.dE12 + po, x0, y0, p.W, r0, r1, tid);
  if constexpr (VEC) cp_async_wait_all();
  __syncthreads();
  f32x2 w2[11];
#pragma unroll
  for (int k = 0; k < 11; k++) w2[k] = pack2(p.win[k], p.win[k]);
  for (int t = tid; t < LOSS_TH * (LOSS_BX / 4); t += 256) {
    int r, cg;
    loss_htask(t, r, cg);
    f32x2 v[16];
    float u[20];
    load16_pairs(&s01[r][cg * 4 + 2], v);  // taps of output column 4cg+j: v[1+j .. 11+j]
    load20(&s2[r][cg * 4], u);             //                              u[3+j .. 13+j]
    f32x2 o01[4];
    float o2[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      f32x2 a = pack2(0.0f, 0.0f);
      float c = 0.0f;
#pragma unroll
      for (int k = 0; k < 11; k++) {
        a = fma2(w2[k], v[1 + j + k], a);
        c = fmaf(p.win[k], u[3 + j + k], c);
      }
      o01[j] = a; o2[j] = c;
    }
    store4_pairs(&hb01[r][cg * 4], o01);
    *reinterpret_cast<float4*>(&hb2[r][(cg ^ ((r & 1) << 2)) * 4]) = make_float4(o2[0], o2[1], o2[2], o2[3]);
  }
  __syncthreads();
  {
    const int c = tid % LOSS_BX, rg = tid / LOSS_BX;
    const int gx = x0 + c;
    // The staged partials are dead now: this thread's four (pred, gt) values are copied into that space asynchronously
    // (each thread copies, waits for and reads only its own words), so their DRAM latency overlaps the vertical filter.
    float* const xy = reinterpret_cast<float*>(&s01[0][0]) + tid;   // word k of this thread at xy[256 k] (conflict-free)
    {
      const float* __restrict__ X = p.pred + po;
      const float* __restrict__ Y = p.gt + po;
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int gy = y0 + rg * 4 + j;
        if (gx >= p.W || gy >= p.H) continue;
        const uint32_t o = (uint32_t)gy * (uint32_t)p.W + (uint32_t)gx;
        cp_async4(smem_addr(xy + 256 * (2 * j)), X + o);
        cp_async4(smem_addr(xy + 256 * (2 * j + 1)), Y + o);
      }
      cp_async_commit();
    }
    f32x2 m01[4];
    float m2[4];
    {
      f32x2 in[14];
#pragma unroll
      for (int k = 0; k < 14; k++) { const float2 t = hb01[rg * 4 + k][c]; in[k] = pack2(t.x, t.y); }
#pragma unroll
      for (int j = 0; j < 4; j++) {
        f32x2 a = pack2(0.0f, 0.0f);
#pragma unroll
        for (int k = 0; k < 11; k++) a = fma2(w2[k], in[j + k], a);
        m01[j] = a;
      }
      float ins[14];
#pragma unroll
      for (int k = 0; k < 14; k++) ins[k] = hb2[rg * 4 + k][c ^ ((k & 1) << 4)];   // (row parity = k parity)
#pragma unroll
      for (int j = 0; j < 4; j++) {
        float a = 0.0f;
#pragma unroll
        for (int k = 0; k < 11; k++) a = fmaf(p.win[k], ins[j + k], a);
        m2[j] = a;
      }
    }
    cp_async_wait_all();
    float* __restrict__ O = p.dpred + po;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int gy = y0 + rg * 4 + j;
      if (gx >= p.W || gy >= p.H) continue;
      const uint32_t o = (uint32_t)gy * (uint32_t)p.W + (uint32_t)gx;
      const float xv = xy[256 * (2 * j)], yv = xy[256 * (2 * j + 1)];
      float g0, g1;
      unpack2(m01[j], g0, g1);
      const float dssim = g0 + 2.0f * xv * g1 + yv * m2[j];
      const float d = xv - yv;
      float sgn = (d > 0.0f) ? 1.0f : ((d < 0.0f) ? -1.0f : 0.0f);   // torch: sign(0) = 0
      if (gy < r0 || gy >= r1) sgn = 0.0f;                            // the L1 term of this row belongs to another rank
      O[o] = (p.lambda * sgn - (1.0f - p.lambda) * dssim) * p.inv_n;
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

// Scratch: the three per-pixel partials (36 B / pixel) of all V views + one double per CTA.  ncu measured 3.57 GB of DRAM
// traffic per step for this two-launch order (0.94 GB of partials written, 1.65 GB re-read) against 0.99 GB of algorithmic
// bytes.  EXPERIMENT (IGS_LOSS_BY_VIEW=1): the two passes view by view over a single-view scratch (49 MB at 1352x1014), so
// that the partials never leave the 126 MB L2 — measured SLOWER (2.16 vs 1.86 ms at C3): the kernels are issue / latency
// bound, not DRAM bound, and 40 launches of 3.5 waves each lose more in their tails than the L2 hits win.  Off by default.
static bool loss_by_view() {
  static const bool on = [] { const char* e = getenv("IGS_LOSS_BY_VIEW"); return e && atoi(e) != 0; }();
  return on;
}

size_t refine_loss_scratch_bytes(int V, int H, int W) {
  size_t planes = (size_t)(loss_by_view() ? 1 : V) * 3, N = (size_t)H * W;
  size_t nblk = (size_t)V * 3 * ((W + LOSS_BX - 1) / LOSS_BX) * ((H + LOSS_BY - 1) / LOSS_BY);
  return 3 * planes * N * sizeof(float) + 256 + nblk * sizeof(double) + 256;
}

void launch_refine_loss(const float* pred, const float* gt, int V, int H, int W, float lambda, float grad_scale,
                        const int* loss_rows_host, float* loss_sum,
                        float* dpred, char* scratch, cudaStream_t s) {
  LossParams p;
  p.has_rows = loss_rows_host != nullptr ? 1 : 0;
  double owned_rows = (double)V * H;   // sum over views of the rows whose loss terms are computed here
  if (p.has_rows) {
    owned_rows = 0.0;
    for (int v = 0; v < V; v++) {
      int a = loss_rows_host[2 * v], b = loss_rows_host[2 * v + 1];
      a = a < 0 ? 0 : (a > H ? H : a);
      b = b < a ? a : (b > H ? H : b);
      p.rows[2 * v] = a; p.rows[2 * v + 1] = b;
      owned_rows += (double)(b - a);
    }
  }
  {
    float gf[11], fsum = 0;
    for (int x = 0; x < 11; x++) gf[x] = (float)exp(-(double)((x - 5) * (x - 5)) / (2.0 * 1.5 * 1.5));
    for (int x = 0; x < 11; x++) fsum += gf[x];  // torch: float32 tensor divided by its float32 sum
    for (int x = 0; x < 11; x++) p.win[x] = gf[x] / fsum;
  }
  const bool by_view = loss_by_view();
  p.planes = V * 3; p.H = H; p.W = W; p.plane_per_view = 3; p.lambda = lambda;
  p.inv_n = grad_scale / (3.0f * (float)H * (float)W);  // only dL_dpred uses it; the loss value is finished in double
  p.pred = pred; p.gt = gt; p.dpred = dpred;
  const size_t N = (size_t)H * W, planes = (size_t)(by_view ? 1 : V) * 3;
  p.d01 = reinterpret_cast<float2*>(scratch);
  p.dE12 = reinterpret_cast<float*>(scratch) + 2 * planes * N;
  size_t off = (3 * planes * N * sizeof(float) + 255) & ~size_t(255);
  p.partial = reinterpret_cast<double*>(scratch + off);
  if ((size_t)H * W >= (1ull << 32)) return;  // in-plane offsets are 32-bit (igs_refine_loss rejects such sizes)
  // 16-byte staging loads need W % 4 == 0 (every float4 entirely inside or outside the image) and aligned planes
  const bool vec = (W % 4 == 0) && ((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(gt) |
                                     reinterpret_cast<uintptr_t>(p.d01)) % 16 == 0) && ((size_t)H * W < (1ull << 31));
  const dim3 grid_all((W + LOSS_BX - 1) / LOSS_BX, (H + LOSS_BY - 1) / LOSS_BY, V * 3);
  const size_t nblk = (size_t)grid_all.x * grid_all.y * grid_all.z;
  if (!by_view) {
    count_launch(2);
    if (vec) {
      loss_pass1_kernel<true><<<grid_all, 256, 0, s>>>(p);
      loss_pass2_kernel<true><<<grid_all, 256, 0, s>>>(p);
    } else {
      loss_pass1_kernel<false><<<grid_all, 256, 0, s>>>(p);
      loss_pass2_kernel<false><<<grid_all, 256, 0, s>>>(p);
    }
  } else {
    const dim3 grid(grid_all.x, grid_all.y, 3);
    const size_t nblk_view = (size_t)grid.x * grid.y * 3;
    for (int v = 0; v < V; v++) {
      LossParams q = p;
      q.planes = 3;
      q.pred = pred + (size_t)v * 3 * N; q.gt = gt + (size_t)v * 3 * N; q.dpred = dpred + (size_t)v * 3 * N;
      q.partial = p.partial + (size_t)v * nblk_view;
      if (p.has_rows) { q.rows[0] = p.rows[2 * v]; q.rows[1] = p.rows[2 * v + 1]; }
      count_launch(2);
      if (vec) {
        loss_pass1_kernel<true><<<grid, 256, 0, s>>>(q);
        loss_pass2_kernel<true><<<grid, 256, 0, s>>>(q);
      } else {
        loss_pass1_kernel<false><<<grid, 256, 0, s>>>(q);
        loss_pass2_kernel<false><<<grid, 256, 0, s>>>(q);
      }
    }
  }
  count_launch(1);
  // sum_v (1 - lambda) * mean(1 - ssim) = (1 - lambda) * (owned rows / H) - (1 - lambda) * sum(ssim) / (3 H W)
  loss_final_kernel<<<1, 256, 0, s>>>(p.partial, nblk, 1.0 / (3.0 * (double)H * (double)W), owned_rows / (double)H * (1.0 - (double)lambda),
                                      loss_sum);
}

}  // namespace igs
