// preprocess.cu — per-Gaussian forward preprocessing.
//
// Two kernels replace the reference's single preprocessCUDA<3,false> (DGR/cuda_rasterizer/forward.cu:307-423):
//   gauss_setup_kernel : view-INDEPENDENT part, once per Gaussian — cov3D from scale/quaternion
//                        (forward.cu:270-304) and the 3x3 eigen-decomposition + Vrk_inv that
//                        computeCov2D needs for the plane/normal terms (forward.cu:135-155).
//   project_kernel     : view-DEPENDENT part, once per (view, Gaussian) — near cull, projection,
//                        EWA cov2D / conic / low-pass coef, radius, tile rect, depth-plane and
//                        normal (forward.cu:78-264), SH -> RGB (forward.cu:23-74); writes the packed
//                        head/tail splat record and the (view | depth) sort key.
#include "common.cuh"
#include "kernels.cuh"
#include "eigen3.cuh"

namespace igs {

// ---------------------------------------------------------------------------------------------
// cov3D, bit-exact with the reference build (operation order from its sm_100a SASS).
// q = (r, x, y, z) is used un-normalised (reference forward.cu:279).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void cov3d_exact(float sx, float sy, float sz, float mod,
                                            float r, float x, float y, float z, float* c) {
  sx = fmul(mod, sx); sy = fmul(mod, sy); sz = fmul(mod, sz);
  float xz = fmul(x, z), zz = fmul(z, z), yy = fmul(y, y), rx = fmul(r, x), rz = fmul(r, z);
  float xz_p_ry = ffma(r, y, xz);
  float xz_m_ry = ffma(-r, y, xz);
  float yz_m_rx = ffma(y, z, -rx);
  float yz_p_rx = ffma(y, z, rx);
  float xy_m_rz = ffma(x, y, -rz);
  float xy_p_rz = ffma(x, y, rz);
  float yy_zz = fadd(yy, zz);
  float xx_zz = ffma(x, x, zz);
  float xx_yy = ffma(x, x, yy);
  // R in glm column-major: Rcr = R[col c][row r]
  float R00 = fsub(1.0f, fadd(yy_zz, yy_zz));
  float R01 = fadd(xy_m_rz, xy_m_rz);
  float R02 = fadd(xz_p_ry, xz_p_ry);
  float R10 = fadd(xy_p_rz, xy_p_rz);
  float R11 = fsub(1.0f, fadd(xx_zz, xx_zz));
  float R12 = fadd(yz_m_rx, yz_m_rx);
  float R20 = fadd(xz_m_ry, xz_m_ry);
  float R21 = fadd(yz_p_rx, yz_p_rx);
  float R22 = fsub(1.0f, fadd(xx_yy, xx_yy));
  // M = S * R : M[c][r] = s_r * R[c][r]
  float M00 = fmul(sx, R00), M01 = fmul(sy, R01), M02 = fmul(sz, R02);
  float M10 = fmul(sx, R10), M11 = fmul(sy, R11), M12 = fmul(sz, R12);
  float M20 = fmul(sx, R20), M21 = fmul(sy, R21), M22 = fmul(sz, R22);
  // Sigma = M^T M : Sigma[c][r] = sum_k M[r][k] * M[c][k]
  c[0] = dot3_ref(M00, M00, M01, M01, M02, M02);
  c[1] = dot3_ref(M10, M00, M11, M01, M12, M02);
  c[2] = dot3_ref(M20, M00, M21, M01, M22, M02);
  c[3] = dot3_ref(M10, M10, M11, M11, M12, M12);
  c[4] = dot3_ref(M20, M10, M21, M11, M22, M12);
  c[5] = dot3_ref(M20, M20, M21, M21, M22, M22);
}

__global__ void __launch_bounds__(256)
gauss_setup_kernel(int P, const float* __restrict__ scales, float mod,
                   const float* __restrict__ rotations, const float* __restrict__ cov3D_precomp,
                   float* __restrict__ cov3D_out, GaussEig* __restrict__ eig_out) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= P) return;
  float c[6];
  if (cov3D_precomp != nullptr) {
#pragma unroll
    for (int i = 0; i < 6; i++) c[i] = cov3D_precomp[6 * g + i];
  } else {
    float4 q = reinterpret_cast<const float4*>(rotations)[g];
    cov3d_exact(scales[3 * g], scales[3 * g + 1], scales[3 * g + 2], mod, q.x, q.y, q.z, q.w, c);
  }
#pragma unroll
  for (int i = 0; i < 6; i++) cov3D_out[6 * g + i] = c[i];

  float w[3], z[3][3];
  bool ok = eigen_sym3(c, w, z);
  GaussEig E;
  unsigned min_id = w[0] > w[1] ? (w[1] > w[2] ? 2u : 1u) : (w[0] > w[2] ? 2u : 0u);
  float wmin = min_id == 0 ? w[0] : (min_id == 1 ? w[1] : w[2]);
  bool well = (double)wmin > 0.00000001;
  if (well) {
    float i0 = 1.0f / w[0], i1 = 1.0f / w[1], i2 = 1.0f / w[2];
    // vinv[c][r] = sum_k (z[r][k] * inv_k) * z[c][k]
#pragma unroll
    for (int cc = 0; cc < 3; cc++)
#pragma unroll
      for (int rr = 0; rr < 3; rr++)
        E.vinv[cc * 3 + rr] = (z[rr][0] * i0) * z[cc][0] + (z[rr][1] * i1) * z[cc][1] + (z[rr][2] * i2) * z[cc][2];
  } else {
    float e0 = min_id == 0 ? z[0][0] : (min_id == 1 ? z[0][1] : z[0][2]);
    float e1 = min_id == 0 ? z[1][0] : (min_id == 1 ? z[1][1] : z[1][2]);
    float e2 = min_id == 0 ? z[2][0] : (min_id == 1 ? z[2][1] : z[2][2]);
    float ev[3] = {e0, e1, e2};
#pragma unroll
    for (int cc = 0; cc < 3; cc++)
#pragma unroll
      for (int rr = 0; rr < 3; rr++) E.vinv[cc * 3 + rr] = ev[rr] * ev[cc];
  }
#pragma unroll
  for (int i = 0; i < 3; i++) {
    E.eval[i] = w[i];
#pragma unroll
    for (int k = 0; k < 3; k++) E.evec[i * 3 + k] = z[k][i];
  }
  E.info = (ok ? 1u : 0u) | (well ? 2u : 0u) | (min_id << 4);
  E.pad[0] = E.pad[1] = 0.0f;
  // 96-byte record written as 6 x float4
  const float4* src = reinterpret_cast<const float4*>(&E);
  float4* dst = reinterpret_cast<float4*>(eig_out + g);
#pragma unroll
  for (int i = 0; i < 6; i++) dst[i] = src[i];
}

// SH basis constants (reference auxiliary.h:35-52)
__device__ __constant__ float kSH_C0 = 0.28209479177387814f;
__device__ __constant__ float kSH_C1 = 0.4886025119029199f;
__device__ __constant__ float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                           -1.0925484305920792f, 0.5462742152960396f};
__device__ __constant__ float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                           0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                           -0.5900435899266435f};

// SH -> RGB for one Gaussian and one camera (reference forward.cu:23-74). sh = [M][3].
__device__ __forceinline__ void sh_to_rgb(int deg, const float* __restrict__ sh, float dx, float dy, float dz,
                                          float rgb[3], uint32_t& clamped) {
  float len = sqrtf(dx * dx + dy * dy + dz * dz);
  float x = dx / len, y = dy / len, z = dz / len;
  float res[3];
#pragma unroll
  for (int c = 0; c < 3; c++) res[c] = kSH_C0 * sh[c];
  if (deg > 0) {
#pragma unroll
    for (int c = 0; c < 3; c++)
      res[c] = res[c] - kSH_C1 * y * sh[3 + c] + kSH_C1 * z * sh[6 + c] - kSH_C1 * x * sh[9 + c];
    if (deg > 1) {
      float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
#pragma unroll
      for (int c = 0; c < 3; c++)
        res[c] = res[c] + kSH_C2[0] * xy * sh[12 + c] + kSH_C2[1] * yz * sh[15 + c] +
                 kSH_C2[2] * (2.0f * zz - xx - yy) * sh[18 + c] + kSH_C2[3] * xz * sh[21 + c] +
                 kSH_C2[4] * (xx - yy) * sh[24 + c];
      if (deg > 2) {
#pragma unroll
        for (int c = 0; c < 3; c++)
          res[c] = res[c] + kSH_C3[0] * y * (3.0f * xx - yy) * sh[27 + c] + kSH_C3[1] * xy * z * sh[30 + c] +
                   kSH_C3[2] * y * (4.0f * zz - xx - yy) * sh[33 + c] +
                   kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[36 + c] +
                   kSH_C3[4] * x * (4.0f * zz - xx - yy) * sh[39 + c] + kSH_C3[5] * z * (xx - yy) * sh[42 + c] +
                   kSH_C3[6] * x * (xx - 3.0f * yy) * sh[45 + c];
      }
    }
  }
  clamped = 0;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    res[c] += 0.5f;
    if (res[c] < 0.0f) clamped |= (1u << c);
    rgb[c] = fmaxf(res[c], 0.0f);
  }
}


#define IGS_PROJECT_THREADS 128
__global__ void __launch_bounds__(IGS_PROJECT_THREADS)
project_kernel(ProjectParams p, const Camera* __restrict__ cams, const float* __restrict__ cov3D,
               const GaussEig* __restrict__ eigs, SplatHead* __restrict__ head, SplatTail* __restrict__ tail,
               float* __restrict__ depth_out, int* __restrict__ radii_geom, int* __restrict__ radii_out,
               uint32_t* __restrict__ tiles_out, uint8_t* __restrict__ clamped_out,
               uint64_t* __restrict__ skey, uint32_t* __restrict__ sval) {
  // View index fastest in the launch order: the V CTAs that project one chunk of 256 Gaussians run together, so the
  // per-Gaussian inputs (192 B of SH coefficients, covariance, eigen-frame) come from DRAM once and from L2 V-1 times.
  int chunk, v;
  if (gridDim.y == 1) { v = blockIdx.x % p.V; chunk = blockIdx.x / p.V; } else { chunk = blockIdx.x; v = blockIdx.y; }
  // SH coefficients of the block's Gaussians staged with coalesced loads: read per thread from the [P][M][3] array,
  // every scalar load touches 32 cache lines and the kernel runs at the L1 request limit (ncu: l1tex 94 % of peak).
  __shared__ float s_sh[IGS_PROJECT_THREADS * 49];
  const int g0 = chunk * IGS_PROJECT_THREADS;
  const int shrow = p.M * 3;
  const bool staged = p.shs != nullptr && p.colors_precomp == nullptr && shrow <= 48;
  if (staged) {
    const int nrows = min(IGS_PROJECT_THREADS, p.P - g0);
    const float* src = p.shs + (size_t)g0 * shrow;
    for (int f = threadIdx.x; f < nrows * shrow; f += IGS_PROJECT_THREADS) {
      const int r = f / shrow;
      s_sh[r * 49 + (f - r * shrow)] = src[f];
    }
    __syncthreads();
  }
  int g = g0 + threadIdx.x;
  if (g >= p.P) return;
  size_t gid = (size_t)v * p.P + g;
  const Camera& cam = cams[v];

  // defaults for an invisible Gaussian (reference forward.cu:348-349)
  int radius_i = 0;
  uint32_t ntiles = 0;
  uint32_t depth_bits = 0xFFFFFFFFu;

  float mx = p.means3D[3 * g], my = p.means3D[3 * g + 1], mz = p.means3D[3 * g + 2];
  float tx = xform_row(cam.view, 0, mx, my, mz);
  float ty = xform_row(cam.view, 1, mx, my, mz);
  float tz = xform_row(cam.view, 2, mx, my, mz);

  if (tz > IGS_NEAR_Z) {
    float hx = xform_row(cam.proj, 0, mx, my, mz);
    float hy = xform_row(cam.proj, 1, mx, my, mz);
    float hw = xform_row(cam.proj, 3, mx, my, mz);
    float p_w = frcp(fadd(hw, 0.0000001f));
    float projx = fmul(hx, p_w), projy = fmul(hy, p_w);

    const float* c3 = cov3D + 6 * (size_t)g;
    float c0 = c3[0], c1 = c3[1], c2 = c3[2], c3_ = c3[3], c4 = c3[4], c5 = c3[5];

    // ---- EWA projection, reference computeCov2D (forward.cu:85-124) ----
    const float limx = fmul(1.3f, p.tan_fovx), limy = fmul(1.3f, p.tan_fovy);
    float txtz = fdiv(tx, tz), tytz = fdiv(ty, tz);
    float cx_ = fmul(tz, fminf(limx, fmaxf(-limx, txtz)));
    float cy_ = fmul(tz, fminf(limy, fmaxf(-limy, tytz)));
    txtz = fdiv(cx_, tz);
    tytz = fdiv(cy_, tz);
    float tz2 = fmul(tz, tz);
    float J00 = fdiv(p.focal_x, tz);
    float J02 = fdiv(fmul(p.focal_x, -cx_), tz2);
    float J11 = fdiv(p.focal_y, tz);
    float J12 = fdiv(fmul(p.focal_y, -cy_), tz2);
    const float* Vm = cam.view;
    // T = W * J  (only the first two columns are non-zero); Tcr = T[col c][row r]
    float T00 = ffma(Vm[2], J02, fmul(Vm[0], J00));
    float T01 = ffma(Vm[6], J02, fmul(Vm[4], J00));
    float T02 = ffma(J02, Vm[10], fmul(Vm[8], J00));
    float T10 = ffma(Vm[2], J12, fmul(J11, Vm[1]));
    float T11 = ffma(Vm[6], J12, fmul(J11, Vm[5]));
    float T12 = ffma(J12, Vm[10], fmul(J11, Vm[9]));
    // (T^T Vrk^T) rows
    float u0 = dot3_ref(T00, c0, T01, c1, T02, c2);
    float u1 = dot3_ref(T00, c1, T01, c3_, T02, c4);
    float u2 = dot3_ref(T00, c2, T01, c4, T02, c5);
    float w0 = dot3_ref(T10, c0, T11, c1, T12, c2);
    float w1 = dot3_ref(T10, c1, T11, c3_, T12, c4);
    float w2 = dot3_ref(T10, c2, T11, c4, T12, c5);
    float cov_a = dot3_ref(T00, u0, T01, u1, T02, u2);
    float cov_b = dot3_ref(T00, w0, T01, w1, T02, w2);
    float cov_c = dot3_ref(T10, w0, T11, w1, T12, w2);
    float bb = fmul(cov_b, cov_b);
    float det = ffma(cov_a, cov_c, -bb);
    // low-pass compensation in double
    float det_0 = (float)fmax(1e-6, (double)det);
    float det_1 = (float)fmax(1e-6, (double)ffma(fadd(p.kernel_size, cov_a), fadd(p.kernel_size, cov_c), -bb));
    float coef = (float)sqrt((double)det_0 / ((double)det_1 + 1e-6) + 1e-6);
    if ((double)det_0 <= 1e-6 || (double)det_1 <= 1e-6) coef = 0.0f;

    // ---- depth plane / normal (forward.cu:157-262), tolerance-level arithmetic ----
    GaussEig E = eigs[g];
    float cp[6] = {0, 0, 0, 0, 0, 0};
    float nrm[3] = {0, 0, 0};
    float rp[2] = {0, 0};
    {
      // cov_cam_inv = W^T * Vrk_inv * W ; uvh_m = cov_cam_inv * uvh
      // W[c][r] = view[r*4 + c]  (W = mat3(v0,v4,v8 | v1,v5,v9 | v2,v6,v10))
      float Wm[3][3];  // Wm[c][r]
#pragma unroll
      for (int c = 0; c < 3; c++)
#pragma unroll
        for (int r = 0; r < 3; r++) Wm[c][r] = Vm[r * 4 + c];
      float uvh[3] = {txtz, tytz, 1.0f};
      // a = W * uvh
      float a[3], b[3], m3[3];
#pragma unroll
      for (int r = 0; r < 3; r++) a[r] = Wm[0][r] * uvh[0] + Wm[1][r] * uvh[1] + Wm[2][r] * uvh[2];
      // b = Vrk_inv * a
#pragma unroll
      for (int r = 0; r < 3; r++) b[r] = E.vinv[0 * 3 + r] * a[0] + E.vinv[1 * 3 + r] * a[1] + E.vinv[2 * 3 + r] * a[2];
      // m3 = W^T * b : (W^T)[c][r] = W[r][c]
#pragma unroll
      for (int r = 0; r < 3; r++) m3[r] = Wm[r][0] * b[0] + Wm[r][1] * b[1] + Wm[r][2] * b[2];
      float inv_len = 1.0f / sqrtf(m3[0] * m3[0] + m3[1] * m3[1] + m3[2] * m3[2]);
      float mn[3] = {m3[0] * inv_len, m3[1] * inv_len, m3[2] * inv_len};
      if (!(isnan(mn[0]) || (E.info & 1u) == 0u)) {
        float u2_ = txtz * txtz, v2_ = tytz * tytz, uv = txtz * tytz;
        float l = sqrtf(cx_ * cx_ + cy_ * cy_ + tz * tz);
        float vbn = mn[0] * uvh[0] + mn[1] * uvh[1] + mn[2] * uvh[2];
        float nl = u2_ + v2_ + 1.0f;
        float factor_normal = l / nl;
        float dv = fmaxf(vbn, 0.0000001f);
        float q0 = mn[0] / dv, q1 = mn[1] / dv, q2 = mn[2] / dv;
        // plane = nJ_inv * q, nJ_inv columns: (v2+1,-uv,0) (-uv,u2+1,0) (-txtz,-tytz,0)
        float pl0 = (v2_ + 1.0f) * q0 + (-uv) * q1 + (-txtz) * q2;
        float pl1 = (-uv) * q0 + (u2_ + 1.0f) * q1 + (-tytz) * q2;
        cp[0] = (-(v2_ + 1.0f) * tz + pl0 * cx_) / nl / p.focal_x;
        cp[1] = (uv * tz + pl1 * cx_) / nl / p.focal_y;
        cp[2] = (uv * tz + pl0 * cy_) / nl / p.focal_x;
        cp[3] = (-(u2_ + 1.0f) * tz + pl1 * cy_) / nl / p.focal_y;
        cp[4] = (cx_ + pl0 * tz) / nl / p.focal_x;
        cp[5] = (cy_ + pl1 * tz) / nl / p.focal_y;
        rp[0] = pl0 * l / nl / p.focal_x;
        rp[1] = pl1 * l / nl / p.focal_y;
        // normal = normalize(nJ * ray_normal), nJ columns: (1/tz,0,-tx/tz^2) (0,1/tz,-ty/tz^2) (tx/l,ty/l,tz/l)
        float rn0 = -pl0 * factor_normal, rn1 = -pl1 * factor_normal, rn2 = -1.0f;
        float n0 = (1.0f / tz) * rn0 + (cx_ / l) * rn2;
        float n1 = (1.0f / tz) * rn1 + (cy_ / l) * rn2;
        float n2 = (-(cx_) / (tz * tz)) * rn0 + (-(cy_) / (tz * tz)) * rn1 + (tz / l) * rn2;
        float il = 1.0f / sqrtf(n0 * n0 + n1 * n1 + n2 * n2);
        nrm[0] = n0 * il; nrm[1] = n1 * il; nrm[2] = n2 * il;
      }
    }
    float ts = fsqrt(ffma(tz, tz, ffma(tx, tx, fmul(ty, ty))));

    if (det != 0.0f) {
      float det_inv = frcp(det);
      float conx = fmul(cov_c, det_inv), cony = fmul(det_inv, -cov_b), conz = fmul(cov_a, det_inv);
      float mid = fmul(fadd(cov_a, cov_c), 0.5f);
      float sq = fsqrt(fmaxf(ffma(mid, mid, -det), 0.1f));
      float lam = fmaxf(fadd(mid, sq), fsub(mid, sq));
      float my_radius = ceilf(fmul(fsqrt(lam), 3.0f));
      // ndc2Pix in double, reference auxiliary.h:57-60 as compiled: fma(v+1, S, -1) * 0.5
      float px = (float)(__dmul_rn(__fma_rn(__dadd_rn((double)projx, 1.0), (double)p.W, -1.0), 0.5));
      float py = (float)(__dmul_rn(__fma_rn(__dadd_rn((double)projy, 1.0), (double)p.H, -1.0), 0.5));
      int ri = (int)my_radius;
      float rf = (float)ri;
      int rminx = min(p.tiles_x, max(0, (int)fmul(fsub(px, rf), 0.0625f)));
      // rows are clipped to the view's tile-row window (the whole image unless the caller renders a band)
      int rminy = min(cam.row1, max(cam.row0, (int)fmul(fsub(py, rf), 0.0625f)));
      int rmaxx = min(p.tiles_x, max(0, (int)fmul(fadd(fadd(fadd(px, rf), 16.0f), -1.0f), 0.0625f)));
      int rmaxy = min(cam.row1, max(cam.row0, (int)fmul(fadd(fadd(fadd(py, rf), 16.0f), -1.0f), 0.0625f)));
      uint32_t area = (uint32_t)(rmaxx - rminx) * (uint32_t)(rmaxy - rminy);
      if (area != 0) {
        float rgb[3];
        uint32_t clamped = 0;
        if (p.colors_precomp != nullptr) {
          rgb[0] = p.colors_precomp[3 * g]; rgb[1] = p.colors_precomp[3 * g + 1]; rgb[2] = p.colors_precomp[3 * g + 2];
        } else {
          const float* sh = staged ? &s_sh[threadIdx.x * 49] : p.shs + (size_t)g * p.M * 3;
          sh_to_rgb(p.D, sh, mx - cam.campos[0], my - cam.campos[1], mz - cam.campos[2], rgb, clamped);
        }
        SplatHead hd;
        hd.a = make_float4(px, py, conx, cony);
        hd.b = make_float4(conz, fmul(p.opacities[g], coef), rgb[0], rgb[1]);
        SplatTail tl;
        tl.a = make_float4(rgb[2], ts, rp[0], rp[1]);
        tl.b = make_float4(nrm[0], nrm[1], nrm[2], tx);
        tl.c = make_float4(ty, tz, cp[0], cp[1]);
        tl.d = make_float4(cp[2], cp[3], cp[4], cp[5]);
        head[gid] = hd;
        tail[gid] = tl;
        clamped_out[gid] = (uint8_t)clamped;
        radius_i = ri;
        ntiles = area;
        depth_bits = __float_as_uint(tz);
      }
    }
  }
  depth_out[gid] = __uint_as_float(depth_bits == 0xFFFFFFFFu ? 0u : depth_bits);
  radii_geom[gid] = radius_i;
  if (radii_out != nullptr) radii_out[gid] = radius_i;
  tiles_out[gid] = ntiles;
  // depth bits of a visible Gaussian are those of a float > 0.2 (sign bit clear); invisible ones sort last with the
  // all-ones 31-bit pattern.  31 + bits(V) key bits: 36 at V = 20, i.e. four 9-bit radix passes instead of five.
  skey[gid] = ((uint64_t)v << 31) | (uint64_t)(depth_bits & 0x7FFFFFFFu);
  sval[gid] = (uint32_t)gid;
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                                    uint8_t* __restrict__ present) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= P) return;
  float tz = xform_row(view, 2, means3D[3 * g], means3D[3 * g + 1], means3D[3 * g + 2]);
  present[g] = tz > IGS_NEAR_Z ? 1 : 0;
}

// -------- host launchers --------
void launch_gauss_setup(int P, const float* scales, float mod, const float* rotations,
                        const float* cov3D_precomp, float* cov3D_out, GaussEig* eig, cudaStream_t s) {
  if (P <= 0) return;
  count_launch();
  gauss_setup_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, scales, mod, rotations, cov3D_precomp, cov3D_out, eig);
}

void launch_project(const ProjectParams& p, const GeomBuf& g, int* radii_out, cudaStream_t s) {
  if (p.P <= 0) return;
  const unsigned chunks = (unsigned)((p.P + IGS_PROJECT_THREADS - 1) / IGS_PROJECT_THREADS);
  dim3 grid(chunks, p.V);
  if ((size_t)chunks * p.V < 0x7fffffffu) grid = dim3(chunks * p.V, 1);  // view-fastest linear order
  count_launch();
  project_kernel<<<grid, IGS_PROJECT_THREADS, 0, s>>>(p, g.cams, g.cov3D, g.eig, g.head, g.tail, g.depth, g.radii, radii_out,
                                      g.tiles, g.clamped, g.skey[0], g.sval[0]);
}

void launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s) {
  if (P <= 0) return;
  count_launch();
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, view, present);
}

}  // namespace igs
