// preprocess_bwd.cu — per-Gaussian backward, replacing the reference's computeCov2DCUDA
// (DGR/cuda_rasterizer/backward.cu:145-488) and backward preprocessCUDA (:560-628, SH :21-140,
// cov3D :492-555).
//
//   gather_bwd_kernel : one thread per (view, Gaussian), in depth-sorted order.  Sums the dense per-instance records that
//                       blend_bwd wrote for this Gaussian's consecutive emission slots (this replaces
//                       the reference's 25 atomics per pixel pair), then runs the view-dependent
//                       chain: conic/coef -> cov2D -> (cov3D, mean), plane/normal -> (Vrk, mean),
//                       mean2D -> mean, view point + ray length -> mean.  Emits a 16-float partial.
//   finish_bwd_kernel : one thread per Gaussian.  Sums the partials over the V views, runs the SH
//                       backward per view (direction depends on the camera) while accumulating
//                       dL/dSH in registers, and the view-independent cov3D -> (scale, quaternion)
//                       backward once.  Every output element is written (no pre-zeroing needed).
//
// Bug compatibility: the reference passes dL_dconic where conic_opacity is expected
// (rasterizer_impl.cu:569), so computeCov2DCUDA's `combined_opacity` is dL_dconic[idx].w
// (SURVEY T19).  IGS_STRICT_REF_BUGS (default 1) reproduces that; 0 uses the real opacity.
#include "common.cuh"
#include <cstdlib>
#include "kernels.cuh"
#include "mat3.cuh"

#ifndef IGS_STRICT_REF_BUGS
#define IGS_STRICT_REF_BUGS 1
#endif

namespace igs {


// Sum of each (view, Gaussian)'s gradient records.  Records of one Gaussian are consecutive (emission slots), and
// positions are walked in depth-sorted order, so neighbouring threads read neighbouring memory: NS/4 lanes per
// position, one 16-byte part each.  Kept apart from gather_bwd_kernel (128 registers, 4 CTAs of 128 threads per SM) so that this
// bandwidth-bound half runs at full occupancy; the k-order of the additions is the list order (deterministic).
template <int NS>   // record stride in floats: 12 (two-pass colour kernel), 16 (first-generation colour kernel), 32 (geometry)
__global__ void __launch_bounds__(256)
sum_records_kernel(size_t VP, const uint32_t* __restrict__ inst_off, const float* __restrict__ grec,
                   float* __restrict__ sums, uint32_t capacity) {
  constexpr int LANES = NS / 4;                // 16-byte parts per record
  constexpr int PARTS = NS <= 16 ? 3 : 7;      // parts that carry values (10 / 25 floats)
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t pos = t / LANES;
  const int part = (int)(t % LANES);
  if (pos >= VP || part >= PARTS) return;
  // (capacity mode: records exist for emission slots < capacity only; an overflowing call is discarded by the caller)
  const uint32_t off = min(inst_off[pos], capacity);
  const uint32_t n = min(inst_off[pos + 1], capacity) - off;
  const float4* r = reinterpret_cast<const float4*>(grec) + (size_t)off * LANES + part;
  float4 acc = make_float4(0, 0, 0, 0);
  uint32_t k = 0;
  for (; k + 4 <= n; k += 4) {
    const float4 a = r[(size_t)k * LANES], b = r[(size_t)(k + 1) * LANES], c = r[(size_t)(k + 2) * LANES],
                 d = r[(size_t)(k + 3) * LANES];
    acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
    acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w;
    acc.x += c.x; acc.y += c.y; acc.z += c.z; acc.w += c.w;
    acc.x += d.x; acc.y += d.y; acc.z += d.z; acc.w += d.w;
  }
  for (; k < n; k++) {
    const float4 a = r[(size_t)k * LANES];
    acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
  }
  reinterpret_cast<float4*>(sums)[pos * PARTS + part] = acc;
}

#ifndef IGS_GATHER_MINB
#define IGS_GATHER_MINB 4
#endif
// Colour-only records (16 floats, the refine loop): the cotangents of the normal / plane / ray-point / depth outputs are
// identically zero, so the block that turns them into dL/dVrk, dL/dnJ, dL/du,v,l — and the eigen-data gather it needs,
// the largest of the kernel — is skipped AT RUN TIME (p.skip_geo; every value it would produce is an exact zero).
// A separate colour-only BUILD with that block compiled out was 0.2 ms faster at C3 but is not used: with the block's
// uses gone nvcc contracts the shared covariance chain differently, and on a few large, strongly anisotropic Gaussians
// — where dL/dcov3D is a small residual of large terms — the scale / rotation gradients moved by 1e-3 relative, five
// times further from the reference than this build, whose expressions mirror the reference's (tools/grad_noise_floor.py).
__global__ void __launch_bounds__(128, IGS_GATHER_MINB)
gather_bwd_kernel(GatherBwdParams p) {
  // Threads are indexed by DEPTH-SORTED position: emission slots of neighbouring positions are adjacent in grec, so
  // the record reads of a warp fall into one contiguous region (the small per-Gaussian inputs become the gathers).
  // (Indexing by (view, Gaussian) instead, with the record sum as the one gather, measured 3 % slower.)
  const size_t pos = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pos >= (size_t)p.V * p.P) return;
  const size_t gid = p.sorted_gid[pos];
  const int v = (int)(gid / (size_t)p.P);
  const int g = (int)(gid - (size_t)v * p.P);
  float4* part = reinterpret_cast<float4*>(p.partial + gid * 16);
  float* dm2 = p.dL_dmeans2D + gid * 3;
  if (!(p.radii[gid] > 0)) {
    part[0] = part[1] = part[2] = part[3] = make_float4(0, 0, 0, 0);
    dm2[0] = dm2[1] = dm2[2] = 0.0f;
    return;
  }
  // ---- the instance records of this (view, Gaussian), already summed by sum_records_kernel ----
  float s[25];
#pragma unroll
  for (int i = 0; i < 25; i++) s[i] = 0.0f;
  if (p.ns == 16) {
    const float4* r = reinterpret_cast<const float4*>(p.sums) + pos * 3;
    const float4 a = r[0], b = r[1], c = r[2];
    s[0] = a.x; s[1] = a.y; s[2] = a.z; s[3] = a.w; s[4] = b.x; s[5] = b.y; s[6] = b.z; s[7] = b.w; s[8] = c.x; s[9] = c.y;
  } else {
    const float4* r = reinterpret_cast<const float4*>(p.sums) + pos * 7;
#pragma unroll
    for (int q = 0; q < 6; q++) {
      const float4 a = r[q];
      s[q * 4 + 0] = a.x; s[q * 4 + 1] = a.y; s[q * 4 + 2] = a.z; s[q * 4 + 3] = a.w;
    }
    s[24] = r[6].x;
  }
  const float dL_dopacity_in = s[9];
  dm2[0] = s[3]; dm2[1] = s[4]; dm2[2] = s[5];
  // A visible Gaussian that contributed to no pixel (hidden behind opaque ones, or below 1/255 everywhere) has an
  // all-zero record sum, and every output below is linear in it: its partial is exactly zero (the |grad| accumulator
  // s[5] >= 0 is zero only if every value is).  Skipping the chain for them matters because this kernel is latency
  // bound (128 registers).
  {
    float nz = 0.0f;
#pragma unroll
    for (int i = 0; i < 25; i++) nz += fabsf(s[i]);
    if (nz == 0.0f) {
      part[0] = part[1] = part[2] = part[3] = make_float4(0, 0, 0, 0);
      return;
    }
  }

  const Camera& cam = p.cams[v];
  const float* view = cam.view;
  const float* proj = cam.proj;
  const float h_x = p.focal_x, h_y = p.focal_y, kernel_size = p.kernel_size;
  const V3 mean = v3(p.means3D[3 * g], p.means3D[3 * g + 1], p.means3D[3 * g + 2]);
  const float* c3 = p.cov3D + 6 * (size_t)g;

  // ======================= computeCov2DCUDA (backward.cu:162-487) =======================
  const V3 dL_dconic = v3(s[6], s[7], s[8]);
  const V3 dL_dnormal = v3(s[22], s[23], s[24]);
  const float cp0x = s[13], cp0y = s[14], cp1x = s[15], cp1y = s[16], cp2x = s[17], cp2y = s[18];
  const float dL_drp_x = s[20], dL_drp_y = s[21];
#if IGS_STRICT_REF_BUGS
  const float combined_opacity = s[8];  // == dL_dconic[idx].w in the reference's float4 view
#else
  const float combined_opacity = p.head[gid].b.y;
#endif

  V3 t = v3(view[0] * mean.x + view[4] * mean.y + view[8] * mean.z + view[12],
            view[1] * mean.x + view[5] * mean.y + view[9] * mean.z + view[13],
            view[2] * mean.x + view[6] * mean.y + view[10] * mean.z + view[14]);
  const float limx = 1.3f * p.tan_fovx, limy = 1.3f * p.tan_fovy;
  float txtz = t.x / t.z, tytz = t.y / t.z;
  t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
  t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
  const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.0f : 1.0f;
  const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.0f : 1.0f;
  txtz = t.x / t.z;
  tytz = t.y / t.z;

  const M3 J = m3_cols(h_x / t.z, 0.0f, -(h_x * t.x) / (t.z * t.z), 0.0f, h_y / t.z, -(h_y * t.y) / (t.z * t.z), 0, 0, 0);
  const M3 Wm = m3_cols(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
  const M3 Vrk = m3_cols(c3[0], c3[1], c3[2], c3[1], c3[3], c3[4], c3[2], c3[4], c3[5]);
  const M3 T = Wm * J;
  const M3 cov2D = transpose(T) * transpose(Vrk) * T;

  const float det_0 = (float)fmax(1e-6, (double)(cov2D.c[0][0] * cov2D.c[1][1] - cov2D.c[0][1] * cov2D.c[0][1]));
  const float det_1 = (float)fmax(1e-6, (double)((cov2D.c[0][0] + kernel_size) * (cov2D.c[1][1] + kernel_size) -
                                                 cov2D.c[0][1] * cov2D.c[0][1]));
  const float coef = (float)sqrt((double)det_0 / ((double)det_1 + 1e-6) + 1e-6);

  const float u2 = txtz * txtz, v2 = tytz * tytz, uv = txtz * tytz;

  M3 dL_dVrk = m3_zero(), dL_dnJ = m3_zero();
  V3 plane = v3(0, 0, 0);
  float dL_du = 0, dL_dv = 0, dL_dl = 0, l = 1, nl = 1;
  if (!p.skip_geo) {
  const GaussEig E = p.eig[g];
  M3 Vrk_inv, evecs;
#pragma unroll
  for (int cc = 0; cc < 3; cc++)
#pragma unroll
    for (int rr = 0; rr < 3; rr++) {
      Vrk_inv.c[cc][rr] = E.vinv[cc * 3 + rr];
      evecs.c[cc][rr] = E.evec[cc * 3 + rr];
    }
  const bool solver_ok = (E.info & 1u) != 0u;
  const bool well_conditioned = (E.info & 2u) != 0u;
  const int min_id = (int)((E.info >> 4) & 3u);
  const V3 eigenvector_min = col(evecs, min_id);

  const M3 cov_cam_inv = transpose(Wm) * Vrk_inv * Wm;
  const V3 uvh = v3(txtz, tytz, 1.0f);
  const V3 uvh_m = cov_cam_inv * uvh;
  const V3 uvh_mn = uvh_m * (1.0f / sqrtf(dot(uvh_m, uvh_m)));
  if (!(isnan(uvh_mn.x) || !solver_ok)) {
    const float vb = dot(uvh_m, uvh);
    const float vbn = dot(uvh_mn, uvh);
    l = sqrtf(t.x * t.x + t.y * t.y + t.z * t.z);
    const M3 nJ = m3_cols(1 / t.z, 0.0f, -(t.x) / (t.z * t.z), 0.0f, 1 / t.z, -(t.y) / (t.z * t.z), t.x / l, t.y / l, t.z / l);
    const M3 nJ_inv = m3_cols(v2 + 1, -uv, 0, -uv, u2 + 1, 0, -txtz, -tytz, 0);
    const float clamp_vb = fmaxf(vb, 0.0000001f);
    const float clamp_vbn = fmaxf(vbn, 0.0000001f);
    nl = u2 + v2 + 1;
    const float factor_normal = l / nl;
    const V3 uvh_m_vb = uvh_mn / clamp_vbn;
    plane = nJ_inv * uvh_m_vb;

    const float c0x = (-(v2 + 1) * t.z + plane.x * t.x) / nl, c0y = (uv * t.z + plane.y * t.x) / nl;
    const float c1x = (uv * t.z + plane.x * t.y) / nl, c1y = (-(u2 + 1) * t.z + plane.y * t.y) / nl;
    const float c2x = (t.x + plane.x * t.z) / nl, c2y = (t.y + plane.y * t.z) / nl;
    const float rpx = plane.x * factor_normal, rpy = plane.y * factor_normal;
    const V3 ray_normal_vector = v3(-plane.x * factor_normal, -plane.y * factor_normal, -1);

    const V3 cam_normal_vector = nJ * ray_normal_vector;
    const float lv = sqrtf(dot(cam_normal_vector, cam_normal_vector));
    const V3 normal_vector = cam_normal_vector * (1.0f / lv);
    const V3 dL_dnormal_lv = dL_dnormal / lv;
    const V3 dL_dcam_normal_vector = dL_dnormal_lv - normal_vector * dot(normal_vector, dL_dnormal_lv);
    const V3 dL_dray_normal_vector = transpose(nJ) * dL_dcam_normal_vector;
    dL_dnJ = outer(dL_dcam_normal_vector, ray_normal_vector);
    dL_dl = (-plane.x * dL_dray_normal_vector.x - plane.y * dL_dray_normal_vector.y + plane.x * dL_drp_x +
             plane.y * dL_drp_y) / nl;

    const float dL_dplane_x = (t.x * cp0x + t.y * cp1x + t.z * cp2x - l * dL_dray_normal_vector.x + dL_drp_x * l) / nl;
    const float dL_dplane_y = (t.x * cp0y + t.y * cp1y + t.z * cp2y - l * dL_dray_normal_vector.y + dL_drp_y * l) / nl;
    const V3 dL_dplane_append = v3(dL_dplane_x, dL_dplane_y, 0);

    const float dL_dnl = (-cp0x * c0x - cp0y * c0y - cp1x * c1x - cp1y * c1y - cp2x * c2x - cp2y * c2y -
                          dL_dray_normal_vector.x * ray_normal_vector.x - dL_dray_normal_vector.y * ray_normal_vector.y -
                          dL_drp_x * rpx - dL_drp_y * rpy) / nl;

    const float tmp = dL_dplane_x * plane.x + dL_dplane_y * plane.y;
    const V3 W_uvh = Wm * uvh;

    if (well_conditioned) {
      const V3 lhs = Vrk_inv * W_uvh;
      const V3 rhs = (Vrk_inv * (1.0f / clamp_vb)) * (W_uvh * (-tmp) + Wm * (transpose(nJ_inv) * dL_dplane_append));
      dL_dVrk = outer(lhs, rhs) * (-1.0f);
    } else {
      const float dL_dvb = -tmp / clamp_vb;
      const V3 nJ_inv_dL_dplane = transpose(nJ_inv) * v3(dL_dplane_x / clamp_vb, dL_dplane_y / clamp_vb, 0);
      const M3 dL_dVrk_inv = outer(W_uvh, W_uvh * dL_dvb + Wm * nJ_inv_dL_dplane);
      const V3 dL_dvv = (dL_dVrk_inv + transpose(dL_dVrk_inv)) * eigenvector_min;
#pragma unroll
      for (int j = 0; j < 3; j++) {
        if (j != min_id) {
          float scale = dot(col(evecs, j), dL_dvv) / fminf(E.eval[min_id] - E.eval[j], -0.0000001f);
          dL_dVrk = dL_dVrk + outer(col(evecs, j) * scale, eigenvector_min);
        }
      }
    }

    const V3 dL_duvh = uvh_m_vb * (2 * (-tmp)) + (cov_cam_inv * (1.0f / clamp_vb)) * (transpose(nJ_inv) * dL_dplane_append);
    const M3 dL_dnJ_inv = outer(dL_dplane_append, uvh_m_vb);

    dL_du = dL_dnl * 2 * txtz + dL_duvh.x + (dL_dnJ_inv.c[0][1] + dL_dnJ_inv.c[1][0]) * (-tytz) +
            2 * dL_dnJ_inv.c[1][1] * txtz - dL_dnJ_inv.c[2][0] + (cp0y * t.y + cp1x * t.y + cp1y * (-2 * t.x)) / nl;
    dL_dv = dL_dnl * 2 * tytz + dL_duvh.y + (dL_dnJ_inv.c[0][1] + dL_dnJ_inv.c[1][0]) * (-txtz) +
            2 * dL_dnJ_inv.c[0][0] * tytz - dL_dnJ_inv.c[2][1] + (cp0x * (-2 * t.y) + cp0y * t.x + cp1x * t.x) / nl;
  }
  }  // !skip_geo

  // ---- low-pass coefficient (double arithmetic where the reference has it) ----
  const float opacity = (float)((double)combined_opacity / ((double)coef + 1e-6));
  const float dL_dcoef = dL_dopacity_in * opacity;
  const float dL_dsqrtcoef = (float)((double)dL_dcoef * 0.5 * 1. / ((double)coef + 1e-6));
  const float dL_ddet0 = (float)((double)dL_dsqrtcoef / ((double)det_1 + 1e-6));
  const float dL_ddet1 = (float)((double)(dL_dsqrtcoef * det_0) * (-1.0 / ((double)(det_1 * det_1) + 1e-6)));
  const float dcoef_da = dL_ddet0 * cov2D.c[1][1] + dL_ddet1 * (cov2D.c[1][1] + kernel_size);
  const float dcoef_db = (float)((double)dL_ddet0 * (-2. * (double)cov2D.c[0][1]) + (double)dL_ddet1 * (-2. * (double)cov2D.c[0][1]));
  const float dcoef_dc = dL_ddet0 * cov2D.c[0][0] + dL_ddet1 * (cov2D.c[0][0] + kernel_size);

  const float a = cov2D.c[0][0] + kernel_size, b = cov2D.c[0][1], c = cov2D.c[1][1] + kernel_size;
  const float denom = a * c - b * b;
  float dL_da = 0, dL_db = 0, dL_dc = 0;
  const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
  float dL_dopacity_out = dL_dopacity_in;
  float dcov[6];
  if (denom2inv != 0) {
    dL_da = denom2inv * (-c * c * dL_dconic.x + 2 * b * c * dL_dconic.y + (denom - a * c) * dL_dconic.z);
    dL_dc = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * dL_dconic.y + (denom - a * c) * dL_dconic.x);
    dL_db = denom2inv * 2 * (b * c * dL_dconic.x - (denom + 2 * b * b) * dL_dconic.y + a * b * dL_dconic.z);
    if ((double)det_0 <= 1e-6 || (double)det_1 <= 1e-6) {
      dL_dopacity_out = 0;
    } else {
      dL_da += dcoef_da;
      dL_dc += dcoef_dc;
      dL_db += dcoef_db;
      dL_dopacity_out = dL_dopacity_in * coef;
    }
    dcov[0] = (T.c[0][0] * T.c[0][0] * dL_da + T.c[0][0] * T.c[1][0] * dL_db + T.c[1][0] * T.c[1][0] * dL_dc);
    dcov[3] = (T.c[0][1] * T.c[0][1] * dL_da + T.c[0][1] * T.c[1][1] * dL_db + T.c[1][1] * T.c[1][1] * dL_dc);
    dcov[5] = (T.c[0][2] * T.c[0][2] * dL_da + T.c[0][2] * T.c[1][2] * dL_db + T.c[1][2] * T.c[1][2] * dL_dc);
    dcov[1] = 2 * T.c[0][0] * T.c[0][1] * dL_da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][1] * dL_dc;
    dcov[2] = 2 * T.c[0][0] * T.c[0][2] * dL_da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][2] * dL_dc;
    dcov[4] = 2 * T.c[0][2] * T.c[0][1] * dL_da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_db + 2 * T.c[1][1] * T.c[1][2] * dL_dc;
  } else {
#pragma unroll
    for (int i = 0; i < 6; i++) dcov[i] = 0;
  }
  dcov[0] += dL_dVrk.c[0][0];
  dcov[3] += dL_dVrk.c[1][1];
  dcov[5] += dL_dVrk.c[2][2];
  dcov[1] += dL_dVrk.c[0][1] + dL_dVrk.c[1][0];
  dcov[2] += dL_dVrk.c[0][2] + dL_dVrk.c[2][0];
  dcov[4] += dL_dVrk.c[1][2] + dL_dVrk.c[2][1];

  const float dL_dT00 = 2 * (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_da +
                        (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_db;
  const float dL_dT01 = 2 * (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_da +
                        (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_db;
  const float dL_dT02 = 2 * (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_da +
                        (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_db;
  const float dL_dT10 = 2 * (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_dc +
                        (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_db;
  const float dL_dT11 = 2 * (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_dc +
                        (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_db;
  const float dL_dT12 = 2 * (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_dc +
                        (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_db;

  const float dL_dJ00 = Wm.c[0][0] * dL_dT00 + Wm.c[0][1] * dL_dT01 + Wm.c[0][2] * dL_dT02;
  const float dL_dJ02 = Wm.c[2][0] * dL_dT00 + Wm.c[2][1] * dL_dT01 + Wm.c[2][2] * dL_dT02;
  const float dL_dJ11 = Wm.c[1][0] * dL_dT10 + Wm.c[1][1] * dL_dT11 + Wm.c[1][2] * dL_dT12;
  const float dL_dJ12 = Wm.c[2][0] * dL_dT10 + Wm.c[2][1] * dL_dT11 + Wm.c[2][2] * dL_dT12;

  const float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
  const float l3 = l * l * l;
  const float dL_dtx = x_grad_mul * (-h_x * tz2 * dL_dJ02 + dL_du * tz - dL_dnJ.c[0][2] * tz2 +
                                     dL_dnJ.c[2][0] * (1 / l - t.x * t.x / l3) + dL_dnJ.c[2][1] * (-t.x * t.y / l3) +
                                     dL_dnJ.c[2][2] * (-t.x * t.z / l3) + (cp0x * plane.x + cp0y * plane.y + cp2x) / nl +
                                     dL_dl * t.x / l);
  const float dL_dty = y_grad_mul * (-h_y * tz2 * dL_dJ12 + dL_dv * tz - dL_dnJ.c[1][2] * tz2 +
                                     dL_dnJ.c[2][0] * (-t.x * t.y / l3) + dL_dnJ.c[2][1] * (1 / l - t.y * t.y / l3) +
                                     dL_dnJ.c[2][2] * (-t.y * t.z / l3) + (cp1x * plane.x + cp1y * plane.y + cp2y) / nl +
                                     dL_dl * t.y / l);
  const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 +
                       (2 * h_y * t.y) * tz3 * dL_dJ12 - (dL_du * t.x + dL_dv * t.y) * tz2 +
                       (dL_dnJ.c[0][0] + dL_dnJ.c[1][1]) * (-tz2) + dL_dnJ.c[0][2] * (2 * t.x * tz3) +
                       dL_dnJ.c[1][2] * (2 * t.y * tz3) + (dL_dnJ.c[2][0] * t.x + dL_dnJ.c[2][1] * t.y) * (-t.z / l3) +
                       dL_dnJ.c[2][2] * (1 / l - t.z * t.z / l3) +
                       (cp0x * (-(v2 + 1)) + cp0y * uv + cp1x * uv + cp1y * (-(u2 + 1)) + cp2x * plane.x + cp2y * plane.y) / nl +
                       dL_dl * t.z / l;

  // dL_dmean from the covariance path: view^T (3x3) applied (transformVec4x3Transpose)
  V3 dmean = v3(view[0] * dL_dtx + view[1] * dL_dty + view[2] * dL_dtz,
                view[4] * dL_dtx + view[5] * dL_dty + view[6] * dL_dtz,
                view[8] * dL_dtx + view[9] * dL_dty + view[10] * dL_dtz);

  // ======================= backward preprocessCUDA, mean part (backward.cu:587-619) =======================
  {
    const V3 m = mean;
    const float m_hom_w = proj[3] * m.x + proj[7] * m.y + proj[11] * m.z + proj[15];
    const float m_w = 1.0f / (m_hom_w + 0.0000001f);
    const float mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * m_w * m_w;
    const float gx = s[3], gy = s[4];
    V3 d1;
    d1.x = (proj[0] * m_w - proj[3] * mul1) * gx + (proj[1] * m_w - proj[3] * mul2) * gy;
    d1.y = (proj[4] * m_w - proj[7] * mul1) * gx + (proj[5] * m_w - proj[7] * mul2) * gy;
    d1.z = (proj[8] * m_w - proj[11] * mul1) * gx + (proj[9] * m_w - proj[11] * mul2) * gy;
    V3 d2 = v3(0, 0, 0);
    if (!p.skip_geo) {
      const V3 mv = v3(view[0] * m.x + view[4] * m.y + view[8] * m.z + view[12],
                       view[1] * m.x + view[5] * m.y + view[9] * m.z + view[13],
                       view[2] * m.x + view[6] * m.y + view[10] * m.z + view[14]);
      const float tlen = sqrtf(mv.x * mv.x + mv.y * mv.y + mv.z * mv.z);
      const float dL_dt = s[19];
      const V3 q = v3(s[10] + mv.x / tlen * dL_dt, s[11] + mv.y / tlen * dL_dt, s[12] + mv.z / tlen * dL_dt);
      d2 = v3(view[0] * q.x + view[1] * q.y + view[2] * q.z, view[4] * q.x + view[5] * q.y + view[6] * q.z,
              view[8] * q.x + view[9] * q.y + view[10] * q.z);
    }
    dmean = dmean + v3(d1.x + d2.x, d1.y + d2.y, d1.z + d2.z);
  }

  const uint32_t cl = p.clamped[gid];
  float4 o0 = make_float4(dmean.x, dmean.y, dmean.z, dcov[0]);
  float4 o1 = make_float4(dcov[1], dcov[2], dcov[3], dcov[4]);
  float4 o2 = make_float4(dcov[5], (cl & 1u) ? 0.0f : s[0], (cl & 2u) ? 0.0f : s[1], (cl & 4u) ? 0.0f : s[2]);
  float4 o3 = make_float4(dL_dopacity_out, s[0], s[1], s[2]);
  part[0] = o0; part[1] = o1; part[2] = o2; part[3] = o3;
}

// SH constants again (device constant memory is per translation unit)
__device__ __constant__ float bSH_C0 = 0.28209479177387814f;
__device__ __constant__ float bSH_C1 = 0.4886025119029199f;
__device__ __constant__ float bSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                           -1.0925484305920792f, 0.5462742152960396f};
__device__ __constant__ float bSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                           0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                           -0.5900435899266435f};


// MAXC = number of SH coefficients held in registers (1, 4, 9 or 16)
template <int MAXC>
__global__ void __launch_bounds__(128)
finish_bwd_kernel(FinishBwdParams p) {
  // The SH coefficients of the block's 128 Gaussians are staged in shared memory with coalesced loads (and dL/dSH
  // leaves the same way): read per thread straight from the [P][M][3] array, every scalar load touches 32 cache
  // lines and the kernel sits at the L1 request limit (ncu: l1tex 84 % of peak, lg_throttle stalls).
  constexpr int SROW = MAXC * 3 + 1;   // odd row stride: conflict-free per-thread rows
  __shared__ float s_sh[MAXC > 1 ? 128 * SROW : 1];
  const int g0 = blockIdx.x * blockDim.x;
  const int g_raw = g0 + threadIdx.x;
  const bool live = g_raw < p.P;
  const int g = live ? g_raw : p.P - 1;
  const int row = p.M * 3;             // floats per Gaussian in shs / dL_dsh
  const bool staged = MAXC > 1 && p.shs != nullptr && row <= MAXC * 3;
  if (staged) {
    const int nrows = min(128, p.P - g0);
    const float* src = p.shs + (size_t)g0 * row;
    for (int f = threadIdx.x; f < nrows * row; f += 128) {
      const int r = f / row;
      s_sh[r * SROW + (f - r * row)] = src[f];
    }
    __syncthreads();
  }
  float dmx = 0, dmy = 0, dmz = 0, dop = 0;
  float dcov[6] = {0, 0, 0, 0, 0, 0};
  float draw[3] = {0, 0, 0};
  float dsh[MAXC * 3];
#pragma unroll
  for (int i = 0; i < MAXC * 3; i++) dsh[i] = 0.0f;
  const float mx = p.means3D[3 * g], my = p.means3D[3 * g + 1], mz = p.means3D[3 * g + 2];
  const float* sh = p.shs ? (staged ? &s_sh[threadIdx.x * SROW] : p.shs + (size_t)g * p.M * 3) : nullptr;

  for (int v = 0; v < p.V; v++) {
    size_t gid = (size_t)v * p.P + g;
    if (!(p.radii[gid] > 0)) continue;
    const float4* part = reinterpret_cast<const float4*>(p.partial + gid * 16);
    float4 o0 = part[0], o1 = part[1], o2 = part[2], o3 = part[3];
    dmx += o0.x; dmy += o0.y; dmz += o0.z;
    dcov[0] += o0.w; dcov[1] += o1.x; dcov[2] += o1.y; dcov[3] += o1.z; dcov[4] += o1.w; dcov[5] += o2.x;
    dop += o3.x;
    draw[0] += o3.y; draw[1] += o3.z; draw[2] += o3.w;
    if (sh != nullptr) {
      // SH backward for this view (reference backward.cu:21-140)
      const float r0 = o2.y, r1 = o2.z, r2 = o2.w;  // dL_dRGB with the clamp mask applied
      const Camera& cam = p.cams[v];
      const float ox = mx - cam.campos[0], oy = my - cam.campos[1], oz = mz - cam.campos[2];
      const float len = sqrtf(ox * ox + oy * oy + oz * oz);
      const float x = ox / len, y = oy / len, z = oz / len;
      float dRdx[3] = {0, 0, 0}, dRdy[3] = {0, 0, 0}, dRdz[3] = {0, 0, 0};
      const float rr[3] = {r0, r1, r2};
#define ACC(k, coef_)                                  \
  {                                                    \
    float cf = (coef_);                                \
    dsh[(k) * 3 + 0] += cf * r0;                       \
    dsh[(k) * 3 + 1] += cf * r1;                       \
    dsh[(k) * 3 + 2] += cf * r2;                       \
  }
      ACC(0, bSH_C0);
      if (MAXC > 1 && p.D > 0) {
        ACC(1, -bSH_C1 * y);
        ACC(2, bSH_C1 * z);
        ACC(3, -bSH_C1 * x);
#pragma unroll
        for (int c = 0; c < 3; c++) {
          dRdx[c] = -bSH_C1 * sh[9 + c];
          dRdy[c] = -bSH_C1 * sh[3 + c];
          dRdz[c] = bSH_C1 * sh[6 + c];
        }
        if (MAXC > 4 && p.D > 1) {
          const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
          ACC(4, bSH_C2[0] * xy);
          ACC(5, bSH_C2[1] * yz);
          ACC(6, bSH_C2[2] * (2.f * zz - xx - yy));
          ACC(7, bSH_C2[3] * xz);
          ACC(8, bSH_C2[4] * (xx - yy));
#pragma unroll
          for (int c = 0; c < 3; c++) {
            dRdx[c] += bSH_C2[0] * y * sh[12 + c] + bSH_C2[2] * 2.f * -x * sh[18 + c] + bSH_C2[3] * z * sh[21 + c] +
                       bSH_C2[4] * 2.f * x * sh[24 + c];
            dRdy[c] += bSH_C2[0] * x * sh[12 + c] + bSH_C2[1] * z * sh[15 + c] + bSH_C2[2] * 2.f * -y * sh[18 + c] +
                       bSH_C2[4] * 2.f * -y * sh[24 + c];
            dRdz[c] += bSH_C2[1] * y * sh[15 + c] + bSH_C2[2] * 2.f * 2.f * z * sh[18 + c] + bSH_C2[3] * x * sh[21 + c];
          }
          if (MAXC > 9 && p.D > 2) {
            ACC(9, bSH_C3[0] * y * (3.f * xx - yy));
            ACC(10, bSH_C3[1] * xy * z);
            ACC(11, bSH_C3[2] * y * (4.f * zz - xx - yy));
            ACC(12, bSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
            ACC(13, bSH_C3[4] * x * (4.f * zz - xx - yy));
            ACC(14, bSH_C3[5] * z * (xx - yy));
            ACC(15, bSH_C3[6] * x * (xx - 3.f * yy));
#pragma unroll
            for (int c = 0; c < 3; c++) {
              dRdx[c] += (bSH_C3[0] * sh[27 + c] * 3.f * 2.f * xy + bSH_C3[1] * sh[30 + c] * yz +
                          bSH_C3[2] * sh[33 + c] * -2.f * xy + bSH_C3[3] * sh[36 + c] * -3.f * 2.f * xz +
                          bSH_C3[4] * sh[39 + c] * (-3.f * xx + 4.f * zz - yy) + bSH_C3[5] * sh[42 + c] * 2.f * xz +
                          bSH_C3[6] * sh[45 + c] * 3.f * (xx - yy));
              dRdy[c] += (bSH_C3[0] * sh[27 + c] * 3.f * (xx - yy) + bSH_C3[1] * sh[30 + c] * xz +
                          bSH_C3[2] * sh[33 + c] * (-3.f * yy + 4.f * zz - xx) + bSH_C3[3] * sh[36 + c] * -3.f * 2.f * yz +
                          bSH_C3[4] * sh[39 + c] * -2.f * xy + bSH_C3[5] * sh[42 + c] * -2.f * yz +
                          bSH_C3[6] * sh[45 + c] * -3.f * 2.f * xy);
              dRdz[c] += (bSH_C3[1] * sh[30 + c] * xy + bSH_C3[2] * sh[33 + c] * 4.f * 2.f * yz +
                          bSH_C3[3] * sh[36 + c] * 3.f * (2.f * zz - xx - yy) + bSH_C3[4] * sh[39 + c] * 4.f * 2.f * xz +
                          bSH_C3[5] * sh[42 + c] * (xx - yy));
            }
          }
        }
      }
#undef ACC
      const float ddx = dRdx[0] * rr[0] + dRdx[1] * rr[1] + dRdx[2] * rr[2];
      const float ddy = dRdy[0] * rr[0] + dRdy[1] * rr[1] + dRdy[2] * rr[2];
      const float ddz = dRdz[0] * rr[0] + dRdz[1] * rr[1] + dRdz[2] * rr[2];
      // dnormvdv (reference auxiliary.h:123-133)
      const float sum2 = ox * ox + oy * oy + oz * oz;
      const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
      dmx += ((+sum2 - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * invsum32;
      dmy += (-ox * oy * ddx + (sum2 - oy * oy) * ddy - oz * oy * ddz) * invsum32;
      dmz += (-ox * oz * ddx - oy * oz * ddy + (sum2 - oz * oz) * ddz) * invsum32;
    }
  }

  if (p.dL_dsh != nullptr) {
    if (staged) {   // through the staging tile: coalesced stores
      __syncthreads();
#pragma unroll
      for (int i = 0; i < MAXC * 3; i++) s_sh[threadIdx.x * SROW + i] = dsh[i];
      __syncthreads();
      const int nrows = min(128, p.P - g0);
      float* dst = p.dL_dsh + (size_t)g0 * row;
      for (int f = threadIdx.x; f < nrows * row; f += 128) {
        const int r = f / row;
        dst[f] = s_sh[r * SROW + (f - r * row)];
      }
    } else if (live) {
      float* o = p.dL_dsh + (size_t)g * p.M * 3;
      for (int i = 0; i < p.M * 3; i++) o[i] = 0.0f;
#pragma unroll
      for (int i = 0; i < MAXC * 3; i++)
        if (i < p.M * 3) o[i] = dsh[i];
    }
  }
  if (!live) return;
  p.dL_dmeans3D[3 * g] = dmx; p.dL_dmeans3D[3 * g + 1] = dmy; p.dL_dmeans3D[3 * g + 2] = dmz;
#pragma unroll
  for (int i = 0; i < 6; i++) p.dL_dcov3D[6 * g + i] = dcov[i];
  p.dL_dopacity[g] = dop;
  if (p.dL_dcolors) { p.dL_dcolors[3 * g] = draw[0]; p.dL_dcolors[3 * g + 1] = draw[1]; p.dL_dcolors[3 * g + 2] = draw[2]; }
  if (p.scales != nullptr && p.dL_dscales != nullptr) {
    // cov3D -> scale / quaternion (reference backward.cu:492-555); quaternion un-normalised
    const float4 q = reinterpret_cast<const float4*>(p.rotations)[g];
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    const M3 R = m3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                         2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                         2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
    const V3 sc = v3(p.scale_modifier * p.scales[3 * g], p.scale_modifier * p.scales[3 * g + 1],
                     p.scale_modifier * p.scales[3 * g + 2]);
    M3 S = m3_zero();
    S.c[0][0] = sc.x; S.c[1][1] = sc.y; S.c[2][2] = sc.z;
    const M3 Mm = S * R;
    const M3 dL_dSigma = m3_cols(dcov[0], 0.5f * dcov[1], 0.5f * dcov[2], 0.5f * dcov[1], dcov[3], 0.5f * dcov[4],
                                 0.5f * dcov[2], 0.5f * dcov[4], dcov[5]);
    const M3 dL_dM = (Mm * 2.0f) * dL_dSigma;
    const M3 Rt = transpose(R);
    M3 dL_dMt = transpose(dL_dM);
    bool invisible_everywhere = true;
    for (int v = 0; v < p.V; v++) invisible_everywhere = invisible_everywhere && !(p.radii[(size_t)v * p.P + g] > 0);
    if (invisible_everywhere) {
      // the reference's kernels return before writing (backward.cu:584), leaving the zero-initialised tensors
      p.dL_dscales[3 * g] = p.dL_dscales[3 * g + 1] = p.dL_dscales[3 * g + 2] = 0.0f;
      reinterpret_cast<float4*>(p.dL_drotations)[g] = make_float4(0, 0, 0, 0);
    } else {
      p.dL_dscales[3 * g] = dot(col(Rt, 0), col(dL_dMt, 0));
      p.dL_dscales[3 * g + 1] = dot(col(Rt, 1), col(dL_dMt, 1));
      p.dL_dscales[3 * g + 2] = dot(col(Rt, 2), col(dL_dMt, 2));
#pragma unroll
      for (int i = 0; i < 3; i++) { dL_dMt.c[0][i] *= sc.x; dL_dMt.c[1][i] *= sc.y; dL_dMt.c[2][i] *= sc.z; }
      float4 dq;
      dq.x = 2 * z * (dL_dMt.c[0][1] - dL_dMt.c[1][0]) + 2 * y * (dL_dMt.c[2][0] - dL_dMt.c[0][2]) + 2 * x * (dL_dMt.c[1][2] - dL_dMt.c[2][1]);
      dq.y = 2 * y * (dL_dMt.c[1][0] + dL_dMt.c[0][1]) + 2 * z * (dL_dMt.c[2][0] + dL_dMt.c[0][2]) + 2 * r * (dL_dMt.c[1][2] - dL_dMt.c[2][1]) - 4 * x * (dL_dMt.c[2][2] + dL_dMt.c[1][1]);
      dq.z = 2 * x * (dL_dMt.c[1][0] + dL_dMt.c[0][1]) + 2 * r * (dL_dMt.c[2][0] - dL_dMt.c[0][2]) + 2 * z * (dL_dMt.c[1][2] + dL_dMt.c[2][1]) - 4 * y * (dL_dMt.c[2][2] + dL_dMt.c[0][0]);
      dq.w = 2 * r * (dL_dMt.c[0][1] - dL_dMt.c[1][0]) + 2 * x * (dL_dMt.c[2][0] + dL_dMt.c[0][2]) + 2 * y * (dL_dMt.c[1][2] + dL_dMt.c[2][1]) - 4 * z * (dL_dMt.c[1][1] + dL_dMt.c[0][0]);
      reinterpret_cast<float4*>(p.dL_drotations)[g] = dq;
    }
  }
}

void launch_gather_bwd(const GatherBwdParams& p, cudaStream_t s) {
  if (p.P <= 0) return;
  const size_t VP = (size_t)p.V * p.P;
  count_launch(2);
  if (p.ns == 16 && p.rec_floats == 12) sum_records_kernel<12><<<(unsigned)((VP * 3 + 255) / 256), 256, 0, s>>>(VP, p.inst_off, p.grec, p.sums, p.capacity);
  else if (p.ns == 16) sum_records_kernel<16><<<(unsigned)((VP * 4 + 255) / 256), 256, 0, s>>>(VP, p.inst_off, p.grec, p.sums, p.capacity);
  else sum_records_kernel<32><<<(unsigned)((VP * 8 + 255) / 256), 256, 0, s>>>(VP, p.inst_off, p.grec, p.sums, p.capacity);
  GatherBwdParams q = p;
  q.skip_geo = p.ns == 16 ? 1 : 0;   // colour-only records: every geometry cotangent is zero
  gather_bwd_kernel<<<(unsigned)((VP + 127) / 128), 128, 0, s>>>(q);
}

void launch_finish_bwd(const FinishBwdParams& p, cudaStream_t s) {
  if (p.P <= 0) return;
  int nb = (p.P + 127) / 128;
  int need = p.shs ? p.M : 1;
  count_launch();
  if (need <= 1) finish_bwd_kernel<1><<<nb, 128, 0, s>>>(p);
  else if (need <= 4) finish_bwd_kernel<4><<<nb, 128, 0, s>>>(p);
  else if (need <= 9) finish_bwd_kernel<9><<<nb, 128, 0, s>>>(p);
  else finish_bwd_kernel<16><<<nb, 128, 0, s>>>(p);
}

}  // namespace igs
