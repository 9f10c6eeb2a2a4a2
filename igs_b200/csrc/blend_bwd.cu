// blend_bwd.cu — backward tile blending, replacing the reference's backward renderCUDA
// (DGR/cuda_rasterizer/backward.cu:631-1016).
//
// The reference issues 25 global float atomicAdd per contributing (pixel, Gaussian) pair.  Here:
//   * traversal is back-to-front exactly like the reference (T is rebuilt by division);
//   * instances are culled per tile and per 8x4 warp block with the same conservative alpha-extent
//     test as the forward kernel, and instances behind every pixel's last contributor are skipped;
//   * per pair, the (up to 25) per-Gaussian partial derivatives are reduced across the 32 pixels of
//     a warp with a butterfly reduce-scatter (31 shuffles for 32 slots instead of 5 per value),
//     accumulated per batch slot in shared memory, and written ONCE per (tile, instance) as a
//     dense gradient record Grec[e] indexed by the instance's emission slot e.  No global atomics:
//     the per-Gaussian sum over its tiles is a contiguous read in gather_bwd (preprocess_bwd.cu),
//     because emission slots of one Gaussian are consecutive.
//   * GEOGRAD=false is the refine-loop case (only the colour map has a cotangent; autograd hands
//     None for the six geometry maps): 10 instead of 25 values per pair, identical results because
//     every omitted term is multiplied by an exactly-zero cotangent in the reference.
#include "common.cuh"
#include "kernels.cuh"

namespace igs {


#define BWD_BATCH 128

__device__ __forceinline__ bool alpha_extent_b(float A, float B, float C, float o, float& ex, float& ey) {
  if (o < 0.00392156862f) return false;
  float det = A * C - B * B;
  float tau = __logf(255.0f * o) * 1.0005f + 1e-3f;
  ex = ey = 3.0e38f;
  if (det > 0.0f && A > 0.0f && C > 0.0f) {
    float k = 2.0f * tau / det;
    float x2 = k * C, y2 = k * A;
    if (x2 < 1e30f && y2 < 1e30f) {
      ex = sqrtf(x2) * 1.0005f + 0.01f;
      ey = sqrtf(y2) * 1.0005f + 0.01f;
    }
  }
  return true;
}

// Butterfly reduce-scatter of NS per-lane values over the 32 lanes of a warp.
// NS == 32: afterwards lane L holds sum over lanes of v[L] in v[0].
// NS == 16: afterwards lanes 2s and 2s+1 hold the sum of v[s] in v[0].
template <int NS>
__device__ __forceinline__ void warp_reduce_scatter(float (&v)[NS], int lane) {
  int n = NS;
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    if (n > 1) {
      const int h = n / 2;
      const bool upper = (lane & off) != 0;
#pragma unroll
      for (int i = 0; i < NS / 2; i++) {
        if (i < h) {
          float keep = upper ? v[i + h] : v[i];
          float send = upper ? v[i] : v[i + h];
          v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
      }
      n = h;
    } else {
      v[0] += __shfl_xor_sync(0xffffffffu, v[0], off);
    }
  }
}

template <bool COORD, bool DEPTH, bool NORMAL, bool GEOGRAD>
__global__ void __launch_bounds__(256)
blend_bwd_kernel(BlendBwdParams p) {
  constexpr int NS = GEOGRAD ? 32 : 16;
  constexpr int NV = GEOGRAD ? 25 : 10;
  constexpr bool GEO = GEOGRAD && (COORD || DEPTH || NORMAL);
  __shared__ float4 s_ha[BWD_BATCH];
  __shared__ float4 s_hb[BWD_BATCH];
  __shared__ float2 s_ext[BWD_BATCH];
  __shared__ uint32_t s_gid[BWD_BATCH];
  __shared__ uint32_t s_pos[BWD_BATCH];
  __shared__ uint32_t s_slot[BWD_BATCH];
  __shared__ uint32_t s_e[BWD_BATCH];
  __shared__ float4 s_t[GEO ? BWD_BATCH * 4 : BWD_BATCH];
  __shared__ float s_grad[BWD_BATCH * NS];
  __shared__ uint32_t s_wcount[8];
  __shared__ uint32_t s_red[8];

  const int tiles = p.tiles_x * p.tiles_y;
  const int gt = blockIdx.x;
  const int v = gt / tiles;
  const int t = gt - v * tiles;
  const int tyi = t / p.tiles_x, txi = t - tyi * p.tiles_x;
  const int w = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int wx0 = txi * IGS_TILE_X + (w & 1) * 8;
  const int wy0 = tyi * IGS_TILE_Y + (w >> 1) * 4;
  const int pxi = wx0 + (lane & 7), pyi = wy0 + (lane >> 3);
  const bool inside = pxi < p.W && pyi < p.H;
  const size_t N = (size_t)p.W * p.H;
  const size_t pix_id = (size_t)p.W * pyi + pxi;
  const float pixfx = (float)pxi, pixfy = (float)pyi;
  const float tx0 = (float)(txi * IGS_TILE_X), ty0 = (float)(tyi * IGS_TILE_Y);
  const float tx1 = (float)(min(txi * IGS_TILE_X + IGS_TILE_X, p.W) - 1);
  const float ty1 = (float)(min(tyi * IGS_TILE_Y + IGS_TILE_Y, p.H) - 1);
  const float bx0 = (float)wx0, by0 = (float)wy0;
  const float bx1 = (float)(wx0 + 7), by1 = (float)(wy0 + 3);

  const uint2 range = p.ranges[gt];
  const uint32_t last_contributor = inside ? p.n_contrib[(size_t)v * 2 * N + pix_id] : 0u;
  const uint32_t max_contributor = inside ? p.n_contrib[(size_t)v * 2 * N + N + pix_id] : 0u;

  // per-tile and per-warp depth of traversal
  uint32_t wlast = last_contributor;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) wlast = max(wlast, __shfl_xor_sync(0xffffffffu, wlast, o));
  if (lane == 0) s_red[w] = wlast;
  __syncthreads();
  uint32_t tile_last = 0;
#pragma unroll
  for (int ww = 0; ww < 8; ww++) tile_last = max(tile_last, s_red[ww]);
  const uint32_t end = range.x + tile_last;  // instances [end, range.y) were never blended

  // zero records of never-visited instances
  for (uint32_t f = (end - range.x) * (NS / 4) + threadIdx.x; f < (range.y - range.x) * (NS / 4); f += 256) {
    uint32_t j = range.x + f / (NS / 4);
    reinterpret_cast<float4*>(p.grec + (size_t)p.point_e[j] * NS)[f % (NS / 4)] = make_float4(0, 0, 0, 0);
  }

  const float w_final = inside ? p.alphas[(size_t)v * N + pix_id] : 0.0f;
  const float T_final = inside ? (1.0f - w_final) : 0.0f;
  float T = T_final;

  float dLp0 = 0, dLp1 = 0, dLp2 = 0, dL_dalpha = 0;
  float dLc0 = 0, dLc1 = 0, dLc2 = 0, dLm0 = 0, dLm1 = 0, dLm2 = 0;
  float dL_dpixel_t = 0, dL_dpixel_mt = 0;
  float dLn0 = 0, dLn1 = 0, dLn2 = 0;
  if (inside) {
    const float* g = p.dL_color + (size_t)v * 3 * N;
    dLp0 = g[pix_id]; dLp1 = g[N + pix_id]; dLp2 = g[2 * N + pix_id];
    if (p.dL_alpha) dL_dalpha = p.dL_alpha[(size_t)v * N + pix_id];
    if constexpr (GEO) {
      const float pnx = (pixfx - p.W / 2.f) / p.focal_x, pny = (pixfy - p.H / 2.f) / p.focal_y;
      const float ln = sqrtf(pnx * pnx + pny * pny + 1.0f);
      const float ww2 = w_final * w_final;
      if constexpr (COORD) {
        const float* ac = p.accum_coord + (size_t)v * 3 * N;
        float a0 = 0, a1 = 0, a2 = 0;
        if (p.dL_coord) {
          const float* q = p.dL_coord + (size_t)v * 3 * N;
          a0 = q[pix_id]; a1 = q[N + pix_id]; a2 = q[2 * N + pix_id];
        }
        dL_dalpha -= a0 * ac[pix_id] / ww2;
        dL_dalpha -= a1 * ac[N + pix_id] / ww2;
        dL_dalpha -= a2 * ac[2 * N + pix_id] / ww2;
        dLc0 = a0 / w_final; dLc1 = a1 / w_final; dLc2 = a2 / w_final;
        if (p.dL_mcoord) {
          const float* q = p.dL_mcoord + (size_t)v * 3 * N;
          dLm0 = q[pix_id]; dLm1 = q[N + pix_id]; dLm2 = q[2 * N + pix_id];
        }
      }
      if constexpr (DEPTH) {
        float dd = p.dL_depth ? p.dL_depth[(size_t)v * N + pix_id] : 0.0f;
        dL_dalpha -= dd * p.accum_depth[(size_t)v * N + pix_id] / ww2;
        dL_dpixel_t = dd / w_final / ln;
        dL_dpixel_mt = (p.dL_mdepth ? p.dL_mdepth[(size_t)v * N + pix_id] : 0.0f) / ln;
      }
      if constexpr (NORMAL) {
        float g0 = 0, g1 = 0, g2 = 0;
        if (p.dL_normal) {
          const float* q = p.dL_normal + (size_t)v * 3 * N;
          g0 = q[pix_id]; g1 = q[N + pix_id]; g2 = q[2 * N + pix_id];
        }
        const float* nm = p.normalmap + (size_t)v * 3 * N;
        float n0 = nm[pix_id], n1 = nm[N + pix_id], n2 = nm[2 * N + pix_id];
        float nlen = p.normal_length[(size_t)v * N + pix_id];
        if (nlen < IGS_NORMALIZE_EPS) {
          dLn0 = g0 / IGS_NORMALIZE_EPS; dLn1 = g1 / IGS_NORMALIZE_EPS; dLn2 = g2 / IGS_NORMALIZE_EPS;
        } else {
          float dt = g0 * n0 + g1 * n1 + g2 * n2;
          dLn0 = (g0 - dt * n0) / nlen; dLn1 = (g1 - dt * n1) / nlen; dLn2 = (g2 - dt * n2) / nlen;
        }
      }
    }
  }
  const float bg_dot_dpixel = p.bg[0] * dLp0 + p.bg[1] * dLp1 + p.bg[2] * dLp2;
  const float ddelx_dx = 0.5f * p.W, ddely_dy = 0.5f * p.H;

  float last_alpha = 0, accum_alpha_rec = 0;
  float ar0 = 0, ar1 = 0, ar2 = 0, lc0 = 0, lc1 = 0, lc2 = 0;      // colour recurrence
  float cr0 = 0, cr1 = 0, cr2 = 0, lco0 = 0, lco1 = 0, lco2 = 0;  // coord recurrence
  float tr = 0, lt = 0;                                            // depth recurrence
  float nr0 = 0, nr1 = 0, nr2 = 0, ln0 = 0, ln1 = 0, ln2 = 0;      // normal recurrence

  for (uint32_t hi = end; hi > range.x;) {
    const uint32_t lo = (hi - range.x > BWD_BATCH) ? hi - BWD_BATCH : range.x;
    const uint32_t nb = hi - lo;
    // ---- 1. fetch (back to front), tile cull, ordered compaction ----
    for (uint32_t f = threadIdx.x; f < BWD_BATCH * NS; f += 256) s_grad[f] = 0.0f;
    bool keep = false;
    float4 ha, hb;
    float ex = 0, ey = 0;
    uint32_t gid = 0, j = 0;
    if (threadIdx.x < nb) {
      j = hi - 1 - threadIdx.x;
      gid = p.point_list[j];
      s_e[threadIdx.x] = p.point_e[j];
      ha = p.head[gid].a;
      hb = p.head[gid].b;
      keep = alpha_extent_b(ha.z, ha.w, hb.x, hb.y, ex, ey);
      keep = keep && (ha.x + ex >= tx0) && (ha.x - ex <= tx1) && (ha.y + ey >= ty0) && (ha.y - ey <= ty1);
    }
    uint32_t bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_wcount[w] = __popc(bal);
    __syncthreads();
    uint32_t woff = 0, count = 0;
#pragma unroll
    for (int ww = 0; ww < 8; ww++) {
      uint32_t c = s_wcount[ww];
      if (ww < w) woff += c;
      count += c;
    }
    if (keep) {
      uint32_t k = woff + __popc(bal & ((1u << lane) - 1u));
      s_ha[k] = ha;
      s_hb[k] = hb;
      s_ext[k] = make_float2(ex, ey);
      s_gid[k] = gid;
      s_pos[k] = j - range.x + 1;
      s_slot[k] = threadIdx.x;
    }
    __syncthreads();
    if constexpr (GEO) {
      for (uint32_t f = threadIdx.x; f < count * 4; f += 256) {
        uint32_t k = f >> 2, q = f & 3;
        s_t[f] = reinterpret_cast<const float4*>(p.tail + s_gid[k])[q];
      }
    } else {
      if (threadIdx.x < count) s_t[threadIdx.x].x = p.tail[s_gid[threadIdx.x]].a.x;
    }
    __syncthreads();
    // ---- 2. per-warp traversal ----
    for (uint32_t base = 0; base < count; base += 32) {
      uint32_t k = base + lane;
      bool hit = false;
      if (k < count) {
        float4 a = s_ha[k];
        float2 e = s_ext[k];
        hit = (s_pos[k] <= wlast) && (a.x + e.x >= bx0) && (a.x - e.x <= bx1) && (a.y + e.y >= by0) &&
              (a.y - e.y <= by1);
      }
      uint32_t m = __ballot_sync(0xffffffffu, hit);
      while (m) {
        int b = __ffs(m) - 1;
        m &= m - 1;
        const uint32_t kk = base + b;
        const uint32_t pos = s_pos[kk];
        const float4 a = s_ha[kk];
        const float4 bq = s_hb[kk];
        const float dx = fsub(a.x, pixfx), dy = fsub(a.y, pixfy);
        const float power = splat_power(dx, dy, a.z, a.w, bq.x);
        const float G = expf(power);
        const float alpha = fminf(0.99f, fmul(bq.y, G));
        const bool active = (pos <= last_contributor) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
        if (!__any_sync(0xffffffffu, active)) continue;

        float val[NS];
#pragma unroll
        for (int i = 0; i < NS; i++) val[i] = 0.0f;
        if (active) {
          T = T / (1.0f - alpha);
          const float aT = alpha * T;
          float dL_dopa = 0.0f;
          float blue;
          if constexpr (GEO) blue = s_t[kk * 4].x; else blue = s_t[kk].x;
          // colour
          ar0 = last_alpha * lc0 + (1.0f - last_alpha) * ar0; lc0 = bq.z; dL_dopa += (bq.z - ar0) * dLp0;
          ar1 = last_alpha * lc1 + (1.0f - last_alpha) * ar1; lc1 = bq.w; dL_dopa += (bq.w - ar1) * dLp1;
          ar2 = last_alpha * lc2 + (1.0f - last_alpha) * ar2; lc2 = blue; dL_dopa += (blue - ar2) * dLp2;
          val[0] = aT * dLp0; val[1] = aT * dLp1; val[2] = aT * dLp2;
          float dL_ddelx_geo = 0.0f, dL_ddely_geo = 0.0f;
          if constexpr (GEO) {
            const float4 t0 = s_t[kk * 4 + 0], t1 = s_t[kk * 4 + 1], t2 = s_t[kk * 4 + 2], t3 = s_t[kk * 4 + 3];
            const bool median = (pos == max_contributor);
            if constexpr (COORD) {
              float c0 = t1.w + t2.z * dx + t2.w * dy;
              float c1 = t2.x + t3.x * dx + t3.y * dy;
              float c2 = t2.y + t3.z * dx + t3.w * dy;
              cr0 = last_alpha * lco0 + (1.0f - last_alpha) * cr0; lco0 = c0; dL_dopa += (c0 - cr0) * dLc0;
              cr1 = last_alpha * lco1 + (1.0f - last_alpha) * cr1; lco1 = c1; dL_dopa += (c1 - cr1) * dLc1;
              cr2 = last_alpha * lco2 + (1.0f - last_alpha) * cr2; lco2 = c2; dL_dopa += (c2 - cr2) * dLc2;
              float g0 = aT * dLc0, g1 = aT * dLc1, g2 = aT * dLc2;
              if (median) { g0 += dLm0; g1 += dLm1; g2 += dLm2; }
              val[10] = g0; val[11] = g1; val[12] = g2;
              val[13] = g0 * dx / p.focal_x; val[14] = g0 * dy / p.focal_y;
              val[15] = g1 * dx / p.focal_x; val[16] = g1 * dy / p.focal_y;
              val[17] = g2 * dx / p.focal_x; val[18] = g2 * dy / p.focal_y;
              dL_ddelx_geo += g0 * t2.z + g1 * t3.x + g2 * t3.z;
              dL_ddely_geo += g0 * t2.w + g1 * t3.y + g2 * t3.w;
            }
            if constexpr (DEPTH) {
              float tt = t0.y + (t0.z * dx + t0.w * dy);
              tr = last_alpha * lt + (1.0f - last_alpha) * tr; lt = tt;
              dL_dopa += (tt - tr) * dL_dpixel_t;
              float dL_dt = aT * dL_dpixel_t;
              if (median) dL_dt += dL_dpixel_mt;
              val[19] = dL_dt;
              val[20] = dL_dt * dx / p.focal_x; val[21] = dL_dt * dy / p.focal_y;
              dL_ddelx_geo += dL_dt * t0.z;
              dL_ddely_geo += dL_dt * t0.w;
            }
            if constexpr (NORMAL) {
              nr0 = last_alpha * ln0 + (1.0f - last_alpha) * nr0; ln0 = t1.x; dL_dopa += (t1.x - nr0) * dLn0;
              nr1 = last_alpha * ln1 + (1.0f - last_alpha) * nr1; ln1 = t1.y; dL_dopa += (t1.y - nr1) * dLn1;
              nr2 = last_alpha * ln2 + (1.0f - last_alpha) * nr2; ln2 = t1.z; dL_dopa += (t1.z - nr2) * dLn2;
              val[22] = aT * dLn0; val[23] = aT * dLn1; val[24] = aT * dLn2;
            }
          }
          accum_alpha_rec = last_alpha + (1.0f - last_alpha) * accum_alpha_rec;
          dL_dopa += (1.0f - accum_alpha_rec) * dL_dalpha;
          dL_dopa *= T;
          last_alpha = alpha;
          dL_dopa += (-T_final / (1.0f - alpha)) * bg_dot_dpixel;

          const float dL_dG = bq.y * dL_dopa;
          const float gdx = G * dx, gdy = G * dy;
          const float dG_ddelx = -gdx * a.z - gdy * a.w;
          const float dG_ddely = -gdy * bq.x - gdx * a.w;
          const float ax = dL_dG * dG_ddelx, ay = dL_dG * dG_ddely;
          val[3] = (ax + dL_ddelx_geo) * ddelx_dx;
          val[4] = (ay + dL_ddely_geo) * ddely_dy;
          val[5] = fabsf(ax * ddelx_dx) + fabsf(ay * ddely_dy);
          val[6] = -0.5f * gdx * dx * dL_dG;
          val[7] = -0.5f * gdx * dy * dL_dG;
          val[8] = -0.5f * gdy * dy * dL_dG;
          val[9] = G * dL_dopa;
        }
        warp_reduce_scatter<NS>(val, lane);
        const int slot = GEOGRAD ? lane : (lane >> 1);
        if (slot < NV && (GEOGRAD || (lane & 1) == 0))
          atomicAdd(&s_grad[s_slot[kk] * NS + slot], val[0]);
      }
    }
    __syncthreads();
    // ---- 3. one dense record per instance of this batch ----
    for (uint32_t f = threadIdx.x; f < nb * (NS / 4); f += 256) {
      uint32_t sl = f / (NS / 4), q = f % (NS / 4);
      reinterpret_cast<float4*>(p.grec + (size_t)s_e[sl] * NS)[q] = reinterpret_cast<const float4*>(s_grad + sl * NS)[q];
    }
    __syncthreads();
    hi = lo;
  }
}

int blend_bwd_record_floats(bool geograd) { return geograd ? 32 : 16; }

void launch_blend_bwd(const BlendBwdParams& p, int V, bool coord, bool depth, bool geograd, cudaStream_t s) {
  int nblocks = V * p.tiles_x * p.tiles_y;
  if (nblocks <= 0) return;
  count_launch();
  if (!geograd || (!coord && !depth)) {
    blend_bwd_kernel<false, false, false, false><<<nblocks, 256, 0, s>>>(p);
  } else if (coord && depth) {
    blend_bwd_kernel<true, true, true, true><<<nblocks, 256, 0, s>>>(p);
  } else if (coord) {
    blend_bwd_kernel<true, false, true, true><<<nblocks, 256, 0, s>>>(p);
  } else {
    blend_bwd_kernel<false, true, true, true><<<nblocks, 256, 0, s>>>(p);
  }
}

}  // namespace igs
