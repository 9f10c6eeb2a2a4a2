// blend_bwd.cu — backward tile blending, replacing the reference's backward renderCUDA
// (DGR/cuda_rasterizer/backward.cu:631-1016).
//
// The reference issues 25 global float atomicAdd per contributing (pixel, Gaussian) pair.  Here:
//   * traversal is back-to-front exactly like the reference (T is rebuilt by division);
//   * instances are culled per tile and per 8x4 warp block with the same conservative alpha-extent
//     test as the forward kernel, and instances behind every pixel's last contributor are skipped;
//   * per pair, the (up to 25) per-Gaussian partial derivatives are reduced across the 32 pixels of
//     a warp with a butterfly reduce-scatter (16 / 31 shuffles for all values together instead of 5 per value),
//     accumulated per batch slot in WARP-PRIVATE shared-memory accumulators (plain load/add/store: float
//     atomics on shared memory compile to CAS loops), and written ONCE per (tile, instance) as a dense
//     gradient record Grec[e] indexed by the instance's emission slot e.  No atomics anywhere: the
//     per-Gaussian sum over its tiles is a contiguous read in gather_bwd (preprocess_bwd.cu), because
//     emission slots of one Gaussian are consecutive;
//   * per-warp culling uses the exact minimum of the splat's quadratic form over the warp's 8x4 block
//     (blend_common.cuh), so instances that cannot reach alpha >= 1/255 on any of its pixels are skipped;
//   * GEOGRAD=false is the refine-loop case (only the colour map has a cotangent; autograd hands
//     None for the six geometry maps): 10 instead of 25 values per pair, identical results because
//     every omitted term is multiplied by an exactly-zero cotangent in the reference.
#include "common.cuh"
#include "blend_common.cuh"
#include <cstdlib>
#include "kernels.cuh"

namespace igs {

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

// instances staged per round: sized so that MINB CTAs fit the SM's shared memory
__host__ __device__ constexpr int bwd_batch(bool geograd, int minb) { return geograd ? 64 : (minb >= 4 ? 96 : 128); }

template <bool COORD, bool DEPTH, bool NORMAL, bool GEOGRAD, int MINB>
__global__ void __launch_bounds__(256, MINB)
blend_bwd_kernel(BlendBwdParams p) {
  constexpr int NS = GEOGRAD ? 32 : 16;         // floats per gradient record in global memory
  constexpr int NV = GEOGRAD ? 25 : 10;         // values actually produced per pair
  constexpr int AS = GEOGRAD ? 25 : 12;         // accumulator stride per batch slot (colour-only: 10 values, a sink word,
                                                // one pad word: 16-byte rows for the vectorised record write-out)
  constexpr int BATCH = bwd_batch(GEOGRAD, MINB);
  constexpr bool GEO = GEOGRAD && (COORD || DEPTH || NORMAL);
  // survivors: [0,B) head.a {x, y, conic.x, conic.y}; [B,2B) head.b {conic.z, opacity, r, g};
  // [2B,3B) {blue, 1.0f, list position (bits), batch slot (bits)}  ((blue, 1) is an fp32x2 operand of the recurrences)
  __shared__ float4 s_rec[BATCH * 3];
  __shared__ float s_tau[BATCH];
  __shared__ float4 s_t[GEO ? BATCH * 4 : 1];
  __shared__ uint32_t s_gid[BATCH];
  __shared__ uint32_t s_e[BATCH];
  __shared__ uint32_t s_wcount[8];
  __shared__ uint32_t s_red[8];
  // per-warp private gradient accumulators [8][BATCH][NV]: no shared-memory atomics (float atomics on shared
  // memory are CAS loops), summed over the 8 warps when the batch's records are written
  extern __shared__ __align__(16) float s_acc[];
  // colour-only path: per-warp [10][36] transpose staging for the cross-pixel reduction (columns 32..35 stay zero)
  __shared__ __align__(16) float s_stage[GEOGRAD ? 4 : 8 * 10 * 36];

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
  const float bx1 = (float)min(wx0 + 7, p.W - 1), by1 = (float)min(wy0 + 3, p.H - 1);
  const uint32_t rec_base = smem_addr(s_rec);
  const uint32_t tail_base = smem_addr(s_t);
  const uint32_t acc_base = smem_addr(s_acc) + (uint32_t)w * (BATCH * AS * 4);
  for (uint32_t f = threadIdx.x; f < 8 * BATCH * AS; f += 256) s_acc[f] = 0.0f;
  if constexpr (!GEOGRAD) {
    for (uint32_t f = threadIdx.x; f < 8 * 10 * 36; f += 256) s_stage[f] = 0.0f;
  }
  const uint32_t stage_base = smem_addr(s_stage) + (GEOGRAD ? 0u : (uint32_t)w * (10 * 36 * 4));
  // reduction role of this lane in the colour-only path: value rv = lane / 3, segment rs = lane % 3 (12 pixels each)
  const uint32_t rv = (uint32_t)lane / 3u, rs = (uint32_t)lane - rv * 3u;
  const uint32_t stage_wr = stage_base + (uint32_t)lane * 4;
  const bool red_owner = (rs == 0u) && (lane < 30);
  const uint32_t stage_rd = stage_base + (lane < 30 ? (rv * 36 + rs * 12) * 4 : 0u);  // lanes 30, 31: any valid row
  // accumulator address of this lane for batch slot s is red_base + s * AS * 4: the 10 owner lanes hit the slot's 10
  // values, every other lane the slot's 11th word (one broadcast address in the bank next to the owners': no conflicts)
  const uint32_t red_base = acc_base + (red_owner ? rv : 10u) * 4;

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
  const f32x2 grad01 = pack2(dLp0, dLp1), grad2a = pack2(dLp2, dL_dalpha);
  const f32x2 pix2 = pack2(pixfx, pixfy), half2 = pack2(p.half_w, p.half_h);

  float last_alpha = 0;
  // colour + alpha recurrences as two fp32x2 pairs: acc01 = (accum_rec r, g), acc2a = (accum_rec b, accum_alpha_rec);
  // last01 = (last colour r, g), last2o = (last colour b, 1);  grad01 = (dL/dpixel r, g), grad2a = (dL/dpixel b, dL/dalpha)
  f32x2 acc01 = pack2(0.0f, 0.0f), acc2a = pack2(0.0f, 0.0f), last01 = pack2(0.0f, 0.0f), last2o = pack2(0.0f, 1.0f);
  float cr0 = 0, cr1 = 0, cr2 = 0, lco0 = 0, lco1 = 0, lco2 = 0;  // coord recurrence
  float tr = 0, lt = 0;                                            // depth recurrence
  float nr0 = 0, nr1 = 0, nr2 = 0, ln0 = 0, ln1 = 0, ln2 = 0;      // normal recurrence

  for (uint32_t hi = end; hi > range.x;) {
    const uint32_t lo = (hi - range.x > BATCH) ? hi - BATCH : range.x;
    const uint32_t nb = hi - lo;
    // ---- 1. fetch (back to front), tile cull, ordered compaction ----
    bool keep = false;
    float4 ha, hb;
    float ex = 0, ey = 0, tau = 0;
    uint32_t gid = 0, j = 0;
    if (threadIdx.x < nb) {
      j = hi - 1 - threadIdx.x;
      gid = p.point_list[j];
      s_e[threadIdx.x] = p.point_e[j];
      ha = p.head[gid].a;
      hb = p.head[gid].b;
      keep = alpha_extent(ha.z, ha.w, hb.x, hb.y, ex, ey, tau);
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
      s_rec[k] = ha;
      s_rec[BATCH + k] = hb;
      s_rec[2 * BATCH + k] = make_float4(p.tail[gid].a.x, 1.0f, __uint_as_float(j - range.x + 1), __uint_as_float(threadIdx.x));
      s_tau[k] = tau;
      s_gid[k] = gid;
    }
    __syncthreads();
    if constexpr (GEO) {
      for (uint32_t f = threadIdx.x; f < count * 4; f += 256) {
        uint32_t k = f >> 2, q = f & 3;
        s_t[f] = reinterpret_cast<const float4*>(p.tail + s_gid[k])[q];
      }
      __syncthreads();
    }
    // ---- 2. per-warp traversal ----
    for (uint32_t base = 0; base < count; base += 32) {
      uint32_t k = base + lane;
      bool hit = false;
      if (k < count) {
        const uint32_t ak = rec_base + k * 16;
        const float4 a = lds128<0>(ak);
        const float4 b = lds128<BATCH * 16>(ak);
        const float4 c = lds128<BATCH * 32>(ak);
        hit = (__float_as_uint(c.z) <= wlast) && block_can_contribute(a.x, a.y, a.z, a.w, b.x, s_tau[k], bx0, by0, bx1, by1);
      }
      uint32_t m = __ballot_sync(0xffffffffu, hit);
      while (m) {
        const uint32_t kk = base + (__ffs(m) - 1);
        m &= m - 1;
        const uint32_t ak = rec_base + kk * 16;
        const float4 a = lds128<0>(ak);
        const float4 bq = lds128<BATCH * 16>(ak);
        const float4 cq = lds128<BATCH * 32>(ak);  // {blue, 1, position, slot}
        const uint32_t pos = __float_as_uint(cq.z);
        const f32x2 d2 = sub2(pack2(a.x, a.y), pix2);  // (dx, dy), each rounded like fsub()
        float dx, dy;
        unpack2(d2, dx, dy);
        const float power = splat_power(dx, dy, a.z, a.w, bq.x);
        const float G = expf(power);
        const float alpha = fminf(0.99f, fmul(bq.y, G));
        const bool active = (pos <= last_contributor) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
        if (!__any_sync(0xffffffffu, active)) continue;

        float val[GEOGRAD ? NS : NV];
#pragma unroll
        for (int i = 0; i < (GEOGRAD ? NS : NV); i++) val[i] = 0.0f;
        if (active) {
          // reference: T = T / (1 - alpha) and (-T_final / (1 - alpha)); one reciprocal serves both
          float rcp;
          asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rcp) : "f"(1.0f - alpha));  // 1 - alpha is in [0.01, 1]
          T = T * rcp;
          const float aT = alpha * T;
          // colour and alpha:  accum_rec = last_alpha * last_colour + (1 - last_alpha) * accum_rec  (alpha: last "colour" 1),
          // dL_dopa = sum_c (colour_c - accum_rec_c) * dL_dpixel_c + (1 - accum_alpha_rec) * dL_dalpha
          const f32x2 la2 = pack2(last_alpha, last_alpha), om2 = pack2(1.0f - last_alpha, 1.0f - last_alpha);
          const f32x2 col01 = pack2(bq.z, bq.w), col2o = pack2(cq.x, cq.y);
          acc01 = fma2(acc01, om2, mul2(last01, la2));
          acc2a = fma2(acc2a, om2, mul2(last2o, la2));
          last01 = col01;
          last2o = col2o;
          float dL_dopa = hsum2(fma2(sub2(col2o, acc2a), grad2a, mul2(sub2(col01, acc01), grad01)));
          unpack2(mul2(grad01, pack2(aT, aT)), val[0], val[1]);
          val[2] = aT * dLp2;
          float dL_ddelx_geo = 0.0f, dL_ddely_geo = 0.0f;
          if constexpr (GEO) {
            const uint32_t at = tail_base + kk * 64;
            const float4 t0 = lds128<0>(at), t1 = lds128<16>(at), t2 = lds128<32>(at), t3 = lds128<48>(at);
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
          dL_dopa *= T;
          last_alpha = alpha;
          dL_dopa += (-T_final * rcp) * bg_dot_dpixel;

          const float dL_dG = bq.y * dL_dopa;
          const f32x2 gd2 = mul2(d2, pack2(G, G));  // (G dx, G dy)
          float gdx, gdy;
          unpack2(gd2, gdx, gdy);
          const float dG_ddelx = -gdx * a.z - gdy * a.w;
          const float dG_ddely = -gdy * bq.x - gdx * a.w;
          const f32x2 axy = mul2(pack2(dG_ddelx, dG_ddely), pack2(dL_dG, dL_dG));
          if constexpr (GEO) {
            float ax, ay;
            unpack2(axy, ax, ay);
            val[3] = (ax + dL_ddelx_geo) * p.half_w;
            val[4] = (ay + dL_ddely_geo) * p.half_h;
            val[5] = fabsf(ax * p.half_w) + fabsf(ay * p.half_h);
          } else {
            unpack2(mul2(axy, half2), val[3], val[4]);
            val[5] = fabsf(val[3]) + fabsf(val[4]);
          }
          const f32x2 h2 = mul2(gd2, pack2(-0.5f, -0.5f));
          float hx, hy;
          unpack2(h2, hx, hy);
          unpack2(mul2(mul2(h2, d2), pack2(dL_dG, dL_dG)), val[6], val[8]);  // -0.5 G dx dx dL_dG, -0.5 G dy dy dL_dG
          val[7] = hx * dy * dL_dG;
          val[9] = G * dL_dopa;
        }
        if constexpr (GEOGRAD) {
          warp_reduce_scatter<NS>(val, lane);  // afterwards lane L holds the warp sum of value L
          if (lane < NV) {
            const uint32_t aa = acc_base + (__float_as_uint(cq.w) * AS + lane) * 4;
            sts32<0>(aa, lds32<0>(aa) + val[0]);
          }
        } else {
          // transpose through shared memory: 10 conflict-free stores, then lane (rv, rs) sums 12 pixels of value
          // rv with three 16-byte loads and the three segments are combined with two shuffles
          sts32<0 * 144>(stage_wr, val[0]); sts32<1 * 144>(stage_wr, val[1]); sts32<2 * 144>(stage_wr, val[2]);
          sts32<3 * 144>(stage_wr, val[3]); sts32<4 * 144>(stage_wr, val[4]); sts32<5 * 144>(stage_wr, val[5]);
          sts32<6 * 144>(stage_wr, val[6]); sts32<7 * 144>(stage_wr, val[7]); sts32<8 * 144>(stage_wr, val[8]);
          sts32<9 * 144>(stage_wr, val[9]);
          __syncwarp();
          const float4 x0 = lds128<0>(stage_rd), x1 = lds128<16>(stage_rd), x2 = lds128<32>(stage_rd);
          float sum = hsum2(add2(add2(add2(pack2(x0.x, x0.y), pack2(x0.z, x0.w)), add2(pack2(x1.x, x1.y), pack2(x1.z, x1.w))),
                                 add2(pack2(x2.x, x2.y), pack2(x2.z, x2.w))));
          sum += __shfl_down_sync(0xffffffffu, sum, 1) + __shfl_down_sync(0xffffffffu, sum, 2);
          const uint32_t aa = red_base + __float_as_uint(cq.w) * (uint32_t)(AS * 4);
          sts32<0>(aa, lds32<0>(aa) + sum);
          __syncwarp();
        }
      }
    }
    __syncthreads();
    // ---- 3. one dense record per instance of this batch: sum (and clear) the 8 warps' accumulators ----
    if constexpr (GEOGRAD) {
      for (uint32_t f = threadIdx.x; f < nb * NS; f += 256) {
        const uint32_t sl = f / NS, q = f % NS;
        float sum = 0.0f;
        if (q < NV) {
#pragma unroll
          for (int ww = 0; ww < 8; ww++) {
            float* pa = s_acc + ((size_t)ww * BATCH + sl) * AS + q;
            sum += *pa;
            *pa = 0.0f;
          }
        }
        p.grec[(size_t)s_e[sl] * NS + q] = sum;
      }
    } else {
      // 16 bytes of a record per thread: three quarters come from the accumulators (values 0-9; words 10, 11 are the
      // sink / pad and are ignored downstream), the fourth is zero so that every 32-byte sector is written whole
      for (uint32_t f = threadIdx.x; f < nb * 4; f += 256) {
        const uint32_t sl = f >> 2, q4 = f & 3;
        float4 sum = make_float4(0, 0, 0, 0);
        if (q4 < 3) {
#pragma unroll
          for (int ww = 0; ww < 8; ww++) {
            float4* pa = reinterpret_cast<float4*>(s_acc + ((size_t)ww * BATCH + sl) * AS) + q4;
            const float4 x = *pa;
            sum.x += x.x; sum.y += x.y; sum.z += x.z; sum.w += x.w;
            *pa = make_float4(0, 0, 0, 0);
          }
        }
        reinterpret_cast<float4*>(p.grec + (size_t)s_e[sl] * NS)[q4] = sum;
      }
    }
    __syncthreads();
    hi = lo;
  }
}


// ---------------------------------------------------------------------------------------------
// Colour-only backward, two-pass form (the refine loop's kernel; selected by launch_blend_bwd).
//
// The kernel above reduces every (8x4 block, instance) pair's ten partial derivatives across the 32 pixels of
// the warp right away: 10 STS + 3 LDS.128 + a read-modify-write per pair, which put it on the shared-memory pipe
// (ncu r01: l1tex 87 %, 27 shared wavefronts and 31 of 113 instructions per pair for the reduction alone).
// Here the two roles of a lane are separated:
//   pass 1  lanes = PIXELS.  For one hit instance the per-pixel recurrences of the reference (T rebuilt by
//           division, the accumulated colour behind the instance; backward.cu:842-1013) collapse, for a colour
//           cotangent, into two scalars per pixel — the blend weight aT = alpha*T and q = G * dL/dopacity_term —
//           because every one of the ten per-Gaussian derivatives is LINEAR in (aT, q) with coefficients that depend
//           only on the pixel position and the instance.  The two scalars are staged in a per-warp [16 hits][32 px]
//           shared-memory tile (one conflict-free 8-byte store per pair).
//   pass 2  lanes = INSTANCES.  After 16 hits, lane (j, h) walks the 16 pixels of half h of the block for hit j and
//           accumulates the ten sums in registers (no cross-lane reduction at all; the two halves are combined
//           with one shuffle per value), then writes them to the warp's private accumulator row of the instance's
//           batch slot — once, with plain stores.
// Measured at C3: shared-memory wavefronts per pair 41 -> 18.5, warp instructions per pair 139 -> 90.
// Same traversal order, same culling, same records as the kernel above; per-pixel activity tests use the same
// bit-pinned power / expf / alpha chain as the forward pass, so exactly the forward's contributors are visited.
// ---------------------------------------------------------------------------------------------
#define BWD2_GROUP 16
#define BWD2_STRIDE 33  // staged float2 per hit: 32 pixels + 1 pad (conflict-free for both access patterns)

__device__ __forceinline__ float2 lds64_at(uint32_t addr) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ float4 lds128_at(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts64_at(uint32_t addr, float a, float b) {
  asm volatile("st.shared.v2.f32 [%0], {%1,%2};" ::"r"(addr), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void sts128_at(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// CAP = accumulator rows per warp (a multiple of the group size).  A warp whose queue for one batch is longer than CAP
// works it off in rounds of CAP hits with a write-out after each (later rounds add to the record in global memory);
// with CAP = BATCH there is exactly one round.  A smaller CAP buys a larger BATCH at the same shared-memory footprint.
// MASK: the forward pass left one byte per instance with the blocks of its tile it can reach (BlendBwdParams::hitmask).
// Then neither the tile-level cull (alpha extent against the tile) nor the eight per-block tests are repeated here: an
// instance survives when its byte is non-zero and warp w queues it when bit w is set — the same pairs the tests below
// select (the forward evaluates them for every instance at or before a block's last contributor), at a fraction of the
// instructions (the per-warp cull was 11.5 % of this kernel's).  MASK = false keeps the self-contained cull.
template <int BATCH, int MINB, int CAP, bool MASK>
__global__ void __launch_bounds__(256, MINB)
blend_bwd2_kernel(BlendBwdParams p) {
  constexpr int NS = 12;  // floats per gradient record in global memory: the ten sums + 2 pad (48 bytes; was 64)
  constexpr int AS = 12;  // accumulator stride per batch slot (10 values + 2 pad words: 16-byte rows)
  constexpr int G = BWD2_GROUP;
  static_assert(BATCH <= 255 && BATCH % 32 == 0, "hit queue entries are bytes; whole 32-entry segments");
  static_assert(CAP % BWD2_GROUP == 0 && CAP <= BATCH, "groups do not straddle rounds");
  __shared__ float4 s_rec[BATCH * 3];     // survivors: {x, y, conic.x, conic.y} {conic.z, opacity, r, g} {blue, position, tau | hit byte, slot}
  __shared__ float4 s_rawh[BATCH * 2];    // cp.async landing zone of the NEXT batch: heads as fetched (entry t: thread t's)
  __shared__ float s_rawb[BATCH];         //   blue
  __shared__ uint32_t s_rawe[BATCH];      //   emission slot
  __shared__ uint32_t s_e[2 * BATCH];     // emission slots of the batch, double-buffered (read by the previous batch's write-out)
  __shared__ uint8_t s_queue[8 * BATCH];  // per warp: shared-memory indices of the survivors that hit the warp's block
  // inverse of the queues, double-buffered like s_e: byte [slot][w] = position of batch entry `slot` in warp w's queue
  // (0xFF: the entry does not touch that warp's block).  The accumulator rows are indexed by QUEUE POSITION, so only
  // rows that were written are ever read: no zero-fill, no clearing, and the write-out reads ~2 rows per entry, not 8.
  __shared__ __align__(16) uint8_t s_inv[2 * BATCH * 8];
  __shared__ uint32_t s_wcount[8];
  __shared__ uint32_t s_red[8];
  extern __shared__ __align__(16) float s_dyn[];
  float* s_acc = s_dyn;                                             // [8 warps][CAP queue positions of a round][AS]
  float* s_stage = s_dyn + 8 * CAP * AS;                            // [8 warps][G][33] float2 (aT, q)
  float4* s_pix = reinterpret_cast<float4*>(s_stage + 8 * G * BWD2_STRIDE * 2);  // [8 warps][34] {dLp r, g, b, 1}

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
  const float bx1 = (float)min(wx0 + 7, p.W - 1), by1 = (float)min(wy0 + 3, p.H - 1);
  const uint32_t rec_base = smem_addr(s_rec);
  const uint32_t acc_base = smem_addr(s_acc) + (uint32_t)w * (CAP * AS * 4);
  const uint32_t stage_base = smem_addr(s_stage) + (uint32_t)w * (G * BWD2_STRIDE * 8);
  const uint32_t stage_wr = stage_base + (uint32_t)lane * 8;  // + i * 33 * 8 for hit i of the group
  // pass-2 role of this lane: hit j of the group, pixel rows {2h, 2h+1} of the block
  const uint32_t j = (uint32_t)lane & 15u, h = (uint32_t)lane >> 4;
  const uint32_t stage_rd = stage_base + (j * BWD2_STRIDE + 16u * h) * 8;  // + s * 8 for pixel 16h + s
  // pixel records are stored at index p + (p >> 4): the pad keeps the two halves' 16-byte reads in different banks
  const uint32_t pix_base = smem_addr(s_pix) + (uint32_t)w * (34 * 16);
  const uint32_t pix_rd = pix_base + (16u * h + h) * 16;  // + s * 16
  uint8_t* const queue = s_queue + w * BATCH;
  const uint32_t rawh_addr = smem_addr(s_rawh) + threadIdx.x * 32;
  const uint32_t rawb_addr = smem_addr(s_rawb) + threadIdx.x * 4;
  const uint32_t rawe_addr = smem_addr(s_rawe) + threadIdx.x * 4;

  const uint2 range = p.ranges[gt];
  const uint32_t last_contributor = inside ? p.n_contrib[(size_t)v * 2 * N + pix_id] : 0u;

  uint32_t wlast = last_contributor;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) wlast = max(wlast, __shfl_xor_sync(0xffffffffu, wlast, o));
  if (lane == 0) s_red[w] = wlast;

  const float w_final = inside ? p.alphas[(size_t)v * N + pix_id] : 0.0f;
  const float T_final = inside ? (1.0f - w_final) : 0.0f;
  float dLp0 = 0, dLp1 = 0, dLp2 = 0, dL_dalpha = 0;
  if (inside) {
    const float* g = p.dL_color + (size_t)v * 3 * N;
    dLp0 = g[pix_id]; dLp1 = g[N + pix_id]; dLp2 = g[2 * N + pix_id];
    if (p.dL_alpha) dL_dalpha = p.dL_alpha[(size_t)v * N + pix_id];
  }
  sts128_at(pix_base + ((uint32_t)lane + ((uint32_t)lane >> 4)) * 16, dLp0, dLp1, dLp2, 1.0f);   // (b, 1): multiplicand of (aT, q)
  // The alpha cotangent and the background enter dL/dopacity of instance i only as
  //   T_final / (1 - alpha_i) * (dL/dalpha - bg . dL/dpixel):
  // the reference's accumulated-alpha recurrence (backward.cu:972-975) telescopes, 1 - accum_alpha_rec = T_final / T_behind.
  const float Kpix = T_final * (dL_dalpha - (p.bg[0] * dLp0 + p.bg[1] * dLp1 + p.bg[2] * dLp2));
  const f32x2 pix2 = pack2(pixfx, pixfy);
  __syncthreads();
  uint32_t tile_last = 0;
#pragma unroll
  for (int ww = 0; ww < 8; ww++) tile_last = max(tile_last, s_red[ww]);
  const uint32_t end = range.x + tile_last;  // instances [end, range.y) were never blended

  // ---- software pipeline of the batch fetch (north_star (3): asynchronous shared-memory staging) ----
  // The list entry -> head gather is two dependent global loads.  They are taken off the critical path:
  //   * the Gaussian id of a batch is loaded into a register TWO batches ahead,
  //   * its head / blue / emission slot are copied global -> shared with cp.async ONE batch ahead,
  // both while the warps traverse the current batch; a batch therefore starts from data already in shared memory.
  // Every thread copies, waits for and reads only ITS OWN landing-zone entry, so no barrier is needed for it.
  auto issue_copies = [&](uint32_t hi_b, uint32_t gid_b, uint32_t mask_b) {
    const uint32_t lo_b = (hi_b - range.x > BATCH) ? hi_b - BATCH : range.x;
    if (threadIdx.x < hi_b - lo_b) {
      if (!MASK || mask_b != 0u) {   // (MASK: an instance that reaches no block is never looked at again)
        const float4* hp = reinterpret_cast<const float4*>(p.head + gid_b);
        cp_async16(rawh_addr, hp);
        cp_async16(rawh_addr + 16, hp + 1);
        cp_async4(rawb_addr, &p.tail[gid_b].a.x);
      }
      cp_async4(rawe_addr, p.point_e + (hi_b - 1 - threadIdx.x));   // every instance gets a record
    }
    cp_async_commit();
  };
  auto load_gid = [&](uint32_t hi_b) -> uint32_t {  // list entry of this thread in the batch that ends at hi_b
    if (hi_b <= range.x) return 0u;
    const uint32_t lo_b = (hi_b - range.x > BATCH) ? hi_b - BATCH : range.x;
    return (threadIdx.x < hi_b - lo_b) ? p.point_list[hi_b - 1 - threadIdx.x] : 0u;
  };
  auto load_mask = [&](uint32_t hi_b) -> uint32_t {  // its block-hit byte (MASK)
    if (!MASK || hi_b <= range.x) return 0u;
    const uint32_t lo_b = (hi_b - range.x > BATCH) ? hi_b - BATCH : range.x;
    return (threadIdx.x < hi_b - lo_b) ? (uint32_t)p.hitmask[hi_b - 1 - threadIdx.x] : 0u;
  };
  uint32_t gid_next = 0, mask_cur = 0, mask_next = 0;
  if (end > range.x) {
    mask_cur = load_mask(end);
    issue_copies(end, load_gid(end), mask_cur);
    const uint32_t hi2 = (end - range.x > BATCH) ? end - BATCH : range.x;
    gid_next = load_gid(hi2);
    mask_next = load_mask(hi2);
  }

  // zero records of never-visited instances
  for (uint32_t f = (end - range.x) * (NS / 4) + threadIdx.x; f < (range.y - range.x) * (NS / 4); f += 256) {
    uint32_t jj = range.x + f / (NS / 4);
    reinterpret_cast<float4*>(p.grec + (size_t)p.point_e[jj] * NS)[f % (NS / 4)] = make_float4(0, 0, 0, 0);
  }

  // per-pixel state of the back-to-front traversal: T (transmittance in front of the last visited contributor) and
  // A = sum_c accum_rec_c * dL/dpixel_c, the accumulated colour BEHIND the next instance dotted with the cotangent
  float T = T_final, A = 0.0f;
  uint32_t parity = 0;

  for (uint32_t hi = end; hi > range.x;) {
    const uint32_t lo = (hi - range.x > BATCH) ? hi - BATCH : range.x;
    const uint32_t nb = hi - lo;
    // ---- 1. this batch's entries (already in shared memory), tile cull, compaction ----
    // Survivors are compacted INSIDE each 32-entry segment (warp w's entries go to s_rec[32 w ...]), so no cross-warp
    // prefix and no barrier is needed before the writes; list order is segment order, then order inside the segment.
    cp_async_wait_all();
    bool keep = false;
    float4 ha, hb;
    float ex = 0, ey = 0, tau = 0, blue = 0;
    uint32_t* const e_cur = s_e + (parity ? BATCH : 0);
    uint8_t* const inv_cur = s_inv + (parity ? BATCH * 8 : 0);
    if (threadIdx.x < BATCH * 2) reinterpret_cast<uint32_t*>(inv_cur)[threadIdx.x] = 0xFFFFFFFFu;
    if (threadIdx.x < nb) {
      ha = lds128_at(rawh_addr);
      hb = lds128_at(rawh_addr + 16);
      blue = s_rawb[threadIdx.x];
      e_cur[threadIdx.x] = s_rawe[threadIdx.x];
      if constexpr (MASK) {
        keep = mask_cur != 0u;
        tau = __uint_as_float(mask_cur);
      } else {
        keep = alpha_extent(ha.z, ha.w, hb.x, hb.y, ex, ey, tau);
        keep = keep && (ha.x + ex >= tx0) && (ha.x - ex <= tx1) && (ha.y + ey >= ty0) && (ha.y - ey <= ty1);
      }
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_wcount[w] = __popc(bal);
    if (keep) {
      const uint32_t k = (uint32_t)w * 32u + __popc(bal & ((1u << lane) - 1u));
      s_rec[k] = ha;
      s_rec[BATCH + k] = hb;
      s_rec[2 * BATCH + k] = make_float4(blue, __uint_as_float(hi - threadIdx.x - range.x), tau, __uint_as_float(threadIdx.x));
    }
    // next batch: global -> shared copies now, its successor's list entries into a register
    if (lo > range.x) {
      issue_copies(lo, gid_next, mask_next);
      mask_cur = mask_next;
      const uint32_t hi2 = (lo - range.x > BATCH) ? lo - BATCH : range.x;
      gid_next = load_gid(hi2);
      mask_next = load_mask(hi2);
    }
    __syncthreads();
    // ---- 2. per-warp traversal ----
    // 2a. survivors against this warp's 8x4 block -> queue of hits (in list order)
    uint32_t qn = 0;
    for (uint32_t seg = 0; seg * 32 < nb; seg++) {
      const uint32_t k = seg * 32 + lane;
      bool hit = false;
      uint32_t slot_of_k = 0;
      if ((uint32_t)lane < s_wcount[seg]) {
        const uint32_t ak = rec_base + k * 16;
        const float4 c = lds128<BATCH * 32>(ak);
        slot_of_k = __float_as_uint(c.w);
        if constexpr (MASK) {
          hit = (__float_as_uint(c.y) <= wlast) && ((__float_as_uint(c.z) >> w) & 1u);
        } else {
          const float4 a = lds128<0>(ak);
          const float4 b = lds128<BATCH * 16>(ak);
          hit = (__float_as_uint(c.y) <= wlast) && block_can_contribute(a.x, a.y, a.z, a.w, b.x, c.z, bx0, by0, bx1, by1);
        }
      }
      const uint32_t m = __ballot_sync(0xffffffffu, hit);
      if (hit) {
        const uint32_t pos = qn + __popc(m & ((1u << lane) - 1u));
        queue[pos] = (uint8_t)k;
        inv_cur[slot_of_k * 8u + (uint32_t)w] = (uint8_t)pos;
      }
      qn += __popc(m);
    }
    __syncwarp();
    for (uint32_t r0 = 0;; r0 += CAP) {
    const uint32_t rend = min(qn, r0 + (uint32_t)CAP);
    for (uint32_t g0 = r0; g0 < rend; g0 += G) {
      const uint32_t ng = min((uint32_t)G, rend - g0);
      // ---- pass 1: lanes = pixels, one hit per iteration ----
#pragma unroll 8
      for (uint32_t i = 0; i < ng; i++) {
        const uint32_t kk = queue[g0 + i];
        const uint32_t ak = rec_base + kk * 16;
        const float4 a = lds128<0>(ak);
        const float4 bq = lds128<BATCH * 16>(ak);
        const float2 cq = lds64_at(ak + BATCH * 32);  // {blue, position}: one 8-byte load
        const uint32_t pos = __float_as_uint(cq.y);
        float dx, dy;
        unpack2(sub2(pack2(a.x, a.y), pix2), dx, dy);  // each rounded like fsub()
        const float power = splat_power(dx, dy, a.z, a.w, bq.x);
        const float Gv = expf(power);
        const float alpha = fminf(0.99f, fmul(bq.y, Gv));
        const bool active = (pos <= last_contributor) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
        // reference: T = T / (1 - alpha) and (-T_final / (1 - alpha)); one reciprocal serves both.  Evaluated on every
        // lane and selected (1 - alpha is in [0.01, 1] wherever the result is kept).
        float rcp;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rcp) : "f"(1.0f - alpha));
        const float Tn = T * rcp;
        const float cd = fmaf(cq.x, dLp2, fmaf(bq.w, dLp1, bq.z * dLp0));  // colour . dL/dpixel
        const float diff = cd - A;
        const float qv = Gv * fmaf(diff, Tn, Kpix * rcp);  // G * ("dL_dopa" of the reference)
        sts64_at(stage_wr + i * (BWD2_STRIDE * 8), active ? alpha * Tn : 0.0f, active ? qv : 0.0f);
        T = active ? Tn : T;
        A = active ? fmaf(alpha, diff, A) : A;  // accumulated colour behind the NEXT instance
      }
      // ---- pass 2: lanes = (hit j of the group, half h of the block) ----
      __syncwarp();
      const uint32_t kj = queue[g0 + (j < ng ? j : 0u)];
      const uint32_t ak = rec_base + kj * 16;
      const float4 a = lds128<0>(ak);
      const float4 bq = lds128<BATCH * 16>(ak);
      // Every per-Gaussian sum is linear in (aT, q) with coefficients that are polynomials of the pixel offset
      // d = centre - pixel.  With the pixel at column c, row r of this lane's half (d = (gx - c, gy - r)) the q-sums
      // expand into MOMENTS of q over the half, which need no per-pixel geometry at all:
      //   sum q dx   = gx S0 - Sx                  sum q dx^2  = gx^2 S0 - 2 gx Sx + Sxx
      //   sum q dy   = gy S0 - Sy                  sum q dx dy = gx gy S0 - gx Sy - gy Sx + Sxy
      //                                            sum q dy^2  = gy^2 S0 - 2 gy Sy + Syy
      // (S0 = sum q, Sx = sum q c, Sxx = sum q c^2, ...).  The moments are taken about the CENTRE of the half (c' = c - 3.5,
      // r' = r - 0.5 = -+0.5, g' = g - (3.5, 0.5)): for a Gaussian inside or next to the block the expansion about the
      // block corner cancels (gx^2 S0 - 2 gx Sx + Sxx with every term ~10x the result), about the centre Sx', Sy' are
      // small.  (Checked against a build with direct per-pixel sums: the scale / rotation gradients moved by less than
      // the reference's own run-to-run spread, tools/grad_noise_floor.py.)
      // With two rows: Sy' = (row1 - row0)/2, Syy' = S0/4, Sxy' = (Sx'_row1 - Sx'_row0)/2.
      // Only the |.| sums of dL/dmean2D still need d per pixel.
      const float gx = a.x - bx0;
      const float gy = (a.y - by0) - (float)(2u * h);
      const float gxc = gx - 3.5f, gyc = gy - 0.5f;
      // The kernel is issue bound: the sums are advanced in fp32x2 pairs — (v0, v1) += aT (dLr, dLg),
      // (v2, S0) += (aT, q) (dLb, 1) straight from the staged (aT, q) pair, (u, w) = (ub, wb) - c (cx, cy) — 7 instead of 10
      // arithmetic instructions per pixel; every element is still one fused multiply-add.
      float b3 = 0, b4 = 0;
      f32x2 V01 = pack2(0.0f, 0.0f);
      f32x2 VR[2];      // per pixel row: (sum aT dLb, sum q)
      float R1[2], R2[2];
      const f32x2 ncxy = pack2(-a.z, -a.w);
#pragma unroll
      for (int r = 0; r < 2; r++) {
        const float ddy = gy - (float)r;
        const f32x2 ubwb = pack2(fmaf(a.z, gx, a.w * ddy), fmaf(a.w, gx, bq.x * ddy));  // cx dx + cy dy, cy dx + cz dy at c = 0
        f32x2 vr = pack2(0.0f, 0.0f);
        float r1 = 0, r2 = 0;
#pragma unroll
        for (int c = 0; c < 8; c++) {
          const int s = r * 8 + c;
          const float2 st = lds64_at(stage_rd + s * 8);
          const float4 dl = lds128_at(pix_rd + s * 16);
          V01 = fma2(pack2(st.x, st.x), pack2(dl.x, dl.y), V01);
          vr = fma2(pack2(st.x, st.y), pack2(dl.z, dl.w), vr);
          r1 = fmaf(st.y, (float)c - 3.5f, r1);
          r2 = fmaf(st.y, ((float)c - 3.5f) * ((float)c - 3.5f), r2);
          float u, wv;
          unpack2(fma2(ncxy, pack2((float)c, (float)c), ubwb), u, wv);
          b3 = fmaf(fabsf(st.y), fabsf(u), b3);   // |q u| = |q| |u|: the absolute values ride on the FFMA's operands
          b4 = fmaf(fabsf(st.y), fabsf(wv), b4);
        }
        VR[r] = vr; R1[r] = r1; R2[r] = r2;
      }
      float v0, v1, v2, R0[2];
      unpack2(V01, v0, v1);
      {
        float va, vb;
        unpack2(VR[0], va, R0[0]);
        unpack2(VR[1], vb, R0[1]);
        v2 = va + vb;
      }
      const float S0 = R0[0] + R0[1], Sx = R1[0] + R1[1], Sxx = R2[0] + R2[1];
      const float Sy = 0.5f * (R0[1] - R0[0]), Sxy = 0.5f * (R1[1] - R1[0]);
      float v9 = S0;
      float a3 = fmaf(gxc, S0, -Sx);
      float a4 = fmaf(gyc, S0, -Sy);
      float s6 = fmaf(gxc, fmaf(gxc, S0, -2.0f * Sx), Sxx);
      float s7 = fmaf(gxc, fmaf(gyc, S0, -Sy), fmaf(-gyc, Sx, Sxy));
      float s8 = fmaf(gyc, fmaf(gyc, S0, -2.0f * Sy), 0.25f * S0);
      v0 += __shfl_xor_sync(0xffffffffu, v0, 16); v1 += __shfl_xor_sync(0xffffffffu, v1, 16);
      v2 += __shfl_xor_sync(0xffffffffu, v2, 16); v9 += __shfl_xor_sync(0xffffffffu, v9, 16);
      s6 += __shfl_xor_sync(0xffffffffu, s6, 16); s7 += __shfl_xor_sync(0xffffffffu, s7, 16);
      s8 += __shfl_xor_sync(0xffffffffu, s8, 16); a3 += __shfl_xor_sync(0xffffffffu, a3, 16);
      a4 += __shfl_xor_sync(0xffffffffu, a4, 16); b3 += __shfl_xor_sync(0xffffffffu, b3, 16);
      b4 += __shfl_xor_sync(0xffffffffu, b4, 16);
      if (h == 0 && j < ng) {
        // dL/dG = opacity * dL_dopa;  dG/ddelx = -G (cx dx + cy dy), dG/ddely = -G (cy dx + cz dy);
        // dL/dconic = -0.5 G d d^T dL/dG;  |grad| accumulates the per-pixel absolute values
        const float o = bq.y;
        const float mo = -o, ho = -0.5f * o;
        const uint32_t aa = acc_base + (g0 - r0 + j) * (uint32_t)(AS * 4);
        sts128_at(aa, v0, v1, v2, mo * fmaf(a.z, a3, a.w * a4) * p.half_w);
        sts128_at(aa + 16, mo * fmaf(a.w, a3, bq.x * a4) * p.half_h, o * fmaf(p.half_w, b3, p.half_h * b4), ho * s6, ho * s7);
        sts64_at(aa + 32, ho * s8, v9);
      }
      __syncwarp();
    }
    const int more = __syncthreads_or(qn > r0 + (uint32_t)CAP);
    // ---- 3. one dense record per instance of this batch: sum the 8 warps' accumulator rows of this round ----
    // 16 bytes of a record per thread (values 0-9; words 10, 11 are pad)
    for (uint32_t f = threadIdx.x; f < nb * 3; f += 256) {
      const uint32_t sl = f / 3u, q4 = f - sl * 3u;
      float4 sum = make_float4(0, 0, 0, 0);
      {
        const uint2 iv = *reinterpret_cast<const uint2*>(inv_cur + sl * 8u);
#pragma unroll
        for (int ww = 0; ww < 8; ww++) {   // warp order, as before: absent rows were zeros then and are skipped now
          const uint32_t qp = ((ww < 4 ? iv.x : iv.y) >> (8 * (ww & 3))) & 0xFFu;
          if (qp - r0 < (uint32_t)CAP) {   // (0xFF and positions of other rounds fall outside)
            const float4 x = reinterpret_cast<const float4*>(s_acc + ((size_t)ww * CAP + (qp - r0)) * AS)[q4];
            sum.x += x.x; sum.y += x.y; sum.z += x.z; sum.w += x.w;
          }
        }
      }
      float4* const rec = reinterpret_cast<float4*>(p.grec + (size_t)e_cur[sl] * NS) + q4;
      if (r0 != 0) {   // a later round of an over-long queue: the same thread wrote this part in the round before
        const float4 old = *rec;
        sum.x += old.x; sum.y += old.y; sum.z += old.z; sum.w += old.w;
      }
      *rec = sum;
    }
    if (!more) break;
    __syncthreads();   // the accumulator rows are free for the next round
    }
    // no barrier here: the next batch writes the OTHER halves of s_e / s_inv, and s_rec / s_acc only change hands at its barrier
    hi = lo;
    parity ^= 1u;
  }
}

static int colour_bwd_impl() {   // IGS_BWD_IMPL=1 selects the first-generation colour-only kernel (reduction per pair); default: two-pass kernel
  static const int impl = [] { const char* e = getenv("IGS_BWD_IMPL"); return e ? atoi(e) : 2; }();
  return impl;
}

// floats per gradient record written by the kernel launch_blend_bwd selects (the stride of BlendBwdParams::grec)
int blend_bwd_record_floats(bool geograd) { return geograd ? 32 : (colour_bwd_impl() >= 2 ? 12 : 16); }

template <bool C, bool D, bool N, bool G, int MINB>
static void launch_one(const BlendBwdParams& p, int nblocks, cudaStream_t s) {
  constexpr int bytes = 8 * bwd_batch(G, MINB) * (G ? 25 : 12) * 4;  // per-warp accumulators (dynamic shared memory)
  // > 48 KB of dynamic shared memory needs the opt-in; it is per device, so it is (cheaply) set on every launch
  cudaFuncSetAttribute(blend_bwd_kernel<C, D, N, G, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  blend_bwd_kernel<C, D, N, G, MINB><<<nblocks, 256, bytes, s>>>(p);
}

template <int BATCH, int MINB, int CAP, bool MASK>
static void launch_two_pass_m(const BlendBwdParams& p, int nblocks, cudaStream_t s) {
  // dynamic shared memory: per-warp accumulators + per-warp (aT, q) staging tiles + per-warp pixel records
  constexpr int bytes = 8 * CAP * 12 * 4 + 8 * BWD2_GROUP * BWD2_STRIDE * 8 + 8 * 34 * 16;
  cudaFuncSetAttribute(blend_bwd2_kernel<BATCH, MINB, CAP, MASK>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  cudaFuncSetAttribute(blend_bwd2_kernel<BATCH, MINB, CAP, MASK>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
  blend_bwd2_kernel<BATCH, MINB, CAP, MASK><<<nblocks, 256, bytes, s>>>(p);
}

template <int BATCH, int MINB, int CAP>
static void launch_two_pass(const BlendBwdParams& p, int nblocks, cudaStream_t s) {
  if (p.hitmask != nullptr) launch_two_pass_m<BATCH, MINB, CAP, true>(p, nblocks, s);
  else launch_two_pass_m<BATCH, MINB, CAP, false>(p, nblocks, s);
}

void launch_blend_bwd(const BlendBwdParams& p, int V, bool coord, bool depth, bool geograd, cudaStream_t s) {
  int nblocks = V * p.tiles_x * p.tiles_y;
  if (nblocks <= 0) return;
  count_launch();
  // IGS_BWD_OCC=4 selects the 64-register / 96-instance-batch build of the colour-only kernel (4 CTAs/SM)
  static const int occ = [] { const char* e = getenv("IGS_BWD_OCC"); return e ? atoi(e) : 4; }();
  if (!geograd || (!coord && !depth)) {
    const int impl = colour_bwd_impl();
    static const int cfg = [] { const char* e = getenv("IGS_BWD2_CFG"); return e ? atoi(e) : 0; }();
    if (impl >= 2) {
      // default: 128-instance batches with 48 accumulator rows per warp (3 CTAs/SM): half the block barriers of the
      // 64-instance build and fuller pass-2 groups (C3: 9.40 -> 8.63 ms); IGS_BWD2_CFG selects the other builds
      if (cfg == 1) launch_two_pass<64, 3, 64>(p, nblocks, s);
      else if (cfg == 2) launch_two_pass<96, 3, 48>(p, nblocks, s);
      else if (cfg == 3) launch_two_pass<96, 3, 64>(p, nblocks, s);
      else if (cfg == 4) launch_two_pass<128, 2, 128>(p, nblocks, s);
      else if (cfg == 5) launch_two_pass<128, 3, 32>(p, nblocks, s);
      else launch_two_pass<128, 3, 48>(p, nblocks, s);
    } else if (occ >= 4) launch_one<false, false, false, false, 4>(p, nblocks, s);
    else launch_one<false, false, false, false, 3>(p, nblocks, s);
  } else if (coord && depth) launch_one<true, true, true, true, 2>(p, nblocks, s);
  else if (coord) launch_one<true, false, true, true, 2>(p, nblocks, s);
  else launch_one<false, true, true, true, 2>(p, nblocks, s);
}

}  // namespace igs
