// blend_fwd.cu — forward tile blending (colour / expected+median coord / expected+median depth /
// normal / weight), replacing the reference's forward renderCUDA (DGR/cuda_rasterizer/forward.cu:428-693).
//
// One CTA per (view, 16x16 tile); each of its 8 warps owns an 8x4 pixel block.  Per batch of up to
// 256 instances of the tile's depth-sorted list:
//   1. every thread fetches one instance's 32-B head and tests it against the tile with a
//      conservative bounding box of the alpha >= 1/255 ellipse; survivors are compacted IN ORDER
//      into shared memory together with their list position (the reference's `contributor` counts
//      every visited instance, SURVEY T12, so positions — not compacted indices — are kept);
//   2. survivors' 64-B tails are fetched;
//   3. each warp tests 32 survivors at a time against its own 8x4 block (one lane per survivor),
//      ballots, and only the hits are evaluated per pixel.
// The per-pixel arithmetic that decides alpha / T / termination is pinned to the reference's
// compiled operation order (common.cuh splat_power; expf, not __expf), so n_contrib is bit-exact.
#include "common.cuh"
#include "blend_common.cuh"
#include <cstdlib>
#include "kernels.cuh"

#ifndef IGS_FWD2_BATCH
#define IGS_FWD2_BATCH 128
#endif

namespace igs {

#define BLEND_BATCH 256

// COUNT = true is the instrumentation build behind igs_count_blend_pairs: the same traversal, no image is written, and
// the work counters of bench.py's roofline are accumulated instead.
template <bool COORD, bool DEPTH, bool NORMAL, int MINB, bool COUNT = false>
__global__ void __launch_bounds__(256, MINB)
blend_fwd_kernel(BlendFwdParams p) {
  constexpr bool GEO = COORD || DEPTH || NORMAL;
  // survivors of the current batch: [0,256) head.a {x, y, conic.x, conic.y}; [256,512) head.b {conic.z, opacity, r, g};
  // [512,768) {blue, 1.0f, list position (bits), tau}  ((blue, 1) is the fp32x2 multiplicand of (C2, weight))
  __shared__ float4 s_rec[BLEND_BATCH * 3];
  __shared__ float4 s_t[GEO ? BLEND_BATCH * 4 : 1];  // tails of the survivors
  __shared__ uint32_t s_gid[BLEND_BATCH];
  __shared__ uint32_t s_wcount[8];
  // Hit masks for the backward pass: s_ball[w][seg] = ballot of "survivor 32 seg + lane can reach warp w's 8x4 block".
  // After a batch every survivor's eight bits are gathered into ONE byte per instance (p.hitmask, by list position), so the
  // backward neither repeats the tile-level cull nor the eight per-block tests: it walks exactly these pairs.
  __shared__ uint32_t s_ball[8][BLEND_BATCH / 32];

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
  // tile bounds in pixel-centre coordinates (clipped to the image)
  const float tx0 = (float)(txi * IGS_TILE_X), ty0 = (float)(tyi * IGS_TILE_Y);
  const float tx1 = (float)(min(txi * IGS_TILE_X + IGS_TILE_X, p.W) - 1);
  const float ty1 = (float)(min(tyi * IGS_TILE_Y + IGS_TILE_Y, p.H) - 1);
  const float bx0 = (float)wx0, by0 = (float)wy0;
  const float bx1 = (float)min(wx0 + 7, p.W - 1), by1 = (float)min(wy0 + 3, p.H - 1);
  const uint32_t rec_base = smem_addr(s_rec);
  const uint32_t tail_base = smem_addr(s_t);

  const uint2 range = p.ranges[gt];
  int done = inside ? 0 : 1;
  uint32_t n_hits = 0, n_eval = 0, n_contr = 0, n_staged = 0;  // COUNT only

  float T = 1.0f;
  uint32_t last_contributor = 0, max_contributor = 0xFFFFFFFFu;
  // Accumulators kept as fp32x2 pairs: one FFMA2 updates two of them (the kernel is issue bound); every element is
  // still one fused multiply-add with the rounding of the scalar instruction it replaces.
  f32x2 C01 = pack2(0.0f, 0.0f), C2w = pack2(0.0f, 0.0f);    // (C0, C1), (C2, weight)
  f32x2 Co0D = pack2(0.0f, 0.0f), Co12 = pack2(0.0f, 0.0f);  // (Co0, Depth), (Co1, Co2)
  f32x2 N01 = pack2(0.0f, 0.0f);
  float N2 = 0;
  const f32x2 pix2 = pack2(pixfx, pixfy);

  // one byte per survivor of the batch just blended (all eight warps' ballots are complete after a block barrier)
  uint32_t prev_count = 0;
  auto write_hit_masks = [&]() {
    if (threadIdx.x < prev_count) {
      const uint32_t seg = threadIdx.x >> 5;
      uint32_t mk = 0;
#pragma unroll
      for (int ww = 0; ww < 8; ww++) mk |= ((s_ball[ww][seg] >> lane) & 1u) << ww;
      const uint32_t pos = __float_as_uint(s_rec[2 * BLEND_BATCH + threadIdx.x].z);   // 1-based list position
      p.hitmask[range.x + pos - 1] = (uint8_t)mk;
    }
    prev_count = 0;
  };
  const bool masks = !COUNT && p.hitmask != nullptr;
  for (uint32_t start = range.x; start < range.y; start += BLEND_BATCH) {
    const int ndone = __syncthreads_count(done != 0);
    if (masks && prev_count) write_hit_masks();
    if (ndone == 256) break;
    // ---- 1. fetch + tile-level cull + ordered compaction ----
    uint32_t j = start + threadIdx.x;
    bool keep = false;
    float4 ha, hb;
    float ex = 0, ey = 0, tau = 0;
    uint32_t gid = 0;
    if (j < range.y) {
      if constexpr (COUNT) n_staged++;
      gid = p.point_list[j];
      ha = p.head[gid].a;
      hb = p.head[gid].b;
      keep = alpha_extent(ha.z, ha.w, hb.x, hb.y, ex, ey, tau);
      keep = keep && (ha.x + ex >= tx0) && (ha.x - ex <= tx1) && (ha.y + ey >= ty0) && (ha.y - ey <= ty1);
    }
    uint32_t bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_wcount[w] = __popc(bal);
    __syncthreads();
    // (every thread has finished gathering the previous batch's ballots; the next ballot is stored after another barrier)
    if (masks && threadIdx.x < 8 * (BLEND_BATCH / 32)) (&s_ball[0][0])[threadIdx.x] = 0u;
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
      s_rec[BLEND_BATCH + k] = hb;
      // blue lives in the tail record; 1-based list position = the reference's `contributor`
      s_rec[2 * BLEND_BATCH + k] = make_float4(p.tail[gid].a.x, 1.0f, __uint_as_float(j - range.x + 1), tau);
      s_gid[k] = gid;
    }
    __syncthreads();
    // ---- 2. tails of the survivors ----
    if constexpr (GEO) {
      for (uint32_t f = threadIdx.x; f < count * 4; f += 256) {
        uint32_t k = f >> 2, q = f & 3;
        s_t[f] = reinterpret_cast<const float4*>(p.tail + s_gid[k])[q];
      }
      __syncthreads();
    }
    // ---- 3. per-warp cull + blend ----
    if (!__all_sync(0xffffffffu, done != 0)) {
      for (uint32_t base = 0; base < count; base += 32) {
        uint32_t k = base + lane;
        bool hit = false;
        if (k < count) {
          const uint32_t ak = rec_base + k * 16;
          float4 a = lds128<0>(ak);
          float4 b = lds128<BLEND_BATCH * 16>(ak);
          float tk = lds32<BLEND_BATCH * 32 + 12>(ak);
          hit = block_can_contribute(a.x, a.y, a.z, a.w, b.x, tk, bx0, by0, bx1, by1);
        }
        uint32_t m = __ballot_sync(0xffffffffu, hit);
        if (masks && lane == 0) s_ball[w][base >> 5] = m;
        while (m) {
          const uint32_t kk = base + (__ffs(m) - 1);
          m &= m - 1;
          // The tests of the reference (done / power > 0 / alpha < 1/255 / T underflow) are evaluated for every lane
          // without branching — a warp instruction costs the same for 1 or 32 active lanes — and folded into ONE
          // predicate, so the loop control stays warp-uniform and the body has a single divergent region.
          const uint32_t ak = rec_base + kk * 16;
          const float4 a = lds128<0>(ak);
          const float4 bq = lds128<BLEND_BATCH * 16>(ak);
          float dx, dy;
          unpack2(sub2(pack2(a.x, a.y), pix2), dx, dy);  // each rounded like fsub()
          const float power = splat_power(dx, dy, a.z, a.w, bq.x);
          const float alpha = fminf(0.99f, fmul(bq.y, expf(power)));
          float aT, test_T;
          unpack2(mul2(pack2(alpha, fsub(1.0f, alpha)), pack2(T, T)), aT, test_T);  // fmul(alpha, T), fmul(T, 1 - alpha)
          const bool cand = (done == 0) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
          const bool term = cand && (test_T < 0.0001f);
          if constexpr (COUNT) {
            n_hits += (lane == 0);
            n_eval += (done == 0) && !(power > 0.0f);
            n_contr += cand && !term;
          }
          if (term) done = 1;
          if (!cand || term) continue;
          const float4 cq = lds128<BLEND_BATCH * 32>(ak);  // {blue, 1, position, tau}
          const uint32_t contributor = __float_as_uint(cq.z);
          const bool before_median = T > 0.5f;
          const f32x2 aT2 = pack2(aT, aT);
          if constexpr (GEO) {
            const uint32_t at = tail_base + kk * 64;
            const float4 t0 = lds128<0>(at), t1 = lds128<16>(at), t2 = lds128<32>(at), t3 = lds128<48>(at);
            float c0 = 0.0f, tt = 0.0f;
            if constexpr (COORD) {
              c0 = ffma(dy, t2.w, ffma(dx, t2.z, t1.w));
              const float c1 = ffma(dy, t3.y, ffma(dx, t3.x, t2.x));
              const float c2 = ffma(dy, t3.w, ffma(dx, t3.z, t2.y));
              Co12 = fma2(aT2, pack2(c1, c2), Co12);
            }
            if constexpr (DEPTH) tt = fadd(t0.y, ffma(t0.z, dx, fmul(t0.w, dy)));
            if constexpr (COORD || DEPTH) Co0D = fma2(aT2, pack2(c0, tt), Co0D);
            if constexpr (NORMAL) {
              N01 = fma2(aT2, pack2(t1.x, t1.y), N01);
              N2 = ffma(t1.z, aT, N2);
            }
            if (before_median) max_contributor = contributor;
          }
          C01 = fma2(aT2, pack2(bq.z, bq.w), C01);
          C2w = fma2(aT2, pack2(cq.x, cq.y), C2w);  // C2 += aT * blue, weight += aT * 1
          T = test_T;
          last_contributor = contributor;
        }
        if (__all_sync(0xffffffffu, done != 0)) break;
      }
    }
    prev_count = count;
  }
  if (masks && prev_count) {   // the list ended with the last batch: no loop-top barrier came after its blend
    __syncthreads();
    write_hit_masks();
  }

  if constexpr (COUNT) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      n_hits += __shfl_xor_sync(0xffffffffu, n_hits, o);
      n_eval += __shfl_xor_sync(0xffffffffu, n_eval, o);
      n_contr += __shfl_xor_sync(0xffffffffu, n_contr, o);
      n_staged += __shfl_xor_sync(0xffffffffu, n_staged, o);
    }
    if (lane == 0) {
      atomicAdd(p.counters + 0, (unsigned long long)n_hits);
      atomicAdd(p.counters + 1, (unsigned long long)n_eval);
      atomicAdd(p.counters + 2, (unsigned long long)n_contr);
      atomicAdd(p.counters + 3, (unsigned long long)n_staged);
    }
    return;
  }
  float C0, C1, C2, weight, Co0, Co1, Co2, Depth, N0, N1;
  unpack2(C01, C0, C1);
  unpack2(C2w, C2, weight);
  unpack2(Co0D, Co0, Depth);
  unpack2(Co12, Co1, Co2);
  unpack2(N01, N0, N1);
  // Median coordinate / depth = the values of the LAST contributor blended while T > 0.5 (max_contributor).  Carrying them
  // through the traversal cost four selects per (block, instance) hit; they are re-evaluated here, once per pixel, from
  // that one instance's records with the same rounded operations as in the loop.
  float mCo0 = 0, mCo1 = 0, mCo2 = 0, mDepth = 0;
  if constexpr (GEO) {
    if (inside && max_contributor != 0xFFFFFFFFu) {
      const uint32_t gid = p.point_list[range.x + max_contributor - 1];
      const float4 ha = p.head[gid].a;
      const float4* tp = reinterpret_cast<const float4*>(p.tail + gid);
      float dx, dy;
      unpack2(sub2(pack2(ha.x, ha.y), pix2), dx, dy);
      if constexpr (COORD) {
        const float4 t1 = tp[1], t2 = tp[2], t3 = tp[3];
        mCo0 = ffma(dy, t2.w, ffma(dx, t2.z, t1.w));
        mCo1 = ffma(dy, t3.y, ffma(dx, t3.x, t2.x));
        mCo2 = ffma(dy, t3.w, ffma(dx, t3.z, t2.y));
      }
      if constexpr (DEPTH) {
        const float4 t0 = tp[0];
        mDepth = fadd(t0.y, ffma(t0.z, dx, fmul(t0.w, dy)));
      }
    }
  }
  if (inside) {
    p.n_contrib[(size_t)v * 2 * N + pix_id] = last_contributor;
    p.n_contrib[(size_t)v * 2 * N + N + pix_id] = max_contributor;
    float* oc = p.out_color + (size_t)v * 3 * N;
    oc[pix_id] = C0 + T * p.bg[0];
    oc[N + pix_id] = C1 + T * p.bg[1];
    oc[2 * N + pix_id] = C2 + T * p.bg[2];
    p.out_alpha[(size_t)v * N + pix_id] = weight;
    if constexpr (GEO) {
      const float pnx = (pixfx - p.W / 2.f) / p.focal_x, pny = (pixfy - p.H / 2.f) / p.focal_y;
      const float ln = sqrtf(pnx * pnx + pny * pny + 1.0f);
      if constexpr (COORD) {
        float* o = p.out_coord + (size_t)v * 3 * N;
        float* om = p.out_mcoord + (size_t)v * 3 * N;
        float* ac = p.accum_coord + (size_t)v * 3 * N;
        if (last_contributor) {
          o[pix_id] = Co0 / weight; o[N + pix_id] = Co1 / weight; o[2 * N + pix_id] = Co2 / weight;
        } else {
          o[pix_id] = 0; o[N + pix_id] = 0; o[2 * N + pix_id] = 0;
        }
        ac[pix_id] = Co0; ac[N + pix_id] = Co1; ac[2 * N + pix_id] = Co2;
        om[pix_id] = mCo0; om[N + pix_id] = mCo1; om[2 * N + pix_id] = mCo2;
      }
      if constexpr (DEPTH) {
        float depth_ln = Depth / ln;
        p.accum_depth[(size_t)v * N + pix_id] = depth_ln;
        p.out_depth[(size_t)v * N + pix_id] = last_contributor ? depth_ln / weight : 0.0f;
        p.out_mdepth[(size_t)v * N + pix_id] = mDepth / ln;
      }
      if constexpr (NORMAL) {
        float* on = p.out_normal + (size_t)v * 3 * N;
        if (last_contributor) {
          float len = sqrtf(N0 * N0 + N1 * N1 + N2 * N2);
          p.normal_length[(size_t)v * N + pix_id] = len;
          len = fmaxf(len, IGS_NORMALIZE_EPS);
          on[pix_id] = N0 / len; on[N + pix_id] = N1 / len; on[2 * N + pix_id] = N2 / len;
        } else {
          p.normal_length[(size_t)v * N + pix_id] = 1.0f;
          on[pix_id] = 0; on[N + pix_id] = 0; on[2 * N + pix_id] = 0;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Forward blend, second generation: asynchronous double-buffered staging (north_star subsystem 3).
//
// The kernel above stages a batch with synchronous loads and four block barriers (instance id -> head -> cull ->
// compaction -> tails), the shape of the reference's staging loop (forward.cu:511-542), and then EVERY warp tests every
// survivor against its own 8x4 block.  Here
//   * the two dependent global loads are off the critical path: the instance ids of batch k+2 are loaded into a
//     register and the heads / tails of batch k+1 are copied global -> shared with cp.async (LDGSTS, no register
//     staging) while the warps blend batch k; the landing zone is double buffered and blended IN PLACE (survivors are
//     not moved);
//   * culling is done ONCE per instance instead of once per (warp, instance): the thread that owns an entry tests it
//     against the eight 8x4 blocks of the tile (exact minimum of the quadratic form over each block, blend_common.cuh)
//     and the warp ballots become eight 32-bit hit masks per 32-entry segment; a blending warp just walks the set bits
//     of its own masks;
//   * ONE block barrier per batch (it publishes batch k and its masks, counts the finished pixels and frees the buffer
//     of batch k-1 for the next copies) instead of four.
// Per-pixel arithmetic is unchanged (the pinned chain of common.cuh), so every output is bit-identical to the kernel
// above; tails are copied for every instance of a list, not only for survivors — the extra L2 traffic (64 B per
// instance) is irrelevant for a kernel at 4 % of HBM bandwidth.
// ---------------------------------------------------------------------------------------------
template <bool COORD, bool DEPTH, bool NORMAL, int MINB, int B>
__global__ void __launch_bounds__(256, MINB)
blend_fwd2_kernel(BlendFwdParams p) {
  constexpr bool GEO = COORD || DEPTH || NORMAL;
  static_assert(B == 128 || B == 256, "batch of 128 (two threads per entry) or 256 (one)");
  constexpr int SPLIT = 256 / B;                       // threads per entry
  constexpr int SEGS = B / 32;
  constexpr int BUF_BYTES = B * 32 + (GEO ? B * 64 : B * 4);   // [head.a][head.b][tails | blue]
  extern __shared__ __align__(16) unsigned char s_land[];   // 2 * BUF_BYTES (dynamic: 48.5 KB at B = 256)
  __shared__ uint32_t s_hit[2][8][SEGS];               // [buffer][8x4 block = blending warp][segment]

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
  const uint32_t land_base = smem_addr(s_land);
  // role of this thread in staging / culling: entry e of the batch; with two threads per entry, half 0 copies the head
  // (+ first tail quarter / blue) and culls block column 0, half 1 copies the other tail quarters and culls column 1
  const uint32_t e = threadIdx.x & (B - 1), half = threadIdx.x / B;
  const uint32_t seg = e >> 5;

  const uint2 range = p.ranges[gt];
  const uint32_t nbatch = (range.y - range.x + B - 1) / B;
  int done = inside ? 0 : 1;

  auto load_gid = [&](uint32_t k) -> uint32_t {
    const uint32_t j = range.x + k * B + e;
    return (k < nbatch && j < range.y) ? p.point_list[j] : 0u;
  };
  auto issue_copies = [&](uint32_t k, uint32_t gid) {
    const uint32_t j = range.x + k * B + e;
    const uint32_t bb = land_base + (k & 1u) * BUF_BYTES;
    if (j < range.y) {
      const float4* hp = reinterpret_cast<const float4*>(p.head + gid);
      const float4* tp = reinterpret_cast<const float4*>(p.tail + gid);
      if (half == 0) {
        cp_async16(bb + e * 16, hp);
        cp_async16(bb + B * 16 + e * 16, hp + 1);
        if constexpr (GEO) cp_async16(bb + B * 32 + e * 64, tp);
        else cp_async4(bb + B * 32 + e * 4, &p.tail[gid].a.x);
      }
      if constexpr (GEO) {
        if (half == SPLIT - 1) {
          cp_async16(bb + B * 32 + e * 64 + 16, tp + 1);
          cp_async16(bb + B * 32 + e * 64 + 32, tp + 2);
          cp_async16(bb + B * 32 + e * 64 + 48, tp + 3);
        }
      }
    }
    cp_async_commit();
  };

  float T = 1.0f;
  uint32_t last_contributor = 0, max_contributor = 0xFFFFFFFFu;
  f32x2 C01 = pack2(0.0f, 0.0f), C2w = pack2(0.0f, 0.0f);    // (C0, C1), (C2, weight)
  f32x2 Co0D = pack2(0.0f, 0.0f), Co12 = pack2(0.0f, 0.0f);  // (Co0, Depth), (Co1, Co2)
  f32x2 N01 = pack2(0.0f, 0.0f);
  float N2 = 0;
  float mCo0 = 0, mCo1 = 0, mCo2 = 0, mDepth = 0;
  const f32x2 pix2 = pack2(pixfx, pixfy);

  uint32_t gid_next = 0;
  if (nbatch > 0) {
    issue_copies(0, load_gid(0));
    gid_next = load_gid(1);
  }
  for (uint32_t k = 0; k < nbatch; k++) {
    const uint32_t bb = land_base + (k & 1u) * BUF_BYTES;
    // ---- 1. this batch has landed; cull the own entry against the tile and its 8x4 blocks, in place ----
    cp_async_wait_all();
    if constexpr (SPLIT == 2) __syncthreads();   // the head was copied by the entry's other thread
    {
      bool keep = false;
      float4 ha = make_float4(0, 0, 0, 0);
      float cz = 0.0f, tau = 0.0f;
      const uint32_t j = range.x + k * B + e;
      if (j < range.y) {
        ha = lds128<0>(bb + e * 16);
        const float2 hb = lds64<B * 16>(bb + e * 16);
        cz = hb.x;
        float ex, ey;
        keep = alpha_extent(ha.z, ha.w, hb.x, hb.y, ex, ey, tau);
        keep = keep && (ha.x + ex >= tx0) && (ha.x - ex <= tx1) && (ha.y + ey >= ty0) && (ha.y - ey <= ty1);
      }
#pragma unroll
      for (int c = 0; c < 2; c++) {
        if (SPLIT == 2 && c != (int)half) continue;
        const float cx0 = tx0 + 8.0f * c, cx1 = fminf(cx0 + 7.0f, tx1);
#pragma unroll
        for (int r = 0; r < 4; r++) {
          const float ry0 = ty0 + 4.0f * r, ry1 = fminf(ry0 + 3.0f, ty1);
          const bool hit = keep && block_can_contribute(ha.x, ha.y, ha.z, ha.w, cz, tau, cx0, ry0, cx1, ry1);
          const uint32_t m = __ballot_sync(0xffffffffu, hit);
          if (lane == 0) s_hit[k & 1u][c + 2 * r][seg] = m;
        }
      }
    }
    // the one full barrier of the batch: publishes batch k and its hit masks, counts finished pixels, frees the buffer
    // of batch k-1
    if (__syncthreads_count(done != 0) == 256) break;
    if (k + 1 < nbatch) {
      issue_copies(k + 1, gid_next);
      gid_next = load_gid(k + 2);
    }
    if (__all_sync(0xffffffffu, done != 0)) continue;
    // ---- 2. blend the hits of this warp's block, in list order ----
    const uint32_t tb = bb + B * 32;
#pragma unroll 1
    for (uint32_t sg = 0; sg < SEGS; sg++) {
      uint32_t m = s_hit[k & 1u][w][sg];
      while (m) {
        const uint32_t kk = sg * 32 + (__ffs(m) - 1);
        m &= m - 1;
        const uint32_t ak = bb + kk * 16;
        const float4 a = lds128<0>(ak);
        const float4 bq = lds128<B * 16>(ak);
        float dx, dy;
        unpack2(sub2(pack2(a.x, a.y), pix2), dx, dy);  // each rounded like fsub()
        const float power = splat_power(dx, dy, a.z, a.w, bq.x);
        const float alpha = fminf(0.99f, fmul(bq.y, expf(power)));
        float aT, test_T;
        unpack2(mul2(pack2(alpha, fsub(1.0f, alpha)), pack2(T, T)), aT, test_T);  // fmul(alpha, T), fmul(T, 1 - alpha)
        const bool cand = (done == 0) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
        const bool term = cand && (test_T < 0.0001f);
        if (term) done = 1;
        if (!cand || term) continue;
        const uint32_t contributor = k * B + kk + 1;   // 1-based list position = the reference's `contributor`
        const bool before_median = T > 0.5f;
        const f32x2 aT2 = pack2(aT, aT);
        float blue;
        if constexpr (GEO) {
          const uint32_t at = tb + kk * 64;
          const float4 t0 = lds128<0>(at), t1 = lds128<16>(at), t2 = lds128<32>(at), t3 = lds128<48>(at);
          blue = t0.x;
          float c0 = 0.0f, tt = 0.0f;
          if constexpr (COORD) {
            c0 = ffma(dy, t2.w, ffma(dx, t2.z, t1.w));
            const float c1 = ffma(dy, t3.y, ffma(dx, t3.x, t2.x));
            const float c2 = ffma(dy, t3.w, ffma(dx, t3.z, t2.y));
            Co12 = fma2(aT2, pack2(c1, c2), Co12);
            if (before_median) { mCo0 = c0; mCo1 = c1; mCo2 = c2; }
          }
          if constexpr (DEPTH) {
            tt = t0.y + (t0.z * dx + t0.w * dy);
            if (before_median) mDepth = tt;
          }
          if constexpr (COORD || DEPTH) Co0D = fma2(aT2, pack2(c0, tt), Co0D);
          if constexpr (NORMAL) {
            N01 = fma2(aT2, pack2(t1.x, t1.y), N01);
            N2 = ffma(t1.z, aT, N2);
          }
          if (before_median) max_contributor = contributor;
        } else {
          blue = lds32<0>(tb + kk * 4);
        }
        C01 = fma2(aT2, pack2(bq.z, bq.w), C01);
        C2w = fma2(aT2, pack2(blue, 1.0f), C2w);  // C2 += aT * blue, weight += aT * 1
        T = test_T;
        last_contributor = contributor;
      }
      if (__all_sync(0xffffffffu, done != 0)) break;
    }
  }
  cp_async_wait_all();

  float C0, C1, C2, weight, Co0, Co1, Co2, Depth, N0, N1;
  unpack2(C01, C0, C1);
  unpack2(C2w, C2, weight);
  unpack2(Co0D, Co0, Depth);
  unpack2(Co12, Co1, Co2);
  unpack2(N01, N0, N1);
  if (inside) {
    p.n_contrib[(size_t)v * 2 * N + pix_id] = last_contributor;
    p.n_contrib[(size_t)v * 2 * N + N + pix_id] = max_contributor;
    float* oc = p.out_color + (size_t)v * 3 * N;
    oc[pix_id] = C0 + T * p.bg[0];
    oc[N + pix_id] = C1 + T * p.bg[1];
    oc[2 * N + pix_id] = C2 + T * p.bg[2];
    p.out_alpha[(size_t)v * N + pix_id] = weight;
    if constexpr (GEO) {
      const float pnx = (pixfx - p.W / 2.f) / p.focal_x, pny = (pixfy - p.H / 2.f) / p.focal_y;
      const float ln = sqrtf(pnx * pnx + pny * pny + 1.0f);
      if constexpr (COORD) {
        float* o = p.out_coord + (size_t)v * 3 * N;
        float* om = p.out_mcoord + (size_t)v * 3 * N;
        float* ac = p.accum_coord + (size_t)v * 3 * N;
        if (last_contributor) {
          o[pix_id] = Co0 / weight; o[N + pix_id] = Co1 / weight; o[2 * N + pix_id] = Co2 / weight;
        } else {
          o[pix_id] = 0; o[N + pix_id] = 0; o[2 * N + pix_id] = 0;
        }
        ac[pix_id] = Co0; ac[N + pix_id] = Co1; ac[2 * N + pix_id] = Co2;
        om[pix_id] = mCo0; om[N + pix_id] = mCo1; om[2 * N + pix_id] = mCo2;
      }
      if constexpr (DEPTH) {
        float depth_ln = Depth / ln;
        p.accum_depth[(size_t)v * N + pix_id] = depth_ln;
        p.out_depth[(size_t)v * N + pix_id] = last_contributor ? depth_ln / weight : 0.0f;
        p.out_mdepth[(size_t)v * N + pix_id] = mDepth / ln;
      }
      if constexpr (NORMAL) {
        float* on = p.out_normal + (size_t)v * 3 * N;
        if (last_contributor) {
          float len = sqrtf(N0 * N0 + N1 * N1 + N2 * N2);
          p.normal_length[(size_t)v * N + pix_id] = len;
          len = fmaxf(len, IGS_NORMALIZE_EPS);
          on[pix_id] = N0 / len; on[N + pix_id] = N1 / len; on[2 * N + pix_id] = N2 / len;
        } else {
          p.normal_length[(size_t)v * N + pix_id] = 1.0f;
          on[pix_id] = 0; on[N + pix_id] = 0; on[2 * N + pix_id] = 0;
        }
      }
    }
  }
}

void launch_blend_count(const BlendFwdParams& p, int V, unsigned long long* counts, cudaStream_t s) {
  int nblocks = V * p.tiles_x * p.tiles_y;
  if (nblocks <= 0) return;
  BlendFwdParams q = p;
  q.counters = counts;
  count_launch();
  blend_fwd_kernel<false, false, false, 4, true><<<nblocks, 256, 0, s>>>(q);
}

template <bool C, bool D, bool NRM, int MINB, int B>
static void launch_fwd2_one(const BlendFwdParams& p, int nblocks, cudaStream_t s) {
  constexpr int bytes = 2 * (B * 32 + ((C || D || NRM) ? B * 64 : B * 4));
  if (bytes > 40 * 1024)   // (static masks come on top) opt-in is per device, so it is (cheaply) set on every launch
    cudaFuncSetAttribute(blend_fwd2_kernel<C, D, NRM, MINB, B>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  blend_fwd2_kernel<C, D, NRM, MINB, B><<<nblocks, 256, bytes, s>>>(p);
}
template <int MINB, int B>
static void launch_fwd2_cfg(const BlendFwdParams& p, int nblocks, bool coord, bool depth, cudaStream_t s) {
  if (coord && depth) launch_fwd2_one<true, true, true, MINB, B>(p, nblocks, s);
  else if (coord && !depth) launch_fwd2_one<true, false, true, MINB, B>(p, nblocks, s);
  else if (!coord && depth) launch_fwd2_one<false, true, true, MINB, B>(p, nblocks, s);
  else launch_fwd2_one<false, false, false, MINB, B>(p, nblocks, s);
}
static void launch_fwd2(const BlendFwdParams& p, int nblocks, bool coord, bool depth, bool occ4, bool big, cudaStream_t s) {
  if (occ4) { if (big) launch_fwd2_cfg<4, 256>(p, nblocks, coord, depth, s); else launch_fwd2_cfg<4, 128>(p, nblocks, coord, depth, s); }
  else { if (big) launch_fwd2_cfg<3, 256>(p, nblocks, coord, depth, s); else launch_fwd2_cfg<3, 128>(p, nblocks, coord, depth, s); }
}

bool blend_fwd_writes_hitmask() {
  static const bool on = [] {
    const char* e = getenv("IGS_HITMASK");
    const char* i = getenv("IGS_FWD_IMPL");
    return !(e && atoi(e) == 0) && !(i && atoi(i) >= 2);   // the cp.async experiment kernel does not write them
  }();
  return on;
}

void launch_blend_fwd(const BlendFwdParams& p, int V, bool coord, bool depth, cudaStream_t s) {
  int nblocks = V * p.tiles_x * p.tiles_y;
  if (nblocks <= 0) return;
  count_launch();
  // 64-register build, 4 CTAs/SM (measured 10% faster than the spill-free 3 CTAs/SM build, IGS_FWD_OCC=3)
  static const int occ = [] { const char* e = getenv("IGS_FWD_OCC"); return e ? atoi(e) : 4; }();
  // IGS_FWD_IMPL=2 selects the cp.async double-buffered kernel (measured 5 % SLOWER at C3: 7.56 vs 7.20 ms);
  // default: the first-generation kernel
  static const int impl = [] { const char* e = getenv("IGS_FWD_IMPL"); return e ? atoi(e) : 1; }();
  if (impl >= 2) {
    static const int batch = [] { const char* e = getenv("IGS_FWD2_BATCH"); return e ? atoi(e) : IGS_FWD2_BATCH; }();
    launch_fwd2(p, nblocks, coord, depth, occ >= 4, batch >= 256, s);
    return;
  }
  if (occ >= 4) {
    if (coord && depth) blend_fwd_kernel<true, true, true, 4><<<nblocks, 256, 0, s>>>(p);
    else if (coord && !depth) blend_fwd_kernel<true, false, true, 4><<<nblocks, 256, 0, s>>>(p);
    else if (!coord && depth) blend_fwd_kernel<false, true, true, 4><<<nblocks, 256, 0, s>>>(p);
    else blend_fwd_kernel<false, false, false, 4><<<nblocks, 256, 0, s>>>(p);
  } else {
    if (coord && depth) blend_fwd_kernel<true, true, true, 3><<<nblocks, 256, 0, s>>>(p);
    else if (coord && !depth) blend_fwd_kernel<true, false, true, 3><<<nblocks, 256, 0, s>>>(p);
    else if (!coord && depth) blend_fwd_kernel<false, true, true, 3><<<nblocks, 256, 0, s>>>(p);
    else blend_fwd_kernel<false, false, false, 3><<<nblocks, 256, 0, s>>>(p);
  }
}

}  // namespace igs
