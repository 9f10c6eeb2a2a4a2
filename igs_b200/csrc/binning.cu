// binning.cu — instance generation, stable radix sort, tile ranges.
//
// Replaces reference K2-K5 (DGR/cuda_rasterizer/rasterizer_impl.cu:350-391: cub InclusiveSum,
// duplicateWithKeys :70-111, cub SortPairs on 64-bit (tile|depth) keys :376-381, identifyTileRanges
// :151-173).  The reference sorts R instances on ~45-bit keys (6 radix passes over 12-byte pairs).
// Here the same permutation is produced by a two-level stable sort:
//   A. sort the V*P Gaussians by (view, depth bits)               — 4-5 passes over V*P pairs
//   B. emit instances in that order (row-major tiles per Gaussian, exactly the reference's emission
//      order inside one Gaussian) and stable-sort them by global tile id only — 2-3 passes over R
// A stable sort by tile of a depth-ordered list equals a stable sort by (tile, depth) with ties in
// ascending Gaussian index, which is what CUB's stable LSD sort yields for the reference
// (SURVEY T8/T9); tests compare the decoded 64-bit keys / point_list / ranges bit-for-bit.
#include "common.cuh"
#include "kernels.cuh"
#include <cooperative_groups.h>
#include <cstdlib>

namespace igs {

// ---------------------------------------------------------------------------------------------
// Exclusive scan (u32), two launches: block sums, then scan-with-lookback-by-summation.
// ---------------------------------------------------------------------------------------------
#define SCAN_THREADS 256
#define SCAN_ITEMS 8
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)  // 2048, must match scan_blocks()

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane_id() >= (uint32_t)o) v += t;
  }
  return v;
}

__device__ __forceinline__ uint32_t block_reduce_sum(uint32_t v, uint32_t* smem /*>=8*/) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane_id() == 0) smem[threadIdx.x >> 5] = v;
  __syncthreads();
  uint32_t t = 0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; w++) t += smem[w];
  return t;
}

__device__ __forceinline__ void scan_block_sums_body(const uint32_t* __restrict__ in, size_t n, uint32_t* __restrict__ sums,
                                                     unsigned blk, uint32_t* sm /*>= 8 words*/) {
  size_t base = (size_t)blk * SCAN_TILE;
  uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) {
    size_t i = base + (size_t)k * SCAN_THREADS + threadIdx.x;
    if (i < n) s += in[i];
  }
  s = block_reduce_sum(s, sm);
  if (threadIdx.x == 0) sums[blk] = s;
}

__global__ void __launch_bounds__(SCAN_THREADS)
scan_block_sums_kernel(const uint32_t* __restrict__ in, size_t n, uint32_t* __restrict__ sums) {
  __shared__ uint32_t sm[8];
  scan_block_sums_body(in, n, sums, blockIdx.x, sm);
}

__device__ __forceinline__ void scan_apply_body(const uint32_t* in, size_t n, const uint32_t* __restrict__ sums, uint32_t* out,
                                                uint32_t* __restrict__ total, unsigned blk, unsigned nblk, uint32_t* sm /*>= 8*/,
                                                uint32_t* wsum /*>= 8*/) {
  // base = sum of all previous block sums
  uint32_t b = 0;
  for (uint32_t j = threadIdx.x; j < blk; j += SCAN_THREADS) b += sums[j];
  b = block_reduce_sum(b, sm);

  size_t base = (size_t)blk * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS];
  uint32_t tsum = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) {
    size_t i = base + k;
    v[k] = (i < n) ? in[i] : 0u;
    tsum += v[k];
  }
  uint32_t incl = warp_incl_scan(tsum);
  __syncthreads();   // wsum may still be read by the previous tile of a persistent caller
  if (lane_id() == 31) wsum[threadIdx.x >> 5] = incl;
  __syncthreads();
  uint32_t woff = 0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; w++)
    if (w < (int)(threadIdx.x >> 5)) woff += wsum[w];
  uint32_t run = b + woff + incl - tsum;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) {
    size_t i = base + k;
    if (i < n) out[i] = run;
    run += v[k];
  }
  if (total != nullptr && blk == nblk - 1 && threadIdx.x == SCAN_THREADS - 1) *total = run;
}

// out may alias in.  If total != nullptr the grand total is written by the last block.
__global__ void __launch_bounds__(SCAN_THREADS)
scan_apply_kernel(const uint32_t* in, size_t n, const uint32_t* __restrict__ sums, uint32_t* out,
                  uint32_t* __restrict__ total) {
  __shared__ uint32_t sm[8];
  __shared__ uint32_t wsum[SCAN_THREADS / 32];
  scan_apply_body(in, n, sums, out, total, blockIdx.x, gridDim.x, sm, wsum);
}

void launch_exclusive_scan(const uint32_t* in, uint32_t* out, size_t n, uint32_t* block_sums,
                           uint32_t* total, cudaStream_t s) {
  if (n == 0) return;
  unsigned nb = (unsigned)scan_blocks(n);
  count_launch(2);
  scan_block_sums_kernel<<<nb, SCAN_THREADS, 0, s>>>(in, n, block_sums);
  scan_apply_kernel<<<nb, SCAN_THREADS, 0, s>>>(in, n, block_sums, out, total);
}

// ---------------------------------------------------------------------------------------------
// Stable LSD radix sort, 8-bit digits, 256 threads x 16 keys per block.
// ---------------------------------------------------------------------------------------------
// Effective element count of a launch that was sized for a CAPACITY: when n_dev is given, the true count lives in device
// memory (the instance total written by the scan, never read back by the host in capacity mode) and is clamped to n.
__device__ __forceinline__ size_t effective_n(size_t n, const uint32_t* __restrict__ n_dev) {
  if (n_dev == nullptr) return n;
  const size_t t = (size_t)*n_dev;
  return t < n ? t : n;
}

template <typename K>
__device__ __forceinline__ uint32_t digit_of(K key, int shift, uint32_t mask) {
  return (uint32_t)(key >> shift) & mask;
}

template <typename K, int BITS>
__device__ __forceinline__ void radix_hist_body(const K* __restrict__ keys, size_t n, int shift, uint32_t mask,
                                                uint32_t* __restrict__ hist, unsigned nblocks, unsigned tile, uint32_t* h) {
  constexpr int NB = 1 << BITS;
  __syncthreads();   // h may still be read by the previous tile of a persistent caller
  for (int d = threadIdx.x; d < NB; d += IGS_SORT_THREADS) h[d] = 0;
  __syncthreads();
  size_t base = (size_t)tile * IGS_SORT_TILE;
#pragma unroll
  for (int k = 0; k < IGS_SORT_ITEMS; k++) {
    size_t i = base + (size_t)k * IGS_SORT_THREADS + threadIdx.x;
    if (i < n) atomicAdd(&h[digit_of<K>(keys[i], shift, mask)], 1u);
  }
  __syncthreads();
  for (int d = threadIdx.x; d < NB; d += IGS_SORT_THREADS) hist[(size_t)d * nblocks + tile] = h[d];
}

template <typename K, int BITS>
__global__ void __launch_bounds__(IGS_SORT_THREADS)
radix_hist_kernel(const K* __restrict__ keys, size_t n, const uint32_t* __restrict__ n_dev, int shift, uint32_t mask,
                  uint32_t* __restrict__ hist, unsigned nblocks) {
  __shared__ uint32_t h[1 << BITS];
  radix_hist_body<K, BITS>(keys, effective_n(n, n_dev), shift, mask, hist, nblocks, blockIdx.x, h);
}

// Each warp ranks a contiguous chunk of 512 keys in 16 rounds of 32 consecutive keys, so that the rank order equals
// the index order (stability).  The ranked keys are then REORDERED IN SHARED MEMORY by digit and written out in index
// order of that staging buffer, so every warp store covers contiguous runs of the destination: scattering 4-byte keys
// straight from the ranking loop costs one L2 sector transaction per key and value (33 M instances x 2 per pass at C3
// was the bound of the pass, ncu: 1.3 TB/s of algorithmic traffic), the staged write-out one per 32 bytes.
template <typename K, int BITS>
struct ScatterSmem {
  static constexpr int NB = 1 << BITS;
  static constexpr int WARPS = IGS_SORT_THREADS / 32;
  static constexpr size_t whist_bytes = (size_t)WARPS * (NB + 1) * 4;
  static constexpr size_t dig_bytes = (size_t)NB * 4 * 2;
  static constexpr size_t key_off = (whist_bytes + dig_bytes + 64 + 15) & ~size_t(15);   // + wsum[8] (+ pad)
  static constexpr size_t val_off = key_off + (size_t)IGS_SORT_TILE * sizeof(K);
  static constexpr size_t bytes = val_off + (size_t)IGS_SORT_TILE * 4;
};

template <typename K, int BITS>
__device__ __forceinline__ void radix_scatter_body(const K* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                   K* __restrict__ keys_out, uint32_t* __restrict__ vals_out, size_t n, int shift,
                                                   uint32_t mask, const uint32_t* __restrict__ hist_scanned, unsigned nblocks,
                                                   unsigned tile, unsigned char* sort_smem) {
  using L = ScatterSmem<K, BITS>;
  constexpr int NB = 1 << BITS;
  constexpr int WARPS = IGS_SORT_THREADS / 32;
  constexpr int CHUNK = IGS_SORT_TILE / WARPS;      // 512 keys per warp
  constexpr int ROUNDS = CHUNK / 32;                // 16
  constexpr int PER = NB / IGS_SORT_THREADS;        // digits per thread in the block-level scan (1 or 2)
  static_assert(NB % IGS_SORT_THREADS == 0, "digit count must be a multiple of the block size");
  uint32_t (*whist)[NB + 1] = reinterpret_cast<uint32_t (*)[NB + 1]>(sort_smem);
  uint32_t* dig_local = reinterpret_cast<uint32_t*>(sort_smem + L::whist_bytes);   // first staged slot of each digit
  uint32_t* dig_global = dig_local + NB;                                             // destination of slot 0 of it
  uint32_t* wsum = dig_global + NB;                                                  // [WARPS]
  K* s_key = reinterpret_cast<K*>(sort_smem + L::key_off);
  uint32_t* s_val = reinterpret_cast<uint32_t*>(sort_smem + L::val_off);
  const int w = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  __syncthreads();   // the staging buffers may still be read by the previous tile of a persistent caller
  for (int i = threadIdx.x; i < WARPS * (NB + 1); i += IGS_SORT_THREADS) (&whist[0][0])[i] = 0;

  const size_t bbase = (size_t)tile * IGS_SORT_TILE;
  const size_t wbase = bbase + (size_t)w * CHUNK;
  K key[ROUNDS];
  uint32_t rnk2[ROUNDS / 2];
#pragma unroll
  for (int r = 0; r < ROUNDS; r++) {  // all loads first: 16 independent requests in flight
    const size_t i = wbase + (size_t)r * 32 + lane;
    key[r] = i < n ? keys_in[i] : K(0);
  }
  __syncthreads();
  uint32_t* const myhist = whist[w];
  const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < ROUNDS; r++) {
    const bool valid = wbase + (size_t)r * 32 + lane < n;
    const uint32_t d = valid ? digit_of<K>(key[r], shift, mask) : (uint32_t)NB;
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    const int leader = __ffs(peers) - 1;
    uint32_t old = 0;
    if ((int)lane == leader) {
      old = myhist[d];
      myhist[d] = old + __popc(peers);
    }
    old = __shfl_sync(0xffffffffu, old, leader);
    const uint32_t rk = old + __popc(peers & lt_mask);
    if (r & 1) rnk2[r >> 1] |= rk << 16; else rnk2[r >> 1] = rk;
    __syncwarp();
  }
  __syncthreads();
  // per digit: exclusive scan over the warps (in place) and the block's count; then an exclusive scan of the counts
  // over the digits gives each digit's first staged slot
  uint32_t cnt[PER], tsum = 0;
#pragma unroll
  for (int q = 0; q < PER; q++) {
    const uint32_t d = threadIdx.x * PER + q;
    uint32_t run = 0;
#pragma unroll
    for (int ww = 0; ww < WARPS; ww++) {
      const uint32_t c = whist[ww][d];
      whist[ww][d] = run;
      run += c;
    }
    cnt[q] = run;
    tsum += run;
  }
  const uint32_t incl = warp_incl_scan(tsum);
  if (lane == 31) wsum[w] = incl;
  __syncthreads();
  uint32_t start = incl - tsum;
#pragma unroll
  for (int ww = 0; ww < WARPS; ww++)
    if (ww < w) start += wsum[ww];
#pragma unroll
  for (int q = 0; q < PER; q++) {
    const uint32_t d = threadIdx.x * PER + q;
    dig_local[d] = start;
    dig_global[d] = hist_scanned[(size_t)d * nblocks + tile] - start;   // modular: slot j of digit d -> dig_global[d] + j
    start += cnt[q];
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < ROUNDS; r++) {
    const size_t i = wbase + (size_t)r * 32 + lane;
    if (i < n) {
      const uint32_t rk = (r & 1) ? (rnk2[r >> 1] >> 16) : (rnk2[r >> 1] & 0xffffu);
      const uint32_t d = digit_of<K>(key[r], shift, mask);
      const uint32_t slot = dig_local[d] + myhist[d] + rk;
      s_key[slot] = key[r];
      s_val[slot] = vals_in[i];
    }
  }
  __syncthreads();
  // (a tile that lies entirely past the element count exists only in capacity mode: it moves nothing)
  const uint32_t nvalid = bbase >= n ? 0u : (uint32_t)(n - bbase < (size_t)IGS_SORT_TILE ? n - bbase : (size_t)IGS_SORT_TILE);
  for (uint32_t j = threadIdx.x; j < nvalid; j += IGS_SORT_THREADS) {
    const K k = s_key[j];
    const uint32_t pos = dig_global[digit_of<K>(k, shift, mask)] + j;
    keys_out[pos] = k;
    vals_out[pos] = s_val[j];
  }
}

template <typename K, int BITS>
__global__ void __launch_bounds__(IGS_SORT_THREADS, sizeof(K) == 4 ? 4 : 3)
radix_scatter_kernel(const K* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, K* __restrict__ keys_out,
                     uint32_t* __restrict__ vals_out, size_t n, const uint32_t* __restrict__ n_dev, int shift, uint32_t mask,
                     const uint32_t* __restrict__ hist_scanned, unsigned nblocks) {
  extern __shared__ __align__(16) unsigned char sort_smem[];
  radix_scatter_body<K, BITS>(keys_in, vals_in, keys_out, vals_out, effective_n(n, n_dev), shift, mask, hist_scanned, nblocks,
                              blockIdx.x, sort_smem);
}

// The whole sort in ONE cooperative launch: persistent CTAs loop over the tiles of each phase (histogram, the two scan
// phases, scatter) with a grid barrier between phases, for every digit pass.  A multi-launch sort is 4 launches per
// pass — 24 launches per rasterizer forward for the two sorts — and at small sizes (one 512x512 or 1280x720 view, or
// the 2.5 views a rank renders at N = 8) those launches, not the kernels, were the cost of binning.
template <typename K, int BITS>
__global__ void __launch_bounds__(IGS_SORT_THREADS, sizeof(K) == 4 ? 4 : 3)
radix_sort_coop_kernel(K* k0, K* k1, uint32_t* v0, uint32_t* v1, size_t n_cap, const uint32_t* n_dev, int end_bit,
                       uint32_t* hist, unsigned ntiles) {
  constexpr int NB = 1 << BITS;
  const size_t n = effective_n(n_cap, n_dev);
  extern __shared__ __align__(16) unsigned char sort_smem[];
  uint32_t* sm32 = reinterpret_cast<uint32_t*>(sort_smem);
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  const size_t entries = (size_t)NB * ntiles;
  const unsigned nchunks = (unsigned)scan_blocks(entries);
  uint32_t* sums = hist + entries;
  K* kin = k0; K* kout = k1;
  uint32_t* vin = v0; uint32_t* vout = v1;
  for (int shift = 0; shift < end_bit; shift += BITS) {
    const int bits = end_bit - shift < BITS ? end_bit - shift : BITS;
    const uint32_t mask = (1u << bits) - 1u;
    for (unsigned t = blockIdx.x; t < ntiles; t += gridDim.x) radix_hist_body<K, BITS>(kin, n, shift, mask, hist, ntiles, t, sm32);
    grid.sync();
    for (unsigned c = blockIdx.x; c < nchunks; c += gridDim.x) scan_block_sums_body(hist, entries, sums, c, sm32);
    grid.sync();
    for (unsigned c = blockIdx.x; c < nchunks; c += gridDim.x)
      scan_apply_body(hist, entries, sums, hist, nullptr, c, nchunks, sm32, sm32 + 8);
    grid.sync();
    for (unsigned t = blockIdx.x; t < ntiles; t += gridDim.x)
      radix_scatter_body<K, BITS>(kin, vin, kout, vout, n, shift, mask, hist, ntiles, t, sort_smem);
    grid.sync();
    K* tk = kin; kin = kout; kout = tk;
    uint32_t* tv = vin; vin = vout; vout = tv;
  }
}

template <typename K, int BITS>
static int coop_capacity() {   // co-resident CTAs of the cooperative sort kernel on the current device (0: unsupported)
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (dev == cached_dev) return cached;
  int coop = 0, sms = 0, per_sm = 0;
  cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaFuncSetAttribute(radix_sort_coop_kernel<K, BITS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       (int)ScatterSmem<K, BITS>::bytes);
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, radix_sort_coop_kernel<K, BITS>, IGS_SORT_THREADS,
                                                    ScatterSmem<K, BITS>::bytes) != cudaSuccess) per_sm = 0;
  (void)cudaGetLastError();
  cached_dev = dev;
  cached = coop ? per_sm * sms : 0;
  return cached;
}

template <typename K, int BITS>
static int radix_sort_passes(K* keys[2], uint32_t* vals[2], size_t n, const uint32_t* n_dev, int end_bit, uint32_t* hist,
                             cudaStream_t s) {
  constexpr int NB = 1 << BITS;
  unsigned nb = (unsigned)sort_blocks(n);
  const int passes = (end_bit + BITS - 1) / BITS;
  static const bool coop_off = [] { const char* e = getenv("IGS_SORT_COOP"); return e && atoi(e) == 0; }();
  const int cap = coop_off ? 0 : coop_capacity<K, BITS>();
  // Only when every tile gets its own resident CTA: with more tiles than that the persistent loop measured slower
  // than separate launches (C3, 20 views: depth sort 0.40 -> 0.47 ms, tile sort 0.70 -> 0.88 ms), and at that size the
  // launches are not the cost.
  if (cap > 0 && passes > 0 && nb <= (unsigned)cap) {
    unsigned grid = nb;
    K* k0 = keys[0]; K* k1 = keys[1];
    uint32_t* v0 = vals[0]; uint32_t* v1 = vals[1];
    void* args[] = {&k0, &k1, &v0, &v1, &n, &n_dev, &end_bit, &hist, &nb};
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)radix_sort_coop_kernel<K, BITS>, dim3(grid), dim3(IGS_SORT_THREADS),
                                                args, ScatterSmem<K, BITS>::bytes, s);
    if (e == cudaSuccess) {
      count_launch(1);
      return passes & 1;
    }
    (void)cudaGetLastError();   // fall through to the multi-launch sort
  }
  int cur = 0;
  // > 48 KB of dynamic shared memory needs the opt-in; it is per device, so it is (cheaply) set on every sort
  cudaFuncSetAttribute(radix_scatter_kernel<K, BITS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       (int)ScatterSmem<K, BITS>::bytes);
  for (int shift = 0; shift < end_bit; shift += BITS) {
    int bits = end_bit - shift < BITS ? end_bit - shift : BITS;
    uint32_t mask = (1u << bits) - 1u;
    count_launch(2);
    radix_hist_kernel<K, BITS><<<nb, IGS_SORT_THREADS, 0, s>>>(keys[cur], n, n_dev, shift, mask, hist, nb);
    // in-place exclusive scan over the bin-major histogram (its block sums live behind the histogram)
    launch_exclusive_scan(hist, hist, (size_t)NB * nb, hist + (size_t)NB * nb, nullptr, s);
    radix_scatter_kernel<K, BITS><<<nb, IGS_SORT_THREADS, ScatterSmem<K, BITS>::bytes, s>>>(
        keys[cur], vals[cur], keys[cur ^ 1], vals[cur ^ 1], n, n_dev, shift, mask, hist, nb);
    cur ^= 1;
  }
  return cur;
}

// Sorts n (key,value) pairs on key bits [0, end_bit).  Buffers ping-pong; returns the index (0/1)
// of the buffer pair that holds the sorted result.  9-bit digits are used when they save a pass
// (e.g. 17 tile-id bits at C3: 2 passes instead of 3).  n_dev (optional): the launches are sized for n, the number of
// pairs actually present is min(*n_dev, n), read on the device (capacity mode: no host read-back of the count).
template <typename K>
int radix_sort_pairs(K* keys[2], uint32_t* vals[2], size_t n, int end_bit, uint32_t* hist,
                     const uint32_t* n_dev, cudaStream_t s) {
  if (n == 0) return 0;
  const int p8 = (end_bit + 7) / 8, p9 = (end_bit + 8) / 9;
  if (p9 < p8) return radix_sort_passes<K, 9>(keys, vals, n, n_dev, end_bit, hist, s);
  return radix_sort_passes<K, 8>(keys, vals, n, n_dev, end_bit, hist, s);
}

template int radix_sort_pairs<uint64_t>(uint64_t* keys[2], uint32_t* vals[2], size_t, int, uint32_t*, const uint32_t*, cudaStream_t);
template int radix_sort_pairs<uint32_t>(uint32_t* keys[2], uint32_t* vals[2], size_t, int, uint32_t*, const uint32_t*, cudaStream_t);

// ---------------------------------------------------------------------------------------------
// After sort A: tiles in depth order + inverse permutation
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
gather_rank_kernel(size_t VP, const uint32_t* __restrict__ sorted_gid, const uint32_t* __restrict__ tiles,
                   uint32_t* __restrict__ tiles_sorted /*[VP+1]*/, uint32_t* __restrict__ rank) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i > VP) return;
  if (i == VP) { tiles_sorted[i] = 0; return; }
  uint32_t gid = sorted_gid[i];
  tiles_sorted[i] = tiles[gid];
  rank[gid] = (uint32_t)i;
}

void launch_gather_rank(size_t VP, const uint32_t* sorted_gid, const uint32_t* tiles, uint32_t* tiles_sorted,
                        uint32_t* rank, cudaStream_t s) {
  unsigned nb = (unsigned)((VP + 1 + 255) / 256);
  count_launch();
  gather_rank_kernel<<<nb, 256, 0, s>>>(VP, sorted_gid, tiles, tiles_sorted, rank);
}

// ---------------------------------------------------------------------------------------------
// Instance emission (reference duplicateWithKeys, rasterizer_impl.cu:70-111, getRect auxiliary.h:62-72)
// One thread per depth-sorted Gaussian; emission slot e = inst_off[i] + row-major tile index.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void tile_rect(float px, float py, int radius, int tiles_x, int row0, int row1, int& x0,
                                          int& y0, int& x1, int& y1) {
  float rf = (float)radius;
  x0 = min(tiles_x, max(0, (int)fmul(fsub(px, rf), 0.0625f)));
  y0 = min(row1, max(row0, (int)fmul(fsub(py, rf), 0.0625f)));
  x1 = min(tiles_x, max(0, (int)fmul(fadd(fadd(fadd(px, rf), 16.0f), -1.0f), 0.0625f)));
  y1 = min(row1, max(row0, (int)fmul(fadd(fadd(fadd(py, rf), 16.0f), -1.0f), 0.0625f)));
}

// One WARP per 32 consecutive depth-sorted Gaussians.  The warp's instances form one contiguous range of emission
// slots; lanes walk that range slot by slot (fully coalesced 128-byte stores to the three arrays), finding each slot's
// Gaussian with a 5-step binary search over the 32 lane-held offsets.  (One thread per Gaussian writing its own run of
// slots — the reference's duplicateWithKeys loop — makes every store instruction touch up to 32 sectors.)
__global__ void __launch_bounds__(256)
emit_instances_kernel(size_t VP, uint32_t P, int tiles_x, int tiles_y, const Camera* __restrict__ cams,
                      const uint32_t* __restrict__ sorted_gid,
                      const uint32_t* __restrict__ inst_off, const SplatHead* __restrict__ head,
                      const int* __restrict__ radii, uint32_t* __restrict__ tkey, uint32_t* __restrict__ tval,
                      uint32_t* __restrict__ inst_gid, uint32_t capacity) {
  const size_t wid = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t lane = lane_id();
  const size_t i0 = wid * 32;
  if (i0 >= VP) return;
  const size_t i = i0 + lane;
  const bool live = i < VP;
  // lanes past the end take the end offset of the last Gaussian, so they own nothing
  const size_t il = live ? i : VP;
  const uint32_t off = inst_off[il];
  const uint32_t n = live ? inst_off[i + 1] - off : 0u;
  uint32_t gid = 0, xy = 0, width = 1, tbase = 0;
  if (n != 0) {
    gid = sorted_gid[i];
    const uint32_t v = gid / P;
    const float4 h = head[gid].a;
    int x0, y0, x1, y1;
    tile_rect(h.x, h.y, radii[gid], tiles_x, cams[v].row0, cams[v].row1, x0, y0, x1, y1);
    xy = (uint32_t)x0 | ((uint32_t)y0 << 16);
    width = (uint32_t)(x1 - x0);
    tbase = v * (uint32_t)(tiles_x * tiles_y);
  }
  const uint32_t begin = __shfl_sync(0xffffffffu, off, 0);
  const uint32_t end = __shfl_sync(0xffffffffu, off + n, 31);
  for (uint32_t e0 = begin; e0 < end; e0 += 32) {     // warp-uniform trip count
    const uint32_t e = e0 + lane;
    // owner = last lane whose offset is <= e (Gaussians without instances share the offset of their successor)
    uint32_t lo = 0;
#pragma unroll
    for (int step = 16; step > 0; step >>= 1) {
      const uint32_t oc = __shfl_sync(0xffffffffu, off, lo + step);
      if (oc <= e) lo += step;
    }
    const uint32_t o_off = __shfl_sync(0xffffffffu, off, lo);
    const uint32_t o_gid = __shfl_sync(0xffffffffu, gid, lo);
    const uint32_t o_xy = __shfl_sync(0xffffffffu, xy, lo);
    const uint32_t o_w = __shfl_sync(0xffffffffu, width, lo);
    const uint32_t o_tb = __shfl_sync(0xffffffffu, tbase, lo);
    if (e < end && e < capacity) {   // capacity mode: instances past the binning buffer are dropped (overflow is flagged)
      const uint32_t local = e - o_off;
      const uint32_t ry = local / o_w, rx = local - ry * o_w;
      tkey[e] = o_tb + ((o_xy >> 16) + ry) * (uint32_t)tiles_x + (o_xy & 0xffffu) + rx;
      tval[e] = e;
      inst_gid[e] = o_gid;
    }
  }
}

void launch_emit_instances(size_t VP, uint32_t P, int tiles_x, int tiles_y, const Camera* cams,
                           const uint32_t* sorted_gid,
                           const uint32_t* inst_off, const SplatHead* head, const int* radii, uint32_t* tkey,
                           uint32_t* tval, uint32_t* inst_gid, uint32_t capacity, cudaStream_t s) {
  if (VP == 0) return;
  unsigned nb = (unsigned)((((VP + 31) / 32) * 32 + 255) / 256);   // one warp per 32 Gaussians, 8 warps per block
  count_launch();
  emit_instances_kernel<<<nb, 256, 0, s>>>(VP, P, tiles_x, tiles_y, cams, sorted_gid, inst_off, head, radii, tkey, tval,
                                           inst_gid, capacity);
}

// ---------------------------------------------------------------------------------------------
// After sort B: tile ranges (reference identifyTileRanges) + sorted Gaussian-id list
// ---------------------------------------------------------------------------------------------
// Four consecutive instances per thread: 16-byte loads / stores of the sorted arrays and four independent gathers of
// the Gaussian id in flight (one element per thread left the kernel latency bound on the dependent gather).
__device__ __forceinline__ void mark_boundary(uint2* __restrict__ ranges, uint32_t tp, uint32_t t, uint32_t j) {
  if (t != tp) {
    ranges[tp].y = j;
    ranges[t].x = j;
  }
}

__global__ void __launch_bounds__(256)
finalize_bins_kernel(uint32_t R_cap, const uint32_t* __restrict__ n_dev, const uint32_t* __restrict__ tkey,
                     const uint32_t* __restrict__ point_e, const uint32_t* __restrict__ inst_gid,
                     uint32_t* __restrict__ point_list, uint2* __restrict__ ranges) {
  const uint32_t R = (uint32_t)effective_n(R_cap, n_dev);
  const uint32_t j0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4u;
  if (j0 >= R) return;
  if (j0 + 4 <= R) {
    const uint4 t = *reinterpret_cast<const uint4*>(tkey + j0);
    const uint4 e = *reinterpret_cast<const uint4*>(point_e + j0);
    const uint4 g = make_uint4(inst_gid[e.x], inst_gid[e.y], inst_gid[e.z], inst_gid[e.w]);
    if (j0 == 0) ranges[t.x].x = 0;
    else mark_boundary(ranges, tkey[j0 - 1], t.x, j0);
    mark_boundary(ranges, t.x, t.y, j0 + 1);
    mark_boundary(ranges, t.y, t.z, j0 + 2);
    mark_boundary(ranges, t.z, t.w, j0 + 3);
    if (j0 + 4 == R) ranges[t.w].y = R;
    *reinterpret_cast<uint4*>(point_list + j0) = g;
  } else {
    for (uint32_t j = j0; j < R; j++) {
      const uint32_t t = tkey[j];
      if (j == 0) ranges[t].x = 0;
      else mark_boundary(ranges, tkey[j - 1], t, j);
      if (j == R - 1) ranges[t].y = R;
      point_list[j] = inst_gid[point_e[j]];
    }
  }
}

void launch_finalize_bins(uint32_t R, const uint32_t* n_dev, const uint32_t* tkey, const uint32_t* point_e,
                          const uint32_t* inst_gid, uint32_t* point_list, uint2* ranges, cudaStream_t s) {
  if (R == 0) return;
  count_launch();
  const uint32_t threads = (R + 3) / 4;
  finalize_bins_kernel<<<(threads + 255) / 256, 256, 0, s>>>(R, n_dev, tkey, point_e, inst_gid, point_list, ranges);
}

// Decode helper for tests: 64-bit reference-style keys in sorted order.
__global__ void decode_keys_kernel(uint32_t R, uint32_t tiles_per_view, uint32_t P, const uint32_t* __restrict__ tkey,
                                   const uint32_t* __restrict__ point_list, const float* __restrict__ depth,
                                   uint64_t* __restrict__ keys, uint32_t* __restrict__ plist_local) {
  uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= R) return;
  uint32_t gid = point_list[j];
  if (keys) keys[j] = ((uint64_t)(tkey[j] % tiles_per_view) << 32) | __float_as_uint(depth[gid]);
  if (plist_local) plist_local[j] = gid % P;
}

void launch_decode_keys(uint32_t R, uint32_t tiles_per_view, uint32_t P, const uint32_t* tkey,
                        const uint32_t* point_list, const float* depth, uint64_t* keys, uint32_t* plist_local,
                        cudaStream_t s) {
  if (R == 0) return;
  count_launch();
  decode_keys_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, tiles_per_view, P, tkey, point_list, depth, keys, plist_local);
}

}  // namespace igs
