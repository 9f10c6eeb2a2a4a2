// common.cuh — shared definitions for the B200-native splat rasterizer kernels.
//
// Scratch layout (private, self-describing from (P, V, R, W*H)); replaces the reference's
// GeometryState / BinningState / ImageState (DGR/cuda_rasterizer/rasterizer_impl.h:29-86,
// carve order rasterizer_impl.cu:190-250).  Differences by design:
//   * per-(view,Gaussian) data is packed into a 32-B "head" (what the alpha test needs) and a
//     64-B "tail" (what only contributing pixels need) instead of 9 strided arrays;
//   * view-independent work (cov3D, 3x3 eigen-decomposition) lives in a per-Gaussian block and is
//     shared by all V views of a batched call;
//   * binning is a two-level stable sort (depth per view, then tile) instead of one 64-bit sort.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#define IGS_TILE_X 16
#define IGS_TILE_Y 16
#define IGS_TILE_PIX 256
#define IGS_NEAR_Z 0.2f            // reference auxiliary.h:170
#define IGS_NORMALIZE_EPS 1.0E-12F // reference auxiliary.h:23
#define IGS_MAGIC 0x49475342u      // "IGSB"

namespace igs {

// ---------------------------------------------------------------------------------------------
// Record formats
// ---------------------------------------------------------------------------------------------
// head: {x, y, conic.x, conic.y} {conic.z, opacity*coef, r, g}
// tail: {b, ts, ray_plane.x, ray_plane.y} {n.x, n.y, n.z, vp.x} {vp.y, vp.z, cp0, cp1} {cp2..cp5}
struct __align__(16) SplatHead { float4 a, b; };
struct __align__(16) SplatTail { float4 a, b, c, d; };

// Per-Gaussian, view-independent block (forward.cu:135-155 / backward.cu:221-241 hoisted).
struct __align__(16) GaussEig {
  float vinv[9];   // Vrk_inv, glm column-major [col][row]
  float evec[9];   // eigenvectors, glm column-major: column i = i-th eigenvector
  float eval[3];
  uint32_t info;   // bit0: D != 0 (solver converged), bit1: well_conditioned, bits 4..5: min_id
  float pad[2];
};
static_assert(sizeof(GaussEig) == 96, "GaussEig size");

struct Camera {
  float view[16];
  float proj[16];
  float campos[3];
  int row0, row1;  // tile-row window [row0, row1) rendered for this view (whole image: 0, tiles_y)
};

struct Header {
  uint32_t magic;
  uint32_t P, V, N;       // N = W*H
  uint32_t R;             // total instances (filled after the scan)
  uint32_t tiles_x, tiles_y;
  uint32_t flags;         // bit0 coord, bit1 depth, bit2 normal, bit3 has_scales, bit4 has_shs, bit8 / bit16 which ping-pong
                          // buffer the depth / tile sort finished in, bit17 block-hit bytes written by the forward blend
  uint32_t pad[8];
};
static_assert(sizeof(Header) == 64, "Header size");

// ---------------------------------------------------------------------------------------------
// Bump carving (256-B aligned sub-arrays)
// ---------------------------------------------------------------------------------------------
struct Carver {
  char* base;
  size_t off;
  __host__ __device__ explicit Carver(char* b) : base(b), off(0) {}
  template <typename T>
  __host__ __device__ T* take(size_t count) {
    off = (off + 255) & ~size_t(255);
    T* p = reinterpret_cast<T*>(base + off);
    off += count * sizeof(T);
    return p;
  }
};

struct GeomBuf {
  Header* hdr;
  Camera* cams;            // [V]
  float* cov3D;            // [P][6]
  GaussEig* eig;           // [P]
  SplatHead* head;         // [V*P]
  SplatTail* tail;         // [V*P]
  float* depth;            // [V*P]
  int* radii;              // [V*P]
  uint32_t* tiles;         // [V*P] tiles touched
  uint8_t* clamped;        // [V*P] bit c: SH colour channel c clamped at 0
  uint64_t* skey[2];       // [V*P] (view<<31 | 31 depth bits), ping-pong
  uint32_t* sval[2];       // [V*P] global Gaussian id (v*P+g), ping-pong
  uint32_t* rank;          // [V*P] gid -> position in depth-sorted order
  uint32_t* inst_off;      // [V*P+1] exclusive scan of tiles in depth-sorted order
  uint32_t* scan_tmp;      // block sums for the scan
  uint32_t* hist;          // radix histograms for the depth sort
  size_t bytes;
};

__host__ __device__ inline size_t scan_blocks(size_t n) { return (n + 2047) / 2048; }
// radix sort: 256 threads x 16 keys per block
#define IGS_SORT_ITEMS 16
#define IGS_SORT_THREADS 256
#define IGS_SORT_TILE (IGS_SORT_ITEMS * IGS_SORT_THREADS)
__host__ __device__ inline size_t sort_blocks(size_t n) { return (n + IGS_SORT_TILE - 1) / IGS_SORT_TILE; }
// u32 words needed for a [512][nblocks] radix histogram (9-bit digits) plus the block sums of its scan
__host__ __device__ inline size_t radix_hist_words(size_t n) {
  size_t nb = sort_blocks(n ? n : 1);
  return 512 * nb + scan_blocks(512 * nb) + 8;
}

__host__ __device__ inline GeomBuf carve_geom(char* base, size_t P, size_t V) {
  GeomBuf g;
  Carver c(base);
  size_t VP = V * P;
  g.hdr = c.take<Header>(1);
  g.cams = c.take<Camera>(V);
  g.cov3D = c.take<float>(P * 6);
  g.eig = c.take<GaussEig>(P);
  g.head = c.take<SplatHead>(VP);
  g.tail = c.take<SplatTail>(VP);
  g.depth = c.take<float>(VP);
  g.radii = c.take<int>(VP);
  g.tiles = c.take<uint32_t>(VP);
  g.clamped = c.take<uint8_t>(VP);
  g.skey[0] = c.take<uint64_t>(VP);
  g.skey[1] = c.take<uint64_t>(VP);
  g.sval[0] = c.take<uint32_t>(VP);
  g.sval[1] = c.take<uint32_t>(VP);
  g.rank = c.take<uint32_t>(VP);
  g.inst_off = c.take<uint32_t>(VP + 1);
  g.scan_tmp = c.take<uint32_t>(scan_blocks(VP + 1) + 1);
  g.hist = c.take<uint32_t>(radix_hist_words(VP));
  g.bytes = c.off + 256;
  return g;
}

// 16-byte tag at the start of the binning / image buffers so that backward can reject a buffer that does not
// belong to this call (wrong order, wrong size) instead of reading through it.
struct BufTag { uint32_t magic, kind, a, b; };
#define IGS_TAG_BIN 0x42494E31u  // "BIN1"
#define IGS_TAG_IMG 0x494D4731u  // "IMG1"

struct BinBuf {
  BufTag* tag;
  uint32_t* tkey[2];   // [R] global tile id (v*tiles + tile), ping-pong
  uint32_t* tval[2];   // [R] emission slot e, ping-pong
  uint32_t* inst_gid;  // [R] emission slot -> global Gaussian id
  uint32_t* point_list;// [R] sorted -> global Gaussian id
  uint32_t* point_e;   // [R] sorted -> emission slot (aliases the final tval buffer)
  uint32_t* hist;
  size_t bytes;
};

__host__ __device__ inline BinBuf carve_bin(char* base, size_t R) {
  BinBuf b;
  Carver c(base);
  size_t Rc = R ? R : 1;
  b.tag = c.take<BufTag>(1);
  b.tkey[0] = c.take<uint32_t>(Rc);
  b.tkey[1] = c.take<uint32_t>(Rc);
  b.tval[0] = c.take<uint32_t>(Rc);
  b.tval[1] = c.take<uint32_t>(Rc);
  b.inst_gid = c.take<uint32_t>(Rc);
  b.point_list = c.take<uint32_t>(Rc);
  b.hist = c.take<uint32_t>(radix_hist_words(Rc));
  b.point_e = nullptr;
  b.bytes = c.off + 256;
  return b;
}

struct ImgBuf {
  BufTag* tag;
  uint2* ranges;        // [V*tiles]
  uint32_t* n_contrib;  // [V][2][N]
  float* accum_coord;   // [V][3][N]
  float* accum_depth;   // [V][N]
  float* normal_length; // [V][N]
  size_t bytes;
};

__host__ __device__ inline ImgBuf carve_img(char* base, size_t V, size_t N, size_t tiles) {
  ImgBuf m;
  Carver c(base);
  m.tag = c.take<BufTag>(1);
  m.ranges = c.take<uint2>(V * tiles);
  m.n_contrib = c.take<uint32_t>(V * 2 * N);
  m.accum_coord = c.take<float>(V * 3 * N);
  m.accum_depth = c.take<float>(V * N);
  m.normal_length = c.take<float>(V * N);
  m.bytes = c.off + 256;
  return m;
}

// Number of radix bits needed to represent values in [0, n)
__host__ __device__ inline int bits_for(uint64_t n) {
  int b = 0;
  while (b < 64 && (uint64_t(1) << b) < n) ++b;
  return b < 1 ? 1 : b;
}

// ---------------------------------------------------------------------------------------------
// Exactly-rounded helpers.  The reference is compiled with nvcc's default -fmad=true; the fused /
// unfused choice of every multiply-add on the path that decides tile keys, radii, conics and alpha
// was read from the reference's sm_100a PTX+SASS and is pinned here with _rn intrinsics so that no
// compiler version can re-associate it (DESIGN.md "bit-exact chain").
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float ffma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float frcp(float a) { return __frcp_rn(a); }
__device__ __forceinline__ float fsqrt(float a) { return __fsqrt_rn(a); }

// a0*b0 + a1*b1 + a2*b2 as the reference build evaluates glm's 3-term products:
// fma(a2,b2, fma(a0,b0, a1*b1))
__device__ __forceinline__ float dot3_ref(float a0, float b0, float a1, float b1, float a2, float b2) {
  return ffma(a2, b2, ffma(a0, b0, fmul(a1, b1)));
}

// m[0]*x + m[4]*y + m[8]*z + m[12] as compiled in the reference (auxiliary.h:74-93):
// (fma(z,m8, fma(x,m0, y*m4))) + m12
__device__ __forceinline__ float xform_row(const float* m, int r, float x, float y, float z) {
  return fadd(m[12 + r], ffma(z, m[8 + r], ffma(x, m[r], fmul(y, m[4 + r]))));
}

// Power of the 2D Gaussian at offset d from the centre, reference forward.cu:555 as compiled:
//   s = fma(dx, dx*cx, dy*(dy*cz)); power = fma(s, -0.5, -(dy*(dx*cy)))
__device__ __forceinline__ float splat_power(float dx, float dy, float cx, float cy, float cz) {
  float s = ffma(dx, fmul(dx, cx), fmul(dy, fmul(dy, cz)));
  return ffma(s, -0.5f, -fmul(dy, fmul(dx, cy)));
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }

}  // namespace igs

// error plumbing shared by the host-side translation units
namespace igs {
void set_error(const char* fmt, ...);
void count_launch(int n = 1);  // bumps the counter behind igs_kernel_launches()
// RAII event pair around one stage (no-op unless igs_stage_timing_enable(1) was called)
class StageScope {
 public:
  StageScope(int stage, cudaStream_t s);
  ~StageScope();
 private:
  int stage_;
  cudaStream_t s_;
  void* a_;
};
}
enum {
  IGS_STAGE_GAUSS_SETUP = 0, IGS_STAGE_PROJECT, IGS_STAGE_DEPTH_SORT, IGS_STAGE_SCAN_RANK, IGS_STAGE_EMIT,
  IGS_STAGE_TILE_SORT, IGS_STAGE_FINALIZE, IGS_STAGE_BLEND_FWD, IGS_STAGE_BLEND_BWD, IGS_STAGE_GATHER_BWD,
  IGS_STAGE_FINISH_BWD, IGS_STAGE_ADAM, IGS_STAGE_LOSS, IGS_STAGE_ACTIVATE, IGS_NUM_STAGES
};
#define IGS_CUDA_CHECK(expr)                                                      \
  do {                                                                            \
    cudaError_t _e = (expr);                                                      \
    if (_e != cudaSuccess) {                                                      \
      igs::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),      \
                     __FILE__, __LINE__);                                         \
      return IGS_ERR_CUDA;                                                        \
    }                                                                             \
  } while (0)
