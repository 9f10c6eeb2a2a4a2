// blend_common.cuh — culling tests and shared-memory access helpers shared by the forward and backward blend kernels.
#pragma once
#include "common.cuh"

namespace igs {

// Conservative half-extents of { d : alpha(d) >= 1/255 } for conic (A,B,C) and opacity o (tile-level cull).
// Returns false if the splat can never reach alpha >= 1/255.  Non-PSD / non-finite conics get infinite extents
// (never culled) so the exact per-pixel test decides.  tau = bound on -power of any contributing pixel.
__device__ __forceinline__ bool alpha_extent(float A, float B, float C, float o, float& ex, float& ey, float& tau) {
  tau = 0.0f;
  if (o < 0.00392156862f) return false;  // o*G <= o < 1/255 for every G in [0,1]
  float det = A * C - B * B;
  tau = __logf(255.0f * o) * 1.0005f + 1e-3f;
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

// q(dx,dy) = 0.5*(A dx^2 + C dy^2) + B dx dy  (= -power) minus a bound on its fp32 evaluation error.
__device__ __forceinline__ float q_lower(float A, float B, float C, float dx, float dy) {
  float a = 0.5f * A * dx * dx, c = 0.5f * C * dy * dy, b = B * dx * dy;
  return (a + c + b) - 2e-6f * (a + c + fabsf(b));
}

// Exact (up to a conservative rounding margin) test "some pixel of the block [bx0,bx1]x[by0,by1] can have
// alpha >= 1/255": the minimum of the convex quadratic q over the block is 0 if the centre is inside, otherwise it
// lies on one of the four edges, where it is a clamped 1-D parabola vertex.
__device__ __forceinline__ bool block_can_contribute(float cx, float cy, float A, float B, float C, float tau, float bx0,
                                                     float by0, float bx1, float by1) {
  const float xl = cx - bx1, xh = cx - bx0, yl = cy - by1, yh = cy - by0;  // d = centre - pixel
  if (xl <= 0.0f && xh >= 0.0f && yl <= 0.0f && yh >= 0.0f) return true;
  if (!(A > 0.0f && C > 0.0f && A * C - B * B > 0.0f)) return true;  // not positive definite: per-pixel test decides
  const float nbc = -B / C, nba = -B / A;
  float q = q_lower(A, B, C, xl, fminf(fmaxf(nbc * xl, yl), yh));
  q = fminf(q, q_lower(A, B, C, xh, fminf(fmaxf(nbc * xh, yl), yh)));
  q = fminf(q, q_lower(A, B, C, fminf(fmaxf(nba * yl, xl), xh), yl));
  q = fminf(q, q_lower(A, B, C, fminf(fmaxf(nba * yh, xl), xh), yh));
  return !(q > tau);  // NaN -> keep
}

// 32-bit shared-window address of a __shared__ object, computed once (the compiler otherwise rematerialises the
// window base — S2R SR_CgaCtaId + LEA — inside the hot loop).
__device__ __forceinline__ uint32_t smem_addr(const void* p) {
  uint32_t a = (uint32_t)__cvta_generic_to_shared(p), r;
  asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(a));  // opaque: keeps the address in a register
  return r;
}

template <int OFF>
__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4+%5];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr), "n"(OFF));
  return v;
}
template <int OFF>
__device__ __forceinline__ float2 lds64(uint32_t addr) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0,%1}, [%2+%3];" : "=f"(v.x), "=f"(v.y) : "r"(addr), "n"(OFF));
  return v;
}
template <int OFF>
__device__ __forceinline__ float lds32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v) : "r"(addr), "n"(OFF));
  return v;
}
template <int OFF>
__device__ __forceinline__ void sts32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0+%1], %2;" ::"r"(addr), "n"(OFF), "f"(v) : "memory");
}

// Asynchronous global -> shared copies (LDGSTS): no register staging, completion by cp.async.wait_all / wait_group.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
// 16-byte copy that writes zeros instead when !valid (src-size 0: nothing is read; src must still be a valid address)
__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void* src, bool valid) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(valid ? 16 : 0) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Packed fp32x2 arithmetic (sm_100 FADD2 / FMUL2 / FFMA2): two IEEE round-to-nearest fp32 operations per issued
// instruction.  The blend kernels are issue bound, so pairing independent operations frees issue slots (measured:
// same FLOP rate as scalar FFMA in half the slots, tools/micro/ffma2_rate.cu).  Each lane of a pair rounds exactly
// like the scalar instruction.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
  asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ float hsum2(f32x2 v) {
  float lo, hi;
  unpack2(v, lo, hi);
  return lo + hi;
}

}  // namespace igs
