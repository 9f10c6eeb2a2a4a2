// mat3.cuh — tiny column-major 3x3 helpers for the per-Gaussian backward (tolerance-level math).
// Storage convention matches the formulas being restated (glm): m.c[col][row].
#pragma once
#include "common.cuh"

namespace igs {

struct V3 {
  float x, y, z;
};
struct M3 {
  float c[3][3];
};

__device__ __forceinline__ V3 v3(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float comp(V3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

__device__ __forceinline__ M3 m3_zero() {
  M3 r;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) r.c[i][j] = 0.0f;
  return r;
}
// columns given in order (like glm::mat3(a,b,c, d,e,f, g,h,i))
__device__ __forceinline__ M3 m3_cols(float a, float b, float c, float d, float e, float f, float g, float h, float i) {
  M3 r;
  r.c[0][0] = a; r.c[0][1] = b; r.c[0][2] = c;
  r.c[1][0] = d; r.c[1][1] = e; r.c[1][2] = f;
  r.c[2][0] = g; r.c[2][1] = h; r.c[2][2] = i;
  return r;
}
__device__ __forceinline__ M3 transpose(const M3& a) {
  M3 r;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) r.c[i][j] = a.c[j][i];
  return r;
}
__device__ __forceinline__ M3 operator*(const M3& a, const M3& b) {
  M3 r;
#pragma unroll
  for (int cc = 0; cc < 3; cc++)
#pragma unroll
    for (int rr = 0; rr < 3; rr++)
      r.c[cc][rr] = a.c[0][rr] * b.c[cc][0] + a.c[1][rr] * b.c[cc][1] + a.c[2][rr] * b.c[cc][2];
  return r;
}
__device__ __forceinline__ V3 operator*(const M3& a, V3 v) {
  return V3{a.c[0][0] * v.x + a.c[1][0] * v.y + a.c[2][0] * v.z, a.c[0][1] * v.x + a.c[1][1] * v.y + a.c[2][1] * v.z,
            a.c[0][2] * v.x + a.c[1][2] * v.y + a.c[2][2] * v.z};
}
__device__ __forceinline__ M3 operator*(const M3& a, float s) {
  M3 r;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) r.c[i][j] = a.c[i][j] * s;
  return r;
}
__device__ __forceinline__ M3 operator+(const M3& a, const M3& b) {
  M3 r;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) r.c[i][j] = a.c[i][j] + b.c[i][j];
  return r;
}
// outer(c, r)[col j][row i] = c_i * r_j
__device__ __forceinline__ M3 outer(V3 cv, V3 rv) {
  M3 m;
  float cc[3] = {cv.x, cv.y, cv.z}, rr[3] = {rv.x, rv.y, rv.z};
#pragma unroll
  for (int j = 0; j < 3; j++)
#pragma unroll
    for (int i = 0; i < 3; i++) m.c[j][i] = cc[i] * rr[j];
  return m;
}
__device__ __forceinline__ V3 col(const M3& a, int j) { return V3{a.c[j][0], a.c[j][1], a.c[j][2]}; }

}  // namespace igs
