// eigen3.cuh — symmetric 3x3 eigen-decomposition used for the RaDe-GS plane/normal terms.
//
// Behavioural restatement, specialised to n = 3, of the solver the reference carries in
// DGR/cuda_rasterizer/auxiliary.h:182-402 (Householder tridiagonalisation followed by implicit-shift
// QL, with ABSOLUTE 1e-7 thresholds and a 30-iteration cap).  The thresholds are part of the observable
// behaviour (they decide `well_conditioned` and the D==0 bail-out, forward.cu:142,162), so the
// control flow is kept decision-for-decision; the arithmetic itself only needs to agree to
// rounding (plane/normal outputs are compared at 1e-4, not bit-exactly).
#pragma once
#include "common.cuh"

namespace igs {

__device__ __forceinline__ bool near_zero(float x) { return fabsf(x) <= 1e-7f; }

__device__ __forceinline__ float hypot_ref(float a, float b) {
  float aa = fabsf(a), ab = fabsf(b);
  if (aa > ab) {
    float q = ab / aa;
    return aa * sqrtf(1.0f + q * q);
  }
  if (near_zero(ab)) return 0.0f;
  float q = aa / ab;
  return ab * sqrtf(1.0f + q * q);
}

// m: symmetric matrix given as its 6 unique entries [xx,xy,xz,yy,yz,zz].
// Outputs eigenvalues w[3] and eigenvectors z[k][i] = component k of eigenvector i.
// Returns false when the QL iteration does not converge (reference returns 0).
__device__ inline bool eigen_sym3(const float* m, float w[3], float z[3][3]) {
  float a[3][3] = {{m[0], m[1], m[2]}, {m[1], m[3], m[4]}, {m[2], m[4], m[5]}};
  float d[3], e[3];

  // ---- Householder step for the last row (i = 2, l = 1 in 0-based terms) ----
  {
    float h = 0.0f;
    float scale = fabsf(a[2][0]) + fabsf(a[2][1]);
    if (near_zero(scale)) {
      e[2] = a[2][1];
    } else {
      a[2][0] /= scale;
      a[2][1] /= scale;
      h = a[2][0] * a[2][0] + a[2][1] * a[2][1];
      float f = a[2][1];
      float g = (f >= 0.0f) ? -sqrtf(h) : sqrtf(h);
      e[2] = scale * g;
      h -= f * g;
      a[2][1] = f - g;
      f = 0.0f;
      // j = 0
      a[0][2] = a[2][0] / h;
      g = a[0][0] * a[2][0] + a[1][0] * a[2][1];
      e[0] = g / h;
      f += e[0] * a[2][0];
      // j = 1
      a[1][2] = a[2][1] / h;
      g = a[1][0] * a[2][0] + a[1][1] * a[2][1];
      e[1] = g / h;
      f += e[1] * a[2][1];
      float hh = f / (h + h);
      // j = 0
      f = a[2][0];
      e[0] = g = e[0] - hh * f;
      a[0][0] -= (f * e[0] + g * a[2][0]);
      // j = 1
      f = a[2][1];
      e[1] = g = e[1] - hh * f;
      a[1][0] -= (f * e[0] + g * a[2][0]);
      a[1][1] -= (f * e[1] + g * a[2][1]);
    }
    d[2] = h;
  }
  // ---- i = 1: nothing to reduce ----
  e[1] = a[1][0];
  d[1] = 0.0f;
  d[0] = 0.0f;
  e[0] = 0.0f;

  // ---- accumulate the transformation ----
  // i = 0
  d[0] = a[0][0];
  a[0][0] = 1.0f;
  // i = 1 (d[1] == 0 -> no update)
  d[1] = a[1][1];
  a[1][1] = 1.0f;
  a[0][1] = a[1][0] = 0.0f;
  // i = 2
  if (!near_zero(d[2])) {
#pragma unroll
    for (int j = 0; j < 2; j++) {
      float g = a[2][0] * a[0][j] + a[2][1] * a[1][j];
      a[0][j] -= g * a[0][2];
      a[1][j] -= g * a[1][2];
    }
  }
  d[2] = a[2][2];
  a[2][2] = 1.0f;
  a[0][2] = a[2][0] = 0.0f;
  a[1][2] = a[2][1] = 0.0f;

  // ---- implicit-shift QL on the tridiagonal (d, e) ----
  e[0] = e[1];
  e[1] = e[2];
  e[2] = 0.0f;
  for (int l = 0; l < 3; l++) {
    int iter = 0;
    int mm;
    do {
      for (mm = l; mm < 2; mm++) {
        if (near_zero(fabsf(e[mm]))) break;
      }
      if (mm != l) {
        if (iter++ == 30) return false;
        float g = (d[l + 1] - d[l]) / (2.0f * e[l]);
        float r = hypot_ref(g, 1.0f);
        g = d[mm] - d[l] + e[l] / (g + ((g >= 0.0f) ? fabsf(r) : -fabsf(r)));
        float s = 1.0f, c = 1.0f, p = 0.0f;
        int i;
        for (i = mm - 1; i >= l; i--) {
          float f = s * e[i];
          float b = c * e[i];
          e[i + 1] = r = hypot_ref(f, g);
          if (near_zero(r)) {
            d[i + 1] -= p;
            e[mm] = 0.0f;
            break;
          }
          s = f / r;
          c = g / r;
          g = d[i + 1] - p;
          r = (d[i] - g) * s + 2.0f * c * b;
          d[i + 1] = g + (p = s * r);
          g = c * r - b;
#pragma unroll
          for (int k = 0; k < 3; k++) {
            f = a[k][i + 1];
            a[k][i + 1] = s * a[k][i] + c * f;
            a[k][i] = c * a[k][i] - s * f;
          }
        }
        if (near_zero(r) && i >= l) continue;
        d[l] -= p;
        e[l] = g;
        e[mm] = 0.0f;
      }
    } while (mm != l);
  }
#pragma unroll
  for (int i = 0; i < 3; i++) {
    w[i] = d[i];
#pragma unroll
    for (int k = 0; k < 3; k++) z[k][i] = a[k][i];
  }
  return true;
}

}  // namespace igs
