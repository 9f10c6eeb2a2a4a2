// Issue-rate microbenchmark: scalar FFMA vs packed FFMA2 (fma.rn.f32x2) on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_rate ffma2_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pack(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpack(u64 r, float& a, float& b) { asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(r)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float s, int iters) {
  float x[16];
#pragma unroll
  for (int i = 0; i < 16; i++) x[i] = threadIdx.x * 0.001f + i;
  u64 y[8];
#pragma unroll
  for (int i = 0; i < 8; i++) y[i] = pack(x[2 * i], x[2 * i + 1]);
  u64 ss = pack(s, s);
  unsigned z[8];
#pragma unroll
  for (int i = 0; i < 8; i++) z[i] = threadIdx.x * 7 + i;
  for (int it = 0; it < iters; it++) {
    if (MODE == 0) {  // 16 scalar FFMA
#pragma unroll
      for (int i = 0; i < 16; i++) x[i] = fma1(x[i], s, s);
    } else if (MODE == 1) {  // 8 packed FFMA2 (same flops as MODE 0)
#pragma unroll
      for (int i = 0; i < 8; i++) y[i] = fma2(y[i], ss, ss);
    } else if (MODE == 3) {  // 16 FFMA + 8 LOP3: 24 issue slots
#pragma unroll
      for (int i = 0; i < 16; i++) x[i] = fma1(x[i], s, s);
#pragma unroll
      for (int i = 0; i < 8; i++) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(z[i]) : "r"(z[(i + 1) & 7]), "r"(it));
    } else if (MODE == 4) {  // 8 FFMA2 + 8 LOP3: 16 issue slots if FFMA2 takes one
#pragma unroll
      for (int i = 0; i < 8; i++) y[i] = fma2(y[i], ss, ss);
#pragma unroll
      for (int i = 0; i < 8; i++) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(z[i]) : "r"(z[(i + 1) & 7]), "r"(it));
    } else if (MODE == 5) {  // 8 LOP3 only
#pragma unroll
      for (int i = 0; i < 8; i++) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(z[i]) : "r"(z[(i + 1) & 7]), "r"(it));
    } else {  // 16 packed FFMA2 over 8 registers (2x flops of MODE 0, same instruction count)
#pragma unroll
      for (int r = 0; r < 2; r++)
#pragma unroll
        for (int i = 0; i < 8; i++) y[i] = fma2(y[i], ss, ss);
    }
  }
  float acc = 0;
  for (int i = 0; i < 8; i++) acc += (float)z[i];
  if (MODE == 0 || MODE == 3) { for (int i = 0; i < 16; i++) acc += x[i]; }
  else { for (int i = 0; i < 8; i++) { float a, b; unpack(y[i], a, b); acc += a + b; } }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int MODE> float run(float* out, int iters) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  k<MODE><<<148 * 8, 256>>>(out, 1.0001f, iters);
  cudaEventRecord(a);
  k<MODE><<<148 * 8, 256>>>(out, 1.0001f, iters);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}
int main() {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  int iters = 20000;
  float t0 = run<0>(out, iters), t1 = run<1>(out, iters), t2 = run<2>(out, iters);
  printf("16 FFMA + 8 LOP3    : %.3f ms\n", run<3>(out, iters));
  printf("8 FFMA2 + 8 LOP3    : %.3f ms\n", run<4>(out, iters));
  printf("8 LOP3              : %.3f ms\n", run<5>(out, iters));
  double inst = 148.0 * 8 * 8 * iters;  // warps * iters
  printf("16 FFMA / iter      : %.3f ms  (%.2f warp-instr/clk/SM @1.9GHz)\n", t0, inst * 16 / (t0 * 1e-3) / 1.9e9 / 148);
  printf("8 FFMA2 / iter      : %.3f ms  (%.2f)\n", t1, inst * 8 / (t1 * 1e-3) / 1.9e9 / 148);
  printf("16 FFMA2 / iter     : %.3f ms  (%.2f)\n", t2, inst * 16 / (t2 * 1e-3) / 1.9e9 / 148);
  return 0;
}
