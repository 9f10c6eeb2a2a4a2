// Pipe-rate microbenchmarks for the blend kernels' roofline (SURVEY 8d asks for measured FP32 / MUFU / shared-memory
// peaks next to MEASURED_PEAKS.json): MUFU.EX2 and LDS.128 throughput per SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_rates pipe_rates.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) mufu_k(float* out, int iters) {
  float x[8];
#pragma unroll
  for (int i = 0; i < 8; i++) x[i] = -0.001f * (threadIdx.x + i);
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
  }
  float a = 0;
  for (int i = 0; i < 8; i++) a += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = a;
}

// scalar FFMA: 16 independent chains per thread
__global__ void __launch_bounds__(256) ffma_k(float* out, int iters) {
  float x[16];
#pragma unroll
  for (int i = 0; i < 16; i++) x[i] = 0.001f * (threadIdx.x + i);
  const float a = 1.0000001f, b = 1e-9f;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
  }
  float s = 0;
  for (int i = 0; i < 16; i++) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) lds_k(float* out, int iters) {
  __shared__ float4 buf[1024];
  for (int i = threadIdx.x; i < 1024; i += 256) buf[i] = make_float4(i, 1, 2, 3);
  __syncthreads();
  unsigned addr = (unsigned)__cvta_generic_to_shared(&buf[threadIdx.x]);
  float4 acc = make_float4(0, 0, 0, 0);
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 4; i++) {
      float4 v;
      asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4+%5];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr), "n"(4096 * 0));
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      addr ^= 4096u * ((i & 1) ? 1 : 2) & 0x3000u;
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc.x + acc.y + acc.z + acc.w;
}

template <typename F> float timeit(F f) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}

int main() {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
  const double ghz = pr.clockRate * 1e-6;
  const int iters = 20000;
  float t = timeit([&] { mufu_k<<<148 * 8, 256>>>(out, iters); });
  double winst = 148.0 * 8 * 8 * iters * 8;
  printf("{\"sm_clock_ghz_nominal\": %.3f,\n", ghz);
  printf(" \"mufu_ex2_warp_inst_per_clk_per_sm\": %.3f, \"mufu_ex2_ms\": %.3f,\n", winst / (t * 1e-3) / (ghz * 1e9) / 148, t);
  t = timeit([&] { ffma_k<<<148 * 8, 256>>>(out, iters); });
  winst = 148.0 * 8 * 8 * iters * 16;
  printf(" \"ffma_warp_inst_per_clk_per_sm\": %.3f, \"ffma_tflops\": %.2f, \"ffma_ms\": %.3f,\n",
         winst / (t * 1e-3) / (ghz * 1e9) / 148, winst * 64 / (t * 1e-3) / 1e12, t);
  t = timeit([&] { lds_k<<<148 * 8, 256>>>(out, iters); });
  winst = 148.0 * 8 * 8 * iters * 4;
  printf(" \"lds128_warp_inst_per_clk_per_sm\": %.3f, \"lds128_bytes_per_clk_per_sm\": %.1f, \"lds128_ms\": %.3f}\n",
         winst / (t * 1e-3) / (ghz * 1e9) / 148, winst * 512 / (t * 1e-3) / (ghz * 1e9) / 148, t);
  return 0;
}
