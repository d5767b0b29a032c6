// Issue rate of FFMA (3-register form) against FFMA2 (fma.rn.f32x2) on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_rate ffma2_rate.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) rate(float *out, float a, float b, int iters)
{
  float x[16];
#pragma unroll
  for (int i = 0; i < 16; i++) x[i] = threadIdx.x * 0.001f + i;
  unsigned long long p[8];
#pragma unroll
  for (int i = 0; i < 8; i++) asm("mov.b64 %0, {%1,%2};" : "=l"(p[i]) : "f"(x[2 * i]), "f"(x[2 * i + 1]));
  unsigned long long aa, bb;
  asm("mov.b64 %0, {%1,%1};" : "=l"(aa) : "f"(a));
  asm("mov.b64 %0, {%1,%1};" : "=l"(bb) : "f"(b));
  for (int it = 0; it < iters; it++) {
    if (MODE == 0) {
#pragma unroll
      for (int r = 0; r < 4; r++)
#pragma unroll
        for (int i = 0; i < 16; i++) x[i] = fmaf(x[i], a, b);
    } else {
#pragma unroll
      for (int r = 0; r < 4; r++)
#pragma unroll
        for (int i = 0; i < 8; i++) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(aa), "l"(bb));
    }
  }
  float s = 0.f;
  if (MODE == 0) {
#pragma unroll
    for (int i = 0; i < 16; i++) s += x[i];
  } else {
#pragma unroll
    for (int i = 0; i < 8; i++) { float lo, hi; asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[i])); s += lo + hi; }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main()
{
  float *out; cudaMalloc(&out, 148 * 8 * 256 * sizeof(float));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int mode = 0; mode < 2; mode++) {
    for (int rep = 0; rep < 2; rep++) {
      cudaEventRecord(e0);
      if (mode == 0) rate<0><<<148 * 8, 256>>>(out, 0.999f, 0.001f, iters);
      else rate<1><<<148 * 8, 256>>>(out, 0.999f, 0.001f, iters);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      // FMAs per thread: mode 0: 64 per iteration; mode 1: 32 instructions = 64 FMAs per iteration
      double fma = (double)148 * 8 * 256 * iters * 64;
      printf("mode %d (%s): %.3f ms  %.2f TFMA/s  (%.1f TFLOP/s)\n", mode, mode ? "FFMA2" : "FFMA", ms, fma / ms * 1e-9, 2 * fma / ms * 1e-9);
    }
  }
  printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
