// Read bandwidth of buffers that fit / do not fit the L2 (float4 loads, grid-stride, sum into a sink).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2_bw l2_bw.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) rd(const float4 *__restrict__ src, size_t n4, int reps, float *sink)
{
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int r = 0; r < reps; r++) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n4; i += 4 * stride) {
      float4 a = __ldg(src + i), b = __ldg(src + i + stride), c = __ldg(src + i + 2 * stride), d = __ldg(src + i + 3 * stride);
      acc.x += a.x + b.x + c.x + d.x; acc.y += a.y + b.y + c.y + d.y;
      acc.z += a.z + b.z + c.z + d.z; acc.w += a.w + b.w + c.w + d.w;
    }
    for (; i < n4; i += stride) { float4 a = __ldg(src + i); acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w; }
  }
  if (acc.x + acc.y + acc.z + acc.w == 1234.5f) *sink = acc.x;
}

int main()
{
  const size_t maxb = (size_t)2 << 30;
  float4 *buf; float *sink;
  cudaMalloc(&buf, maxb); cudaMalloc(&sink, 4);
  cudaMemset(buf, 0, maxb);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const size_t sizes[] = {(size_t)8 << 20, (size_t)16 << 20, (size_t)32 << 20, (size_t)48 << 20, (size_t)64 << 20, (size_t)96 << 20, (size_t)128 << 20, (size_t)256 << 20, (size_t)1 << 30, (size_t)2 << 30};
  for (size_t sz : sizes) {
    const size_t n4 = sz / 16;
    const int reps = (int)(((size_t)8 << 30) / sz);
    rd<<<148 * 8, 256>>>(buf, n4, 1, sink);   // warm
    cudaEventRecord(e0);
    rd<<<148 * 8, 256>>>(buf, n4, reps, sink);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%5zu MiB x %4d: %8.3f ms  %7.1f GB/s\n", sz >> 20, reps, ms, (double)sz * reps / ms * 1e-6);
  }
  printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
