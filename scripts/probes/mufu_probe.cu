// MUFU.EX2 throughput per SM on this GPU:  nvcc -arch=sm_100a -O3 -o /tmp/mufu_probe scripts/probes/mufu_probe.cu && /tmp/mufu_probe
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(float* out, int iters, float seed) {
  float x[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) x[j] = seed + 0.001f * (threadIdx.x + j);
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 16; ++j) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[j]));
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 16; ++j) s += x[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
  float* out; cudaMalloc(&out, 148 * 1024 * sizeof(float));
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  for (int threads : {256, 512, 1024}) {
    const int iters = 4096;
    k<<<148, threads>>>(out, 16, -0.5f);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); k<<<148, threads>>>(out, iters, -0.5f); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double ops = 148.0 * threads * iters * 16.0;
    printf("%4d threads/SM: %.3f ms, %.2f ex2 per ns per SM = %.2f per clk at %d MHz (max clock)\n", threads, ms,
           ops / 148.0 / (ms * 1e6), ops / 148.0 / (ms * 1e6) / (khz / 1e6), khz / 1000);
  }
  return 0;
}
