// MUFU.EX2 throughput per SM on this GPU (f32, packed f16x2, packed bf16x2):
//   nvcc -arch=sm_100a -O3 -o /tmp/mufu_probe scripts/probes/mufu_probe.cu && /tmp/mufu_probe
// (cuobjdump -sass shows that one ex2.approx.f16x2 / .bf16x2 becomes TWO MUFU.EX2.F16 / .BF16: no XU saving from packing)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
template <int kMode>
__global__ void k(float* out, int iters, float seed) {
  float x[16];
  uint32_t h[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) { x[j] = seed + 0.001f * (threadIdx.x + j); h[j] = 0xb800b800u + threadIdx.x + j; }
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      if (kMode == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[j]));
      if (kMode == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h[j]));
      if (kMode == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h[j]));
    }
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 16; ++j) s += x[j] + __uint_as_float(h[j]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int kMode>
void run(const char* name, float* out, int khz) {
  for (int threads : {256, 512, 1024}) {
    const int iters = 4096;
    k<kMode><<<148, threads>>>(out, 16, -0.5f);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); k<kMode><<<148, threads>>>(out, iters, -0.5f); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double ops = 148.0 * threads * iters * 16.0;
    printf("%-8s %4d threads/SM: %.3f ms, %.2f instruction lanes per ns per SM = %.2f per clk at %d MHz (max clock)%s\n", name,
           threads, ms, ops / 148.0 / (ms * 1e6), ops / 148.0 / (ms * 1e6) / (khz / 1e6), khz / 1000,
           kMode ? "  (two exponentials per lane)" : "");
  }
}
int main() {
  float* out; cudaMalloc(&out, 148 * 1024 * sizeof(float));
  int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  run<0>("f32", out, khz);
  run<1>("f16x2", out, khz);
  run<2>("bf16x2", out, khz);
  return 0;
}
