// Micro-probe: what does each piece of a tcgen05 kernel's prologue / epilogue cost per launch?
// Variants are timed as 64 dependent launches inside a CUDA graph.  nvcc -arch=sm_100a -o build/launch_probe ...
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include "../../aha_b200/csrc/ptx_sm100.cuh"
using namespace sm100;

// level 0: empty  1: +smem carve/syncthreads  2: +tmem alloc/dealloc  3: +mbarrier init  4: +TMA load 16KB + wait
// 5: + 4 MMAs + commit + wait  6: + tcgen05.ld + store
template <int kSmemBig>
__global__ void __launch_bounds__(256, 1) probe(const __grid_constant__ CUtensorMap tm, int level, float* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 65536);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 4);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (level == 0) return;
  if (level >= 3 && warp == 1 && lane == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); fence_barrier_init(); }
  if (level >= 2 && warp == 2) tmem_alloc<128>(slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tb = level >= 2 ? *slot : 0;
  if (level >= 4 && warp == 0 && elect_one()) {
    mbar_arrive_expect_tx(&bar[0], 32768);
    tma_load_2d(smem, &tm, &bar[0], 0, blockIdx.x * 128);
    tma_load_2d(smem + 16384, &tm, &bar[0], 64, blockIdx.x * 128);
  }
  if (level >= 4 && warp == 1 && elect_one()) {
    mbar_wait(&bar[0], 0);
    if (level >= 5) {
      tc_fence_after();
      const uint32_t idesc = make_idesc_bf16_f32(128, 128);
      for (int k = 0; k < 4; ++k)
        umma_f16_ss(tb, make_kmajor_sw128_desc(smem_u32(smem) + 32 * k), make_kmajor_sw128_desc(smem_u32(smem + 16384) + 32 * k), idesc, k != 0);
      umma_commit(&bar[1]);
    }
  }
  if (level >= 5 && warp >= 4) {
    mbar_wait(&bar[1], 0);
    tc_fence_after();
    if (level >= 6) {
      uint32_t r[16];
      tmem_ld_32x32b_x16(tb + ((uint32_t)((warp & 3) * 32) << 16), r);
      tmem_ld_wait();
      out[(blockIdx.x * 128 + (warp & 3) * 32 + lane)] = __uint_as_float(r[3]);
    }
    tc_fence_before();
  }
  __syncthreads();
  if (level >= 2 && warp == 2) { tc_fence_after(); tmem_dealloc<128>(tb); }
}

typedef CUresult (*PFN)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                        const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  uint16_t* w; cudaMalloc(&w, (size_t)148 * 128 * 128 * 2 * 64); cudaMemset(w, 0, (size_t)148 * 128 * 128 * 2 * 64);
  float* out; cudaMalloc(&out, 148 * 128 * 4 * 4);
  CUtensorMap tm;
  cuuint64_t dims[2] = {128, 148 * 128 * 64}; cuuint64_t str[1] = {256}; cuuint32_t box[2] = {64, 128}; cuuint32_t es[2] = {1, 1};
  ((PFN)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, w, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  cudaStream_t st; cudaStreamCreate(&st);
  const int smem = 65536 + 1024 + 256;
  cudaFuncSetAttribute(probe<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int grid : {1, 148}) {
    for (int cluster : {0, 1}) {
      for (int level = 0; level <= 6; ++level) {
        cudaGraph_t g; cudaGraphExec_t ge;
        cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal);
        for (int i = 0; i < 64; ++i) {
          cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = smem; cfg.stream = st;
          cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim = {1, 1, 1};
          cfg.attrs = at; cfg.numAttrs = cluster;
          cudaLaunchKernelEx(&cfg, probe<1>, tm, level, out);
        }
        cudaStreamEndCapture(st, &g); cudaGraphInstantiate(&ge, g, 0);
        cudaGraphLaunch(ge, st); cudaStreamSynchronize(st);
        cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
        cudaEventRecord(a, st); cudaGraphLaunch(ge, st); cudaEventRecord(b, st); cudaStreamSynchronize(st);
        float ms; cudaEventElapsedTime(&ms, a, b);
        printf("grid=%3d clusterattr=%d level=%d  %.2f us/launch  (%s)\n", grid, cluster, level, ms * 1e3 / 64, cudaGetErrorString(cudaGetLastError()));
      }
    }
  }
  return 0;
}
