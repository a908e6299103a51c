// Micro-probe: HBM streaming bandwidth of TMA box loads, row-major weight slabs vs tile-contiguous weights.
// Each CTA streams `kblocks` boxes of [128 rows x 64 bf16] through a kStages ring (no math).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include "../../aha_b200/csrc/ptx_sm100.cuh"
using namespace sm100;

template <int kStages, int kBoxes>
__global__ void __launch_bounds__(128, 1) stream(const __grid_constant__ CUtensorMap tm, int kblocks, int mode, int ctas_per_slab) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kStages * kBoxes * 16384);
  uint64_t* empty = full + kStages;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { for (int s = 0; s < kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); } fence_barrier_init(); }
  __syncthreads();
  const int slab = blockIdx.x / ctas_per_slab, part = blockIdx.x % ctas_per_slab;
  const int kb0 = part * kblocks;
  if (warp == 0 && elect_one()) {
    for (int it = 0; it < kblocks / kBoxes; ++it) {
      const int s = it % kStages; const uint32_t ph = (it / kStages) & 1;
      mbar_wait(&empty[s], ph ^ 1);
      mbar_arrive_expect_tx(&full[s], kBoxes * 16384);
      for (int b = 0; b < kBoxes; ++b) {
        const int kb = kb0 + it * kBoxes + b;
        if (mode == 0) tma_load_2d(smem + (s * kBoxes + b) * 16384, &tm, &full[s], kb * 64, slab * 128);   // row-major [rows, K]
        else tma_load_2d(smem + (s * kBoxes + b) * 16384, &tm, &full[s], 0, (slab * (kblocks * ctas_per_slab) + kb) * 128);  // tiles contiguous
      }
    }
  } else if (warp == 1 && elect_one()) {
    for (int it = 0; it < kblocks / kBoxes; ++it) {
      const int s = it % kStages; const uint32_t ph = (it / kStages) & 1;
      mbar_wait(&full[s], ph);
      mbar_arrive(&empty[s]);
    }
  }
}

typedef CUresult (*PFN)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                        const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN enc;
static CUtensorMap mk(void* p, uint64_t rows, uint64_t cols) {
  CUtensorMap tm; cuuint64_t dims[2] = {cols, rows}; cuuint64_t str[1] = {cols * 2}; cuuint32_t box[2] = {64, 128}; cuuint32_t es[2] = {1, 1};
  enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, p, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
      CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return tm;
}

template <int kStages, int kBoxes>
void run(const char* name, void* w, int rows, int K, int ctas_per_slab, int mode, cudaStream_t st) {
  const int slabs = rows / 128, kblocks = K / 64 / ctas_per_slab;
  CUtensorMap tm = mode == 0 ? mk(w, rows, K) : mk(w, (uint64_t)rows * K / 64, 64);
  const int smem = kStages * kBoxes * 16384 + 1024 + 256;
  cudaFuncSetAttribute(stream<kStages, kBoxes>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  float best = 1e9;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(a, st);
    stream<kStages, kBoxes><<<slabs * ctas_per_slab, 128, smem, st>>>(tm, kblocks, mode, ctas_per_slab);
    cudaEventRecord(b, st); cudaStreamSynchronize(st);
    float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
  }
  printf("%-34s rows=%5d K=%5d ctas=%4d stages=%d boxes/stage=%d mode=%s: %.1f us  %.0f GB/s (%s)\n", name, rows, K, slabs * ctas_per_slab,
         kStages, kBoxes, mode ? "tile-contiguous" : "row-major", best * 1e3, (double)rows * K * 2 / best / 1e6, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q); enc = (PFN)fn;
  uint16_t* w; const size_t bytes = (size_t)37888 * 3584 * 2 * 2; cudaMalloc(&w, bytes); cudaMemset(w, 1, bytes);
  uint16_t* flush; cudaMalloc(&flush, 256u << 20);
  cudaStream_t st; cudaStreamCreate(&st);
  for (int mode = 0; mode < 2; ++mode) {
    run<4, 1>("gate_up 37888x3584", w, 37888, 3584, 1, mode, st);
    run<8, 1>("gate_up 37888x3584", w, 37888, 3584, 1, mode, st);
    run<12, 1>("gate_up 37888x3584", w, 37888, 3584, 1, mode, st);
    run<3, 4>("gate_up 37888x3584", w, 37888, 3584, 1, mode, st);
    run<8, 1>("down 3584x18944 (split 8)", w, 3584, 18944 - 18944 % 512, 8, mode, st);
    run<8, 1>("o_proj 3584x3584 (split 4)", w, 3584, 3584, 4, mode, st);
    run<8, 1>("o_proj 3584x3584 (split 7)", w, 3584, 3584, 7, mode, st);
    cudaMemsetAsync(flush, 0, 256u << 20, st);
  }
  return 0;
}
