// Shared device/host helpers for the aha_b200 sm_100a kernels.
//
// Rounding convention: PyTorch evaluates every bf16 tensor op in
// fp32 and rounds the result to bf16 (RNE) before the next op.  `bf16r()` is that rounding
// step; kernels that must be bit-faithful to the reference's op order call it after every
// reference-level op.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <string>

typedef __nv_bfloat16 bf16;

#include "aha_b200.h"  // status codes of the C-ABI

extern unsigned long long g_aha_launch_count;  // every kernel launch of this library bumps it

void aha_set_error(const std::string& msg);
int aha_cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define AHA_CUDA(expr)                                                        \
  do {                                                                        \
    cudaError_t _e = (expr);                                                  \
    if (_e != cudaSuccess) return aha_cuda_fail(_e, #expr, __FILE__, __LINE__); \
  } while (0)

#define AHA_CHECK_LAUNCH() AHA_CUDA(cudaGetLastError())

#define AHA_REQUIRE(cond, msg)                         \
  do {                                                 \
    if (!(cond)) {                                     \
      aha_set_error(std::string("ValueError: ") + msg); \
      return AHA_ERR_VALUE;                            \
    }                                                  \
  } while (0)

// ---- programmatic dependent launch (PDL) ----
// Every hot-path kernel is launched with cudaLaunchAttributeProgrammaticStreamSerialization: its prologue (barrier
// init, TMEM allocation, weight prefetch) may start while the previous kernel on the stream is still running; it calls
// pdl_wait() before touching anything an earlier kernel produced (and before its first global write), and
// pdl_launch_dependents() as early as possible so that its own successor can do the same.  AHA_PDL=0 disables it.
bool aha_pdl_enabled();

#ifdef __CUDACC__
template <typename... KArgs, typename... Args>
inline int launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = aha_pdl_enabled() ? 1 : 0;
  AHA_CUDA(cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...));
  ++g_aha_launch_count;
  return AHA_OK;
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif

// ---- weight-stream L2 prefetch ----
// The single-stream decoder step is bound by streaming ~13 GB of weights per frame, yet HBM idles during every kernel's
// ramp-up / tail and during the latency-bound kernels (attention, norms, RoPE).  Each hot-path kernel therefore carries
// up to two byte ranges of the weights that a LATER kernel will stream; one elected thread per CTA requests its slice
// of them into L2 (cp.async.bulk.prefetch.L2) once the CTA's own loads are in flight.  model.cu keeps the rolling
// window (how far ahead of consumption the stream has been requested) bounded by the L2 capacity.
struct L2Prefetch {
  const void* p[2];
  unsigned long long n[2];   // bytes, multiples of 16
};

#ifdef __CUDACC__
__device__ __forceinline__ void l2_prefetch_slice(const L2Prefetch& pf, unsigned cta, unsigned ncta) {
  constexpr unsigned kGran = 8192;
#pragma unroll 1
  for (int s = 0; s < 2; ++s) {
    const unsigned long long n = pf.n[s];
    if (n == 0) continue;
    const unsigned ngran = (unsigned)((n + kGran - 1) / kGran);
    const unsigned per = (ngran + ncta - 1) / ncta;
    const unsigned g1 = min(ngran, (cta + 1) * per);
#pragma unroll 1
    for (unsigned g = cta * per; g < g1; ++g) {
      const unsigned long long off = (unsigned long long)g * kGran;
      const unsigned bytes = (unsigned)min((unsigned long long)kGran, n - off);
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"((const char*)pf.p[s] + off), "r"(bytes) : "memory");
    }
  }
}
#endif

#ifdef __CUDACC__
// bring-up aid (AHA_TS=1): one CTA of a kernel stamps its phases with the SM cycle counter
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void dbg_stamp(unsigned long long* ts, int slot) {
  if (ts) ts[slot] = (unsigned long long)clock64();
}
#endif

#ifdef __CUDACC__

__device__ __forceinline__ float bf16r(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
__device__ __forceinline__ float b2f(bf16 x) { return __bfloat162float(x); }
__device__ __forceinline__ bf16 f2b(float x) { return __float2bfloat16_rn(x); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}

// bf16 rounding of two values at once.  A single-value conversion (F2F.BF16.F32) issues on the XU pipe, 16 lanes per
// clock per SM, and bounds any kernel that rounds after every op; the packed conversion (F2FP) does not.
__device__ __forceinline__ void bf16r2(float& a, float& b) {
  const uint32_t p = pack_bf16x2(a, b);
  a = __uint_as_float(p << 16);
  b = __uint_as_float(p & 0xffff0000u);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum for blockDim.x <= 1024 (multiple of 32); `red` is >= 32 floats of smem.
__device__ __forceinline__ float block_sum(float v, float* red) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : 0.f;
  r = warp_sum(r);
  return r;
}

// 128-bit streaming load / store (read-once data: keep it out of L1).
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_v4(void* p, const uint4& v) {
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// Activation functions in fp32, written the way ATen evaluates them.
__device__ __forceinline__ float gelu_erf_f(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float gelu_tanh_f(float x) {
  const float kBeta = 0.79788456080286535588f;  // sqrt(2/pi)
  const float kKappa = 0.044715f;
  float inner = kBeta * (x + kKappa * x * x * x);
  return 0.5f * x * (1.f + tanhf(inner));
}
__device__ __forceinline__ float silu_f(float x) { return x / (1.f + expf(-x)); }

#endif  // __CUDACC__
