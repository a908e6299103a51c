// C-ABI entry for the standalone GEMM (used by the parity tests and by the Python op wrappers).
#include "common.cuh"
#include "gemm.h"
#include <mutex>

const char* aha_get_error_cstr();

static GemmWorkspace g_ws[16];
static bool g_ws_init[16] = {false};
static std::mutex g_ws_mu;

GemmWorkspace* aha_default_workspace(int* rc_out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 16) {
    *rc_out = aha_cuda_fail(e == cudaSuccess ? cudaErrorInvalidDevice : e, "cudaGetDevice", __FILE__, __LINE__);
    return nullptr;
  }
  std::lock_guard<std::mutex> lk(g_ws_mu);
  if (!g_ws_init[dev]) {
    int rc = gemm_workspace_create(&g_ws[dev], dev);
    if (rc) { *rc_out = rc; return nullptr; }
    g_ws_init[dev] = true;
  }
  *rc_out = AHA_OK;
  return &g_ws[dev];
}

extern "C" {

const char* aha_last_error(void) { return aha_get_error_cstr(); }

unsigned long long aha_launch_count(void) { return g_aha_launch_count; }

int aha_gemm_bf16(const void* x, int tokens, int ldx, const void* w, int features, int ldw, int K, void* out, int ldo,
                  int out_f32, int round_bf16, int act, const void* bias, const void* residual, int ldr, int res_mod,
                  int force_swap, int force_bn, int force_splits, void* stream) {
  int rc;
  GemmWorkspace* ws = aha_default_workspace(&rc);
  if (!ws) return rc;
  GemmCall c{};
  c.x = (const bf16*)x; c.tokens = tokens; c.ldx = ldx;
  c.w = (const bf16*)w; c.features = features; c.ldw = ldw; c.K = K;
  c.out = out; c.ldo = ldo; c.out_f32 = out_f32; c.round_bf16 = round_bf16; c.act = act;
  c.bias = (const bf16*)bias; c.residual = (const bf16*)residual; c.ldr = ldr; c.res_mod = res_mod;
  c.force_swap = force_swap; c.force_bn = force_bn; c.force_splits = force_splits; c.force_deep = -1;
  return gemm_bf16(c, ws, (cudaStream_t)stream);
}

}  // extern "C"
