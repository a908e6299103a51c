// Thread-local error string behind aha_last_error() (include/aha_b200.h).
#include "common.cuh"
#include <string>

static thread_local std::string g_err;

void aha_set_error(const std::string& msg) { g_err = msg; }
const char* aha_get_error_cstr() { return g_err.c_str(); }

int aha_cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  g_err = std::string("RuntimeError: CUDA error '") + cudaGetErrorString(e) + "' at " + file + ":" +
          std::to_string(line) + " in " + what;
  return AHA_ERR_RUNTIME;
}

#include <cstdlib>
bool aha_pdl_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("AHA_PDL"); v = (e && e[0] == '0') ? 0 : 1; }
  return v != 0;
}
