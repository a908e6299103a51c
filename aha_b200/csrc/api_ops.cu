// C-ABI entries of the single operators (include/aha_b200.h, last section) - thin argument checks around ops.h.
#include "ops.h"
#include <cmath>
#include <vector>

extern "C" {

int aha_rmsnorm(const void* x, const void* weight, void* out, int rows, int cols, float eps, void* stream) {
  AHA_REQUIRE(x && weight && out, "aha_rmsnorm: NULL argument");
  return op_rmsnorm((const bf16*)x, (const bf16*)weight, (bf16*)out, rows, cols, eps, (cudaStream_t)stream);
}

int aha_layernorm(const void* x, const void* weight, const void* bias, void* out, int rows, int cols, float eps,
                  void* stream) {
  AHA_REQUIRE(x && weight && bias && out, "aha_layernorm: NULL argument");
  return op_layernorm((const bf16*)x, (const bf16*)weight, (const bf16*)bias, (bf16*)out, rows, cols, eps,
                      (cudaStream_t)stream);
}

int aha_preprocess(const uint8_t* frames_u8, void* out_bf16, long long n, void* stream) {
  AHA_REQUIRE(frames_u8 && out_bf16, "aha_preprocess: NULL argument");
  return op_preprocess(frames_u8, (bf16*)out_bf16, n, (cudaStream_t)stream);
}

int aha_ingest_frames(const uint8_t* frames_hwc, int n_frames, int height, int width, int flavour, uint8_t* out_chw,
                      int out_side, void* stream) {
  AHA_REQUIRE(n_frames == 0 || (frames_hwc && out_chw), "aha_ingest_frames: NULL argument");
  return op_ingest_frames(frames_hwc, n_frames, height, width, flavour, out_chw, out_side, (cudaStream_t)stream);
}

int aha_pool(const void* x, void* out, int frames, int side, int channels, int stride, int mode, void* stream) {
  AHA_REQUIRE(x && out, "aha_pool: NULL argument");
  return op_pool((const bf16*)x, (bf16*)out, frames, side, channels, stride, mode, (cudaStream_t)stream);
}

int aha_rope_table(int pos0, int q, int head_dim, float theta, void* cos_out, void* sin_out, void* stream) {
  AHA_REQUIRE(cos_out && sin_out && q > 0 && head_dim > 0 && head_dim % 2 == 0, "aha_rope_table: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  std::vector<float> inv(head_dim / 2);
  for (int j = 0; j < head_dim / 2; ++j) inv[j] = 1.0f / powf(theta, (float)(2 * j) / (float)head_dim);
  float* dinv = nullptr;
  AHA_CUDA(cudaMallocAsync(&dinv, inv.size() * sizeof(float), st));
  AHA_CUDA(cudaMemcpyAsync(dinv, inv.data(), inv.size() * sizeof(float), cudaMemcpyHostToDevice, st));
  AHA_CUDA(cudaStreamSynchronize(st));  // `inv` is pageable host memory that dies with this frame
  RopeTableArgs a{};
  a.pos0[0] = pos0;
  int rc = op_rope_table(a, 1, q, head_dim, dinv, (bf16*)cos_out, (bf16*)sin_out, st);
  cudaFreeAsync(dinv, st);
  return rc;
}

// q: [kv_heads, nq, group, d] (the layout the fused RoPE kernel writes); k, v: [kv_heads, klen, d]; out: [nq, heads*d]
int aha_attention(const void* q, const void* k, const void* v, void* out, int heads, int kv_heads, int nq, int klen,
                  int head_dim, int off, float scale, void* stream) {
  AHA_REQUIRE(q && k && v && out, "aha_attention: NULL argument");
  AHA_REQUIRE(heads > 0 && kv_heads > 0 && heads % kv_heads == 0 && nq > 0 && klen >= 0, "aha_attention: bad shape");
  AHA_REQUIRE(head_dim >= 64 && head_dim % 16 == 0 && head_dim <= 128, "aha_attention: head_dim must be in {64..128}, multiple of 16");
  const int G = heads / kv_heads;
  AttnProblem ap{};
  ap.q = (const bf16*)q; ap.q_rows_total = kv_heads * nq * G; ap.q_ld = head_dim;
  ap.k = (const bf16*)k; ap.k_rows_total = (long long)kv_heads * klen; ap.k_ld = head_dim;
  ap.v = (const bf16*)v; ap.v_rows_total = (long long)kv_heads * klen; ap.v_ld = head_dim;
  ap.q_block_rows = nq * G; ap.k_head_rows = klen;
  ap.n_blocks = kv_heads; ap.nq_rows = nq * G; ap.group = G; ap.d = head_dim; ap.d_pad = head_dim;
  ap.scale = scale; ap.out = (bf16*)out; ap.out_ld = heads * head_dim;
  ap.n_segs = 1;
  ap.segs[0] = AttnSeg{0, 0, 0, klen, off, 0, 0, 0, 0};
  return op_attention(ap, (cudaStream_t)stream);
}

int aha_find_ticks(const double* scores, int n_streams, int T, double prominence, int distance,
                   const double* thresh_or_null, int32_t* peaks, int max_peaks, int32_t* n_peaks, double* smoothed_out,
                   double* thresh_out, void* stream) {
  return op_find_ticks(scores, n_streams, T, prominence, distance, thresh_or_null, peaks, max_peaks, n_peaks,
                       smoothed_out, thresh_out, (cudaStream_t)stream);
}

}  // extern "C"
