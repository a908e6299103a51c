// Host-side interface of the tcgen05 GEMM (gemm_tcgen05.cu).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <map>
#include <mutex>
#include "common.cuh"

enum GemmAct { GEMM_ACT_NONE = 0, GEMM_ACT_GELU_TANH = 1, GEMM_ACT_GELU_ERF = 2, GEMM_ACT_SILU_MUL = 3 };

// Device-visible parameter block (passed by value).
struct GemmParams {
  int Ma, Nb, K;        // rows of A, rows of B, shared inner dimension
  int swap;             // 0: A = tokens, B = features ; 1: A = features (weights), B = tokens
  int splits;           // split-K factor (gridDim.z)
  int kb_per_split;     // 64-wide k-blocks per split
  void* out;            // out[token * ldo + feature]; bf16 unless out_f32
  int ldo;
  int out_f32;
  int round_bf16;       // with out_f32: round the sum to bf16 first (reference: bf16 Linear then .float())
  int act;              // GemmAct
  const __nv_bfloat16* bias;      // [features] or null
  const __nv_bfloat16* residual;  // residual[(token % res_mod) * ldr + feature] or null
  int ldr;
  int res_mod;          // 0 = no modulo
  int dbg;              // bring-up knob (AHA_GEMM_DBG): 1 = return at entry, 2 = skip epilogue math/stores
  int ring, ring_deep, ring_shallow;   // operand-ring stages actually used (0 = all the kernel variant has)
  float* ws_partial;    // split-K workspace: [tile][split][BN][128] fp32 partial tiles
  int* ws_counters;     // split-K arrival counter per tile, in this launch's counter bank (zero on entry)
  int* ws_zero;         // the OTHER counter bank, ws_zero_n ints: this launch zeroes it for the next split launch
  int ws_zero_n;
  int sk_ranks;         // persistent skinny kernel, stream-K decomposition: CTAs per token tile (0: fixed split); `splits` is then
                        // the workspace slot count per tile (the largest number of ranks that share one tile)
  int* ws_pcounters;    // persistent skinny kernel: self-resetting counter per (tile, column group), zero between launches
  unsigned long long* dbg_ts;   // bring-up: phase stamps of CTA (0,0,0), or null
  unsigned long long* dbg_all;  // bring-up: (accumulator complete, exit) of every CTA on the global ns timer, or null
  L2Prefetch pf;        // weights of LATER kernels to request into L2 once this CTA's own loads are in flight
  // RMSNorm folded around the GEMMs (skinny orientation only, see GemmCall):
  const float* rs_partials;   // consumer: sum over i < rs_n of rs_partials[i * rs_ld + token] = sum of squares of the row
  int rs_n, rs_ld;
  float rs_eps, rs_inv_cols;  // out = rsqrt(sumsq * rs_inv_cols + rs_eps) * (x . W^T) (+ bias ...)
  float* sumsq_out;           // producer: sumsq_out[feature_tile * sumsq_ld + token] = sum of out[token, f]^2 over the tile
  int sumsq_ld;
};

// Description of one GEMM call as the model layers see it:  out[tokens, features] = X[tokens, K] . W[features, K]^T
struct GemmCall {
  const __nv_bfloat16* x; int tokens; int ldx;    // activations, row stride in elements
  const __nv_bfloat16* w; int features; int ldw;  // weights
  int K;
  void* out; int ldo; int out_f32; int round_bf16;
  int act;
  const __nv_bfloat16* bias;
  const __nv_bfloat16* residual; int ldr; int res_mod;
  int force_swap;    // -1 auto, 0 tall, 1 skinny
  int force_bn;      // 0 auto
  int force_splits;  // 0 auto
  int force_deep;    // -1 auto, 0 two CTAs per SM, 1 one CTA per SM with the deep operand ring (skinny only)
  L2Prefetch pf;     // see common.cuh (all zero: nothing to prefetch)
  // RMSNorm without a kernel of its own (tokens <= 256, no GELU: the skinny orientation).  The norm's per-token factor
  // commutes with the GEMM that consumes the normalised row, so the CONSUMER multiplies its accumulators by
  // rsqrt(mean(x^2) + eps) in the epilogue (its weight matrix has the norm weight folded into its columns) and reads the
  // raw residual stream; the sum of squares comes from the PRODUCER of that stream, whose epilogue adds up the squares
  // of the bf16 values it stores, one partial per 128-feature tile (summed in tile order: deterministic).
  const float* rs_partials; int rs_n; int rs_ld; float rs_eps;   // consumer side (rs_partials null: off); mean over K
  float* sumsq_out; int sumsq_ld;                                // producer side (null: off)
  int* sumsq_tiles;                                              // out: partial rows the producer writes per token
};

// Split-K scratch is per CUDA stream (launches on one stream are ordered; two streams must not share counters).
// Arrival counters live in two banks that successive split launches of a stream (GEMM or attention) alternate between:
// a launch counts arrivals in its bank and zeroes the other one for its successor, so no CTA has to find out that it
// is the last one out (that took an atomic with a return value - a full L2 round trip at the very end of every
// split kernel, in front of the dependent launch).  Inside a bank the GEMM uses [0, kCounters), attention the rest.
struct GemmSplitScratch {
  float* partial = nullptr;
  int* counters = nullptr;       // all banks (allocation base)
  int* pcounters = nullptr;      // self-resetting counters of the last-arriver merges (persistent skinny GEMM, long-key attention)
  int* bank = nullptr;           // this launch's bank
  int* zero = nullptr;           // the bank to zero
  int zero_n = 0;
  unsigned long long seq = 0;    // split launches so far on this stream
};
struct GemmWorkspace {
  int sm_count = 148;
  static constexpr size_t kPartialBytes = 64u << 20;
  static constexpr int kCounters = 1024;
  static constexpr int kPCounters = 8192;   // [0, 4096): persistent skinny GEMM, [4096, 8192): attention last-arriver merge
  std::map<cudaStream_t, GemmSplitScratch> scratch;
  std::mutex mu;
};

int gemm_workspace_create(GemmWorkspace* w, int device);
bool aha_sm_shared();   // aha_set_sm_sharing(1) is in effect
// Per-stream split scratch of the default workspace for ONE split launch (every call advances the stream's bank
// sequence: call it only when the split kernel is really launched).  out->bank points at the attention half.
// advance = false: only the partial workspace and the self-resetting counters (attention half) are wanted; no bank switch.
int aha_split_scratch(cudaStream_t st, GemmSplitScratch* out, bool advance = true);
void gemm_workspace_destroy(GemmWorkspace* w);
// Enqueues the GEMM on `stream`. Returns AHA_OK or a negative error code.
int gemm_bf16(const GemmCall& c, GemmWorkspace* w, cudaStream_t stream);
// Cached TMA descriptor of a row-major [rows, K] bf16 matrix (row stride ld elements): box = 64 x box_rows, 128B swizzle.
#include <cuda.h>
int get_tmap(const void* ptr, int rows, int K, int ld, int box_rows, CUtensorMap* out);
// Launch counters (bench.py's gpu_launches claim).
extern unsigned long long g_aha_launch_count;

// Largest token count the skinny (weights = MMA A operand) orientation takes: 4 column tiles of the persistent kernel.
constexpr int kGemmSkinnyMaxTokens = 832;

// Optional per-launch CUDA-event timing (aha_profile_gemm / aha_profile_gemm_read): returns null when disabled.
struct ProfSlot;
ProfSlot* prof_begin(int cls, cudaStream_t st);
void prof_end(ProfSlot* s, cudaStream_t st, double bytes, double flops);
