// Greedy next-token choice for the response decode loop: what fast_greedy_generate
// (models/modeling_live.py:64-90) does per step on the host side of the logits - transformers'
// RepetitionPenaltyLogitsProcessor (score < 0 ? score * p : score / p for every id generated so far, applied once per
// id, on a copy) followed by argmax over the vocabulary - as one kernel launch, the chosen id staying on the device.
#include "ops.h"

namespace {

constexpr int kPickThreads = 1024;

__global__ void __launch_bounds__(kPickThreads) greedy_pick_kernel(const float* __restrict__ logits, int vocab,
                                                                   const long long* __restrict__ generated, int n_generated,
                                                                   float penalty, float* __restrict__ scratch,
                                                                   long long* __restrict__ out_id) {
  __shared__ float s_val[kPickThreads / 32];
  __shared__ int s_idx[kPickThreads / 32];
  const int tid = threadIdx.x;
  const float* src = logits;
  if (n_generated > 0 && penalty != 1.0f) {
    for (int i = tid; i < vocab; i += kPickThreads) scratch[i] = logits[i];
    __syncthreads();
    // gather from the ORIGINAL scores, scatter into the copy: duplicates write the same value (penalised once)
    for (int g = tid; g < n_generated; g += kPickThreads) {
      const long long id = generated[g];
      if (id >= 0 && id < vocab) {
        const float s = logits[id];
        scratch[id] = s < 0.f ? s * penalty : s / penalty;
      }
    }
    __syncthreads();
    src = scratch;
  }
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int i = tid; i < vocab; i += kPickThreads) {
    const float v = src[i];
    if (v > best || bi == 0x7fffffff) { best = v; bi = i; }   // ascending i per thread: ties keep the earlier index
  }
  // first maximum wins ties (torch.argmax on a contiguous row)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if ((tid & 31) == 0) { s_val[tid >> 5] = best; s_idx[tid >> 5] = bi; }
  __syncthreads();
  if (tid < 32) {
    best = s_val[tid]; bi = s_idx[tid];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if (tid == 0) *out_id = bi;
  }
}

}  // namespace

extern "C" int aha_greedy_pick(const float* logits, int vocab, const long long* generated_ids, int n_generated,
                               float repetition_penalty, float* scratch, long long* out_id, void* stream) {
  AHA_REQUIRE(logits && out_id && vocab > 0, "aha_greedy_pick: bad arguments");
  AHA_REQUIRE(n_generated == 0 || (generated_ids && scratch), "aha_greedy_pick: generated ids / scratch missing");
  AHA_REQUIRE(repetition_penalty > 0.f, "`penalty` has to be a strictly positive float");
  greedy_pick_kernel<<<1, kPickThreads, 0, (cudaStream_t)stream>>>(logits, vocab, generated_ids, n_generated,
                                                                   repetition_penalty, scratch, out_id);
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}
