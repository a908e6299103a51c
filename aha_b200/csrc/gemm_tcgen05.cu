// bf16 GEMM on the 5th-gen tensor cores (tcgen05.mma, accumulator in TMEM, operands staged by TMA).
//
//   D[i, j] = sum_k A[i, k] * B[j, k]        A: [Ma, K] row-major bf16, B: [Nb, K] row-major bf16
//
// One CTA computes one 128 x BN tile of D (optionally one K-split of it).  Warp roles:
//   warp 0 : TMA producer (one elected lane) - fills a kStages ring of {A 128x64, B BNx64} bf16 tiles,
//            128-byte swizzled, signalling `full[s]` via mbarrier complete_tx
//   warp 1 : MMA issuer (one elected lane) - 4 x tcgen05.mma (K=16 each) per stage, tcgen05.commit -> `empty[s]`,
//            final commit -> `accum_full`
//   warp 2 : TMEM allocator / deallocator
//   warps 4-7 : epilogue - tcgen05.ld (lane = tile row), bias / activation / residual, bf16 or fp32 store
//
// The same kernel serves two orientations (GemmParams::swap):
//   swap = 0 ("tall"): A = activations [tokens, K], B = weights [features, K]   (SigLIP tower, projector)
//   swap = 1 ("skinny"): A = weights [features, K], B = activations [tokens, K] (Qwen2 prefill with
//            49..256 tokens: the 128-wide MMA M dimension is filled with weight rows, the token count
//            becomes the MMA N dimension, so no tensor work is wasted on padding and every CTA streams a
//            distinct slab of weights - the skinny GEMMs are HBM-bound on weight bytes)
// Output is always out[token * ldo + feature].
//
// Split-K: the gridDim.z K-splits of a tile are ordinary, independently scheduled CTAs (no cluster: co-scheduling 5-8
// CTAs inside one GPC cost more than it saved, profiles/r01_decoder_breakdown.txt).  Every split writes its fp32
// partial tile to an L2-resident workspace and arrives at the tile's counter; once all splits have arrived, split r
// reduces the 16-column chunks c with c mod s == r in split order 0..s-1 (deterministic) and runs their epilogue.
//
// Replaces the cuBLAS calls behind nn.Linear in the reference: SigLIP q/k/v/out/fc1/fc2
// (LLaVA_NeXT/llava/model/multimodal_encoder/siglip_encoder.py:197-255), mm_projector
// (multimodal_projector/builder.py:42-48), Qwen2 q/k/v/o/gate/up/down + lm_head
// (transformers 4.49 modeling_qwen2.py, called at models/live_llava/video_head_live_llava_qwen.py:160-175).
#include "common.cuh"
#include "ptx_sm100.cuh"
#include "gemm.h"

#include <cuda.h>
#include <cstdlib>
#include <deque>
#include <map>
#include <mutex>
#include <tuple>

using namespace sm100;

namespace {

constexpr int kBlockM = 128;  // rows of A per tile == TMEM lanes
constexpr int kBlockK = 64;   // bf16 elements per stage row == one 128-byte swizzle atom
constexpr int kUmmaK = 16;
constexpr int kThreads = 256;

// kDeep = 0: two CTAs per SM share the shared memory (grids of ~2 CTAs per SM: the big weight matrices, the tall GEMMs).
// kDeep = 1: one CTA per SM owns a ~190 KB operand ring (grids of <= 1 CTA per SM: the small / split-K skinny GEMMs);
//            the whole ring of weight tiles is requested before griddepcontrol.wait, i.e. under the previous kernel.
template <int BN, bool kDeep>
struct Cfg {
  static constexpr int kStages = kDeep ? ((BN <= 32) ? 10 : (BN <= 64) ? 8 : (BN <= 128) ? 6 : 4)
                                       : ((BN <= 32) ? 5 : (BN <= 64) ? 4 : (BN <= 128) ? 3 : 4);
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBBytes = BN * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kTmemCols = BN < 32 ? 32 : BN;  // power of two >= 32
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

// ---- epilogue math: fast approximations whose error is far below one bf16 step of the result ----
// MUFU.RCP (about 1 ulp); __frcp_rn would add a branch and a slow-path call per element and serialise the epilogue
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_silu(float x) { return x * rcp_approx(1.f + __expf(-x)); }
// gelu_tanh(x) = 0.5 x (1 + tanh(u)) = x * sigmoid(2u), u = sqrt(2/pi) (x + 0.044715 x^3): one ex2, one rcp and five FP
// operations per element (the tanh form spent eleven; the fc1 GEMM of the SigLIP tower is co-limited by its epilogue).
// Relative error ~1e-7.  Launch time of fc1 relative to fc2 (same FLOPs, no activation) by ncu: 1.26 with the tanh form,
// 1.17 with this one - and 1.17 with tanh.approx.f32, which would save the second MUFU but is only good to 5e-4 absolute,
// ten bf16 steps of the small negative outputs: not used.
__device__ __forceinline__ float fast_gelu_tanh(float x) {
  const float w = x * fmaf(0.044715f * -2.3022082f, x * x, -2.3022082f);   // -2u * log2(e)
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(w));
  return x * rcp_approx(1.f + e);
}
__device__ __forceinline__ float fast_gelu_erf(float x) {
  // erf by Abramowitz-Stegun 7.1.26 (|error| <= 1.5e-7)
  const float z = fabsf(x) * 0.70710678118654752440f;
  const float t = rcp_approx(1.f + 0.3275911f * z);
  const float poly = t * (0.254829592f + t * (-0.284496736f + t * (1.421413741f + t * (-1.453152027f + t * 1.061405429f))));
  const float e = 1.f - poly * __expf(-z * z);
  return 0.5f * x * (1.f + copysignf(e, x));
}
template <int kAct>
__device__ __forceinline__ float apply_act(float x) {   // x is already bf16-rounded; result is rounded by the caller
  if (kAct == GEMM_ACT_GELU_TANH) return fast_gelu_tanh(x);
  if (kAct == GEMM_ACT_GELU_ERF) return fast_gelu_erf(x);
  return x;
}

// Skinny orientation, 16 tokens of one feature row (thread = feature).  Lanes hold consecutive features, so every
// per-token access of a warp is one contiguous 64-byte (bf16) or 128-byte (fp32) segment.
// raw bf16 bits: converting at the load would make the loading thread wait for the data right there
__device__ __forceinline__ void load_residual16(const GemmParams& p, int feat, int tok0, uint16_t (&r)[16]) {
  const int ntok = min(16, p.Nb - tok0);
  const uint16_t* res = reinterpret_cast<const uint16_t*>(p.residual);
#pragma unroll
  for (int t = 0; t < 16; ++t) {
    const int token = tok0 + t;
    const int rrow = p.res_mod > 0 ? (token % p.res_mod) : token;
    r[t] = (t < ntok) ? res[(size_t)rrow * p.ldr + feat] : (uint16_t)0;
  }
}

// `pre_res`: the 16 residual values when the caller has already requested them (split-K: during the arrival wait)
template <int kAct>
__device__ __forceinline__ void epi_swap16(const GemmParams& p, int feat, bool fvalid, float bias, int i, int tok0,
                                           float (&acc)[16], const uint16_t* pre_res = nullptr) {
  const bool round_out = !p.out_f32 || p.round_bf16;
#pragma unroll
  for (int t = 0; t < 16; ++t) acc[t] += bias;
  if (round_out) {
#pragma unroll
    for (int t = 0; t < 16; t += 2) bf16r2(acc[t], acc[t + 1]);
  }
  const int ntok = min(16, p.Nb - tok0);
  if (kAct == GEMM_ACT_SILU_MUL) {
    // rows interleaved: even A row = gate feature f, odd A row = up feature f (the neighbouring lane)
    bf16* o = reinterpret_cast<bf16*>(p.out) + (feat >> 1) + (size_t)tok0 * p.ldo;
    const bool writer = (i & 1) == 0 && fvalid;
#pragma unroll
    for (int t = 0; t < 16; t += 2) {
      const float up0 = __shfl_down_sync(0xffffffffu, acc[t], 1), up1 = __shfl_down_sync(0xffffffffu, acc[t + 1], 1);
      float s0 = fast_silu(acc[t]), s1 = fast_silu(acc[t + 1]);
      bf16r2(s0, s1);
      const uint32_t x = pack_bf16x2(s0 * up0, s1 * up1);
      if (writer && t < ntok) *reinterpret_cast<uint16_t*>(o + (size_t)t * p.ldo) = (uint16_t)(x & 0xffffu);
      if (writer && t + 1 < ntok) *reinterpret_cast<uint16_t*>(o + (size_t)(t + 1) * p.ldo) = (uint16_t)(x >> 16);
    }
    return;
  }
  if (kAct != GEMM_ACT_NONE) {
#pragma unroll
    for (int t = 0; t < 16; t += 2) {
      acc[t] = apply_act<kAct>(acc[t]); acc[t + 1] = apply_act<kAct>(acc[t + 1]);
      bf16r2(acc[t], acc[t + 1]);
    }
  }
  if (!fvalid) return;
  if (p.residual) {
    // all loads first (they may alias the output elements written below, by this same thread only)
    uint16_t raw[16];
    if (pre_res) {
#pragma unroll
      for (int t = 0; t < 16; ++t) raw[t] = pre_res[t];
    } else {
      load_residual16(p, feat, tok0, raw);
    }
    float r[16];
#pragma unroll
    for (int t = 0; t < 16; ++t) r[t] = __uint_as_float((uint32_t)raw[t] << 16);
#pragma unroll
    for (int t = 0; t < 16; t += 2) {
      acc[t] = r[t] + acc[t]; acc[t + 1] = r[t + 1] + acc[t + 1];
      bf16r2(acc[t], acc[t + 1]);
    }
  }
  if (p.out_f32) {
    float* o = reinterpret_cast<float*>(p.out) + feat + (size_t)tok0 * p.ldo;
#pragma unroll
    for (int t = 0; t < 16; ++t)
      if (t < ntok) o[(size_t)t * p.ldo] = acc[t];
  } else {
    uint16_t* o = reinterpret_cast<uint16_t*>(p.out) + feat + (size_t)tok0 * p.ldo;
#pragma unroll
    for (int t = 0; t < 16; t += 2) {
      const uint32_t x = pack_bf16x2(acc[t], acc[t + 1]);
      if (t < ntok) o[(size_t)t * p.ldo] = (uint16_t)(x & 0xffffu);
      if (t + 1 < ntok) o[(size_t)(t + 1) * p.ldo] = (uint16_t)(x >> 16);
    }
  }
}

// Skinny orientation, staged form of the above for bf16 outputs of a fully valid 128-feature tile: the warp's
// [16 tokens][32 features] block (16 features after SiLU*mul) goes through a small shared-memory tile so that it
// leaves as 16-byte pieces - 8x fewer global memory instructions than the 2-byte per-token stores, the residual is
// read the same way.  Same arithmetic and rounding order as epi_swap16.
// `sq_warp` (plain epilogue only): this warp's 16 per-token sums of squares of the values it stores (its 32 features),
// the producer side of the folded RMSNorm - four lanes hold the four 16-byte pieces of a token, two shuffles add them.
template <int kAct>
__device__ __forceinline__ void epi_swap16_staged(const GemmParams& p, uint8_t* stage_warp, int a0, int ew, int lane,
                                                  float bias, int tok0, float (&acc)[16], float* sq_warp = nullptr) {
  constexpr int kFeat = (kAct == GEMM_ACT_SILU_MUL) ? 16 : 32;
  constexpr int kPitch = kFeat * 2 + 16;
  constexpr int kPieces = kFeat / 8;            // 16-byte pieces per token
  // every rounding is a packed conversion of a token pair (see bf16r2)
  if (kAct == GEMM_ACT_SILU_MUL) {
#pragma unroll
    for (int t = 0; t < 16; t += 2) {
      float g0 = acc[t] + bias, g1 = acc[t + 1] + bias;
      bf16r2(g0, g1);
      const float up0 = __shfl_down_sync(0xffffffffu, g0, 1), up1 = __shfl_down_sync(0xffffffffu, g1, 1);
      float s0 = fast_silu(g0), s1 = fast_silu(g1);
      bf16r2(s0, s1);
      const uint32_t x = pack_bf16x2(s0 * up0, s1 * up1);
      if ((lane & 1) == 0) {
        *reinterpret_cast<uint16_t*>(stage_warp + t * kPitch + (lane >> 1) * 2) = (uint16_t)(x & 0xffffu);
        *reinterpret_cast<uint16_t*>(stage_warp + (t + 1) * kPitch + (lane >> 1) * 2) = (uint16_t)(x >> 16);
      }
    }
  } else {
#pragma unroll
    for (int t = 0; t < 16; t += 2) {
      const uint32_t x = pack_bf16x2(acc[t] + bias, acc[t + 1] + bias);
      *reinterpret_cast<uint16_t*>(stage_warp + t * kPitch + lane * 2) = (uint16_t)(x & 0xffffu);
      *reinterpret_cast<uint16_t*>(stage_warp + (t + 1) * kPitch + lane * 2) = (uint16_t)(x >> 16);
    }
  }
  __syncwarp();
  const int ntok = min(16, p.Nb - tok0);
  const int fbase = (kAct == GEMM_ACT_SILU_MUL) ? (a0 >> 1) + ew * 16 : a0 + ew * 32;
  bf16* out = reinterpret_cast<bf16*>(p.out);
#pragma unroll
  for (int idx = lane; idx < 16 * kPieces; idx += 32) {
    const int t = idx / kPieces, pc = idx % kPieces;
    float ss = 0.f;
    if (t < ntok) {
      uint4 u = *reinterpret_cast<const uint4*>(stage_warp + t * kPitch + pc * 16);
      const int token = tok0 + t, f = fbase + pc * 8;
      if (p.residual) {
        const int rrow = p.res_mod > 0 ? (token % p.res_mod) : token;
        const uint4 ru = *reinterpret_cast<const uint4*>(p.residual + (size_t)rrow * p.ldr + f);
        uint32_t w[4] = {u.x, u.y, u.z, u.w};
        const uint32_t rw[4] = {ru.x, ru.y, ru.z, ru.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 a = unpack_bf16x2(w[e]), b = unpack_bf16x2(rw[e]);
          w[e] = pack_bf16x2(b.x + a.x, b.y + a.y);
        }
        u = make_uint4(w[0], w[1], w[2], w[3]);
      }
      st_v4(out + (size_t)token * p.ldo + f, u);
      if (kAct == GEMM_ACT_NONE && sq_warp) {
        const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
        ss = a.x * a.x + a.y * a.y + b.x * b.x + b.y * b.y + c.x * c.x + c.y * c.y + d.x * d.x + d.y * d.y;
      }
    }
    if (kAct == GEMM_ACT_NONE && sq_warp) {      // (kPieces == 4: lanes 4j .. 4j+3 hold token j of this pass)
      ss += __shfl_xor_sync(0xffffffffu, ss, 1);
      ss += __shfl_xor_sync(0xffffffffu, ss, 2);
      if (pc == 0) sq_warp[t] = ss;
    }
  }
  __syncwarp();
}

// Producer side of the folded RMSNorm: acc holds the 16 bf16 values (as floats) this thread has just stored for its
// feature; their squares are summed over the tile's 128 features (shuffles, then the four warps through shared memory)
// and the tile's partial goes to sumsq_out[feature tile][token].  Called by all four epilogue warps together.
__device__ __forceinline__ void emit_sumsq16(const GemmParams& p, float (*sq_red)[16], const float (&acc)[16], bool fvalid,
                                             int ew, int lane, int tok0, int tile, int bar_id = 1) {
  const int ntok = min(16, p.Nb - tok0);
#pragma unroll
  for (int t = 0; t < 16; ++t) {
    const float s = warp_sum(fvalid ? acc[t] * acc[t] : 0.f);
    if (lane == 0) sq_red[ew][t] = s;
  }
  asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
  if (ew == 0 && lane < ntok)
    p.sumsq_out[(size_t)tile * p.sumsq_ld + tok0 + lane] = sq_red[0][lane] + sq_red[1][lane] + sq_red[2][lane] + sq_red[3][lane];
  asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
}

// Finishes the staged form of the sums of squares: the four warps of an epilogue group have each left their 16 per-token
// partials in sq[ew][.]; one barrier, then warp 0 adds them in warp order and writes the tile's partial.  The caller
// alternates between two `sq` buffers from chunk to chunk, so one barrier per chunk is enough.
__device__ __forceinline__ void finish_sumsq16(const GemmParams& p, float (*sq)[16], int ew, int lane, int tok0, int tile,
                                               int bar_id) {
  asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
  if (ew == 0 && lane < min(16, p.Nb - tok0))
    p.sumsq_out[(size_t)tile * p.sumsq_ld + tok0 + lane] = sq[0][lane] + sq[1][lane] + sq[2][lane] + sq[3][lane];
}

// Split-K merge of one 16-column chunk: the partials of all S splits are requested before the first use (one L2 round
// trip whatever S is) and added in split order, which keeps the result deterministic.
template <int S>
__device__ __forceinline__ void reduce_splits16(const float* base, size_t split_stride, float (&acc)[16]) {
  float v[S][16];
#pragma unroll
  for (int sp = 0; sp < S; ++sp)
#pragma unroll
    for (int t = 0; t < 16; ++t) v[sp][t] = __ldcg(base + (size_t)sp * split_stride + t * kBlockM);
#pragma unroll
  for (int sp = 0; sp < S; ++sp)
#pragma unroll
    for (int t = 0; t < 16; ++t) acc[t] += v[sp][t];
}

// the staged skinny epilogue needs a fully valid feature tile and 16-byte aligned bf16 rows
__device__ __forceinline__ bool swap_staged_ok(const GemmParams& p, int a0) {
  if (p.out_f32 || a0 + kBlockM > p.Ma || (p.ldo & 7) || (reinterpret_cast<uintptr_t>(p.out) & 15)) return false;
  if (p.residual && ((p.ldr & 7) || (reinterpret_cast<uintptr_t>(p.residual) & 15))) return false;
  return true;
}

// Tall orientation, direct form: 16 consecutive features of one token row (thread = token).  Used by the split-K merge
// and for fp32 outputs; the unsplit bf16 path goes through the shared-memory staged form below.
template <int kAct>
__device__ __forceinline__ void epi_tall16(const GemmParams& p, int token, int f0, float (&acc)[16]) {
  if (token >= p.Ma || f0 >= p.Nb) return;
  const int nvalid = min(16, p.Nb - f0);
  const bool round_out = !p.out_f32 || p.round_bf16;
  if (p.bias) {
#pragma unroll
    for (int t = 0; t < 16; ++t) acc[t] += (t < nvalid) ? b2f(p.bias[f0 + t]) : 0.f;
  }
  if (round_out) {
#pragma unroll
    for (int t = 0; t < 16; t += 2) bf16r2(acc[t], acc[t + 1]);
  }
  if (kAct == GEMM_ACT_SILU_MUL) {
    uint16_t* o = reinterpret_cast<uint16_t*>(p.out) + (size_t)token * p.ldo + (f0 >> 1);
#pragma unroll
    for (int t = 0; t < 8; t += 2) {
      float s0 = fast_silu(acc[2 * t]), s1 = fast_silu(acc[2 * t + 2]);
      bf16r2(s0, s1);
      const uint32_t x = pack_bf16x2(s0 * acc[2 * t + 1], s1 * acc[2 * t + 3]);
      if (2 * t + 1 < nvalid) o[t] = (uint16_t)(x & 0xffffu);
      if (2 * t + 3 < nvalid) o[t + 1] = (uint16_t)(x >> 16);
    }
    return;
  }
  if (kAct != GEMM_ACT_NONE) {
#pragma unroll
    for (int t = 0; t < 16; t += 2) {
      acc[t] = apply_act<kAct>(acc[t]); acc[t + 1] = apply_act<kAct>(acc[t + 1]);
      bf16r2(acc[t], acc[t + 1]);
    }
  }
  if (p.residual) {
    const int rrow = p.res_mod > 0 ? (token % p.res_mod) : token;
    const bf16* r = p.residual + (size_t)rrow * p.ldr + f0;
    if (nvalid == 16 && (reinterpret_cast<uintptr_t>(r) & 15) == 0) {
      // two 16-byte loads instead of sixteen 2-byte ones (every lane reads a different row: each load instruction costs
      // 32 separate sectors either way)
      const uint4 u0 = *reinterpret_cast<const uint4*>(r), u1 = *reinterpret_cast<const uint4*>(r + 8);
      const uint32_t w[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
#pragma unroll
      for (int t = 0; t < 16; t += 2) {
        const float2 f = unpack_bf16x2(w[t >> 1]);
        acc[t] = f.x + acc[t]; acc[t + 1] = f.y + acc[t + 1];
        bf16r2(acc[t], acc[t + 1]);
      }
    } else {
#pragma unroll
      for (int t = 0; t < 16; t += 2) {
        if (t < nvalid) acc[t] = b2f(r[t]) + acc[t];
        if (t + 1 < nvalid) acc[t + 1] = b2f(r[t + 1]) + acc[t + 1];
        bf16r2(acc[t], acc[t + 1]);
      }
    }
  }
  if (p.out_f32) {
    float* o = reinterpret_cast<float*>(p.out) + (size_t)token * p.ldo + f0;
#pragma unroll
    for (int t = 0; t < 16; ++t)
      if (t < nvalid) o[t] = acc[t];
  } else {
    bf16* o = reinterpret_cast<bf16*>(p.out) + (size_t)token * p.ldo + f0;
    if (nvalid == 16 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
      st_v4(o, make_uint4(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]), pack_bf16x2(acc[4], acc[5]),
                          pack_bf16x2(acc[6], acc[7])));
      st_v4(o + 8, make_uint4(pack_bf16x2(acc[8], acc[9]), pack_bf16x2(acc[10], acc[11]), pack_bf16x2(acc[12], acc[13]),
                              pack_bf16x2(acc[14], acc[15])));
    } else {
#pragma unroll
      for (int t = 0; t < 16; t += 2) {
        const uint32_t x = pack_bf16x2(acc[t], acc[t + 1]);
        if (t < nvalid) reinterpret_cast<uint16_t*>(o)[t] = (uint16_t)(x & 0xffffu);
        if (t + 1 < nvalid) reinterpret_cast<uint16_t*>(o)[t + 1] = (uint16_t)(x >> 16);
      }
    }
  }
}

// Tall orientation, staged form.  Phase 1 (thread = token row): TMEM -> bias / activation -> bf16 -> this warp's
// [32 rows][BN] staging tile in shared memory (row pitch BN*2 + 16 bytes: conflict-free 16-byte accesses).
// Phase 2 (warp-cooperative): 16-byte pieces leave in row order, so a warp instruction covers whole 256-byte row
// segments of the output (and of the residual it adds): fully coalesced HBM traffic.
template <int BN, int kAct>
__device__ __forceinline__ void epi_tall_staged(const GemmParams& p, uint32_t taddr, uint8_t* stage_warp, int a0, int b0,
                                                int ew, int lane, uint64_t* release_bar = nullptr,
                                                uint32_t release_cluster_addr = 0) {
  constexpr int kPitch = BN * 2 + 16;
  constexpr int kOutCols = (kAct == GEMM_ACT_SILU_MUL) ? BN / 2 : BN;    // SiLU*mul halves the feature count
  uint8_t* my = stage_warp + lane * kPitch;
#pragma unroll 1
  for (int jc = 0; jc < BN; jc += 32) {
    uint32_t r[32];
    tmem_ld_32x32b_x32(taddr + jc, r);
    tmem_ld_wait();
    float v[32];
#pragma unroll
    for (int t = 0; t < 32; ++t) v[t] = __uint_as_float(r[t]);
    if (p.bias) {
      const int f0 = b0 + jc;
      if (f0 + 32 <= p.Nb) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const uint4 bu = *reinterpret_cast<const uint4*>(p.bias + f0 + c * 8);   // same address in every lane: one broadcast
          const uint32_t bw[4] = {bu.x, bu.y, bu.z, bu.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 f = unpack_bf16x2(bw[e]);
            v[c * 8 + 2 * e] += f.x; v[c * 8 + 2 * e + 1] += f.y;
          }
        }
      } else {
#pragma unroll
        for (int t = 0; t < 32; ++t) v[t] += (f0 + t < p.Nb) ? b2f(p.bias[f0 + t]) : 0.f;
      }
    }
    if (kAct != GEMM_ACT_NONE) {    // (no activation: the pack below is the rounding of the linear output)
#pragma unroll
      for (int t = 0; t < 32; t += 2) bf16r2(v[t], v[t + 1]);
    }
    if (kAct == GEMM_ACT_SILU_MUL) {
      uint32_t pk[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        float s0 = fast_silu(v[4 * t]), s1 = fast_silu(v[4 * t + 2]);
        bf16r2(s0, s1);
        pk[t] = pack_bf16x2(s0 * v[4 * t + 1], s1 * v[4 * t + 3]);
      }
      *reinterpret_cast<uint4*>(my + jc) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      *reinterpret_cast<uint4*>(my + jc + 16) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
    } else {
      uint32_t pk[16];
#pragma unroll
      for (int t = 0; t < 16; ++t) pk[t] = pack_bf16x2(apply_act<kAct>(v[2 * t]), apply_act<kAct>(v[2 * t + 1]));
#pragma unroll
      for (int c = 0; c < 4; ++c)
        *reinterpret_cast<uint4*>(my + jc * 2 + c * 16) = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
    }
  }
  if (release_bar) {   // persistent kernel: the accumulator buffer is free again as soon as it has been read
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(release_bar);
  }
  if (release_cluster_addr) {   // CTA-pair kernel: the accumulator's "drained" barrier lives in the leader CTA
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive_cluster(release_cluster_addr);
  }
  __syncwarp();
  constexpr int kChunksPerRow = kOutCols / 8;
  const int ob0 = (kAct == GEMM_ACT_SILU_MUL) ? (b0 >> 1) : b0;
  const int n_out = (kAct == GEMM_ACT_SILU_MUL) ? (p.Nb >> 1) : p.Nb;
  bf16* out = reinterpret_cast<bf16*>(p.out);
  // Interior tiles (all 32 rows and all columns valid, 16-byte aligned rows): branch-free, kBatch pieces per lane in
  // flight - the residual read is a dependent global round trip and with four epilogue warps per SM only the number of
  // loads in flight per lane hides it (the residual GEMMs of the SigLIP tower were epilogue-bound at 4 in flight).
  const bool interior = a0 + ew * 32 + 32 <= p.Ma && ob0 + kOutCols <= n_out && (p.ldo & 7) == 0 &&
                        (reinterpret_cast<uintptr_t>(out) & 15) == 0 && (ob0 & 7) == 0 &&
                        (!p.residual || ((p.ldr & 7) == 0 && (reinterpret_cast<uintptr_t>(p.residual) & 15) == 0));
  if (interior) {
    constexpr int kPieces = 32 * kChunksPerRow / 32;     // pieces per lane
    constexpr int kB = (kPieces % 8 == 0) ? 8 : (kPieces % 4 == 0) ? 4 : kPieces;
    static_assert(kPieces % kB == 0 && kB <= 12, "piece count per lane");
#pragma unroll 1
    for (int base = 0; base < kPieces; base += kB) {
      uint4 u[kB], ru[kB];
      size_t off[kB];
#pragma unroll
      for (int j = 0; j < kB; ++j) {
        const int idx = (base + j) * 32 + lane;
        const int rr_w = idx / kChunksPerRow, ch = idx % kChunksPerRow;
        const int token = a0 + ew * 32 + rr_w;
        u[j] = *reinterpret_cast<const uint4*>(stage_warp + rr_w * kPitch + ch * 16);
        off[j] = (size_t)token * p.ldo + ob0 + ch * 8;
        if (p.residual) {
          const int rr = p.res_mod > 0 ? (token % p.res_mod) : token;
          ru[j] = *reinterpret_cast<const uint4*>(p.residual + (size_t)rr * p.ldr + ob0 + ch * 8);
        }
      }
#pragma unroll
      for (int j = 0; j < kB; ++j) {
        if (p.residual) {
          uint32_t w[4] = {u[j].x, u[j].y, u[j].z, u[j].w};
          const uint32_t rw[4] = {ru[j].x, ru[j].y, ru[j].z, ru[j].w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 a = unpack_bf16x2(w[e]), b = unpack_bf16x2(rw[e]);
            w[e] = pack_bf16x2(b.x + a.x, b.y + a.y);
          }
          u[j] = make_uint4(w[0], w[1], w[2], w[3]);
        }
        st_v4(out + off[j], u[j]);
      }
    }
    return;
  }
#pragma unroll 4
  for (int idx = lane; idx < 32 * kChunksPerRow; idx += 32) {
    const int rrow_in_warp = idx / kChunksPerRow, ch = idx % kChunksPerRow;
    const int token = a0 + ew * 32 + rrow_in_warp;
    const int f = ob0 + ch * 8;
    if (token >= p.Ma || f >= n_out) continue;
    uint4 u = *reinterpret_cast<const uint4*>(stage_warp + rrow_in_warp * kPitch + ch * 16);
    bf16* o = out + (size_t)token * p.ldo + f;
    const bool full = (f + 8 <= n_out) && ((reinterpret_cast<uintptr_t>(o) & 15) == 0);
    if (p.residual) {
      const int rr = p.res_mod > 0 ? (token % p.res_mod) : token;
      const bf16* rp = p.residual + (size_t)rr * p.ldr + f;
      uint32_t w[4] = {u.x, u.y, u.z, u.w};
      if (full && ((reinterpret_cast<uintptr_t>(rp) & 15) == 0)) {
        const uint4 ru = *reinterpret_cast<const uint4*>(rp);
        const uint32_t rw[4] = {ru.x, ru.y, ru.z, ru.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 a = unpack_bf16x2(w[e]), b = unpack_bf16x2(rw[e]);
          w[e] = pack_bf16x2(b.x + a.x, b.y + a.y);
        }
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 a = unpack_bf16x2(w[e]);
          const float r0 = (f + 2 * e < n_out) ? b2f(rp[2 * e]) : 0.f, r1 = (f + 2 * e + 1 < n_out) ? b2f(rp[2 * e + 1]) : 0.f;
          w[e] = pack_bf16x2(r0 + a.x, r1 + a.y);
        }
      }
      u = make_uint4(w[0], w[1], w[2], w[3]);
    }
    if (full) {
      st_v4(o, u);
    } else {
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int e = 0; e < 8; ++e)
        if (f + e < n_out) reinterpret_cast<uint16_t*>(o)[e] = (uint16_t)(w[e >> 1] >> ((e & 1) * 16));
    }
  }
}

template <int BN, bool kSwap, int kAct, bool kDeep>
__global__ void __launch_bounds__(kThreads, (BN <= 128 && !kDeep) ? 2 : 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const GemmParams p) {
  using C = Cfg<BN, kDeep>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ float rstd_s[kSwap ? BN : 1];    // folded RMSNorm: per-token factor of this CTA's token columns
  __shared__ float sq_red[4][16];             // folded RMSNorm: cross-warp sum of squares
  // SWIZZLE_128B tiles need 1024-byte alignment.
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + C::kStages * C::kABytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* empty_bar = full_bar + C::kStages;
  uint64_t* accum_bar = empty_bar + C::kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int a0 = blockIdx.x * kBlockM;
  const int b0 = blockIdx.y * BN;
  const int split = blockIdx.z;
  const int kb_total = (p.K + kBlockK - 1) / kBlockK;
  const int kb_begin = split * p.kb_per_split;
  const int kb_end = min(kb_total, kb_begin + p.kb_per_split);
  const int num_kb = kb_end - kb_begin;  // host guarantees >= 1
  if (p.dbg == 1) return;
  unsigned long long* ts = (p.dbg_ts && threadIdx.x == 128 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) ? p.dbg_ts : nullptr;
  // bring-up (AHA_TS=1): every CTA's "accumulator complete" and "exit" on the global nanosecond timer (tail skew of a grid)
  unsigned long long* ts_all = (p.dbg_all && threadIdx.x == 128)
      ? p.dbg_all + 2 * (((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) : nullptr;
  dbg_stamp(ts, 0);

  // The weight operand (A when skinny, B when tall) does not depend on any earlier kernel: its first ring of tiles is
  // requested before griddepcontrol.wait, i.e. while the previous kernel on the stream may still be running (PDL).
  const int nst = (p.ring > 0 && p.ring < C::kStages) ? p.ring : C::kStages;   // stages actually used (bytes in flight)
  const int pre = min(nst, num_kb);
  const uint64_t pol_a = kSwap ? l2_policy_evict_first() : l2_policy_evict_last();
  const uint64_t pol_b = kSwap ? l2_policy_evict_last() : l2_policy_evict_first();
  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmap_a);
    prefetch_tensormap(&tmap_b);
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(accum_bar, 1);
    fence_barrier_init();
    for (int kb = 0; kb < pre; ++kb) {
      mbar_arrive_expect_tx(&full_bar[kb], C::kStageBytes);
      const int kc = (kb_begin + kb) * kBlockK;
      if (kSwap) tma_load_2d_hint(smem_a + kb * C::kABytes, &tmap_a, &full_bar[kb], kc, a0, pol_a);
      else tma_load_2d_hint(smem_b + kb * C::kBBytes, &tmap_b, &full_bar[kb], kc, b0, pol_b);
    }
  }
  if (warp == 2) tmem_alloc<C::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      pdl_wait();   // activations (and everything else written by earlier kernels) are visible from here on
      for (int kb = 0; kb < pre; ++kb) {
        const int kc = (kb_begin + kb) * kBlockK;
        if (kSwap) tma_load_2d_hint(smem_b + kb * C::kBBytes, &tmap_b, &full_bar[kb], kc, b0, pol_b);
        else tma_load_2d_hint(smem_a + kb * C::kABytes, &tmap_a, &full_bar[kb], kc, a0, pol_a);
      }
      for (int kb = pre; kb < num_kb; ++kb) {
        const int s = kb % nst;
        const uint32_t ph = (kb / nst) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1);
        mbar_arrive_expect_tx(&full_bar[s], C::kStageBytes);
        const int kc = (kb_begin + kb) * kBlockK;
        tma_load_2d_hint(smem_a + s * C::kABytes, &tmap_a, &full_bar[s], kc, a0, pol_a);
        tma_load_2d_hint(smem_b + s * C::kBBytes, &tmap_b, &full_bar[s], kc, b0, pol_b);
      }
      // every load of this CTA is in flight: keep HBM busy through the tail (reduction, epilogue, the next kernel's
      // ramp-up) by requesting this CTA's slice of the weights a later kernel will stream
      l2_prefetch_slice(p.pf, (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x,
                        gridDim.x * gridDim.y * gridDim.z);
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16_f32(kBlockM, BN);
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % nst;
        const uint32_t ph = (kb / nst) & 1;
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        const uint64_t da = make_kmajor_sw128_desc(smem_u32(smem_a + s * C::kABytes));
        const uint64_t db = make_kmajor_sw128_desc(smem_u32(smem_b + s * C::kBBytes));
#pragma unroll
        for (int k = 0; k < kBlockK / kUmmaK; ++k) {
          // advance 16 bf16 = 32 bytes inside the swizzle atom: +2 in the (addr >> 4) field
          umma_f16_ss(tmem_base, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0);
        }
        umma_commit(&empty_bar[s]);  // frees the smem slot once these MMAs have read it
      }
      umma_commit(accum_bar);  // accumulator complete
    }
  } else if (warp >= 4) {
    // ===== epilogue: warp (w & 3) owns TMEM lanes 32*(w&3) .. +31 =====
    const int ew = warp & 3;
    const int i = ew * 32 + lane;  // tile row
    dbg_stamp(ts, 1);
    pdl_wait();   // the epilogue reads residuals and writes outputs that earlier kernels may still be using
    dbg_stamp(ts, 2);
    if (kSwap && p.rs_partials) {
      // per-token RMSNorm factor of this CTA's token columns, while the main loop streams the weights
      for (int t = threadIdx.x - 128; t < BN; t += 128) {
        const int token = b0 + t;
        float sum = 0.f;
        if (token < p.Nb)
          for (int r = 0; r < p.rs_n; ++r) sum += __ldcg(p.rs_partials + (size_t)r * p.rs_ld + token);
        rstd_s[t] = rsqrtf(sum * p.rs_inv_cols + p.rs_eps);
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
    }
    mbar_wait(accum_bar, 0);
    tc_fence_after();
    dbg_stamp(ts, 3);
    if (ts_all) ts_all[0] = global_timer_ns();
    const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16);
    if (p.dbg == 2) {
    } else if (p.splits == 1) {
      if (kSwap) {
        const int feat = a0 + i;
        const bool fvalid = feat < p.Ma;
        const float bias = (p.bias && fvalid) ? b2f(p.bias[feat]) : 0.f;
        const bool staged = swap_staged_ok(p, a0) && !p.sumsq_out;
        // the operand ring is idle by now (accum_bar fired after the last MMA read it): 2 KB of it per warp stage the stores
        uint8_t* stage_warp = smem + ew * 2048;
#pragma unroll 1
        for (int jc = 0; jc < BN; jc += 16) {
          if (b0 + jc >= p.Nb) break;
          uint32_t r[16];
          tmem_ld_32x32b_x16(taddr + jc, r);
          tmem_ld_wait();
          dbg_stamp(ts, 8 + (jc >> 4) * 2);
          float acc[16];
#pragma unroll
          for (int t = 0; t < 16; ++t) acc[t] = __uint_as_float(r[t]);
          if (p.rs_partials) {
#pragma unroll
            for (int t = 0; t < 16; ++t) acc[t] *= rstd_s[jc + t];
          }
          if (staged) epi_swap16_staged<kAct>(p, stage_warp, a0, ew, lane, bias, b0 + jc, acc);
          else epi_swap16<kAct>(p, feat, fvalid, bias, i, b0 + jc, acc);
          if (p.sumsq_out) emit_sumsq16(p, sq_red, acc, fvalid, ew, lane, b0 + jc, blockIdx.x);
          dbg_stamp(ts, 9 + (jc >> 4) * 2);
        }
      } else if (p.out_f32) {
#pragma unroll 1
        for (int jc = 0; jc < BN; jc += 16) {
          uint32_t r[16];
          tmem_ld_32x32b_x16(taddr + jc, r);
          tmem_ld_wait();
          float acc[16];
#pragma unroll
          for (int t = 0; t < 16; ++t) acc[t] = __uint_as_float(r[t]);
          epi_tall16<kAct>(p, a0 + i, b0 + jc, acc);
        }
      } else {
        // the operand ring is idle by now (accum_bar fired after the last MMA read it): reuse it as the staging tile
        epi_tall_staged<BN, kAct>(p, taddr, smem + ew * 32 * (BN * 2 + 16), a0, b0, ew, lane);
      }
    } else {
      // Split-K: park this split's fp32 partial tile in the workspace ([column][128 rows]: a warp store is one
      // 128-byte line), then the last CTA to arrive at the tile's counter reduces in split order.
      const int tile_lin = blockIdx.y * gridDim.x + blockIdx.x;
      const int ncols = min(BN, p.Nb - b0);
      {   // zero this CTA's slice of the other counter bank (the next split launch on the stream counts in it)
        const int ncta = gridDim.x * gridDim.y * gridDim.z, cta = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        for (int z = cta * 128 + (int)threadIdx.x - 128; z < p.ws_zero_n; z += ncta * 128) p.ws_zero[z] = 0;
      }
      float* part = p.ws_partial + ((size_t)tile_lin * p.splits + split) * (size_t)(BN * kBlockM) + i;
#pragma unroll 1
      for (int jc = 0; jc < ncols; jc += 16) {
        uint32_t r[16];
        tmem_ld_32x32b_x16(taddr + jc, r);
        tmem_ld_wait();
#pragma unroll
        for (int t = 0; t < 16; ++t) __stcg(part + (jc + t) * kBlockM, __uint_as_float(r[t]));
      }
      dbg_stamp(ts, 4);
      const int feat = a0 + i;
      const bool fvalid = feat < p.Ma;
      uint16_t res[16];
      const bool pre_res = kSwap && p.residual != nullptr && fvalid && split * 16 < ncols;
      __threadfence();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      int* cnt = p.ws_counters + tile_lin;
      if (threadIdx.x == 128) atomicAdd(cnt, 1);   // arrive (result unused: a fire-and-forget reduction)
      // the residual of this split's first chunk does not depend on the other splits: its loads fly during the wait
      // (after the arrival - the fence in front of it would have waited for them)
      if (pre_res) load_residual16(p, feat, b0 + split * 16, res);
      if (threadIdx.x == 128) {
        // wait for the other splits of this tile.  All CTAs of a split grid are co-resident (the host keeps
        // tiles*splits within the device's resident-CTA capacity), so the wait cannot deadlock; it is bounded anyway.
        uint32_t spins = 0;
        while (*reinterpret_cast<volatile int*>(cnt) < p.splits) {
          if (++spins > (1u << 24)) { printf("aha_b200: split-K arrival wait timed out (tile %d)\n", tile_lin); __trap(); }
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      __threadfence();
      dbg_stamp(ts, 5);
      // 16-column chunk c is reduced (in split order 0..s-1: deterministic) and finished by split (c mod splits)
      const float* base = p.ws_partial + (size_t)tile_lin * p.splits * (size_t)(BN * kBlockM) + i;
      const float bias = (kSwap && p.bias && fvalid) ? b2f(p.bias[feat]) : 0.f;
      constexpr size_t kSplitStride = (size_t)BN * kBlockM;
#pragma unroll 1
      for (int jc = split * 16; jc < ncols; jc += 16 * p.splits) {
        float acc[16];
#pragma unroll
        for (int t = 0; t < 16; ++t) acc[t] = 0.f;
        const float* pp = base + jc * kBlockM;
        switch (p.splits) {
          case 2: reduce_splits16<2>(pp, kSplitStride, acc); break;
          case 3: reduce_splits16<3>(pp, kSplitStride, acc); break;
          case 4: reduce_splits16<4>(pp, kSplitStride, acc); break;
          case 5: reduce_splits16<5>(pp, kSplitStride, acc); break;
          case 6: reduce_splits16<6>(pp, kSplitStride, acc); break;
          default:
#pragma unroll 1
            for (int sp = 0; sp < p.splits; sp += 4) {     // four splits in flight at a time
              if (sp + 4 <= p.splits) reduce_splits16<4>(pp + (size_t)sp * kSplitStride, kSplitStride, acc);
              else for (int q = sp; q < p.splits; ++q) reduce_splits16<1>(pp + (size_t)q * kSplitStride, kSplitStride, acc);
            }
        }
        if (kSwap && p.rs_partials) {
#pragma unroll
          for (int t = 0; t < 16; ++t) acc[t] *= rstd_s[jc + t];
        }
        if (kSwap) epi_swap16<kAct>(p, feat, fvalid, bias, i, b0 + jc, acc, (pre_res && jc == split * 16) ? res : nullptr);
        else epi_tall16<kAct>(p, a0 + i, b0 + jc, acc);
        if (kSwap && p.sumsq_out) emit_sumsq16(p, sq_red, acc, fvalid, ew, lane, b0 + jc, blockIdx.x);
      }
    }
    tc_fence_before();
    dbg_stamp(ts, 6);
  }

  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<C::kTmemCols>(tmem_base);
  }
  dbg_stamp(ts, 7);
  if (ts_all) ts_all[1] = global_timer_ns();
}

// ---------------------------------------------------------------------------------------------------------------------
// Persistent tall GEMM (tokens >> 256: the SigLIP tower and the projector).  One CTA per SM walks over 128 x BN output
// tiles (n fastest, so the CTAs in flight share a few activation row blocks and the whole weight matrix in L2):
//   * the TMA producer and the MMA issuer run ahead across tile boundaries - the operand ring never drains,
//   * two accumulator buffers in TMEM (2 x BN columns): the epilogue warps drain tile t while tile t+1 is being
//     multiplied, so bias / GELU / residual / stores cost no tensor-pipe time,
//   * a 128 x 256 tile reads 48 KB of shared memory per 512 tensor-pipe cycles (96 B/clk) where the 128 x 128 tiles of
//     the one-shot kernel sit exactly at the 128 B/clk shared-memory limit.
// ---------------------------------------------------------------------------------------------------------------------
template <int BN>
struct PCfg {
  // 128 x 192 tiles consume a k-block in 384 tensor-pipe cycles: three stages are 0.6 us of look-ahead, less than an L2
  // round trip under load (the tower's qkv GEMM ran at 0.57 of the sustained peak where the 256-wide fc2 reaches 0.99)
  static constexpr int kStages = BN <= 192 ? 4 : 3;
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBBytes = BN * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  // BN = 128 is the variant that shares the SM with a weight-streaming decoder CTA of another stream: 256 TMEM columns,
  // ~115 KB of shared memory (its epilogue stages 64 columns at a time)
  // EIGHT epilogue warps: warps w and w + 4 share the TMEM lanes 32*(w & 3).. and take one half of the tile's columns
  // each (with four warps - one per scheduler, nothing to hide their latencies - the GELU and residual GEMMs of the
  // SigLIP tower were epilogue-bound: 8-9 us of epilogue against 5-8 us of main loop per tile)
  static constexpr int kEpiWarps = 8;
  static constexpr int kThreads = 128 + 32 * kEpiWarps;
  static constexpr int kEpiBN = BN / 2;
  static constexpr int kEpiWarpBytes = 32 * (kEpiBN * 2 + 16);
  static constexpr int kTmemCols = (BN <= 128) ? 256 : 512;   // two accumulators of BN columns, power of two
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiWarps * kEpiWarpBytes + 1024 + 256;
};

template <int BN, int kAct>
__global__ void __launch_bounds__(PCfg<BN>::kThreads, 1)
gemm_tall_persistent_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                            const GemmParams p, const int tiles_m, const int tiles_n) {
  using C = PCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + C::kStages * C::kABytes;
  uint8_t* smem_epi = smem + C::kStages * C::kStageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_epi + C::kEpiWarps * C::kEpiWarpBytes);
  uint64_t* empty_bar = full_bar + C::kStages;
  uint64_t* tmem_full = empty_bar + C::kStages;     // [2]
  uint64_t* tmem_empty = tmem_full + 2;             // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kb_total = (p.K + kBlockK - 1) / kBlockK;
  const int n_tiles = tiles_m * tiles_n;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmap_a);
    prefetch_tensormap(&tmap_b);
    for (int s = 0; s < C::kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], C::kEpiWarps); }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<C::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait();   // activations, residuals and the output buffer belong to earlier kernels until here

  if (warp == 0) {
    // ===== TMA producer: one continuous stream of k-blocks over all tiles of this CTA =====
    if (lane == 0) {
      const uint64_t pol_a = l2_policy_evict_first(), pol_b = l2_policy_evict_last();
      uint32_t it = 0;
      for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        const int a0 = (t / tiles_n) * kBlockM, b0 = (t % tiles_n) * BN;
        for (int kb = 0; kb < kb_total; ++kb, ++it) {
          const int s = it % C::kStages;
          mbar_wait(&empty_bar[s], ((it / C::kStages) & 1) ^ 1);
          mbar_arrive_expect_tx(&full_bar[s], C::kStageBytes);
          tma_load_2d_hint(smem_a + s * C::kABytes, &tmap_a, &full_bar[s], kb * kBlockK, a0, pol_a);
          tma_load_2d_hint(smem_b + s * C::kBBytes, &tmap_b, &full_bar[s], kb * kBlockK, b0, pol_b);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16_f32(kBlockM, BN);
      uint32_t it = 0, tl = 0;
      for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++tl) {
        const uint32_t buf = tl & 1;
        mbar_wait(&tmem_empty[buf], ((tl >> 1) & 1) ^ 1);      // the epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t tacc = tmem_base + buf * BN;
        for (int kb = 0; kb < kb_total; ++kb, ++it) {
          const int s = it % C::kStages;
          mbar_wait(&full_bar[s], (it / C::kStages) & 1);
          tc_fence_after();
          const uint64_t da = make_kmajor_sw128_desc(smem_u32(smem_a + s * C::kABytes));
          const uint64_t db = make_kmajor_sw128_desc(smem_u32(smem_b + s * C::kBBytes));
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k)
            umma_f16_ss(tacc, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0);
          umma_commit(&empty_bar[s]);
        }
        umma_commit(&tmem_full[buf]);
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue: warp (w & 3) owns TMEM lanes 32*(w&3) .. +31 of both accumulators, column half (w - 4) / 4 =====
    const int ew = warp & 3, half = (warp - 4) >> 2;
    uint32_t tl = 0;
    unsigned long long* ts = (p.dbg_ts && threadIdx.x == 128 && blockIdx.x == 0) ? p.dbg_ts : nullptr;
    dbg_stamp(ts, 0);
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++tl) {
      const int a0 = (t / tiles_n) * kBlockM, b0 = (t % tiles_n) * BN;
      const uint32_t buf = tl & 1;
      mbar_wait(&tmem_full[buf], (tl >> 1) & 1);
      tc_fence_after();
      if (tl < 7) dbg_stamp(ts, 1 + 2 * tl);
      const uint32_t taddr = tmem_base + buf * BN + half * C::kEpiBN + ((uint32_t)(ew * 32) << 16);
      epi_tall_staged<C::kEpiBN, kAct>(p, taddr, smem_epi + (warp - 4) * C::kEpiWarpBytes, a0, b0 + half * C::kEpiBN, ew, lane,
                                       &tmem_empty[buf]);
      if (tl < 7) dbg_stamp(ts, 2 + 2 * tl);
    }
    tc_fence_before();
  }

  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<C::kTmemCols>(tmem_base);
  }
}


// ---------------------------------------------------------------------------------------------------------------------
// Persistent skinny ("swap") GEMM for batched decoder steps: 256 < tokens <= 832 (e.g. 8 streams x 49 frame tokens = 392).
// A = weights (128 feature rows per tile), B = tokens in column tiles of 208 (392 tokens = 2 tiles, 94 % of the MMA N
// dimension used; the tall orientation would pad 392 token rows to 512).  One CTA per SM walks over work items
// (feature tile, k split, token tile) - the token tiles of a feature tile are neighbours in the item order, so the CTAs
// that share a weight tile pull it at the same time and HBM sees it once.  Two 208-column accumulators in TMEM: the
// eight epilogue warps (two groups of four: columns [0,112) and [112,208)) drain item t under the main loop of item t+1.
// Split-K items park their fp32 partial in the L2 workspace and bump the counter of their (tile, column group); the
// LAST arriver reduces all splits in split order (deterministic) and runs the epilogue - nobody waits, so the items of a
// tile need not be co-resident.  The counters reset themselves.  Same epilogues as the one-shot skinny kernel: bias,
// SiLU*mul on interleaved gate|up rows, residual, folded RMSNorm on both sides (GemmCall).
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kPsBN = 208;
// BN = 208: two token tiles cover 392 tokens (the big matrices: 296 feature tiles of gate|up fill the device four times);
// BN = 112: four token tiles - the matrices with 28 / 36 feature tiles (o_proj, down, qkv) then have 112 / 144 work items
// WITHOUT splitting K, i.e. without parking and merging partial tiles (in the step those merges cost more than the main
// loops they balanced: o_proj 35 us, down 69 us, qkv 28 us split 2 ways).
template <int BN, bool kPair = false>
struct PSCfgT {
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  // kPair: the CTA loads only its half of the token rows, the tensor cores read the other half from the partner CTA
  static constexpr int kBRows = kPair ? BN / 2 : BN;
  static constexpr int kBBytes = kBRows * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = kPair ? (BN > 128 ? 7 : 9) : (BN > 128 ? 5 : 7);
  static constexpr int kEpiWarps = 8;
  static constexpr int kThreads = 128 + 32 * kEpiWarps;
  static constexpr int kEpiWarpBytes = 1280;
  static constexpr int kMaxTokens = 896;                  // 4 column tiles of 208, 8 of 112 (832 tokens)
  static constexpr int kAccStride = 256;                  // TMEM columns between the two accumulators
  static constexpr int kGroupCols = BN > 128 ? 112 : 64;  // columns of epilogue group 0 (group 1: the rest)
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiWarps * kEpiWarpBytes + kMaxTokens * 4 + 256 + 1024;
  static_assert(kBBytes % 1024 == 0 && BN % 16 == 0 && kSmemBytes <= 232448 - 1100, "tile shape");
};
using PSCfg = PSCfgT<kPsBN>;

__device__ __forceinline__ void reduce_splits_dyn(const float* pp, size_t stride, int splits, float (&acc)[16]) {
  switch (splits) {
    case 2: reduce_splits16<2>(pp, stride, acc); break;
    case 3: reduce_splits16<3>(pp, stride, acc); break;
    case 4: reduce_splits16<4>(pp, stride, acc); break;
    case 5: reduce_splits16<5>(pp, stride, acc); break;
    case 6: reduce_splits16<6>(pp, stride, acc); break;
    default:
#pragma unroll 1
      for (int sp = 0; sp < splits; sp += 4) {     // four splits in flight at a time
        if (sp + 4 <= splits) reduce_splits16<4>(pp + (size_t)sp * stride, stride, acc);
        else for (int q = sp; q < splits; ++q) reduce_splits16<1>(pp + (size_t)q * stride, stride, acc);
      }
  }
}

template <int kAct, int BN, bool kPair>
__global__ void __launch_bounds__(PSCfgT<BN, kPair>::kThreads, 1)
gemm_swap_persistent_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                            const GemmParams p, const int tiles_a, const int tiles_b) {
  using C = PSCfgT<BN, kPair>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ float sq_red[2][2][4][16];      // [epilogue group][chunk parity][warp][token]
  __shared__ int last_flag[2];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + C::kStages * C::kABytes;
  uint8_t* smem_epi = smem + C::kStages * C::kStageBytes;
  float* rstd_all = reinterpret_cast<float*>(smem_epi + C::kEpiWarps * C::kEpiWarpBytes);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(rstd_all + C::kMaxTokens);
  uint64_t* empty_bar = full_bar + C::kStages;
  uint64_t* tmem_full = empty_bar + C::kStages;     // [2]
  uint64_t* tmem_empty = tmem_full + 2;             // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kb_total = (p.K + kBlockK - 1) / kBlockK;
  // kPair: the two CTAs of a cluster (one TPC) take two neighbouring feature tiles and the SAME token tile; `tiles_a` then
  // counts PAIRS of feature tiles, work items are dealt to clusters, the leader (rank 0) issues tcgen05.mma.cta_group::2.
  const uint32_t rank = kPair ? cluster_ctarank() : 0u;
  const int wg_id = kPair ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;        // work-group (CTA or CTA pair) index
  const int wg_n = kPair ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  // Work items of this CTA, in order.  Two decompositions:
  //  * fixed split (p.sk_ranks == 0): item -> (feature tile, split, token tile), token tile fastest (the readers of one
  //    weight tile are neighbours), items dealt round-robin over the CTAs;
  //  * stream-K (p.sk_ranks = R): the CTAs of one token tile (blockIdx.x % tiles_b) share the tiles_a x kb_total k-block
  //    units of that token tile in R equal contiguous ranges, whatever the tile boundaries: every SM carries the same number
  //    of k-blocks (down at 392 tokens has 56 tiles of 296 k-blocks - any fixed split leaves a third of the SMs idle or
  //    runs a second, nearly empty round).  Rank r of EVERY token tile walks the same range at the same time, so a weight
  //    tile is still pulled by neighbours together.  A tile that spans several ranks is merged by its last arriver; the
  //    rank's position among the tile's contributors is its split index (fixed summation order).
  struct Item { int a0, b0, kb0, kb1, tile_lin, split, nsplit; };
  const int sk_R = p.sk_ranks;
  const long long sk_U = (long long)tiles_a * kb_total;
  const int sk_r = wg_id / tiles_b, sk_tb = wg_id % tiles_b;     // (stream-K is not used with pairs)
  const long long sk_u0 = sk_R ? (sk_r * sk_U) / sk_R : 0, sk_u1 = sk_R ? ((sk_r + 1) * sk_U) / sk_R : 0;
  auto sk_rank_of = [&](long long u) { return (int)(((u + 1) * sk_R + sk_U - 1) / sk_U) - 1; };
  auto get_item = [&](int i, Item& it) -> bool {
    if (sk_R) {
      if (sk_u1 <= sk_u0) return false;
      const int f = (int)(sk_u0 / kb_total) + i;
      if (f > (int)((sk_u1 - 1) / kb_total)) return false;
      const long long t0 = (long long)f * kb_total, t1 = t0 + kb_total;
      it.a0 = f * kBlockM; it.b0 = sk_tb * BN;
      it.kb0 = (int)(max(sk_u0, t0) - t0); it.kb1 = (int)(min(sk_u1, t1) - t0);
      it.tile_lin = f * tiles_b + sk_tb;
      const int rf = sk_rank_of(t0), rl = sk_rank_of(t1 - 1);
      it.split = sk_r - rf; it.nsplit = rl - rf + 1;
      return true;
    }
    const int g = wg_id + i * wg_n;
    if (g >= tiles_a * tiles_b * p.splits) return false;
    const int tb = g % tiles_b, r = g / tiles_b;
    it.split = r % p.splits; it.nsplit = p.splits;
    const int f = kPair ? 2 * (r / p.splits) + (int)rank : r / p.splits;
    it.a0 = f * kBlockM; it.b0 = tb * BN;
    it.kb0 = it.split * p.kb_per_split; it.kb1 = min(kb_total, it.kb0 + p.kb_per_split);
    it.tile_lin = f * tiles_b + tb;
    return true;
  };
  const uint64_t pol_a = l2_policy_evict_first(), pol_b = l2_policy_evict_last();
  int pre = 0;
  Item first{};
  const bool have_first = get_item(0, first);
  if (have_first) pre = min(C::kStages, first.kb1 - first.kb0);
  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmap_a);
    prefetch_tensormap(&tmap_b);
    for (int s = 0; s < C::kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    // (pairs: the leader's tmem_empty barriers collect the epilogue warps of both CTAs, its full barriers the bytes of both)
    for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], kPair ? 2 * C::kEpiWarps : C::kEpiWarps); }
    fence_barrier_init();
    // the weight tiles of the first ring do not depend on any earlier kernel: requested before griddepcontrol.wait (PDL)
    if (pre > 0 && !kPair) {
      for (int kb = 0; kb < pre; ++kb) {
        mbar_arrive_expect_tx(&full_bar[kb], C::kStageBytes);
        tma_load_2d_hint(smem_a + kb * C::kABytes, &tmap_a, &full_bar[kb], (first.kb0 + kb) * kBlockK, first.a0, pol_a);
      }
    }
  }
  if (warp == 2) { if (kPair) tmem_alloc_pair<512>(tmem_slot); else tmem_alloc<512>(tmem_slot); }
  tc_fence_before();
  if (kPair) cluster_sync_all();   // both CTAs' barriers are initialised before a remote arrival / TMA completion can reach them
  else __syncthreads();
  tc_fence_after();
  if (kPair && warp == 0 && lane == 0 && pre > 0) {
    // (pairs: the first ring of weight tiles can only be requested once the LEADER's barriers exist - after the cluster sync,
    // still before griddepcontrol.wait)
    for (int kb = 0; kb < pre; ++kb) {
      if (rank == 0) mbar_arrive_expect_tx(&full_bar[kb], 2u * (uint32_t)C::kStageBytes);
      tma_load_2d_pair_hint(smem_a + kb * C::kABytes, &tmap_a, &full_bar[kb], (first.kb0 + kb) * kBlockK, first.a0, pol_a);
    }
  }
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();

  if (warp == 0) {
    // ===== TMA producer: one continuous stream of k-blocks over all items of this CTA =====
    if (lane == 0) {
      pdl_wait();
      uint32_t cnt = 0;
      Item it;
      for (int ii = 0; get_item(ii, it); ++ii) {
        const int a0 = it.a0, b0 = it.b0;
        for (int kb = it.kb0; kb < it.kb1; ++kb, ++cnt) {
          const int s = cnt % C::kStages;
          if (kPair) {
            const int bh = b0 + (int)rank * C::kBRows;       // this CTA's half of the token rows
            if (cnt >= (uint32_t)pre) {
              mbar_wait(&empty_bar[s], ((cnt / C::kStages) & 1) ^ 1);
              if (rank == 0) mbar_arrive_expect_tx(&full_bar[s], 2u * (uint32_t)C::kStageBytes);     // both CTAs' bytes
              tma_load_2d_pair_hint(smem_a + s * C::kABytes, &tmap_a, &full_bar[s], kb * kBlockK, a0, pol_a);
            }
            tma_load_2d_pair_hint(smem_b + s * C::kBBytes, &tmap_b, &full_bar[s], kb * kBlockK, bh, pol_b);
          } else if (cnt < (uint32_t)pre) {           // first ring: the weight tile is already on its way
            tma_load_2d_hint(smem_b + s * C::kBBytes, &tmap_b, &full_bar[s], kb * kBlockK, b0, pol_b);
          } else {
            mbar_wait(&empty_bar[s], ((cnt / C::kStages) & 1) ^ 1);
            mbar_arrive_expect_tx(&full_bar[s], C::kStageBytes);
            tma_load_2d_hint(smem_a + s * C::kABytes, &tmap_a, &full_bar[s], kb * kBlockK, a0, pol_a);
            tma_load_2d_hint(smem_b + s * C::kBBytes, &tmap_b, &full_bar[s], kb * kBlockK, b0, pol_b);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if ((!kPair || rank == 0) && elect_one()) {      // pairs: the leader's elected lane drives the tensor cores of both SMs
      constexpr uint32_t idesc = make_idesc_bf16_f32(kPair ? 2 * kBlockM : kBlockM, BN);
      uint32_t cnt = 0, tl = 0;
      Item it;
      for (int ii = 0; get_item(ii, it); ++ii, ++tl) {
        const int kb0 = it.kb0, kb1 = it.kb1;
        const uint32_t buf = tl & 1;
        mbar_wait(&tmem_empty[buf], ((tl >> 1) & 1) ^ 1);      // the epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t tacc = tmem_base + buf * C::kAccStride;
        for (int kb = kb0; kb < kb1; ++kb, ++cnt) {
          const int s = cnt % C::kStages;
          mbar_wait(&full_bar[s], (cnt / C::kStages) & 1);
          tc_fence_after();
          const uint64_t da = make_kmajor_sw128_desc(smem_u32(smem_a + s * C::kABytes));
          const uint64_t db = make_kmajor_sw128_desc(smem_u32(smem_b + s * C::kBBytes));
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k)
            if (kPair) umma_f16_ss_pair(tacc, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb != kb0 || k != 0) ? 1u : 0u);
            else umma_f16_ss(tacc, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb != kb0 || k != 0) ? 1u : 0u);
          if (kPair) umma_commit_pair(&empty_bar[s], 0b11);    // frees the stage in both CTAs
          else umma_commit(&empty_bar[s]);
        }
        if (kPair) umma_commit_pair(&tmem_full[buf], 0b11);
        else umma_commit(&tmem_full[buf]);
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue: warp (w & 3) owns TMEM lanes 32*(w&3).., group (w - 4) / 4 owns columns [0,112) or [112,208) =====
    const int ew = warp & 3, grp = (warp - 4) >> 2;
    const int i = ew * 32 + lane;
    const int col_begin = grp ? C::kGroupCols : 0, col_end = grp ? BN : C::kGroupCols;
    const int bar_id = 1 + grp;
    uint8_t* stage_warp = smem_epi + (warp - 4) * C::kEpiWarpBytes;
    pdl_wait();   // residuals, norm partials and the output buffer belong to earlier kernels until here
    if (p.rs_partials) {
      for (int t = threadIdx.x - 128; t < C::kMaxTokens; t += 32 * C::kEpiWarps) {
        float sum = 0.f;
        if (t < p.Nb) {
          // partial sums in tile order (the order the one-shot kernel adds them in), eight loads in flight at a time:
          // one dependent L2 round trip per partial would take longer than the first item's main loop
#pragma unroll 1
          for (int r0 = 0; r0 < p.rs_n; r0 += 8) {
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = (r0 + j < p.rs_n) ? __ldcg(p.rs_partials + (size_t)(r0 + j) * p.rs_ld + t) : 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) sum += v[j];
          }
        }
        rstd_all[t] = (t < p.Nb) ? rsqrtf(sum * p.rs_inv_cols + p.rs_eps) : 0.f;
      }
      asm volatile("bar.sync 3, 256;" ::: "memory");
    }
    uint32_t tl = 0, sq_cnt = 0;
    Item it;
    for (int ii = 0; get_item(ii, it); ++ii, ++tl) {
      const int a0 = it.a0, b0 = it.b0, tile_lin = it.tile_lin, split = it.split, nsplit = it.nsplit;
      const uint32_t buf = tl & 1;
      mbar_wait(&tmem_full[buf], (tl >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + buf * C::kAccStride + ((uint32_t)(ew * 32) << 16);
      const int feat = a0 + i;
      const bool fvalid = feat < p.Ma;
      const float bias = (p.bias && fvalid) ? b2f(p.bias[feat]) : 0.f;
      const bool staged = swap_staged_ok(p, a0);
      const int ncols = min(BN, p.Nb - b0);
      const int my_end = min(col_end, ncols);
      if (kPair && a0 >= p.Ma) {     // the partner of an odd last feature tile: zeros in the accumulator, nothing to store
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(mapa_shared(smem_u32(&tmem_empty[buf]), 0));
        continue;
      }
      if (nsplit == 1) {
#pragma unroll 1
        for (int jc = col_begin; jc < my_end; jc += 16) {
          uint32_t r[16];
          tmem_ld_32x32b_x16(taddr + jc, r);
          tmem_ld_wait();
          float acc[16];
#pragma unroll
          for (int t = 0; t < 16; ++t) acc[t] = __uint_as_float(r[t]);
          if (p.rs_partials) {
#pragma unroll
            for (int t = 0; t < 16; ++t) acc[t] *= rstd_all[b0 + jc + t];
          }
          if (staged) {
            float (*sq)[16] = sq_red[grp][sq_cnt++ & 1];     // running parity: consecutive chunks never share a buffer
            epi_swap16_staged<kAct>(p, stage_warp, a0, ew, lane, bias, b0 + jc, acc, p.sumsq_out ? sq[ew] : nullptr);
            if (p.sumsq_out) finish_sumsq16(p, sq, ew, lane, b0 + jc, a0 / kBlockM, bar_id);
          } else {
            epi_swap16<kAct>(p, feat, fvalid, bias, i, b0 + jc, acc);
            if (p.sumsq_out) emit_sumsq16(p, sq_red[grp][0], acc, fvalid, ew, lane, b0 + jc, a0 / kBlockM, bar_id);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) { if (kPair) mbar_arrive_cluster(mapa_shared(smem_u32(&tmem_empty[buf]), 0)); else mbar_arrive(&tmem_empty[buf]); }
      } else {
        constexpr size_t kSplitStride = (size_t)BN * kBlockM;
        float* part = p.ws_partial + ((size_t)tile_lin * p.splits + split) * kSplitStride + i;
#pragma unroll 1
        for (int jc = col_begin; jc < my_end; jc += 16) {
          uint32_t r[16];
          tmem_ld_32x32b_x16(taddr + jc, r);
          tmem_ld_wait();
#pragma unroll
          for (int t = 0; t < 16; ++t) __stcg(part + (jc + t) * kBlockM, __uint_as_float(r[t]));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) { if (kPair) mbar_arrive_cluster(mapa_shared(smem_u32(&tmem_empty[buf]), 0)); else mbar_arrive(&tmem_empty[buf]); }     // the accumulator is free: the rest works on the parked copy
        __threadfence();
        asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
        if (ew == 0 && lane == 0) {
          int* cnt = p.ws_pcounters + tile_lin * 2 + grp;
          const int old = atomicAdd(cnt, 1);
          const int last = (old == nsplit - 1) ? 1 : 0;
          if (last) *cnt = 0;                              // self-resetting: nobody else touches it in this launch
          last_flag[grp] = last;
        }
        asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
        if (last_flag[grp]) {
          __threadfence();
          const float* base = p.ws_partial + (size_t)tile_lin * p.splits * kSplitStride + i;
#pragma unroll 1
          for (int jc = col_begin; jc < my_end; jc += 16) {
            float acc[16];
#pragma unroll
            for (int t = 0; t < 16; ++t) acc[t] = 0.f;
            reduce_splits_dyn(base + jc * kBlockM, kSplitStride, nsplit, acc);
            if (p.rs_partials) {
#pragma unroll
              for (int t = 0; t < 16; ++t) acc[t] *= rstd_all[b0 + jc + t];
            }
            if (staged) {
              float (*sq)[16] = sq_red[grp][sq_cnt++ & 1];     // running parity: consecutive chunks never share a buffer
              epi_swap16_staged<kAct>(p, stage_warp, a0, ew, lane, bias, b0 + jc, acc, p.sumsq_out ? sq[ew] : nullptr);
              if (p.sumsq_out) finish_sumsq16(p, sq, ew, lane, b0 + jc, a0 / kBlockM, bar_id);
            } else {
              epi_swap16<kAct>(p, feat, fvalid, bias, i, b0 + jc, acc);
              if (p.sumsq_out) emit_sumsq16(p, sq_red[grp][0], acc, fvalid, ew, lane, b0 + jc, a0 / kBlockM, bar_id);
            }
          }
        }
      }
    }
    tc_fence_before();
  }

  if (kPair) {
    __syncwarp();
    cluster_sync_all();    // the partner's tensor-core reads of this CTA's shared memory / writes to its TMEM are over
  } else {
    __syncthreads();
  }
  if (warp == 2) {
    tc_fence_after();
    if (kPair) tmem_dealloc_pair<512>(tmem_base); else tmem_dealloc<512>(tmem_base);
  }
}


// ---------------------------------------------------------------------------------------------------------------------
// The same batched decoder GEMM on CTA PAIRS (tcgen05.mma.cta_group::2, M = 256).  The single-CTA persistent kernel above
// is bound by L2 -> SM bandwidth: a 128 x 208 tile pulls 42 KB per 416 tensor-pipe cycles (101 B/clk per SM, 15 KB/clk
// over the chip against the ~6.3 KB/clk the L2 slices deliver) - measured 610 TFLOP/s = 0.43 of the cuBLAS peak.
// A pair of CTAs (the two SMs of a TPC) takes two neighbouring weight tiles and ALL tokens of a 416-token group: every
// CTA loads its own 128 weight rows (16 KB) and only HALF of the token rows of each 208-token MMA (2 x 104 rows, 26 KB)
// per k-block, the tensor cores read the other half from the partner's shared memory: 42 KB per 832 cycles = 50 B/clk.
// Per CTA the accumulator is 128 features x 416 tokens (two 208-column halves, 416 of the 512 TMEM columns): there is no
// second accumulator, so the epilogue of an item is not hidden under the next item's main loop; epilogue group g (four
// warps) drains token half g.  Items = (weight-tile pair, k split, token group); split-K merges are last-arriver as above.
// ---------------------------------------------------------------------------------------------------------------------
struct PPCfg {
  static constexpr int kStages = 5;
  static constexpr int kHalfRows = kPsBN / 2;                       // token rows of one MMA half held by one CTA
  static constexpr int kABytes = kBlockM * kBlockK * 2;             // 16384
  static constexpr int kBHalfBytes = kHalfRows * kBlockK * 2;       // 13312 (13 swizzle atoms)
  static constexpr int kBBytes = 2 * kBHalfBytes;
  static constexpr int kStageBytes = kABytes + kBBytes;             // 43008 per CTA
  static constexpr int kGroupTokens = 2 * kPsBN;                    // 416
  static constexpr int kEpiWarps = 8;
  static constexpr int kThreads = 128 + 32 * kEpiWarps;
  static constexpr int kEpiWarpBytes = 1280;
  static constexpr int kMaxTokens = 832;
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiWarps * kEpiWarpBytes + kMaxTokens * 4 + 256 + 1024;
};

template <int kAct>
__global__ void __launch_bounds__(PPCfg::kThreads, 1)
gemm_swap_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                      const GemmParams p, const int pairs, const int groups) {
  using C = PPCfg;
  extern __shared__ uint8_t smem_raw[];
  __shared__ float sq_red[2][2][4][16];      // [epilogue group][chunk parity][warp][token]
  __shared__ int last_flag[2];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + C::kStages * C::kABytes;
  uint8_t* smem_epi = smem + C::kStages * C::kStageBytes;
  float* rstd_all = reinterpret_cast<float*>(smem_epi + C::kEpiWarps * C::kEpiWarpBytes);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(rstd_all + C::kMaxTokens);   // used in the leader only
  uint64_t* empty_bar = full_bar + C::kStages;
  uint64_t* tmem_full = empty_bar + C::kStages;     // [1]
  uint64_t* tmem_empty = tmem_full + 1;             // [1], counted in the leader only (both CTAs' epilogue warps arrive)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();           // 0 = leader
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const int kb_total = (p.K + kBlockK - 1) / kBlockK;
  const int n_items = pairs * groups * p.splits;
  // item -> (weight-tile pair, split, token group); this CTA's weight tile is 2 * pair + rank
  auto decode = [&](int it, int& a0, int& b0, int& kb0, int& kb1, int& tile_lin, int& split, int& halves) {
    const int g = it % groups, r = it / groups;
    split = r % p.splits;
    const int pr = r / p.splits;
    a0 = (2 * pr + (int)rank) * kBlockM; b0 = g * C::kGroupTokens;
    kb0 = split * p.kb_per_split; kb1 = min(kb_total, kb0 + p.kb_per_split);
    tile_lin = (2 * pr + (int)rank) * groups + g;
    halves = (p.Nb - b0 > kPsBN) ? 2 : 1;
  };
  const uint64_t pol_a = l2_policy_evict_first(), pol_b = l2_policy_evict_last();
  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmap_a);
    prefetch_tensormap(&tmap_b);
    for (int s = 0; s < C::kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(tmem_full, 1);
    mbar_init(tmem_empty, 2 * C::kEpiWarps);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc_pair<512>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();        // both CTAs' barriers are initialised before any remote arrival / TMA completion can reach them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();

  if (warp == 0) {
    // ===== TMA producer (both CTAs): own weight rows + own half of the token rows, counted by the leader's barrier =====
    if (lane == 0) {
      // The weight tiles do not depend on any earlier kernel: the first ring of them (first item only) is requested
      // before griddepcontrol.wait, i.e. while the previous kernel on the stream may still be running (PDL).
      int pre = 0;
      if (cluster_id < n_items) {
        int a0, b0, kb0, kb1, tl_, sp_, halves;
        decode(cluster_id, a0, b0, kb0, kb1, tl_, sp_, halves);
        pre = min(C::kStages, kb1 - kb0);
        const uint32_t stage_tx = 2u * (uint32_t)(C::kABytes + halves * C::kBHalfBytes);     // both CTAs' bytes
        for (int kb = 0; kb < pre; ++kb) {
          if (rank == 0) mbar_arrive_expect_tx(&full_bar[kb], stage_tx);
          tma_load_2d_pair_hint(smem_a + kb * C::kABytes, &tmap_a, &full_bar[kb], (kb0 + kb) * kBlockK, a0, pol_a);
        }
      }
      pdl_wait();
      uint32_t cnt = 0;
      for (int it = cluster_id; it < n_items; it += n_clusters) {
        int a0, b0, kb0, kb1, tl_, sp_, halves;
        decode(it, a0, b0, kb0, kb1, tl_, sp_, halves);
        const uint32_t stage_tx = 2u * (uint32_t)(C::kABytes + halves * C::kBHalfBytes);
        for (int kb = kb0; kb < kb1; ++kb, ++cnt) {
          const int s = cnt % C::kStages;
          if (cnt >= (uint32_t)pre) {
            mbar_wait(&empty_bar[s], ((cnt / C::kStages) & 1) ^ 1);
            if (rank == 0) mbar_arrive_expect_tx(&full_bar[s], stage_tx);
            tma_load_2d_pair_hint(smem_a + s * C::kABytes, &tmap_a, &full_bar[s], kb * kBlockK, a0, pol_a);
          }
          for (int h = 0; h < halves; ++h)
            tma_load_2d_pair_hint(smem_b + s * C::kBBytes + h * C::kBHalfBytes, &tmap_b, &full_bar[s], kb * kBlockK,
                                  b0 + h * kPsBN + (int)rank * C::kHalfRows, pol_b);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the leader's elected lane drives the tensor cores of both SMs =====
    if (rank == 0 && elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16_f32(2 * kBlockM, kPsBN);
      uint32_t cnt = 0, tl = 0;
      for (int it = cluster_id; it < n_items; it += n_clusters, ++tl) {
        int a0, b0, kb0, kb1, tl_, sp_, halves;
        decode(it, a0, b0, kb0, kb1, tl_, sp_, halves);
        mbar_wait(tmem_empty, (tl & 1) ^ 1);       // both CTAs' epilogues have drained the accumulator
        tc_fence_after();
        for (int kb = kb0; kb < kb1; ++kb, ++cnt) {
          const int s = cnt % C::kStages;
          mbar_wait(&full_bar[s], (cnt / C::kStages) & 1);
          tc_fence_after();
          const uint64_t da = make_kmajor_sw128_desc(smem_u32(smem_a + s * C::kABytes));
          for (int h = 0; h < halves; ++h) {
            const uint64_t db = make_kmajor_sw128_desc(smem_u32(smem_b + s * C::kBBytes + h * C::kBHalfBytes));
#pragma unroll
            for (int k = 0; k < kBlockK / kUmmaK; ++k)
              umma_f16_ss_pair(tmem_base + h * kPsBN, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc,
                               (kb != kb0 || k != 0) ? 1u : 0u);
          }
          umma_commit_pair(&empty_bar[s], 0b11);    // frees the stage in both CTAs
        }
        umma_commit_pair(tmem_full, 0b11);
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue (both CTAs): warp (w & 3) owns TMEM lanes 32*(w&3).., group (w - 4) / 4 owns token half g =====
    const int ew = warp & 3, grp = (warp - 4) >> 2;
    const int i = ew * 32 + lane;
    const int bar_id = 1 + grp;
    uint8_t* stage_warp = smem_epi + (warp - 4) * C::kEpiWarpBytes;
    const uint32_t tmem_empty_leader = mapa_shared(smem_u32(tmem_empty), 0);
    pdl_wait();   // residuals, norm partials and the output buffer belong to earlier kernels until here
    if (p.rs_partials) {
      for (int t = threadIdx.x - 128; t < C::kMaxTokens; t += 32 * C::kEpiWarps) {
        float sum = 0.f;
        if (t < p.Nb) {
          // partial sums in tile order (the order the one-shot kernel adds them in), eight loads in flight at a time:
          // one dependent L2 round trip per partial would take longer than the first item's main loop
#pragma unroll 1
          for (int r0 = 0; r0 < p.rs_n; r0 += 8) {
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = (r0 + j < p.rs_n) ? __ldcg(p.rs_partials + (size_t)(r0 + j) * p.rs_ld + t) : 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) sum += v[j];
          }
        }
        rstd_all[t] = (t < p.Nb) ? rsqrtf(sum * p.rs_inv_cols + p.rs_eps) : 0.f;
      }
      asm volatile("bar.sync 3, 256;" ::: "memory");
    }
    uint32_t tl = 0, sq_cnt = 0;
    for (int it = cluster_id; it < n_items; it += n_clusters, ++tl) {
      int a0, b0, kb0, kb1, tile_lin, split, halves;
      decode(it, a0, b0, kb0, kb1, tile_lin, split, halves);
      mbar_wait(tmem_full, tl & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16);
      const int feat = a0 + i;
      const bool fvalid = feat < p.Ma;
      const bool tile_valid = a0 < p.Ma;
      const float bias = (p.bias && fvalid) ? b2f(p.bias[feat]) : 0.f;
      const bool staged = swap_staged_ok(p, a0);
      const int ncols = min(C::kGroupTokens, p.Nb - b0);
      const int col_begin = grp * kPsBN, my_end = min(col_begin + kPsBN, ncols);
      if (p.splits == 1) {
#pragma unroll 1
        for (int jc = col_begin; jc < my_end; jc += 16) {
          uint32_t r[16];
          tmem_ld_32x32b_x16(taddr + jc, r);
          tmem_ld_wait();
          float acc[16];
#pragma unroll
          for (int t = 0; t < 16; ++t) acc[t] = __uint_as_float(r[t]);
          if (p.rs_partials) {
#pragma unroll
            for (int t = 0; t < 16; ++t) acc[t] *= rstd_all[b0 + jc + t];
          }
          if (staged) {
            float (*sq)[16] = sq_red[grp][sq_cnt++ & 1];     // running parity: consecutive chunks never share a buffer
            epi_swap16_staged<kAct>(p, stage_warp, a0, ew, lane, bias, b0 + jc, acc, p.sumsq_out ? sq[ew] : nullptr);
            if (p.sumsq_out) finish_sumsq16(p, sq, ew, lane, b0 + jc, a0 / kBlockM, bar_id);
          } else {
            epi_swap16<kAct>(p, feat, fvalid, bias, i, b0 + jc, acc);
            if (p.sumsq_out && tile_valid) emit_sumsq16(p, sq_red[grp][0], acc, fvalid, ew, lane, b0 + jc, a0 / kBlockM, bar_id);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(tmem_empty_leader);
      } else {
        constexpr size_t kSplitStride = (size_t)C::kGroupTokens * kBlockM;
        float* part = p.ws_partial + ((size_t)tile_lin * p.splits + split) * kSplitStride + i;
#pragma unroll 1
        for (int jc = col_begin; jc < my_end; jc += 16) {
          uint32_t r[16];
          tmem_ld_32x32b_x16(taddr + jc, r);
          tmem_ld_wait();
#pragma unroll
          for (int t = 0; t < 16; ++t) __stcg(part + (jc + t) * kBlockM, __uint_as_float(r[t]));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(tmem_empty_leader);   // the accumulator is free: the rest works on the parked copy
        __threadfence();
        asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
        if (ew == 0 && lane == 0) {
          int* cnt = p.ws_pcounters + tile_lin * 2 + grp;
          const int old = atomicAdd(cnt, 1);
          const int last = (old == p.splits - 1) ? 1 : 0;
          if (last) *cnt = 0;                              // self-resetting: nobody else touches it in this launch
          last_flag[grp] = last;
        }
        asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
        if (last_flag[grp]) {
          __threadfence();
          const float* base = p.ws_partial + (size_t)tile_lin * p.splits * kSplitStride + i;
#pragma unroll 1
          for (int jc = col_begin; jc < my_end; jc += 16) {
            float acc[16];
#pragma unroll
            for (int t = 0; t < 16; ++t) acc[t] = 0.f;
            reduce_splits_dyn(base + jc * kBlockM, kSplitStride, p.splits, acc);
            if (p.rs_partials) {
#pragma unroll
              for (int t = 0; t < 16; ++t) acc[t] *= rstd_all[b0 + jc + t];
            }
            if (staged) {
              float (*sq)[16] = sq_red[grp][sq_cnt++ & 1];     // running parity: consecutive chunks never share a buffer
              epi_swap16_staged<kAct>(p, stage_warp, a0, ew, lane, bias, b0 + jc, acc, p.sumsq_out ? sq[ew] : nullptr);
              if (p.sumsq_out) finish_sumsq16(p, sq, ew, lane, b0 + jc, a0 / kBlockM, bar_id);
            } else {
              epi_swap16<kAct>(p, feat, fvalid, bias, i, b0 + jc, acc);
              if (p.sumsq_out && tile_valid) emit_sumsq16(p, sq_red[grp][0], acc, fvalid, ew, lane, b0 + jc, a0 / kBlockM, bar_id);
            }
          }
        }
      }
    }
    tc_fence_before();
  }

  __syncwarp();
  cluster_sync_all();      // the partner's tensor-core reads of this CTA's shared memory / writes to its TMEM are over
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair<512>(tmem_base);
  }
}


// ---------------------------------------------------------------------------------------------------------------------
// The persistent tall GEMM on CTA PAIRS (tcgen05.mma.cta_group::2, M = 256): the two SMs of a TPC take two neighbouring
// 128-row blocks and the same BN columns.  Every CTA loads its own 128 activation rows (16 KB) and only HALF of the BN
// weight rows of a k-block; the tensor cores read the other half from the partner's shared memory.  Per CTA and k-block
// that is 32 KB instead of 48 KB at BN = 256 (28 instead of 40 KB at BN = 192): a third less L2 -> SM traffic and
// shared-memory fill for the same MMA work, and room for a fourth (BN = 256) / fifth (BN = 192) operand stage.  Same
// two-accumulator pipeline and the same epilogue as the single-CTA kernel; TMA completions of both CTAs are counted by the
// leader's full barrier, tcgen05.commit multicasts "stage free" and "accumulator full" to both CTAs, both CTAs' epilogue
// warps release an accumulator on the leader's barrier.
// ---------------------------------------------------------------------------------------------------------------------
template <int BN>
struct TPairCfg {
  static constexpr int kStages = BN <= 192 ? 5 : 4;
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBHalfRows = BN / 2;
  static constexpr int kBBytes = kBHalfRows * kBlockK * 2;          // this CTA's half of the weight rows
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kEpiWarps = 8;
  static constexpr int kThreads = 128 + 32 * kEpiWarps;
  static constexpr int kEpiBN = BN / 2;
  static constexpr int kEpiWarpBytes = 32 * (kEpiBN * 2 + 16);
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiWarps * kEpiWarpBytes + 1024 + 256;
  static_assert(kBBytes % 1024 == 0 && kSmemBytes <= 232448, "tile shape");
};

template <int BN, int kAct>
__global__ void __launch_bounds__(TPairCfg<BN>::kThreads, 1)
gemm_tall_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                      const GemmParams p, const int tiles_mp, const int tiles_n) {
  using C = TPairCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + C::kStages * C::kABytes;
  uint8_t* smem_epi = smem + C::kStages * C::kStageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_epi + C::kEpiWarps * C::kEpiWarpBytes);   // used in the leader only
  uint64_t* empty_bar = full_bar + C::kStages;
  uint64_t* tmem_full = empty_bar + C::kStages;     // [2]
  uint64_t* tmem_empty = tmem_full + 2;             // [2], counted in the leader only (both CTAs' epilogue warps arrive)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();           // 0 = leader
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const int kb_total = (p.K + kBlockK - 1) / kBlockK;
  const int n_tiles = tiles_mp * tiles_n;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmap_a);
    prefetch_tensormap(&tmap_b);
    for (int s = 0; s < C::kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], 2 * C::kEpiWarps); }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc_pair<512>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();        // both CTAs' barriers are initialised before any remote arrival / TMA completion can reach them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait();   // activations, residuals and the output buffer belong to earlier kernels until here

  if (warp == 0) {
    // ===== TMA producer (both CTAs): own activation rows + own half of the weight rows, counted by the leader's barrier =====
    if (lane == 0) {
      const uint64_t pol_a = l2_policy_evict_first(), pol_b = l2_policy_evict_last();
      uint32_t it = 0;
      for (int t = cluster_id; t < n_tiles; t += n_clusters) {
        const int a0 = ((t / tiles_n) * 2 + (int)rank) * kBlockM;
        const int b0 = (t % tiles_n) * BN + (int)rank * C::kBHalfRows;
        for (int kb = 0; kb < kb_total; ++kb, ++it) {
          const int s = it % C::kStages;
          mbar_wait(&empty_bar[s], ((it / C::kStages) & 1) ^ 1);
          if (rank == 0) mbar_arrive_expect_tx(&full_bar[s], 2u * (uint32_t)C::kStageBytes);     // both CTAs' bytes
          tma_load_2d_pair_hint(smem_a + s * C::kABytes, &tmap_a, &full_bar[s], kb * kBlockK, a0, pol_a);
          tma_load_2d_pair_hint(smem_b + s * C::kBBytes, &tmap_b, &full_bar[s], kb * kBlockK, b0, pol_b);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the leader's elected lane drives the tensor cores of both SMs =====
    if (rank == 0 && elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16_f32(2 * kBlockM, BN);
      uint32_t it = 0, tl = 0;
      for (int t = cluster_id; t < n_tiles; t += n_clusters, ++tl) {
        const uint32_t buf = tl & 1;
        mbar_wait(&tmem_empty[buf], ((tl >> 1) & 1) ^ 1);      // both CTAs' epilogues have drained this accumulator
        tc_fence_after();
        const uint32_t tacc = tmem_base + buf * BN;
        for (int kb = 0; kb < kb_total; ++kb, ++it) {
          const int s = it % C::kStages;
          mbar_wait(&full_bar[s], (it / C::kStages) & 1);
          tc_fence_after();
          const uint64_t da = make_kmajor_sw128_desc(smem_u32(smem_a + s * C::kABytes));
          const uint64_t db = make_kmajor_sw128_desc(smem_u32(smem_b + s * C::kBBytes));
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k)
            umma_f16_ss_pair(tacc, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0);
          umma_commit_pair(&empty_bar[s], 0b11);    // frees the stage in both CTAs
        }
        umma_commit_pair(&tmem_full[buf], 0b11);
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue (both CTAs): warp (w & 3) owns TMEM lanes 32*(w&3) .. +31 of both accumulators, column half (w - 4) / 4 =====
    const int ew = warp & 3, half = (warp - 4) >> 2;
    uint32_t tl = 0;
    for (int t = cluster_id; t < n_tiles; t += n_clusters, ++tl) {
      const int a0 = ((t / tiles_n) * 2 + (int)rank) * kBlockM, b0 = (t % tiles_n) * BN;
      const uint32_t buf = tl & 1;
      mbar_wait(&tmem_full[buf], (tl >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + buf * BN + half * C::kEpiBN + ((uint32_t)(ew * 32) << 16);
      epi_tall_staged<C::kEpiBN, kAct>(p, taddr, smem_epi + (warp - 4) * C::kEpiWarpBytes, a0, b0 + half * C::kEpiBN, ew, lane,
                                       nullptr, mapa_shared(smem_u32(&tmem_empty[buf]), 0));
    }
    tc_fence_before();
  }

  __syncwarp();
  cluster_sync_all();      // the partner's tensor-core reads of this CTA's shared memory / writes to its TMEM are over
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair<512>(tmem_base);
  }
}

}  // namespace

// =====================================================================================
// Host side: tensor-map cache, tiling heuristics, launch.
// =====================================================================================
unsigned long long g_aha_launch_count = 0;

namespace {

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

struct TmapKey {
  const void* ptr; int rows; int K; int ld; int box_rows;
  bool operator<(const TmapKey& o) const {
    return std::tie(ptr, rows, K, ld, box_rows) < std::tie(o.ptr, o.rows, o.K, o.ld, o.box_rows);
  }
};
std::map<TmapKey, CUtensorMap> g_tmaps;
std::mutex g_tmap_mu;

}  // namespace

// K-major [rows, K] bf16 matrix, box = 64 (K) x box_rows, 128-byte swizzle, OOB reads return zero.
int get_tmap(const void* ptr, int rows, int K, int ld, int box_rows, CUtensorMap* out) {
  TmapKey key{ptr, rows, K, ld, box_rows};
  std::lock_guard<std::mutex> lk(g_tmap_mu);
  auto it = g_tmaps.find(key);
  if (it != g_tmaps.end()) { *out = it->second; return AHA_OK; }
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) { aha_set_error("RuntimeError: cuTensorMapEncodeTiled not available from the driver"); return AHA_ERR_RUNTIME; }
  AHA_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "GEMM operand base must be 16-byte aligned");
  AHA_REQUIRE(((size_t)ld * 2) % 16 == 0, "GEMM operand row stride must be a multiple of 8 elements");
  cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)kBlockK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    aha_set_error("RuntimeError: cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
    return AHA_ERR_RUNTIME;
  }
  if (g_tmaps.size() > 65536) g_tmaps.clear();
  g_tmaps[key] = m;
  *out = m;
  return AHA_OK;
}

struct ProfSlot { cudaEvent_t e0, e1; int cls; double bytes, flops; };
namespace {
std::deque<ProfSlot> g_prof;
std::mutex g_prof_mu;
bool g_prof_enabled = false;
}  // namespace

ProfSlot* prof_begin(int cls, cudaStream_t st) {
  if (!g_prof_enabled) return nullptr;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (g_prof.size() >= 200000) return nullptr;
  ProfSlot s{};
  s.cls = cls;
  if (cudaEventCreate(&s.e0) != cudaSuccess || cudaEventCreate(&s.e1) != cudaSuccess) return nullptr;
  cudaEventRecord(s.e0, st);
  g_prof.push_back(s);
  return &g_prof.back();
}
void prof_end(ProfSlot* s, cudaStream_t st, double bytes, double flops) {
  s->bytes = bytes; s->flops = flops;
  cudaEventRecord(s->e1, st);
}

namespace {

template <int BN, bool kSwap, int kAct, bool kDeep>
int launch_bn(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, dim3 grid, cudaStream_t stream) {
  using C = Cfg<BN, kDeep>;
  static bool attr_done[16] = {false};     // function attributes are per device
  int dev = 0;
  AHA_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 16 && !attr_done[dev]) {
    AHA_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, kSwap, kAct, kDeep>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    AHA_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, kSwap, kAct, kDeep>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    attr_done[dev] = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = C::kSmemBytes; cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = aha_pdl_enabled() ? 1 : 0;
  AHA_CUDA(cudaLaunchKernelEx(&cfg, gemm_tcgen05_kernel<BN, kSwap, kAct, kDeep>, ta, tb, p));
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}

}  // namespace

namespace {
template <int BN, bool kSwap, bool kDeep>
int dispatch_act(int act, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, dim3 grid, cudaStream_t st) {
  if (kSwap) {   // skinny orientation: plain or SiLU*mul (GELU callers are routed to the tall orientation)
    if (act == GEMM_ACT_SILU_MUL) return launch_bn<BN, kSwap, GEMM_ACT_SILU_MUL, kDeep>(ta, tb, p, grid, st);
    return launch_bn<BN, kSwap, GEMM_ACT_NONE, kDeep>(ta, tb, p, grid, st);
  }
  switch (act) {
    case GEMM_ACT_GELU_TANH: return launch_bn<BN, false, GEMM_ACT_GELU_TANH, false>(ta, tb, p, grid, st);
    case GEMM_ACT_GELU_ERF: return launch_bn<BN, false, GEMM_ACT_GELU_ERF, false>(ta, tb, p, grid, st);
    case GEMM_ACT_SILU_MUL: return launch_bn<BN, false, GEMM_ACT_SILU_MUL, false>(ta, tb, p, grid, st);
    default: return launch_bn<BN, false, GEMM_ACT_NONE, false>(ta, tb, p, grid, st);
  }
}
int dispatch(int bn, int swap, int deep, int act, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p,
             dim3 grid, cudaStream_t st) {
  if (swap && deep) {
    switch (bn) {
      case 32: return dispatch_act<32, true, true>(act, ta, tb, p, grid, st);
      case 64: return dispatch_act<64, true, true>(act, ta, tb, p, grid, st);
      case 128: return dispatch_act<128, true, true>(act, ta, tb, p, grid, st);
      default: return dispatch_act<256, true, false>(act, ta, tb, p, grid, st);   // BN 256 is one CTA per SM already
    }
  }
  if (swap) {
    switch (bn) {
      case 32: return dispatch_act<32, true, false>(act, ta, tb, p, grid, st);
      case 64: return dispatch_act<64, true, false>(act, ta, tb, p, grid, st);
      case 128: return dispatch_act<128, true, false>(act, ta, tb, p, grid, st);
      default: return dispatch_act<256, true, false>(act, ta, tb, p, grid, st);
    }
  }
  switch (bn) {
    case 64: return dispatch_act<64, false, false>(act, ta, tb, p, grid, st);
    case 128: return dispatch_act<128, false, false>(act, ta, tb, p, grid, st);
    default: return dispatch_act<256, false, false>(act, ta, tb, p, grid, st);
  }
}
}  // namespace

namespace {
template <int BN, int kAct>
int launch_persistent(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int tiles_m, int tiles_n, int sm_count,
                      cudaStream_t stream) {
  using C = PCfg<BN>;
  static bool attr_done[16] = {false};     // function attributes are per device
  int dev = 0;
  AHA_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 16 && !attr_done[dev]) {
    AHA_CUDA(cudaFuncSetAttribute(gemm_tall_persistent_kernel<BN, kAct>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    // same (maximal) shared-memory carve-out as every other kernel of the library: CTAs of two kernels only share an SM
    // when their carve-outs agree
    AHA_CUDA(cudaFuncSetAttribute(gemm_tall_persistent_kernel<BN, kAct>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    attr_done[dev] = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(std::min(tiles_m * tiles_n, sm_count)); cfg.blockDim = dim3(C::kThreads);
  cfg.dynamicSmemBytes = C::kSmemBytes; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = aha_pdl_enabled() ? 1 : 0;
  AHA_CUDA(cudaLaunchKernelEx(&cfg, gemm_tall_persistent_kernel<BN, kAct>, ta, tb, p, tiles_m, tiles_n));
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}
template <int BN>
int dispatch_persistent(int act, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int tiles_m, int tiles_n,
                        int sm_count, cudaStream_t st) {
  switch (act) {
    case GEMM_ACT_GELU_TANH: return launch_persistent<BN, GEMM_ACT_GELU_TANH>(ta, tb, p, tiles_m, tiles_n, sm_count, st);
    case GEMM_ACT_GELU_ERF: return launch_persistent<BN, GEMM_ACT_GELU_ERF>(ta, tb, p, tiles_m, tiles_n, sm_count, st);
    case GEMM_ACT_SILU_MUL: return launch_persistent<BN, GEMM_ACT_SILU_MUL>(ta, tb, p, tiles_m, tiles_n, sm_count, st);
    default: return launch_persistent<BN, GEMM_ACT_NONE>(ta, tb, p, tiles_m, tiles_n, sm_count, st);
  }
}
}  // namespace

namespace {
template <int BN, int kAct>
int launch_tall_pair(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int tiles_mp, int tiles_n, int sm_count,
                     cudaStream_t stream) {
  using C = TPairCfg<BN>;
  static bool attr_done[16] = {false};     // function attributes are per device
  int dev = 0;
  AHA_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 16 && !attr_done[dev]) {
    AHA_CUDA(cudaFuncSetAttribute(gemm_tall_pair_kernel<BN, kAct>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    AHA_CUDA(cudaFuncSetAttribute(gemm_tall_pair_kernel<BN, kAct>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    attr_done[dev] = true;
  }
  const int n_clusters = std::min(tiles_mp * tiles_n, sm_count / 2);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * n_clusters); cfg.blockDim = dim3(C::kThreads);
  cfg.dynamicSmemBytes = C::kSmemBytes; cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = aha_pdl_enabled() ? 2 : 1;
  AHA_CUDA(cudaLaunchKernelEx(&cfg, gemm_tall_pair_kernel<BN, kAct>, ta, tb, p, tiles_mp, tiles_n));
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}
template <int BN>
int dispatch_tall_pair(int act, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int tiles_mp, int tiles_n,
                       int sm_count, cudaStream_t st) {
  switch (act) {
    case GEMM_ACT_GELU_TANH: return launch_tall_pair<BN, GEMM_ACT_GELU_TANH>(ta, tb, p, tiles_mp, tiles_n, sm_count, st);
    case GEMM_ACT_GELU_ERF: return launch_tall_pair<BN, GEMM_ACT_GELU_ERF>(ta, tb, p, tiles_mp, tiles_n, sm_count, st);
    case GEMM_ACT_SILU_MUL: return launch_tall_pair<BN, GEMM_ACT_SILU_MUL>(ta, tb, p, tiles_mp, tiles_n, sm_count, st);
    default: return launch_tall_pair<BN, GEMM_ACT_NONE>(ta, tb, p, tiles_mp, tiles_n, sm_count, st);
  }
}
}  // namespace

namespace {
template <int kAct, int BN, bool kPair = false>
int launch_swap_persistent(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int tiles_a, int tiles_b,
                           int grid_ctas, cudaStream_t stream) {
  using C = PSCfgT<BN, kPair>;
  static bool attr_done[16] = {false};
  int dev = 0;
  AHA_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 16 && !attr_done[dev]) {
    AHA_CUDA(cudaFuncSetAttribute(gemm_swap_persistent_kernel<kAct, BN, kPair>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    AHA_CUDA(cudaFuncSetAttribute(gemm_swap_persistent_kernel<kAct, BN, kPair>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    attr_done[dev] = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid_ctas); cfg.blockDim = dim3(C::kThreads);
  cfg.dynamicSmemBytes = C::kSmemBytes; cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (kPair) {      // (tiles_a counts PAIRS of feature tiles, grid_ctas is even)
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 2; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (aha_pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr; cfg.numAttrs = na;
  AHA_CUDA(cudaLaunchKernelEx(&cfg, gemm_swap_persistent_kernel<kAct, BN, kPair>, ta, tb, p, tiles_a, tiles_b));
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}
}  // namespace

namespace {
template <int kAct>
int launch_swap_pair(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int pairs, int groups, int sm_count,
                     cudaStream_t stream) {
  using C = PPCfg;
  static bool attr_done[16] = {false};
  int dev = 0;
  AHA_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 16 && !attr_done[dev]) {
    AHA_CUDA(cudaFuncSetAttribute(gemm_swap_pair_kernel<kAct>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    AHA_CUDA(cudaFuncSetAttribute(gemm_swap_pair_kernel<kAct>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    attr_done[dev] = true;
  }
  const int n_clusters = std::min(pairs * groups * p.splits, sm_count / 2);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * n_clusters); cfg.blockDim = dim3(C::kThreads);
  cfg.dynamicSmemBytes = C::kSmemBytes; cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = aha_pdl_enabled() ? 2 : 1;
  AHA_CUDA(cudaLaunchKernelEx(&cfg, gemm_swap_pair_kernel<kAct>, ta, tb, p, pairs, groups));
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}
}  // namespace

// ---- bring-up: phase stamps (SM cycles) of CTA (0,0,0) of the most recent launch of each (grid.x, grid.z) class ----
static unsigned long long* g_gemm_ts = nullptr;   // [64 slots][8]
static std::map<std::pair<int, int>, int> g_gemm_ts_slot;
static unsigned long long* g_gemm_ts_all = nullptr;   // [64 slots][kTsAllCtas][2]
constexpr int kTsAllCtas = 512;
static unsigned long long* gemm_ts_all_slot(dim3 grid);
static unsigned long long* gemm_ts_slot(dim3 grid) {
  static const bool want = getenv("AHA_TS") && atoi(getenv("AHA_TS")) == 1;
  if (!want) return nullptr;
  if (!g_gemm_ts) { if (cudaMalloc(&g_gemm_ts, 64 * 16 * 8) != cudaSuccess) return nullptr; cudaMemset(g_gemm_ts, 0, 64 * 8 * 8); }
  auto key = std::make_pair((int)grid.x, (int)grid.z);
  auto it = g_gemm_ts_slot.find(key);
  if (it == g_gemm_ts_slot.end()) { if (g_gemm_ts_slot.size() >= 64) return nullptr; it = g_gemm_ts_slot.emplace(key, (int)g_gemm_ts_slot.size()).first; }
  return g_gemm_ts + it->second * 16;
}
static unsigned long long* gemm_ts_all_slot(dim3 grid) {
  if (!gemm_ts_slot(grid) || (long)grid.x * grid.y * grid.z > kTsAllCtas) return nullptr;
  if (!g_gemm_ts_all) {
    if (cudaMalloc(&g_gemm_ts_all, (size_t)64 * kTsAllCtas * 2 * 8) != cudaSuccess) return nullptr;
    cudaMemset(g_gemm_ts_all, 0, (size_t)64 * kTsAllCtas * 2 * 8);
  }
  return g_gemm_ts_all + (size_t)g_gemm_ts_slot[std::make_pair((int)grid.x, (int)grid.z)] * kTsAllCtas * 2;
}
// per-CTA (accumulator complete, exit) nanosecond stamps of the most recent launch of slot `slot` (aha_debug_gemm_timeline order)
extern "C" int aha_debug_gemm_tails(int slot, unsigned long long* out /*[512*2]*/) {
  if (!g_gemm_ts_all || slot < 0 || slot >= 64) { aha_set_error("ValueError: run with AHA_TS=1"); return AHA_ERR_VALUE; }
  AHA_CUDA(cudaDeviceSynchronize());
  AHA_CUDA(cudaMemcpy(out, g_gemm_ts_all + (size_t)slot * kTsAllCtas * 2, (size_t)kTsAllCtas * 2 * 8, cudaMemcpyDeviceToHost));
  return AHA_OK;
}
extern "C" int aha_debug_gemm_timeline(int* n_out, int* grid_x, int* grid_z, unsigned long long* out /*[64*16]*/) {
  if (!g_gemm_ts) { aha_set_error("ValueError: run with AHA_TS=1"); return AHA_ERR_VALUE; }
  AHA_CUDA(cudaDeviceSynchronize());
  AHA_CUDA(cudaMemcpy(out, g_gemm_ts, 64 * 16 * 8, cudaMemcpyDeviceToHost));
  *n_out = (int)g_gemm_ts_slot.size();
  for (auto& kv : g_gemm_ts_slot) { grid_x[kv.second] = kv.first.first; grid_z[kv.second] = kv.first.second; }
  return AHA_OK;
}

int gemm_workspace_create(GemmWorkspace* w, int device) {
  cudaDeviceProp prop;
  AHA_CUDA(cudaGetDeviceProperties(&prop, device));
  w->sm_count = prop.multiProcessorCount;
  return AHA_OK;
}

void gemm_workspace_destroy(GemmWorkspace* w) {
  for (auto& kv : w->scratch) { cudaFree(kv.second.partial); cudaFree(kv.second.counters); }
  w->scratch.clear();
}

// advance = false: a launch that uses only the partial workspace and the self-resetting counters (no bank switch)
static int split_scratch(GemmWorkspace* w, cudaStream_t st, GemmSplitScratch* out, bool advance = true) {
  std::lock_guard<std::mutex> lk(w->mu);
  auto it = w->scratch.find(st);
  if (it == w->scratch.end()) {
    GemmSplitScratch s;
    AHA_CUDA(cudaMalloc(&s.partial, GemmWorkspace::kPartialBytes));
    const size_t n_cnt = 4 * GemmWorkspace::kCounters + GemmWorkspace::kPCounters;
    AHA_CUDA(cudaMalloc(&s.counters, n_cnt * sizeof(int)));
    AHA_CUDA(cudaMemsetAsync(s.counters, 0, n_cnt * sizeof(int), st));
    s.pcounters = s.counters + 4 * GemmWorkspace::kCounters;
    it = w->scratch.emplace(st, s).first;
  }
  GemmSplitScratch& s = it->second;
  if (!advance) { *out = s; return AHA_OK; }
  const int bank = (int)(s.seq++ & 1);
  *out = s;
  out->bank = s.counters + bank * 2 * GemmWorkspace::kCounters;
  out->zero = s.counters + (bank ^ 1) * 2 * GemmWorkspace::kCounters;
  out->zero_n = 2 * GemmWorkspace::kCounters;
  return AHA_OK;
}

GemmWorkspace* aha_default_workspace(int* rc_out);
int aha_split_scratch(cudaStream_t st, GemmSplitScratch* out, bool advance) {
  int rc;
  GemmWorkspace* w = aha_default_workspace(&rc);
  if (!w) return rc;
  rc = split_scratch(w, st, out, advance);
  if (rc) return rc;
  if (advance) out->bank += GemmWorkspace::kCounters;
  else out->pcounters += GemmWorkspace::kPCounters / 2;     // the attention half of the self-resetting counters
  return AHA_OK;
}

static int g_sm_shared = 0;
extern "C" int aha_set_sm_sharing(int shared) { g_sm_shared = shared != 0; return AHA_OK; }
bool aha_sm_shared() { return g_sm_shared != 0; }

int gemm_bf16(const GemmCall& c, GemmWorkspace* w, cudaStream_t stream) {
  AHA_REQUIRE(c.tokens > 0 && c.features > 0 && c.K > 0, "GEMM with an empty dimension");
  // skinny orientation (weights = MMA A operand): up to 256 tokens in the one-shot kernel, up to kGemmSkinnyMaxTokens in
  // the persistent one (bf16 outputs only)
  int swap = c.force_swap >= 0 ? c.force_swap : ((c.tokens <= 256 || (c.tokens <= kGemmSkinnyMaxTokens && !c.out_f32)) ? 1 : 0);
  if (c.act == GEMM_ACT_GELU_TANH || c.act == GEMM_ACT_GELU_ERF) swap = 0;   // GELU epilogues exist in the tall orientation only
  if (swap && c.tokens > 256 && (c.out_f32 || c.tokens > kGemmSkinnyMaxTokens)) {
    AHA_REQUIRE(c.force_swap < 0, "skinny orientation above 256 tokens: bf16 output and at most 832 tokens");
    swap = 0;
  }
  if (c.act == GEMM_ACT_SILU_MUL) {
    AHA_REQUIRE(!c.out_f32 && !c.residual, "SiLU*mul epilogue writes bf16 without residual");
  }
  GemmParams p{};
  p.K = c.K; p.swap = swap;
  p.out = c.out; p.ldo = c.ldo; p.out_f32 = c.out_f32; p.round_bf16 = c.round_bf16; p.act = c.act;
  p.bias = c.bias; p.residual = c.residual; p.ldr = c.ldr; p.res_mod = c.res_mod;
  if (c.rs_partials || c.sumsq_out) {
    AHA_REQUIRE(swap && c.act != GEMM_ACT_GELU_TANH && c.act != GEMM_ACT_GELU_ERF, "folded RMSNorm needs the skinny orientation (tokens <= 832, bf16 output, no GELU)");
    AHA_REQUIRE(!c.rs_partials || (c.rs_n > 0 && c.rs_ld >= c.tokens), "folded RMSNorm: bad partial-sum layout");
    AHA_REQUIRE(!c.sumsq_out || (c.sumsq_ld >= c.tokens && !c.out_f32 && c.act == GEMM_ACT_NONE), "folded RMSNorm: bad sum-of-squares output");
    p.rs_partials = c.rs_partials; p.rs_n = c.rs_n; p.rs_ld = c.rs_ld; p.rs_eps = c.rs_eps; p.rs_inv_cols = 1.0f / (float)c.K;
    p.sumsq_out = c.sumsq_out; p.sumsq_ld = c.sumsq_ld;
  }
  p.pf = c.pf;
  const void* a; const void* b; int lda, ldb;
  if (swap) { p.Ma = c.features; p.Nb = c.tokens; a = c.w; lda = c.ldw; b = c.x; ldb = c.ldx; }
  else      { p.Ma = c.tokens; p.Nb = c.features; a = c.x; lda = c.ldx; b = c.w; ldb = c.ldw; }

  // Batched decoder steps (256 < tokens <= 832) on CTA pairs: opt-in (AHA_PS_PAIR=1 or force_bn = 416); the default is the
  // single-CTA persistent kernel below
  static const int use_pair = getenv("AHA_PS_PAIR") ? atoi(getenv("AHA_PS_PAIR")) : 0;   // measured slower in the step (its epilogue is not hidden): opt-in
  if (swap && c.tokens > 256 && (use_pair || c.force_bn == 416) && c.force_bn != 208 && c.force_bn != 112) {
    const int tiles_a = (p.Ma + kBlockM - 1) / kBlockM, pairs = (tiles_a + 1) / 2;
    const int groups = (p.Nb + PPCfg::kGroupTokens - 1) / PPCfg::kGroupTokens;
    const int kb_total = (p.K + kBlockK - 1) / kBlockK;
    const int n_clusters = w->sm_count / 2;
    if (c.sumsq_tiles) *c.sumsq_tiles = tiles_a;
    int splits = c.force_splits;
    if (splits <= 0) {
      // every cluster should get the same number of k-blocks; an item pays for its (unhidden) epilogue, a split item for
      // parking / merging its partial tile as well (in units of one k-block = 832 tensor-pipe cycles)
      static const int max_env = getenv("AHA_PS_SPLITS") ? atoi(getenv("AHA_PS_SPLITS")) : 8;
      long best_cost = -1;
      for (int s = 1; s <= max_env && s <= kb_total; ++s) {
        const int kb_per = (kb_total + s - 1) / s;
        if ((kb_total + kb_per - 1) / kb_per != s || (s > 1 && kb_per < 8)) continue;
        const long items = (long)pairs * groups * s;
        const long cost = ((items + n_clusters - 1) / n_clusters) * (kb_per + (s > 1 ? 13 : 7));
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; splits = s; }
      }
    }
    if (splits > kb_total) splits = kb_total;
    int kb_per = (kb_total + splits - 1) / splits;
    splits = (kb_total + kb_per - 1) / kb_per;
    const size_t tile_floats = (size_t)PPCfg::kGroupTokens * kBlockM;
    while (splits > 1 && ((size_t)2 * pairs * groups * splits * tile_floats * sizeof(float) > GemmWorkspace::kPartialBytes ||
                          2 * pairs * groups * 2 > 4096)) {
      --splits;
      kb_per = (kb_total + splits - 1) / splits;
      splits = (kb_total + kb_per - 1) / kb_per;
    }
    p.splits = splits; p.kb_per_split = kb_per; p.dbg = 0;
    if (splits > 1) {
      GemmSplitScratch sc;
      int src = split_scratch(w, stream, &sc, false); if (src) return src;
      p.ws_partial = sc.partial; p.ws_pcounters = sc.pcounters;
    }
    CUtensorMap ta, tb;
    int rc = get_tmap(a, p.Ma, p.K, lda, kBlockM, &ta); if (rc) return rc;
    rc = get_tmap(b, p.Nb, p.K, ldb, PPCfg::kHalfRows, &tb); if (rc) return rc;
    ProfSlot* slot = prof_begin(5, stream);
    rc = c.act == GEMM_ACT_SILU_MUL ? launch_swap_pair<GEMM_ACT_SILU_MUL>(ta, tb, p, pairs, groups, w->sm_count, stream)
                                    : launch_swap_pair<GEMM_ACT_NONE>(ta, tb, p, pairs, groups, w->sm_count, stream);
    if (slot) {
      const double out_bytes = (double)c.tokens * (c.act == GEMM_ACT_SILU_MUL ? c.features / 2 : c.features) * 2;
      prof_end(slot, stream, ((double)c.features * c.K + (double)c.tokens * c.K) * 2.0 + out_bytes,
               2.0 * c.tokens * (double)c.features * c.K);
    }
    return rc;
  }
  // Persistent skinny path, single CTAs (the default; force_bn = 208 selects it explicitly)
  if (swap && c.tokens > 256) {
    const int tiles_a = (p.Ma + kBlockM - 1) / kBlockM;
    // token tile width: 208 for the big matrices, 112 where 208-token tiles would leave most SMs without a work item
    static const int bn_env = getenv("AHA_PS_BN") ? atoi(getenv("AHA_PS_BN")) : 0;
    // CTA pairs (tcgen05.mma.cta_group::2: two feature tiles x one token tile per cluster, each CTA loads half of the token
    // rows): AHA_PS_PAIR2=1, or force_bn = 2112 / 2208 (tests)
    static const int pair2_env = getenv("AHA_PS_PAIR2") ? atoi(getenv("AHA_PS_PAIR2")) : 0;
    const bool pair2 = (pair2_env && c.force_bn == 0 && c.force_splits >= 0) || c.force_bn == 2112 || c.force_bn == 2208;
    const int wg = pair2 ? w->sm_count / 2 : w->sm_count;                 // work groups: CTAs or CTA pairs
    int bn_tok = kPsBN;
    if (c.force_bn == 112 || c.force_bn == 208) bn_tok = c.force_bn;
    else if (c.force_bn == 2112 || c.force_bn == 2208) bn_tok = c.force_bn - 2000;
    else if (bn_env == 112 || bn_env == 208) bn_tok = bn_env;
    // (short K only: with 112-token tiles every weight tile is pulled by four CTAs - down (K = 18 944) measured 87 us
    // unsplit on 112-token tiles, at the L2 -> SM bandwidth, against 69 us on 208-token tiles split two ways)
    else if ((long)tiles_a * ((p.Nb + kPsBN - 1) / kPsBN) * 5 <= (long)w->sm_count * 3 && (p.K + kBlockK - 1) / kBlockK <= 96) bn_tok = 112;
    const int tiles_b = (p.Nb + bn_tok - 1) / bn_tok;
    const int tiles_aw = pair2 ? (tiles_a + 1) / 2 : tiles_a;              // feature tiles (or pairs of them) per token tile
    const int tiles = (pair2 ? 2 * tiles_aw : tiles_a) * tiles_b;           // workspace tiles (an odd last pair has a dummy partner)
    const int tiles_w = tiles_aw * tiles_b;                                 // work items before splitting
    const int kb_total = (p.K + kBlockK - 1) / kBlockK;
    if (c.sumsq_tiles) *c.sumsq_tiles = tiles_a;
    static const int use_sk = getenv("AHA_PS_STREAMK") ? atoi(getenv("AHA_PS_STREAMK")) : 0;   // measured slower in the step: opt-in
    // (force_splits = -1 selects it explicitly)
    int splits = c.force_splits, kb_per = kb_total, grid_ctas = 0;
    p.sk_ranks = 0;
    if (!pair2 && (splits == -1 || (splits == 0 && use_sk))) {
      // stream-K: every CTA carries the same number of k-blocks (see the kernel).  R ranks per token tile.
      const long long U = (long long)tiles_a * kb_total;
      int R = std::max(1, w->sm_count / tiles_b);
      if (U / R < 8) R = (int)std::max<long long>(1, U / 8);        // at least 8 k-blocks per CTA
      auto rank_of = [&](long long u) { return (int)(((u + 1) * R + U - 1) / U) - 1; };
      int cap = 1;
      for (int f = 0; f < tiles_a; ++f)
        cap = std::max(cap, rank_of((long long)(f + 1) * kb_total - 1) - rank_of((long long)f * kb_total) + 1);
      if (cap == 1 || ((size_t)tiles * cap * bn_tok * kBlockM * sizeof(float) <= GemmWorkspace::kPartialBytes && tiles * 2 <= 4096)) {
        p.sk_ranks = R; splits = cap; grid_ctas = R * tiles_b;
      }
    }
    if (!p.sk_ranks) {
      if (splits <= 0) {
        // fixed split: a split item pays for parking / merging its partial tile
        static const int max_env = getenv("AHA_PS_SPLITS") ? atoi(getenv("AHA_PS_SPLITS")) : 8;
        long best_cost = -1;
        for (int s = 1; s <= max_env && s <= kb_total; ++s) {
          const int kbp = (kb_total + s - 1) / s;
          if ((kb_total + kbp - 1) / kbp != s || (s > 1 && kbp < 8)) continue;
          const long items = (long)tiles_w * s;
          // (measured in the batched step: with residual + sum-of-squares epilogues, down runs faster split 2 ways than 5)
          const long cost = ((items + wg - 1) / wg) * (kbp + (s > 1 ? 14 + 6 * s : 2));
          if (best_cost < 0 || cost < best_cost) { best_cost = cost; splits = s; }
        }
      }
      if (splits > kb_total) splits = kb_total;
      kb_per = (kb_total + splits - 1) / splits;
      splits = (kb_total + kb_per - 1) / kb_per;
      while (splits > 1 && ((size_t)tiles * splits * bn_tok * kBlockM * sizeof(float) > GemmWorkspace::kPartialBytes || tiles * 2 > 4096)) {
        --splits;
        kb_per = (kb_total + splits - 1) / splits;
        splits = (kb_total + kb_per - 1) / kb_per;
      }
      grid_ctas = pair2 ? 2 * std::min(tiles_w * splits, wg) : std::min(tiles * splits, w->sm_count);
    }
    if (splits < 1) splits = 1;
    p.splits = splits; p.kb_per_split = kb_per; p.dbg = 0;
    if (splits > 1) {
      GemmSplitScratch sc;
      int src = split_scratch(w, stream, &sc, false); if (src) return src;
      p.ws_partial = sc.partial; p.ws_pcounters = sc.pcounters;
    }
    CUtensorMap ta, tb;
    int rc = get_tmap(a, p.Ma, p.K, lda, kBlockM, &ta); if (rc) return rc;
    rc = get_tmap(b, p.Nb, p.K, ldb, pair2 ? bn_tok / 2 : bn_tok, &tb); if (rc) return rc;
    ProfSlot* slot = prof_begin(5, stream);
    if (pair2 && bn_tok == 112)
      rc = c.act == GEMM_ACT_SILU_MUL ? launch_swap_persistent<GEMM_ACT_SILU_MUL, 112, true>(ta, tb, p, tiles_aw, tiles_b, grid_ctas, stream)
                                      : launch_swap_persistent<GEMM_ACT_NONE, 112, true>(ta, tb, p, tiles_aw, tiles_b, grid_ctas, stream);
    else if (pair2)
      rc = c.act == GEMM_ACT_SILU_MUL ? launch_swap_persistent<GEMM_ACT_SILU_MUL, kPsBN, true>(ta, tb, p, tiles_aw, tiles_b, grid_ctas, stream)
                                      : launch_swap_persistent<GEMM_ACT_NONE, kPsBN, true>(ta, tb, p, tiles_aw, tiles_b, grid_ctas, stream);
    else if (bn_tok == 112)
      rc = c.act == GEMM_ACT_SILU_MUL ? launch_swap_persistent<GEMM_ACT_SILU_MUL, 112>(ta, tb, p, tiles_a, tiles_b, grid_ctas, stream)
                                      : launch_swap_persistent<GEMM_ACT_NONE, 112>(ta, tb, p, tiles_a, tiles_b, grid_ctas, stream);
    else
      rc = c.act == GEMM_ACT_SILU_MUL ? launch_swap_persistent<GEMM_ACT_SILU_MUL, kPsBN>(ta, tb, p, tiles_a, tiles_b, grid_ctas, stream)
                                      : launch_swap_persistent<GEMM_ACT_NONE, kPsBN>(ta, tb, p, tiles_a, tiles_b, grid_ctas, stream);
    if (slot) {
      const double out_bytes = (double)c.tokens * (c.act == GEMM_ACT_SILU_MUL ? c.features / 2 : c.features) * 2;
      prof_end(slot, stream, ((double)c.features * c.K + (double)c.tokens * c.K) * 2.0 + out_bytes,
               2.0 * c.tokens * (double)c.features * c.K);
    }
    return rc;
  }
  // Persistent tall path: bf16 output, enough 128 x BN tiles to give every SM several of them
  {
    static int persist = -1;
    if (persist < 0) { const char* e = getenv("AHA_GEMM_PERSIST"); persist = e ? atoi(e) : 1; }
    // While another stream shares the SMs (vision under the decoder) the one-shot kernel is used instead: a persistent CTA
    // holds its SM for the whole GEMM and the decoder's grids queue behind it (measured 4.80 vs 4.53 ms per frame; the
    // 128-column persistent variant, which leaves room for a decoder CTA on paper, measured 4.77).
    const bool forced = c.force_bn == 192 || c.force_bn == 256 || c.force_bn == 1128 || c.force_bn == 2192 || c.force_bn == 2256;
    if (!swap && persist && !c.out_f32 && c.force_splits <= 1 && (forced || (c.force_bn == 0 && !g_sm_shared))) {
      int pbn = c.force_bn == 1128 ? 128 : c.force_bn >= 2000 ? c.force_bn - 2000 : c.force_bn;   // 1128: the persistent 128-column variant, 2192 / 2256: CTA pairs (tests / experiments)
      if (pbn == 0) {      // fewer wasted columns wins; ties go to the wider tile
        const int w256 = (p.Nb + 255) / 256 * 256, w192 = (p.Nb + 191) / 192 * 192;
        pbn = w192 < w256 ? 192 : 256;
        static const int tall_bn = getenv("AHA_TALL_BN") ? atoi(getenv("AHA_TALL_BN")) : 0;   // experiment knob
        if (tall_bn == 192 || tall_bn == 256) pbn = tall_bn;
      }
      const int tm = (p.Ma + kBlockM - 1) / kBlockM, tn = (p.Nb + pbn - 1) / pbn;
      if ((long)tm * tn >= 2L * w->sm_count || c.force_bn != 0) {
        p.splits = 1; p.kb_per_split = (p.K + kBlockK - 1) / kBlockK; p.dbg = 0; p.dbg_ts = gemm_ts_slot(dim3(148, 1, 1));
        CUtensorMap ta, tb;
        int rc = get_tmap(a, p.Ma, p.K, lda, kBlockM, &ta); if (rc) return rc;
        // CTA pairs (cta_group::2) for the 192- and 256-column tiles: AHA_TALL_PAIR=1, or force_bn 2192 / 2256 (tests)
        static const int tall_pair = getenv("AHA_TALL_PAIR") ? atoi(getenv("AHA_TALL_PAIR")) : 1;
        const bool pair = (tall_pair || c.force_bn == 2192 || c.force_bn == 2256) && pbn != 128;
        rc = get_tmap(b, p.Nb, p.K, ldb, pair ? pbn / 2 : pbn, &tb); if (rc) return rc;
        ProfSlot* slot = prof_begin(1, stream);
        if (pair)
          rc = pbn == 192 ? dispatch_tall_pair<192>(c.act, ta, tb, p, (tm + 1) / 2, tn, w->sm_count, stream)
                          : dispatch_tall_pair<256>(c.act, ta, tb, p, (tm + 1) / 2, tn, w->sm_count, stream);
        else
        rc = pbn == 128 ? dispatch_persistent<128>(c.act, ta, tb, p, tm, tn, w->sm_count, stream)
           : pbn == 192 ? dispatch_persistent<192>(c.act, ta, tb, p, tm, tn, w->sm_count, stream)
                        : dispatch_persistent<256>(c.act, ta, tb, p, tm, tn, w->sm_count, stream);
        if (slot) {
          const double out_bytes = (double)c.tokens * (c.act == GEMM_ACT_SILU_MUL ? c.features / 2 : c.features) * 2;
          prof_end(slot, stream, ((double)c.features * c.K + (double)c.tokens * c.K) * 2.0 + out_bytes,
                   2.0 * c.tokens * (double)c.features * c.K);
        }
        return rc;
      }
    }
  }
  int bn = c.force_bn;
  if (bn == 0) {
    if (swap) bn = p.Nb <= 32 ? 32 : p.Nb <= 64 ? 64 : p.Nb <= 128 ? 128 : 256;
    else bn = p.Nb <= 64 ? 64 : 128;
    // Tall orientation with a short K and too few 128 x 128 tiles for the device (the one-frame SigLIP out_proj: 729 x 1152
    // x 1152 = 54 tiles): 128 x 64 tiles, unsplit, beat 128 x 128 with split-K 4 - 10.6 against 18.3 us (scripts/
    // gpu_gemm_sweep.py vision1) - because the tall split merge costs more than the 18 k-blocks it saves.
    if (!swap && bn == 128 && c.force_splits == 0 && p.K <= 2048 &&
        (long)((p.Ma + kBlockM - 1) / kBlockM) * ((p.Nb + 127) / 128) < w->sm_count)
      bn = 64;
  }
  AHA_REQUIRE(bn == 32 || bn == 64 || bn == 128 || bn == 256, "unsupported GEMM BN tile (one-shot kernel: 32/64/128/256)");
  if (!swap && bn == 32) bn = 64;
  const int tiles_a = (p.Ma + kBlockM - 1) / kBlockM;
  const int tiles_b = (p.Nb + bn - 1) / bn;
  const int tiles = tiles_a * tiles_b;
  if (c.sumsq_tiles) *c.sumsq_tiles = tiles_a;
  const int kb_total = (p.K + kBlockK - 1) / kBlockK;
  int splits = c.force_splits;
  int deep = c.force_deep;
  { static int env = -2; if (env == -2) { const char* e = getenv("AHA_GEMM_DEEP"); env = e ? atoi(e) : -1; } if (env >= 0) deep = env; }
  if (splits == 0 && swap) {
    // Skinny (weight-streaming) orientation: every SM should pull the same number of weight bytes.  With s splits the
    // grid has tiles*s CTAs of ceil(kb/s) k-blocks; CTAs beyond one per SM run as a second resident CTA (or a second
    // wave), so the per-SM critical path is ceil(ctas / SMs) * kb_per_split.  Fewer splits win ties (cheaper merge).
    long best_cost = -1;
    for (int s = 1; s <= 8 && s <= kb_total; ++s) {
      const int kb_per = (kb_total + s - 1) / s;
      const int s_eff = (kb_total + kb_per - 1) / kb_per;
      if (s_eff != s) continue;
      const long ctas = (long)tiles * s;
      if (s > 1 && ctas > (long)w->sm_count * (bn <= 128 ? 2 : 1)) continue;   // split CTAs must all be co-resident
      const long cost = ((ctas + w->sm_count - 1) / w->sm_count) * kb_per + (s - 1);
      if (best_cost < 0 || cost < best_cost) { best_cost = cost; splits = s; }
    }
  } else if (splits == 0) {
    splits = 1;
    // tall orientation (measured on B200, profiles/r01_gemm_sweep.json): keep the grid within ~1.75 CTAs per SM
    const int target = (7 * w->sm_count) / 4;
    if (tiles < target && !(bn == 64 && p.Nb > 64 && p.K <= 2048)) {   // (the short-K 128 x 64 case above stays unsplit)
      splits = target / tiles;
      const int max_by_k = kb_total / 4 > 0 ? kb_total / 4 : 1;  // at least 4 k-blocks per split
      if (splits > max_by_k) splits = max_by_k;
      if (splits > 8) splits = 8;
    }
  }
  // Split CTAs wait for their siblings.  Two grids of waiting CTAs on two streams could starve each other of SM slots, so
  // while the streams share the SMs only the decoder's weight-streaming GEMMs (skinny orientation) may split.
  if (g_sm_shared && !swap && c.force_splits == 0) splits = 1;
  if (splits > kb_total) splits = kb_total;
  int kb_per = (kb_total + splits - 1) / splits;
  splits = (kb_total + kb_per - 1) / kb_per;  // no empty splits
  p.splits = splits; p.kb_per_split = kb_per;
  { static int dbg = -1; if (dbg < 0) { const char* e = getenv("AHA_GEMM_DBG"); dbg = e ? atoi(e) : 0; } p.dbg = dbg; }
  { static int ring = -1, ring2 = -1;
    if (ring < 0) { const char* e = getenv("AHA_GEMM_RING"); ring = e ? atoi(e) : 0; e = getenv("AHA_GEMM_RING2"); ring2 = e ? atoi(e) : 0; }
    p.ring = 0; p.ring_deep = ring; p.ring_shallow = ring2; }
  AHA_REQUIRE(splits <= 16, "split-K factor above 16");
  // the splits of a tile wait for each other: the whole grid has to fit the device at once
  while (splits > 1 && ((long)tiles * splits > (long)w->sm_count * (bn <= 128 ? 2 : 1) ||
                        (size_t)tiles * splits * bn * kBlockM * sizeof(float) > GemmWorkspace::kPartialBytes ||
                        tiles > GemmWorkspace::kCounters)) {
    --splits;
    kb_per = (kb_total + splits - 1) / splits;
    splits = (kb_total + kb_per - 1) / kb_per;
  }
  p.splits = splits; p.kb_per_split = kb_per;
  if (splits > 1) {
    GemmSplitScratch sc;
    int src = split_scratch(w, stream, &sc); if (src) return src;
    p.ws_partial = sc.partial; p.ws_counters = sc.bank; p.ws_zero = sc.zero; p.ws_zero_n = sc.zero_n;
  }
  CUtensorMap ta, tb;
  int rc = get_tmap(a, p.Ma, p.K, lda, kBlockM, &ta); if (rc) return rc;
  rc = get_tmap(b, p.Nb, p.K, ldb, bn, &tb); if (rc) return rc;
  dim3 grid(tiles_a, tiles_b, splits);
  p.dbg_ts = gemm_ts_slot(grid);
  p.dbg_all = gemm_ts_all_slot(grid);
  if (deep < 0) deep = (swap && !g_sm_shared && (long)tiles * splits <= w->sm_count) ? 1 : 0;   // <= one CTA per SM: give it the whole SM
  if (deep && splits > 1 && (long)tiles * splits > w->sm_count) deep = 0;       // split CTAs must stay co-resident
  ProfSlot* slot = prof_begin(swap ? 0 : 1, stream);
  p.ring = swap ? (deep ? p.ring_deep : p.ring_shallow) : 0;
  rc = dispatch(bn, swap, deep, c.act, ta, tb, p, grid, stream);
  if (slot) {
    const double out_bytes = (double)c.tokens * (c.act == GEMM_ACT_SILU_MUL ? c.features / 2 : c.features) * (c.out_f32 ? 4 : 2);
    prof_end(slot, stream, ((double)c.features * c.K + (double)c.tokens * c.K) * 2.0 + out_bytes,
             2.0 * c.tokens * (double)c.features * c.K);
  }
  return rc;
}

// ---- optional per-launch CUDA-event timing of the GEMM launches (bench.py's roofline leg) ----
extern "C" int aha_profile_gemm(int enable) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (auto& s : g_prof) { cudaEventDestroy(s.e0); cudaEventDestroy(s.e1); }
  g_prof.clear();
  g_prof_enabled = enable != 0;
  return AHA_OK;
}
// cls 0 = skinny GEMM (weight-streaming, tokens <= 256), 1 = tall GEMM, 2 = KV re-rotation (eviction).  Synchronises the recorded events.
extern "C" int aha_profile_gemm_read(int cls, double* ms, long long* launches, double* bytes, double* flops) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  double t = 0, b = 0, f = 0; long long n = 0;
  for (auto& s : g_prof) {
    if (s.cls != cls) continue;
    AHA_CUDA(cudaEventSynchronize(s.e1));
    float e = 0.f;
    AHA_CUDA(cudaEventElapsedTime(&e, s.e0, s.e1));
    t += e; b += s.bytes; f += s.flops; ++n;
  }
  if (ms) *ms = t; if (launches) *launches = n; if (bytes) *bytes = b; if (flops) *flops = f;
  return AHA_OK;
}
