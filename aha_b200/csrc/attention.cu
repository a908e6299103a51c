// Fused attention on the 5th-gen tensor cores: S = Q.K^T and O += P.V are tcgen05.mma tiles with fp32 accumulators
// in TMEM, Q/K/V tiles are staged by TMA (128-byte swizzle), the online softmax runs in registers between the two.
//
// One kernel serves both attention shapes of the ingest-and-score path:
//   * frame-chunk versus KV cache (transformers 4.49 Qwen2Attention called at video_head_live_llava_qwen.py:160-169):
//     per (stream, kv head) the q*group query rows (GQA heads folded into the M dimension) against the cache's
//     physical ring slots; the reference's mask rule allowed(i, j) <=> logical(j) <= off + i is
//     evaluated from the ring geometry, so keys are streamed in PHYSICAL order (softmax is order independent) and an
//     eviction never has to move them.
//   * SigLIP self-attention (siglip_encoder.py:197-239): per (frame, head) 729 x 729, head_dim 72 padded to 80 columns.
//
// CTA = one 128-row query tile of one (segment, head block).  Warp roles:
//   warp 0    TMA producer: Q once, then a 2-stage ring of {K tile, V tile} (128 keys each)
//   warp 1    MMA issuer (one elected lane): S[b] = Q.K^T for tile kt+1 is issued before P.V of tile kt, so the
//             tensor pipe works on the next scores while the softmax warps are busy with the current ones
//   warp 2    TMEM allocator (512 columns: S0 | S1 | O)
//   warps 4-7 softmax / correction / epilogue: thread = query row = TMEM lane.  P is written to shared memory in the
//             K-major 128B-swizzled layout the MMA expects; O is rescaled in TMEM only when the running maximum grows by
//             more than 2^8 (lazy rescale), which keeps the correction off the critical path.
#include "ops.h"
#include "gemm.h"
#include "ptx_sm100.cuh"
#include <cstdlib>
#include <cstdio>

using namespace sm100;

namespace {

constexpr int kTile = 128;          // query rows per CTA == keys per tile
constexpr int kAtomBytes = 16384;   // one [128 rows][64 bf16] swizzled box
constexpr int kStagesKV = 2;
constexpr int kAttnThreads = 256;
constexpr uint32_t kColS0 = 0, kColS1 = 128, kColO = 256;
constexpr float kRescaleThreshold = 8.f;  // log2 domain

struct AttnSmem {
  // offsets inside the 1024-aligned dynamic shared memory
  static __host__ __device__ int q_off() { return 0; }
  static __host__ __device__ int k_off(int DB) { return DB * kAtomBytes; }
  static __host__ __device__ int v_off(int DB) { return DB * kAtomBytes * (1 + kStagesKV); }
  static __host__ __device__ int p_off(int DB) { return DB * kAtomBytes * (1 + 2 * kStagesKV); }
  static __host__ __device__ int bar_off(int DB) { return p_off(DB) + 2 * kAtomBytes; }
  static __host__ __device__ int total(int DB) { return bar_off(DB) + 256 + 1024; }
};

__global__ void __launch_bounds__(kAttnThreads, 1)
attention_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                 const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ AttnProblem p, const uint32_t v_lbo,
                 const uint32_t v_sbo) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int DB = (p.d_pad + 63) >> 6;
  uint8_t* sQ = smem + AttnSmem::q_off();
  uint8_t* sK = smem + AttnSmem::k_off(DB);
  uint8_t* sV = smem + AttnSmem::v_off(DB);
  uint8_t* sP = smem + AttnSmem::p_off(DB);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + AttnSmem::bar_off(DB));
  uint64_t* q_full = bars;             // 1
  uint64_t* kv_full = bars + 1;        // kStagesKV
  uint64_t* kv_empty = bars + 3;       // kStagesKV
  uint64_t* s_full = bars + 5;         // 2
  uint64_t* p_full = bars + 7;         // 1 (128 arrivals)
  uint64_t* pv_done = bars + 8;        // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const AttnSeg& seg = p.segs[blockIdx.z];
  const int blk = blockIdx.y;
  const int qt = blockIdx.x;
  const int nkt = (seg.klen + kTile - 1) / kTile;
  const int stage_bytes = DB * kAtomBytes;

  if (warp == 0 && lane == 0) { prefetch_tensormap(&tmap_q); prefetch_tensormap(&tmap_k); prefetch_tensormap(&tmap_v); }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < kStagesKV; ++s) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
    mbar_init(&s_full[0], 1); mbar_init(&s_full[1], 1);
    mbar_init(p_full, 128);
    mbar_init(pv_done, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (nkt > 0) {
    if (warp == 0) {
      // ===================== TMA producer =====================
      if (elect_one()) {
        const int q_row = seg.q_row0 + blk * p.q_block_rows + qt * kTile;
        const int q_col = p.q_col0 + blk * p.q_col_step;
        mbar_arrive_expect_tx(q_full, stage_bytes);
        for (int b = 0; b < DB; ++b) tma_load_2d(sQ + b * kAtomBytes, &tmap_q, q_full, q_col + b * 64, q_row);
        const int k_row = seg.k_row0 + blk * p.k_head_rows, k_col = p.k_col0 + blk * p.k_col_step;
        const int v_row = seg.v_row0 + blk * p.k_head_rows, v_col = p.v_col0 + blk * p.v_col_step;
        for (int kt = 0; kt < nkt; ++kt) {
          const int s = kt % kStagesKV;
          const uint32_t ph = (kt / kStagesKV) & 1;
          mbar_wait(&kv_empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&kv_full[s], 2 * stage_bytes);
          for (int b = 0; b < DB; ++b) {
            tma_load_2d(sK + s * stage_bytes + b * kAtomBytes, &tmap_k, &kv_full[s], k_col + b * 64, k_row + kt * kTile);
            tma_load_2d(sV + s * stage_bytes + b * kAtomBytes, &tmap_v, &kv_full[s], v_col + b * 64, v_row + kt * kTile);
          }
        }
      }
    } else if (warp == 1) {
      // ===================== MMA issuer =====================
      if (elect_one()) {
        const uint32_t idesc_qk = make_idesc_bf16_f32(kTile, kTile);
        const uint32_t idesc_pv = make_idesc_bf16_f32_bmn(kTile, (uint32_t)p.d_pad);
        const int ksteps = p.d_pad >> 4;
        auto issue_qk = [&](int kt) {
          const int s = kt % kStagesKV;
          mbar_wait(&kv_full[s], (kt / kStagesKV) & 1);
          tc_fence_after();
          const uint32_t qa = smem_u32(sQ), ka = smem_u32(sK + s * stage_bytes);
          for (int ks = 0; ks < ksteps; ++ks) {
            const uint32_t o = (uint32_t)(ks >> 2) * kAtomBytes + (uint32_t)(ks & 3) * 32;
            umma_f16_ss(tmem_base + ((kt & 1) ? kColS1 : kColS0), make_kmajor_sw128_desc(qa + o),
                        make_kmajor_sw128_desc(ka + o), idesc_qk, ks != 0);
          }
          umma_commit(&s_full[kt & 1]);
        };
        mbar_wait(q_full, 0);
        issue_qk(0);
        for (int kt = 0; kt < nkt; ++kt) {
          if (kt + 1 < nkt) issue_qk(kt + 1);
          mbar_wait(p_full, kt & 1);
          tc_fence_after();
          const int s = kt % kStagesKV;
          const uint32_t pa = smem_u32(sP), va = smem_u32(sV + s * stage_bytes);
          for (int ks = 0; ks < kTile / 16; ++ks) {
            const uint64_t da = make_kmajor_sw128_desc(pa + (uint32_t)(ks >> 2) * kAtomBytes + (uint32_t)(ks & 3) * 32);
            const uint64_t db = make_mnmajor_sw128_desc(va + (uint32_t)ks * 2048, v_lbo, v_sbo);
            umma_f16_ss(tmem_base + kColO, da, db, idesc_pv, (kt | ks) != 0);
          }
          umma_commit(&kv_empty[s]);
          umma_commit(pv_done);
        }
      }
    } else if (warp >= 4) {
      // ===================== softmax / correction / epilogue =====================
      const int ew = warp & 3;
      const int row = ew * 32 + lane;            // tile row == TMEM lane
      const int m = qt * kTile + row;            // query row inside the (segment, block)
      const bool row_valid = m < p.nq_rows;
      const int tok = m / p.group;
      const int limit = seg.off + tok;           // logical key index must be <= limit
      const uint32_t lane_addr = tmem_base + ((uint32_t)(ew * 32) << 16);
      const float sl2 = p.scale * 1.4426950408889634f;
      const bool all_visible = seg.off >= seg.klen - 1;
      float m_used = -INFINITY, l_sum = 0.f;
      uint8_t* prow = sP + row * 128;
      for (int kt = 0; kt < nkt; ++kt) {
        mbar_wait(&s_full[kt & 1], (kt >> 1) & 1);
        tc_fence_after();
        const uint32_t s_addr = lane_addr + ((kt & 1) ? kColS1 : kColS0);
        const bool need_mask = !(all_visible && (kt * kTile + kTile <= seg.klen));
        auto visible = [&](int c) -> bool {
          const int pslot = kt * kTile + c;
          if (pslot >= seg.klen) return false;
          int j = pslot;
          if (seg.ring_R > 0 && pslot >= seg.ring_S) {
            j = pslot - seg.ring_S - seg.ring_base;
            if (j < 0) j += seg.ring_R;
            j += seg.ring_S;
          }
          return j <= limit;
        };
        // pass 1: tile maximum of this row
        float tmax = -INFINITY;
#pragma unroll 1
        for (int cc = 0; cc < kTile; cc += 32) {
          uint32_t r[32];
          tmem_ld_32x32b_x32(s_addr + cc, r);
          tmem_ld_wait();
#pragma unroll
          for (int t = 0; t < 32; ++t) {
            float x = __uint_as_float(r[t]) * sl2;
            if (need_mask && !visible(cc + t)) x = -INFINITY;
            tmax = fmaxf(tmax, x);
          }
        }
        float alpha = 1.f;
        bool rescale = false;
        if (tmax > m_used + kRescaleThreshold || (m_used == -INFINITY && tmax > -INFINITY)) {
          alpha = (m_used == -INFINITY) ? 0.f : exp2f(m_used - tmax);
          m_used = tmax;
          rescale = true;
        }
        // P smem and O are owned by the tensor pipe until P.V of the previous tile has completed
        if (kt > 0) mbar_wait(pv_done, (kt - 1) & 1);
        if (kt > 0 && __any_sync(0xffffffffu, rescale)) {
          tc_fence_after();
#pragma unroll 1
          for (int cc = 0; cc < p.d_pad; cc += 16) {
            uint32_t r[16];
            tmem_ld_32x32b_x16(lane_addr + kColO + cc, r);
            tmem_ld_wait();
#pragma unroll
            for (int t = 0; t < 16; ++t) r[t] = __float_as_uint(__uint_as_float(r[t]) * alpha);
            tmem_st_32x32b_x16(lane_addr + kColO + cc, r);
          }
          tmem_st_wait();
        }
        l_sum *= alpha;
        // pass 2: probabilities -> bf16 -> swizzled K-major P tile
        const float mref = (m_used == -INFINITY) ? 0.f : m_used;
#pragma unroll 1
        for (int cc = 0; cc < kTile; cc += 32) {
          uint32_t r[32];
          tmem_ld_32x32b_x32(s_addr + cc, r);
          tmem_ld_wait();
          uint32_t pk[16];
#pragma unroll
          for (int t = 0; t < 32; t += 2) {
            float x0 = exp2f(__uint_as_float(r[t]) * sl2 - mref);
            float x1 = exp2f(__uint_as_float(r[t + 1]) * sl2 - mref);
            if (need_mask) {
              if (!visible(cc + t)) x0 = 0.f;
              if (!visible(cc + t + 1)) x1 = 0.f;
            }
            l_sum += x0 + x1;
            pk[t >> 1] = pack_bf16x2(x0, x1);
          }
          uint8_t* base = prow + (cc >> 6) * kAtomBytes;
          const int chunk0 = (cc & 63) >> 3;
#pragma unroll
          for (int ch = 0; ch < 4; ++ch) {
            const int sw = (chunk0 + ch) ^ (row & 7);
            *reinterpret_cast<uint4*>(base + sw * 16) = make_uint4(pk[4 * ch], pk[4 * ch + 1], pk[4 * ch + 2], pk[4 * ch + 3]);
          }
        }
        fence_proxy_async();
        tc_fence_before();
        mbar_arrive(p_full);
      }
      // epilogue: O / l -> bf16 -> out[token, (blk*group + r)*d + c]
      mbar_wait(pv_done, (nkt - 1) & 1);
      tc_fence_after();
      const float inv = l_sum > 0.f ? 1.f / l_sum : 0.f;
      bf16* orow = p.out + (size_t)(seg.out_row0 + tok) * p.out_ld + (size_t)(blk * p.group + (m % p.group)) * p.d;
#pragma unroll 1
      for (int cc = 0; cc < p.d_pad; cc += 16) {
        uint32_t r[16];
        tmem_ld_32x32b_x16(lane_addr + kColO + cc, r);
        tmem_ld_wait();
        if (row_valid) {
          float v[16];
#pragma unroll
          for (int t = 0; t < 16; ++t) v[t] = __uint_as_float(r[t]) * inv;
          if (cc + 16 <= p.d && ((reinterpret_cast<uintptr_t>(orow + cc) & 15) == 0)) {
            st_v4(orow + cc, make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                                        pack_bf16x2(v[6], v[7])));
            st_v4(orow + cc + 8, make_uint4(pack_bf16x2(v[8], v[9]), pack_bf16x2(v[10], v[11]),
                                            pack_bf16x2(v[12], v[13]), pack_bf16x2(v[14], v[15])));
          } else {
#pragma unroll
            for (int t = 0; t < 16; ++t)
              if (cc + t < p.d) orow[cc + t] = f2b(v[t]);
          }
        }
      }
      tc_fence_before();
    }
  } else if (warp >= 4) {
    // no visible keys at all: the reference's softmax over an empty set never happens in practice; write zeros
    const int m = qt * kTile + (warp & 3) * 32 + lane;
    if (m < p.nq_rows) {
      bf16* orow = p.out + (size_t)(seg.out_row0 + m / p.group) * p.out_ld + (size_t)(blk * p.group + (m % p.group)) * p.d;
      for (int c = 0; c < p.d; ++c) orow[c] = f2b(0.f);
    }
  }
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace

int op_attention(const AttnProblem& p, cudaStream_t st) {
  AHA_REQUIRE(p.d_pad % 16 == 0 && p.d_pad >= 16 && p.d_pad <= 128 && p.d <= p.d_pad, "attention: head_dim must pad to <= 128");
  AHA_REQUIRE(p.n_segs > 0 && p.n_segs <= 32 && p.n_blocks > 0 && p.nq_rows > 0 && p.group > 0, "attention: bad problem");
  const int DB = (p.d_pad + 63) / 64;
  CUtensorMap tq, tk, tv;
  int rc = get_tmap(p.q, p.q_rows_total, p.q_ld, p.q_ld, kTile, &tq); if (rc) return rc;
  rc = get_tmap(p.k, (int)p.k_rows_total, p.k_ld, p.k_ld, kTile, &tk); if (rc) return rc;
  rc = get_tmap(p.v, (int)p.v_rows_total, p.v_ld, p.v_ld, kTile, &tv); if (rc) return rc;
  const int smem = AttnSmem::total(DB);
  static int attr_smem = 0;
  if (attr_smem < smem) {
    AHA_CUDA(cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_smem = smem;
  }
  // MN-major V descriptor strides: LBO = bytes between 64-column blocks (one [128][64] box), SBO = 8 keys * 128 B.
  uint32_t v_lbo = kAtomBytes, v_sbo = 1024;
  if (const char* e = getenv("AHA_DEBUG_VDESC")) {  // bring-up aid: "lbo,sbo" in bytes
    unsigned a = 0, b = 0;
    if (sscanf(e, "%u,%u", &a, &b) == 2) { v_lbo = a; v_sbo = b; }
  }
  dim3 grid((p.nq_rows + kTile - 1) / kTile, p.n_blocks, p.n_segs);
  attention_kernel<<<grid, kAttnThreads, smem, st>>>(tq, tk, tv, p, v_lbo, v_sbo);
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}
