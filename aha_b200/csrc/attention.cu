// Fused attention on the 5th-gen tensor cores: S = Q.K^T and O += P.V are tcgen05.mma tiles with fp32 accumulators
// in TMEM, Q/K/V tiles are staged by TMA (128-byte swizzle), the online softmax runs in registers between the two.
//
// One kernel serves both attention shapes of the ingest-and-score path:
//   * frame-chunk versus KV cache (transformers 4.49 Qwen2Attention called at video_head_live_llava_qwen.py:160-169):
//     per (stream, kv head) the q*group query rows (GQA heads folded into the M dimension) against the cache's
//     physical ring slots; the reference's mask rule allowed(i, j) <=> logical(j) <= off + i is
//     evaluated from the ring geometry, so keys are streamed in PHYSICAL order (softmax is order independent) and an
//     eviction never has to move them.  A chunk has only a few query tiles, so the key axis is split over up to 8 CTAs
//     (flash-decoding style); the partial (O, m, l) are parked column-major in an L2-resident workspace, the splits
//     meet at an arrival counter and merge the 16-column blocks c = split (mod splits) in split order (deterministic).
//   * SigLIP self-attention (siglip_encoder.py:197-239): per (frame, head) 729 x 729, head_dim 72 padded to 80 columns.
//
// CTA = one 128-row query tile of one (segment, head block, key split).  Warp roles:
//   warp 0    TMA producer of the K ring (three stages of 128 keys)
//   warp 3    TMA producer of the V ring (a K stage is free after Q.K^T, a V stage only after P.V)
//   warp 1    MMA issuer (one elected lane): S[b] = Q.K^T for tile kt+1 is issued before P.V of tile kt, so the
//             tensor pipe works on the next scores while the softmax warps are busy with the current ones.  BOTH products
//             take their A operand from TENSOR MEMORY: Q is written there once per CTA (64 columns of packed bf16 pairs,
//             straight from global memory), P(kt) overwrites the first 64 columns of S(kt) - shared memory carries only
//             K and V
//   warp 2    TMEM allocator (512 columns: S0 | S1 | O | Q)
//   warps 4-11 softmax / correction / epilogue: TWO threads per query row (lane l of warps w and w + 4 share TMEM lane
//             quadrant w & 3 and own 64 of the tile's 128 key columns each; row maximum and sum are exchanged through
//             shared memory).  O is rescaled in TMEM only when the running maximum grows by more than 2^8 (lazy rescale),
//             which keeps the correction off the critical path.  The unsplit epilogue leaves through a shared-memory tile
//             as 16-byte pieces in row order.
// Measured on B200 (profiles/r02_attention_notes.txt): the loop runs at about 1900 cycles per 128 x 128 tile against the
// 1024 cycles the 16 384 exponentials of a tile need on the XU pipe (16 per clock per SM, measured); what was tried on the
// way: two P tiles, key-tile ping-pong of two softmax groups with two O accumulators, TMEM operands (this version).
#include "ops.h"
#include "gemm.h"
#include "ptx_sm100.cuh"
#include <cstdio>
#include <cstdlib>

using namespace sm100;

namespace {

constexpr int kTile = 128;          // query rows per CTA == keys per tile
constexpr int kAtomBytes = 16384;   // one [128 rows][64 bf16] swizzled box
constexpr int kStagesKV = 3;
constexpr int kPBufs = 2;           // P tiles (tensor memory: the first 64 columns of the two score buffers)
constexpr int kAttnThreads = 384;   // warps 0-3: TMA / MMA / TMEM, warps 4-11: softmax (two threads per query row)
constexpr uint32_t kColS0 = 0, kColS1 = 128, kColO = 256, kColQ = 384;   // Q: 64 columns of packed bf16 pairs
constexpr float kRescaleThreshold = 8.f;  // log2 domain

struct AttnSmem {
  // offsets inside the 1024-aligned dynamic shared memory: only the K and V rings live here - Q and P are tensor-memory
  // operands of their MMAs (see the kernel), which left room for a third stage of each ring
  static __host__ __device__ int k_off(int DB) { return 0; }
  static __host__ __device__ int v_off(int DB) { return DB * kAtomBytes * kStagesKV; }
  static __host__ __device__ int bar_off(int DB) { return DB * kAtomBytes * 2 * kStagesKV; }
  static __host__ __device__ int xch_off(int DB) { return bar_off(DB) + 256; }      // float [2][2][128] row-statistics exchange
  // no alignment slack: the dynamic window of a kernel without static shared memory starts 1024-aligned (checked at entry)
  static __host__ __device__ int total(int DB) { return xch_off(DB) + 2048; }
};

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// bit t of the result: key slot p0 + t is visible to a query whose logical limit is `limit`
__device__ __noinline__ uint32_t visible_mask32(const AttnSeg& seg, int p0, int limit) {
  uint32_t vm = 0;
#pragma unroll 1
  for (int t = 0; t < 32; ++t) {
    const int pslot = p0 + t;
    int j = pslot;
    if (seg.ring_R > 0 && pslot >= seg.ring_S) {
      j = pslot - seg.ring_S - seg.ring_base;
      if (j < 0) j += seg.ring_R;
      j += seg.ring_S;
    }
    if (pslot < seg.klen && j <= limit) vm |= 1u << t;
  }
  return vm;
}

__global__ void __launch_bounds__(kAttnThreads, 1)
attention_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                 const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ AttnProblem p, const int splits) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem_raw) & 1023u) != 0) { if (threadIdx.x == 0) printf("aha_b200: attention shared memory is not 1024-byte aligned\n"); __trap(); }
  const int DB = (p.d_pad + 63) >> 6;
  uint8_t* sK = smem + AttnSmem::k_off(DB);
  uint8_t* sV = smem + AttnSmem::v_off(DB);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + AttnSmem::bar_off(DB));
  // K and V have their own rings: a K stage is free again as soon as its Q.K^T has completed (one whole softmax period
  // before the matching P.V), so its refill has more time to land
  uint64_t* q_full = bars;                          // 1 (256 arrivals: the softmax threads store Q into tensor memory)
  uint64_t* k_full = bars + 1;                      // kStagesKV
  uint64_t* k_empty = k_full + kStagesKV;           // kStagesKV
  uint64_t* s_full = k_empty + kStagesKV;           // 2
  uint64_t* p_full = s_full + 2;                    // kPBufs (256 arrivals each)
  uint64_t* pv_done = p_full + kPBufs;              // kPBufs: P.V of tile t commits to pv_done[t & 1]
  uint64_t* v_full = pv_done + kPBufs;              // kStagesKV
  uint64_t* v_empty = v_full + kStagesKV;           // kStagesKV
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(v_empty + kStagesKV);
  static_assert((7 + 4 * kStagesKV) * 8 + 4 <= 256, "barrier block");
  float* xch = reinterpret_cast<float*>(smem + AttnSmem::xch_off(DB));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned long long* ts = (p.dbg_ts && threadIdx.x == 128 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) ? p.dbg_ts : nullptr;
  dbg_stamp(ts, 0);
  const int seg_idx = blockIdx.z / splits, split = blockIdx.z % splits;
  const AttnSeg& seg = p.segs[seg_idx];
  const int blk = blockIdx.y;
  const int qt = blockIdx.x;
  const int nkt_all = (seg.klen + kTile - 1) / kTile;
  const int per = (nkt_all + splits - 1) / splits;
  const int kt0 = split * per;
  const int nkt = max(0, min(nkt_all, kt0 + per) - kt0);   // key tiles of this CTA: kt0 .. kt0 + nkt - 1
  const int stage_bytes = DB * kAtomBytes;
  // K / V tensor maps: per segment (device memory) when the streams of a batched launch own separate cache slabs
  const void* tmk = seg.k_tmap ? seg.k_tmap : static_cast<const void*>(&tmap_k);
  const void* tmv = seg.v_tmap ? seg.v_tmap : static_cast<const void*>(&tmap_v);

  if (warp == 0 && lane == 0) { prefetch_tensormap(&tmap_q); prefetch_tensormap(tmk); prefetch_tensormap(tmv); }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 256);
    for (int s = 0; s < kStagesKV; ++s) {
      mbar_init(&k_full[s], 1); mbar_init(&k_empty[s], 1); mbar_init(&v_full[s], 1); mbar_init(&v_empty[s], 1);
    }
    mbar_init(&s_full[0], 1); mbar_init(&s_full[1], 1);
    for (int b = 0; b < kPBufs; ++b) { mbar_init(&p_full[b], 256); mbar_init(&pv_done[b], 1); }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  // chunk-versus-cache attention is latency-bound and leaves HBM idle: warp 3 (otherwise unused) requests this CTA's
  // slice of the weights the following GEMMs will stream (independent of every earlier kernel, so before pdl_wait)
  if (warp == 3 && lane == 0 && p.pf_early)
    l2_prefetch_slice(p.pf, (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x,
                      gridDim.x * gridDim.y * gridDim.z);
  // Chunk versus cache (key-split launches): this CTA's K / V tiles are a frame old - in HBM, not in L2 - except for the
  // rows the kernel in front of this one is writing right now.  Requesting the tiles into L2 BEFORE griddepcontrol.wait
  // overlaps their HBM latency with that kernel (the new rows land in the same L2 lines, so the TMA loads after the wait
  // read current data).
  static_assert(kStagesKV >= 1, "");
  if (warp == 3 && lane == 0 && p.kv_prefetch) {
    const int k_row = seg.k_row0 + blk * p.k_head_rows, k_col = p.k_col0 + blk * p.k_col_step;
    const int v_row = seg.v_row0 + blk * p.k_head_rows, v_col = p.v_col0 + blk * p.v_col_step;
    // (at most kv_prefetch tiles ahead: a long key axis is requested in a rolling window by the producers below - all of
    // it at once would not fit the L2 next to the other CTAs' windows)
    for (int it = 0; it < min(nkt, p.kv_prefetch); ++it) {
      const int key0 = (kt0 + it) * kTile;
      for (int b = 0; b < DB; ++b) {
        tma_prefetch_2d(tmk, k_col + b * 64, k_row + key0);
        tma_prefetch_2d(tmv, v_col + b * 64, v_row + key0);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  dbg_stamp(ts, 1);
  pdl_wait();                 // Q, K and V were all written by earlier kernels on the stream
  pdl_launch_dependents();
  dbg_stamp(ts, 2);

  // softmax-thread state (meaningful in warps 4-11).  A query row is shared by two threads, lane l of warps w and w + 4
  // (same TMEM lane quadrant): `half` picks the 64 of the tile's 128 key columns the thread owns, and the 16-column
  // blocks of O it rescales / stores.  With one thread per row the four softmax warps - one per scheduler, every
  // latency exposed - were the critical path of the kernel (2.2 us per key tile against 0.3 us of tensor work).
  const int ew = warp & 3;
  const int half = warp >= 8 ? 1 : 0;
  const int row = ew * 32 + lane;            // tile row == TMEM lane
  const int m = qt * kTile + row;            // query row inside the (segment, block)
  const bool row_valid = m < p.nq_rows;
  const int tok = m / p.group;
  const uint32_t lane_addr = tmem_base + ((uint32_t)(ew * 32) << 16);
  float m_used = -INFINITY, l_sum = 0.f;

  if (nkt > 0) {
    if (warp == 0) {
      // ===================== TMA producer =====================
      if (elect_one()) {
        const int k_row = seg.k_row0 + blk * p.k_head_rows, k_col = p.k_col0 + blk * p.k_col_step;
        for (int it = 0; it < nkt; ++it) {
          const int s = it % kStagesKV;
          const uint32_t ph = (it / kStagesKV) & 1;
          // rolling L2 window: with two ring stages a tile is requested one tile period before it is needed, which is about
          // the HBM latency - the period then stretches to the latency (2.4 us per tile at 8192 keys, tensor pipe 28 % busy)
          if (p.kv_prefetch && it + p.kv_prefetch < nkt)
            for (int b = 0; b < DB; ++b) tma_prefetch_2d(tmk, k_col + b * 64, k_row + (kt0 + it + p.kv_prefetch) * kTile);
          mbar_wait(&k_empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&k_full[s], stage_bytes);
          const int key0 = (kt0 + it) * kTile;
          for (int b = 0; b < DB; ++b)
            tma_load_2d(sK + s * stage_bytes + b * kAtomBytes, tmk, &k_full[s], k_col + b * 64, k_row + key0);
        }
        if (!p.pf_early)
          l2_prefetch_slice(p.pf, (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x,
                            gridDim.x * gridDim.y * gridDim.z);
      }
    } else if (warp == 3) {
      // ===================== TMA producer of the V ring =====================
      if (elect_one()) {
        const int v_row = seg.v_row0 + blk * p.k_head_rows, v_col = p.v_col0 + blk * p.v_col_step;
        for (int it = 0; it < nkt; ++it) {
          const int s = it % kStagesKV;
          const uint32_t ph = (it / kStagesKV) & 1;
          if (p.kv_prefetch && it + p.kv_prefetch < nkt)
            for (int b = 0; b < DB; ++b) tma_prefetch_2d(tmv, v_col + b * 64, v_row + (kt0 + it + p.kv_prefetch) * kTile);
          mbar_wait(&v_empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&v_full[s], stage_bytes);
          const int key0 = (kt0 + it) * kTile;
          for (int b = 0; b < DB; ++b)
            tma_load_2d(sV + s * stage_bytes + b * kAtomBytes, tmv, &v_full[s], v_col + b * 64, v_row + key0);
        }
      }
    } else if (warp == 1) {
      // ===================== MMA issuer =====================
      if (elect_one()) {
        const uint32_t idesc_qk = make_idesc_bf16_f32(kTile, kTile);
        const uint32_t idesc_pv = make_idesc_bf16_f32_bmn(kTile, (uint32_t)p.d_pad);
        const int ksteps = p.d_pad >> 4;
        auto issue_qk = [&](int it) {
          const int s = it % kStagesKV;
          mbar_wait(&k_full[s], (it / kStagesKV) & 1);
          tc_fence_after();
          const uint32_t ka = smem_u32(sK + s * stage_bytes);
          for (int ks = 0; ks < ksteps; ++ks) {
            const uint32_t o = (uint32_t)(ks >> 2) * kAtomBytes + (uint32_t)(ks & 3) * 32;
            // Q is the A operand in tensor memory (8 columns per K = 16 step): the tensor pipe re-read the 32 KB Q tile from
            // shared memory for every key tile, a quarter of the kernel's shared-memory traffic
            umma_f16_ts(tmem_base + ((it & 1) ? kColS1 : kColS0), tmem_base + kColQ + (uint32_t)ks * 8u,
                        make_kmajor_sw128_desc(ka + o), idesc_qk, ks != 0);
          }
          umma_commit(&k_empty[s]);      // the K stage is free once these MMAs have read it
          umma_commit(&s_full[it & 1]);
        };
        mbar_wait(q_full, 0);
        issue_qk(0);
        for (int it = 0; it < nkt; ++it) {
          if (it + 1 < nkt) issue_qk(it + 1);
          const int s = it % kStagesKV;
          mbar_wait(&v_full[s], (it / kStagesKV) & 1);
          mbar_wait(&p_full[it & 1], (it >> 1) & 1);
          tc_fence_after();
          const uint32_t va = smem_u32(sV + s * stage_bytes);
          const uint32_t pt = tmem_base + ((it & 1) ? kColS1 : kColS0);     // P(it) sits in the first 64 columns of S(it)
          for (int ks = 0; ks < kTile / 16; ++ks) {
            // V tile is MN-major: 64-column blocks kAtomBytes apart (LBO), 8-key groups 1024 B apart (SBO)
            const uint64_t db = make_mnmajor_sw128_desc(va + (uint32_t)ks * 2048, kAtomBytes, 1024);
            umma_f16_ts(tmem_base + kColO, pt + (uint32_t)ks * 8u, db, idesc_pv, (it | ks) != 0);
          }
          umma_commit(&v_empty[s]);
          umma_commit(&pv_done[it & 1]);
        }
      }
    } else if (warp >= 4) {
      // ===================== softmax / correction =====================
      {
        // This row of Q goes from global memory straight into tensor memory (no TMA, no shared memory): thread `half` of the
        // row packs elements [64 half, 64 half + 64) into 32 columns; elements beyond d_pad and rows beyond the matrix are 0.
        const long long q_row = (long long)seg.q_row0 + (long long)blk * p.q_block_rows + (long long)qt * kTile + row;
        const bf16* qsrc = p.q + q_row * (long long)p.q_ld + (p.q_col0 + blk * p.q_col_step) + half * 64;
        const bool rv = q_row < (long long)p.q_rows_total;
        uint32_t qv[32];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          uint4 u = make_uint4(0u, 0u, 0u, 0u);
          if (rv && half * 64 + c * 8 < p.d_pad) u = *reinterpret_cast<const uint4*>(qsrc + c * 8);
          qv[4 * c] = u.x; qv[4 * c + 1] = u.y; qv[4 * c + 2] = u.z; qv[4 * c + 3] = u.w;
        }
        const uint32_t qaddr = lane_addr + kColQ + (uint32_t)half * 32u;
        tmem_st_32x32b_x16(qaddr, *reinterpret_cast<uint32_t(*)[16]>(&qv[0]));
        tmem_st_32x32b_x16(qaddr + 16, *reinterpret_cast<uint32_t(*)[16]>(&qv[16]));
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(q_full);
      }
      const int limit = seg.off + tok;           // logical key index must be <= limit
      const float sl2 = p.scale * 1.4426950408889634f;
      const bool all_visible = seg.off >= seg.klen - 1;
      // Two P tiles: the probabilities of tile t go to buffer t & 1, so the exponentials of tile t + 1 are computed and
      // stored while the tensor pipe still multiplies P(t) by V(t).  With one buffer every tile paid for softmax, P.V and
      // two barrier round trips in sequence (4460 cycles per tile at 8192 keys, tensor pipe 28 % busy, XU pipe 28 %).
      for (int it = 0; it < nkt; ++it) {
        const uint32_t p_addr = lane_addr + ((it & 1) ? kColS1 : kColS0);
        const int kt = kt0 + it;
        mbar_wait(&s_full[it & 1], (it >> 1) & 1);
        tc_fence_after();
        dbg_stamp(ts, 3 + 2 * min(it, 1));
        const int col0 = half * 64;
        const uint32_t s_addr = lane_addr + ((it & 1) ? kColS1 : kColS0) + col0;
        uint32_t sv[64];
        tmem_ld_32x32b_x32p(s_addr, sv);
        tmem_ld_32x32b_x32p(s_addr + 32, sv + 32);
        tmem_ld_wait();
        if (all_visible && kt * kTile + kTile > seg.klen) {
          // ragged last tile without a causal limit (SigLIP's 729 keys, a full cache): only slots >= klen are masked
          const int nvalid = seg.klen - kt * kTile - col0;
#pragma unroll
          for (int t = 0; t < 64; ++t)
            if (t >= nvalid) sv[t] = 0xff800000u;   // -inf
        } else if (!all_visible) {
          // slow path (growing cache: causal limit through the ring geometry): knock masked scores out before the statistics
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            const uint32_t vm = visible_mask32(seg, kt * kTile + col0 + c * 32, limit);
#pragma unroll
            for (int t = 0; t < 32; ++t)
              if (!((vm >> t) & 1u)) sv[c * 32 + t] = 0xff800000u;   // -inf
          }
        }
        // four independent maximum chains (one dependent chain of 32 three-input maxima was ~260 cycles per tile, with both
        // warps of the scheduler in it at the same time)
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
        for (int t = 0; t < 16; ++t) {
          mx0 = fmaxf(mx0, __uint_as_float(sv[t]));
          mx1 = fmaxf(mx1, __uint_as_float(sv[16 + t]));
          mx2 = fmaxf(mx2, __uint_as_float(sv[32 + t]));
          mx3 = fmaxf(mx3, __uint_as_float(sv[48 + t]));
        }
        float tmax = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
        {
          // row maximum over both halves: the two threads of a row then take identical rescale decisions.  The slots
          // alternate with the tile parity; the pair barrier of tile it + 1 orders their reuse at tile it + 2.
          float* x = xch + (it & 1) * 256;
          x[half * 128 + row] = tmax;
          asm volatile("bar.sync %0, 64;" ::"r"(2 + ew) : "memory");
          tmax = fmaxf(tmax, x[(half ^ 1) * 128 + row]);
        }
        tmax *= sl2;                                 // scale > 0: max commutes with the scaling
        // one compare decides (m_used = -inf: -inf + 8 = -inf, so any finite tmax passes; tmax = -inf never does); the whole
        // slow path sits behind one warp vote
        const bool rescale = tmax > m_used + kRescaleThreshold;
        float alpha = 1.f;
        if (__any_sync(0xffffffffu, rescale)) {
        if (rescale) {
          alpha = (m_used == -INFINITY) ? 0.f : ex2_approx(m_used - tmax);
          m_used = tmax;
        }
        // (P(it) overwrites the first half of S(it) in tensor memory: both threads of the row have their scores in registers
        // - the pair barrier above - and the P.V product that read P(it - 2) from these columns ran before Q.K^T(it) did)
        if (it > 0) {
          // O may only be rescaled between two P.V products: wait for the previous one (the MMAs complete in order) ...
          mbar_wait(&pv_done[(it - 1) & 1], ((it - 1) >> 1) & 1);
          tc_fence_after();
#pragma unroll 1
          for (int cc = half * 16; cc < p.d_pad; cc += 32) {
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
        }
        // probabilities -> bf16 -> swizzled K-major P tile
        const float mref = (m_used == -INFINITY) ? 0.f : m_used;
        float acc_l = 0.f;
#pragma unroll
        for (int cc = 0; cc < 64; cc += 32) {
          uint32_t pk[16];
#pragma unroll
          for (int t = 0; t < 32; t += 2) {
            const float x0 = ex2_approx(fmaf(__uint_as_float(sv[cc + t]), sl2, -mref));
            const float x1 = ex2_approx(fmaf(__uint_as_float(sv[cc + t + 1]), sl2, -mref));
            acc_l += x0 + x1;
            pk[t >> 1] = pack_bf16x2(x0, x1);
          }
          // keys 64 half + cc .. + 31 of this row = 16 columns of packed bf16 pairs
          tmem_st_32x32b_x16(p_addr + (uint32_t)(half * 32 + (cc >> 1)), pk);
        }
        l_sum += acc_l;
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&p_full[it & 1]);
        dbg_stamp(ts, 4 + 2 * min(it, 1));
      }
      {
        // row sum = half 0's partial + half 1's partial, added in that order by both threads
        float* x = xch + (nkt & 1) * 256;
        x[half * 128 + row] = l_sum;
        asm volatile("bar.sync %0, 64;" ::"r"(2 + ew) : "memory");
        l_sum = x[row] + x[128 + row];
      }
      mbar_wait(&pv_done[(nkt - 1) & 1], ((nkt - 1) >> 1) & 1);   // O complete
      tc_fence_after();
      dbg_stamp(ts, 7);
    }
  }

  if (splits == 1) {
    if (warp >= 4) {
      // epilogue straight from TMEM: O / l -> bf16 -> out[token, (blk*group + r)*d + c]
      const float inv = l_sum > 0.f ? 1.f / l_sum : 0.f;
      bf16* orow = p.out + (size_t)(seg.out_row0 + tok) * p.out_ld + (size_t)(blk * p.group + (m % p.group)) * p.d;
      // Staged form (head_dim a multiple of 8, 16-byte aligned rows): the quadrant's 32 x d tile goes through shared
      // memory (the K ring, idle by now) and leaves as 16-byte pieces in row order - d/8 consecutive pieces per output
      // row.  One 32-byte store per lane straight from TMEM is 32 separate lines per instruction and cost 2.9 us per CTA.
      const bool staged = nkt > 0 && (p.d & 7) == 0 && (p.out_ld & 7) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15) == 0;
      if (staged) {
        const int pitch = p.d_pad * 2 + 16;
        uint8_t* tile = sK + ew * 32 * pitch;
        dbg_stamp(ts, 13);
#pragma unroll 1
        for (int cc = half * 16; cc < p.d_pad; cc += 32) {
          uint32_t r[16];
          tmem_ld_32x32b_x16(lane_addr + kColO + cc, r);
          tmem_ld_wait();
          uint32_t w[8];
#pragma unroll
          for (int t = 0; t < 8; ++t) w[t] = pack_bf16x2(__uint_as_float(r[2 * t]) * inv, __uint_as_float(r[2 * t + 1]) * inv);
          uint4* dst = reinterpret_cast<uint4*>(tile + lane * pitch + cc * 2);
          dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
          dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
        }
        dbg_stamp(ts, 14);
        asm volatile("bar.sync %0, 64;" ::"r"(2 + ew) : "memory");
        const int ppr = p.d >> 3;                      // 16-byte pieces per output row
        for (int i = half * 32 + lane; i < 32 * ppr; i += 64) {
          const int r_in = i / ppr, pc = i - r_in * ppr;
          const int mm = qt * kTile + ew * 32 + r_in;
          if (mm < p.nq_rows) {
            const int tk = mm / p.group;
            bf16* o = p.out + (size_t)(seg.out_row0 + tk) * p.out_ld + (size_t)(blk * p.group + (mm - tk * p.group)) * p.d + pc * 8;
            st_v4(o, *reinterpret_cast<const uint4*>(tile + r_in * pitch + pc * 16));
          }
        }
        dbg_stamp(ts, 11);
        tc_fence_before();
      } else {
#pragma unroll 1
      for (int cc = half * 16; cc < p.d_pad; cc += 32) {
        uint32_t r[16];
        if (cc == 0) dbg_stamp(ts, 13);
        if (nkt > 0) {
          tmem_ld_32x32b_x16(lane_addr + kColO + cc, r);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int t = 0; t < 16; ++t) r[t] = 0u;
        }
        if (cc == 0) dbg_stamp(ts, 14);
        if (cc == 32) dbg_stamp(ts, 15);
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
#pragma unroll 4
            for (int t = 0; t < 16; ++t)
              if (cc + t < p.d) orow[cc + t] = f2b(v[t]);
          }
        }
      }
      dbg_stamp(ts, 11);
      tc_fence_before();
      }
    }
  } else {
    // ---- key-split merge through an L2-resident workspace, deterministic split order ----
    // (DSMEM moved these 72 KB per CTA at ~4 B/clk; L2 does it an order of magnitude faster and needs no cluster.)
    // Every split parks its unnormalised O (fp32, column-major [d_pad][128]) and (m, l), arrives at the counter of its
    // (segment, block, query tile), waits for the other splits - the host keeps the grid co-resident - and then merges
    // the 16-column chunks c with c mod splits == split.
    if (warp >= 4) {
      {   // zero this CTA's slice of the other counter bank (the next split launch on the stream counts in it)
        const int ncta = gridDim.x * gridDim.y * gridDim.z, cta = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        for (int z = cta * 256 + (int)threadIdx.x - 128; z < p.ws_zero_n; z += ncta * 256) p.ws_zero[z] = 0;
      }
      const int group_lin = (seg_idx * gridDim.y + blk) * gridDim.x + qt;
      const size_t cta_floats = (size_t)kTile * p.d_pad + 2 * kTile;
      float* mine = p.ws_partial + ((size_t)group_lin * splits + split) * cta_floats;
#pragma unroll 1
      for (int cc = half * 32; cc < p.d_pad; cc += 64) {   // 32 columns per TMEM round trip (d_pad is a multiple of 16)
        uint32_t r[32];
        const bool two = cc + 32 <= p.d_pad;
        if (nkt > 0) {
          if (two) tmem_ld_32x32b_x32p(lane_addr + kColO + cc, r);
          else { uint32_t h[16]; tmem_ld_32x32b_x16(lane_addr + kColO + cc, h);
#pragma unroll
                 for (int t = 0; t < 16; ++t) r[t] = h[t]; }
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int t = 0; t < 32; ++t) r[t] = 0u;
        }
        // parked column-major ([d_pad][128 rows]): a warp's store (and the merge's load) of one column is one contiguous
        // 128-byte segment; row-major float4 pieces were 32 separate lines per instruction and cost 2.7 us per CTA
        float* dst = mine + (size_t)cc * kTile + row;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j < 16 || two) __stcg(dst + (size_t)j * kTile, __uint_as_float(r[j]));
      }
      if (half == 0) {
        __stcg(mine + (size_t)kTile * p.d_pad + row, m_used);
        __stcg(mine + (size_t)kTile * p.d_pad + kTile + row, l_sum);
      }
      tc_fence_before();
      dbg_stamp(ts, 8);
      __threadfence();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      bool merge_all = false;
      if (p.merge_last) {
        // grids that are not co-resident: nobody waits - the last split to arrive merges every chunk of the group
        int* cnt = p.ws_pcounters + group_lin;
        if (threadIdx.x == 128) {
          const int old = atomicAdd(cnt, 1);
          const int last = (old == splits - 1) ? 1 : 0;
          if (last) *cnt = 0;                       // self-resetting
          xch[0] = __int_as_float(last);
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        merge_all = __float_as_int(xch[0]) != 0;
      } else {
        int* cnt = p.ws_counters + group_lin;
        if (threadIdx.x == 128) {
          atomicAdd(cnt, 1);
          uint32_t spins = 0;
          while (*reinterpret_cast<volatile int*>(cnt) < splits) {
            if (++spins > (1u << 24)) { printf("aha_b200: attention split arrival wait timed out (group %d)\n", group_lin); __trap(); }
          }
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
      }
      if (!p.merge_last || merge_all) {
      const int c_first = merge_all ? 0 : split, c_step = merge_all ? 1 : splits;
      __threadfence();
      dbg_stamp(ts, 9);
      const float* grp = p.ws_partial + (size_t)group_lin * splits * cta_floats;
      // all loads of a step are issued before the first use (splits beyond `splits` alias split 0 with weight 0); the O
      // pieces of this CTA's first chunk are requested together with the statistics: one L2 round trip, not two
      // (a 16-column chunk is shared by the two threads of a row: 8 columns each)
      float v[8][8];
      if (c_first < (p.d_pad >> 4)) {
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          const float* peer = grp + (size_t)(s < splits ? s : 0) * cta_floats + (size_t)(c_first * 16 + half * 8) * kTile + row;
#pragma unroll
          for (int j = 0; j < 8; ++j) v[s][j] = __ldcg(peer + (size_t)j * kTile);
        }
      }
      float ms[8], ls[8];
#pragma unroll
      for (int s = 0; s < 8; ++s) {
        const float* st = grp + (size_t)(s < splits ? s : 0) * cta_floats + (size_t)kTile * p.d_pad;
        ms[s] = __ldcg(st + row);
        ls[s] = __ldcg(st + kTile + row);
      }
      float mx = -INFINITY;
#pragma unroll
      for (int s = 0; s < 8; ++s) if (s < splits) mx = fmaxf(mx, ms[s]);
      float wgt[8];
      float L = 0.f;
#pragma unroll
      for (int s = 0; s < 8; ++s) {
        wgt[s] = (s < splits && ms[s] != -INFINITY) ? ex2_approx(ms[s] - mx) : 0.f;
        L += wgt[s] * ls[s];
      }
      const float inv = L > 0.f ? 1.f / L : 0.f;
      bf16* orow = p.out + (size_t)(seg.out_row0 + tok) * p.out_ld + (size_t)(blk * p.group + (m % p.group)) * p.d;
#pragma unroll 1
      for (int c = c_first; c < (p.d_pad >> 4); c += c_step) {
        const int cc = c * 16 + half * 8;
        if (c != c_first) {
#pragma unroll
          for (int s = 0; s < 8; ++s) {
            const float* peer = grp + (size_t)(s < splits ? s : 0) * cta_floats + (size_t)cc * kTile + row;
#pragma unroll
            for (int j = 0; j < 8; ++j) v[s][j] = __ldcg(peer + (size_t)j * kTile);
          }
        }
        float acc[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) acc[t] = 0.f;
#pragma unroll
        for (int s = 0; s < 8; ++s) {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] += wgt[s] * v[s][j];
        }
        if (row_valid) {
          if (cc + 8 <= p.d && ((reinterpret_cast<uintptr_t>(orow + cc) & 15) == 0)) {
#pragma unroll
            for (int t = 0; t < 8; ++t) acc[t] *= inv;
            st_v4(orow + cc, make_uint4(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]),
                                        pack_bf16x2(acc[4], acc[5]), pack_bf16x2(acc[6], acc[7])));
          } else {
#pragma unroll 4
            for (int t = 0; t < 8; ++t)
              if (cc + t < p.d) orow[cc + t] = f2b(acc[t] * inv);
          }
        }
      }
      dbg_stamp(ts, 10);
      }
    }
  }

  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
  dbg_stamp(ts, 12);
}

}  // namespace

static unsigned long long* g_attn_ts = nullptr;
extern "C" int aha_debug_attention_timeline(unsigned long long* out16) {
  if (!g_attn_ts) { aha_set_error("ValueError: run with AHA_TS=1"); return AHA_ERR_VALUE; }
  AHA_CUDA(cudaDeviceSynchronize());
  AHA_CUDA(cudaMemcpy(out16, g_attn_ts, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  return AHA_OK;
}

int op_attention(const AttnProblem& p_in, cudaStream_t st) {
  // the heavy shapes (a vision batch; frame chunks of several streams against long windows) run on the persistent kernel
  if (!(getenv("AHA_TS") && atoi(getenv("AHA_TS")) == 1)) {
    const int taken = op_attention_persistent(p_in, st);
    if (taken != 0) return taken < 0 ? taken : AHA_OK;
  }
  AttnProblem p = p_in;
  static const bool want_ts = getenv("AHA_TS") && atoi(getenv("AHA_TS")) == 1;
  if (want_ts) {
    if (!g_attn_ts) { AHA_CUDA(cudaMalloc(&g_attn_ts, 16 * sizeof(unsigned long long))); AHA_CUDA(cudaMemset(g_attn_ts, 0, 128)); }
    p.dbg_ts = g_attn_ts;
  }
  AHA_REQUIRE(p.d_pad % 16 == 0 && p.d_pad >= 16 && p.d_pad <= 128 && p.d <= p.d_pad, "attention: head_dim must pad to <= 128");
  AHA_REQUIRE(p.n_segs > 0 && p.n_segs <= 32 && p.n_blocks > 0 && p.nq_rows > 0 && p.group > 0, "attention: bad problem");
  const int DB = (p.d_pad + 63) / 64;
  CUtensorMap tq, tk, tv;
  int rc = get_tmap(p.q, p.q_rows_total, p.q_ld, p.q_ld, kTile, &tq); if (rc) return rc;
  rc = get_tmap(p.k, (int)p.k_rows_total, p.k_ld, p.k_ld, kTile, &tk); if (rc) return rc;
  rc = get_tmap(p.v, (int)p.v_rows_total, p.v_ld, p.v_ld, kTile, &tv); if (rc) return rc;
  const int smem = AttnSmem::total(DB);
  static int attr_smem[16] = {0};          // function attributes and the SM count are per device
  static int sm_counts[16] = {0};
  int dev = 0;
  AHA_CUDA(cudaGetDevice(&dev));
  AHA_REQUIRE(dev >= 0 && dev < 16, "attention: device index out of range");
  if (attr_smem[dev] < smem) {
    AHA_CUDA(cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_smem[dev] = smem;
  }
  // key-axis split: a frame chunk has only a handful of (query tile, kv head) pairs, far fewer than SMs
  const int q_tiles = (p.nq_rows + kTile - 1) / kTile;
  const int ctas = q_tiles * p.n_blocks * p.n_segs;
  int max_tiles = 0;
  for (int i = 0; i < p.n_segs; ++i) max_tiles = max_tiles > (p.segs[i].klen + kTile - 1) / kTile ? max_tiles : (p.segs[i].klen + kTile - 1) / kTile;
  // the splits of a group wait for each other, so the whole grid must be resident at once: one CTA per SM
  if (!sm_counts[dev]) AHA_CUDA(cudaDeviceGetAttribute(&sm_counts[dev], cudaDevAttrMultiProcessorCount, dev));
  const int sm_count = sm_counts[dev];
  int splits = 1;
  if (ctas < sm_count && max_tiles > 1 && !p.no_split) {
    splits = sm_count / ctas;
    if (const char* e = getenv("AHA_ATTN_SPLITS")) { const int v = atoi(e); if (v >= 1 && v < splits) splits = v; }
    if (splits > 8) splits = 8;
    if (splits > max_tiles) splits = max_tiles;
    const int per = (max_tiles + splits - 1) / splits;
    splits = (max_tiles + per - 1) / per;       // no empty splits for the longest segment
  }
  // Long key axes with too few (query tile, head) pairs to fill the device evenly (8 streams x 4 kv heads x 3 query tiles
  // = 96 CTAs of 64 key tiles each at a window of 8192): split further than co-residency allows and let the last split to
  // arrive merge (nobody waits).  Cost model: waves x key tiles per CTA, one tile's worth of time for parking + merging.
  p.merge_last = 0;
  if (!p.no_split && max_tiles >= 16 && ctas <= GemmWorkspace::kPCounters / 2) {
    long best = (long)((ctas * splits + sm_count - 1) / sm_count) * ((max_tiles + splits - 1) / splits) + (splits > 1 ? 1 : 0);
    int best_s = splits;
    for (int s = 2; s <= 8 && s <= max_tiles; ++s) {
      const int per = (max_tiles + s - 1) / s;
      if ((max_tiles + per - 1) / per != s) continue;
      const long cost = (long)((ctas * s + sm_count - 1) / sm_count) * per + 1;
      if (cost < best) { best = cost; best_s = s; }
    }
    if (const char* e = getenv("AHA_ATTN_SPLITS_LONG")) { const int v = atoi(e); if (v >= 1 && v <= 8) best_s = v; }
    if (best_s != splits && ctas * best_s > sm_count) { splits = best_s; p.merge_last = 1; }
  }
  if (splits > 1) {
    const size_t need = (size_t)ctas * splits * ((size_t)kTile * p.d_pad + 2 * kTile) * sizeof(float);
    if (need > GemmWorkspace::kPartialBytes || (!p.merge_last && ctas > GemmWorkspace::kCounters)) { splits = 1; p.merge_last = 0; }
  }
  if (splits > 1) {
    GemmSplitScratch sc;
    rc = aha_split_scratch(st, &sc, !p.merge_last); if (rc) return rc;     // (advances the stream's counter bank: only for a real waiting-split launch)
    p.ws_partial = sc.partial; p.ws_counters = sc.bank; p.ws_zero = sc.zero; p.ws_zero_n = p.merge_last ? 0 : sc.zero_n;
    p.ws_pcounters = sc.pcounters;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(q_tiles, p.n_blocks, p.n_segs * splits);
  cfg.blockDim = dim3(kAttnThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = aha_pdl_enabled() ? 1 : 0;
  // K / V tiles requested into L2 ahead of the TMA ring (0: off); the short key axes of the single-stream step fit whole
  { static const int kvpf = getenv("AHA_KV_PF") ? atoi(getenv("AHA_KV_PF")) : 3; p.kv_prefetch = (splits > 1 || max_tiles >= 16) ? kvpf : 0; }
  // per-launch timing for bench.py's roofline leg: class 3 = many-tile batched attention (SigLIP), 4 = chunk versus cache
  ProfSlot* slot = prof_begin(ctas >= sm_count ? 3 : 4, st);
  AHA_CUDA(cudaLaunchKernelEx(&cfg, attention_kernel, tq, tk, tv, p, splits));
  ++g_aha_launch_count;
  if (slot) {
    double flops = 0, bytes = 0;
    for (int i = 0; i < p.n_segs; ++i) {
      flops += 4.0 * p.n_blocks * (double)p.nq_rows * p.segs[i].klen * p.d;
      bytes += 2.0 * p.n_blocks * (2.0 * p.nq_rows + 2.0 * p.segs[i].klen) * p.d;
    }
    prof_end(slot, st, bytes, flops);
  }
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// aha_attention_noalloc: aha_attention for callers that must not allocate (CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).  Same kernel; the key-split scratch - the parked fp32 partials and the arrival
// counters, which op_attention takes from the per-stream workspace it cudaMallocs on first use - lives in a file-scope
// device array that every GPU loading the library holds until exit.  No counterpart in the
// reference.
//   * limit: kNoAllocPartialBytes (8 MiB) of partials = (query tiles x kv heads x key splits) x (128 x head_dim + 256) x 4
//     bytes, and at most kNoAllocCounters (query tile, kv head) groups; larger problems are rejected with a ValueError.
//   * all calls share the one array, which keeps the previous call's contents: the counters are cleared on the stream
//     before every launch, the partials are fully overwritten before they are read.  Calls that overlap on the GPU
//     (different streams, or several host threads) race on it - serialise them.
//   * no cudaMalloc / cudaFree / allocating library call on this path; the launch is not timed by aha_profile_gemm.
constexpr size_t kNoAllocPartialBytes = 8u << 20;
constexpr int kNoAllocCounters = 1024;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocPartialBytes + kNoAllocCounters * sizeof(int)];

extern "C" int aha_attention_noalloc(const void* q, const void* k, const void* v, void* out, int heads, int kv_heads, int nq,
                                     int klen, int head_dim, int off, float scale, void* stream) {
  AHA_REQUIRE(q && k && v && out, "aha_attention_noalloc: NULL argument");
  AHA_REQUIRE(heads > 0 && kv_heads > 0 && heads % kv_heads == 0 && nq > 0 && klen >= 0, "aha_attention_noalloc: bad shape");
  AHA_REQUIRE(head_dim >= 64 && head_dim % 16 == 0 && head_dim <= 128,
              "aha_attention_noalloc: head_dim must be in {64..128}, multiple of 16");
  cudaStream_t st = (cudaStream_t)stream;
  // per-device set-up with that device current: in host code the array's name is only a stand-in, the kernel needs the
  // address cudaGetSymbolAddress returns for THIS device
  static unsigned char* scratch_dev[16] = {nullptr};
  static int sm_counts[16] = {0};
  static int attr_smem[16] = {0};
  int dev = 0;
  AHA_CUDA(cudaGetDevice(&dev));
  AHA_REQUIRE(dev >= 0 && dev < 16, "aha_attention_noalloc: device index out of range");
  if (!scratch_dev[dev]) {
    void* ptr = nullptr;
    AHA_CUDA(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
    AHA_REQUIRE(ptr != nullptr, "aha_attention_noalloc: no device address for the scratch array");
    AHA_CUDA(cudaDeviceGetAttribute(&sm_counts[dev], cudaDevAttrMultiProcessorCount, dev));
    scratch_dev[dev] = static_cast<unsigned char*>(ptr);
  }
  const int G = heads / kv_heads;
  AttnProblem p{};
  p.q = (const bf16*)q; p.q_rows_total = kv_heads * nq * G; p.q_ld = head_dim;
  p.k = (const bf16*)k; p.k_rows_total = (long long)kv_heads * klen; p.k_ld = head_dim;
  p.v = (const bf16*)v; p.v_rows_total = (long long)kv_heads * klen; p.v_ld = head_dim;
  p.q_block_rows = nq * G; p.k_head_rows = klen;
  p.n_blocks = kv_heads; p.nq_rows = nq * G; p.group = G; p.d = head_dim; p.d_pad = head_dim;
  p.scale = scale; p.out = (bf16*)out; p.out_ld = heads * head_dim;
  p.n_segs = 1;
  p.segs[0] = AttnSeg{0, 0, 0, klen, off, 0, 0, 0, 0, nullptr, nullptr};

  const int DB = (p.d_pad + 63) / 64;
  CUtensorMap tq, tk, tv;
  int rc = get_tmap(p.q, p.q_rows_total, p.q_ld, p.q_ld, kTile, &tq); if (rc) return rc;
  rc = get_tmap(p.k, (int)p.k_rows_total, p.k_ld, p.k_ld, kTile, &tk); if (rc) return rc;
  rc = get_tmap(p.v, (int)p.v_rows_total, p.v_ld, p.v_ld, kTile, &tv); if (rc) return rc;
  const int smem = AttnSmem::total(DB);
  if (attr_smem[dev] < smem) {
    AHA_CUDA(cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_smem[dev] = smem;
  }
  // key-axis split as in op_attention's co-resident case; the merge is always the last arriver's (nobody waits)
  const int q_tiles = (p.nq_rows + kTile - 1) / kTile;
  const int ctas = q_tiles * p.n_blocks;
  const int max_tiles = (klen + kTile - 1) / kTile;
  int splits = 1;
  if (ctas < sm_counts[dev] && max_tiles > 1) {
    splits = sm_counts[dev] / ctas;
    if (splits > 8) splits = 8;
    if (splits > max_tiles) splits = max_tiles;
    const int per = (max_tiles + splits - 1) / splits;
    splits = (max_tiles + per - 1) / per;       // no empty splits
  }
  if (splits > 1) {
    const size_t need = (size_t)ctas * splits * ((size_t)kTile * p.d_pad + 2 * kTile) * sizeof(float);
    if (need > kNoAllocPartialBytes || ctas > kNoAllocCounters) {
      aha_set_error("ValueError: aha_attention_noalloc: this problem needs " + std::to_string(need) + " bytes of key-split scratch for " +
                    std::to_string(ctas) + " (query tile, kv head) groups; the built-in array holds " +
                    std::to_string(kNoAllocPartialBytes) + " bytes and " + std::to_string(kNoAllocCounters) +
                    " groups - use aha_attention");
      return AHA_ERR_VALUE;
    }
    int* counters = reinterpret_cast<int*>(scratch_dev[dev] + kNoAllocPartialBytes);
    AHA_CUDA(cudaMemsetAsync(counters, 0, (size_t)ctas * sizeof(int), st));     // must start at zero, whatever the last call left
    p.merge_last = 1;
    p.ws_partial = reinterpret_cast<float*>(scratch_dev[dev]);
    p.ws_pcounters = counters;
    p.ws_counters = nullptr; p.ws_zero = nullptr; p.ws_zero_n = 0;
    p.kv_prefetch = 3;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(q_tiles, p.n_blocks, splits);
  cfg.blockDim = dim3(kAttnThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cfg.attrs = nullptr; cfg.numAttrs = 0;        // (no programmatic launch: the kernel follows a memset)
  AHA_CUDA(cudaLaunchKernelEx(&cfg, attention_kernel, tq, tk, tv, p, splits));
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}
