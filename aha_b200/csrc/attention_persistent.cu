// Persistent fused attention for the two heavy shapes of the ingest-and-score path (same math, masks and operand
// layouts as attention.cu, which stays the kernel of the small launches):
//   * SigLIP self-attention of a vision batch (siglip_encoder.py:197-239): thousands of (frame, head, query tile) groups
//     of six key tiles each.  One CTA per group spent more than a third of its life outside the key loop (barrier set-up,
//     tensor-memory allocation, the Q -> K -> first-score latency chain, the output epilogue, the launch gap to the next
//     CTA); here a CTA walks over many groups (dealt round-robin, see build_items) and every pipeline - K/V ring, score
//     buffers, Q, output - runs across the group boundaries.
//   * frame chunks of several streams against long KV windows (transformers 4.49 Qwen2Attention called at
//     video_head_live_llava_qwen.py:160-169; BASELINE config 3: 8 streams x 4 kv heads x 3 query tiles x 64 key tiles):
//     one team of CTAs per query tile; inside a team the (stream, kv head, key tile) units are dealt as equal contiguous
//     ranges (stream-K), so every SM gets the same number of key tiles; a (stream, head) shared by several CTAs is merged by
//     the last part to arrive, in part order.
//
// CTA = 512 threads, one per SM, tensor memory: S0 | S1 | O | Q0 | Q1 (512 columns).
//   warp 0      TMA producer of the K ring, warp 3 of the V ring (three stages of 128 keys each), L2 prefetch hints
//   warp 1      MMA issuer: two score tiles ahead - as soon as P(n) arrives it issues O += P(n).V and S(n+2) = Q.K^T back to
//               back, also across item boundaries; Q and P are tensor-memory operands (P(n) overwrites the first 64
//               columns of S(n))
//   warp 2      tensor-memory allocation, builds this CTA's work list in shared memory
//   warps 4-11  online softmax, two threads per query row (64 key columns each); they never touch O except for the rare
//               lazy rescale, and go straight on to the next item's first scores
//   warps 12-15 one thread per query row: load the NEXT item's Q rows from global memory into the spare Q buffer of
//               tensor memory, then finish the current item - O / l -> bf16 -> out, or park the partial (O, m, l) and, as
//               the last part to arrive, merge the group
#include "ops.h"
#include "gemm.h"
#include "ptx_sm100.cuh"
#include <cstdio>
#include <cstdlib>

using namespace sm100;

namespace {

constexpr int kTile = 128;
constexpr int kAtomBytes = 16384;   // one [128 rows][64 bf16] swizzled box
constexpr int kStages = 3;
constexpr int kThreads = 512;
constexpr int kMaxItems = 64;       // work items per CTA (shared-memory table)
constexpr int kMaxParts = 4;        // CTAs sharing one group in stream-K mode
constexpr uint32_t kColS0 = 0, kColS1 = 128, kColO = 256, kColQ = 384;   // Q(i) at kColQ + 64 (i & 1)
constexpr float kRescaleThreshold = 8.f;  // log2 domain

struct Item { int seg, blk, qt, kt0, nkt, parts, part, group_lin; };
static_assert(sizeof(Item) == 32, "Item");

struct PSmem {
  static __host__ __device__ int k_off(int DB) { return 0; }
  static __host__ __device__ int v_off(int DB) { return DB * kAtomBytes * kStages; }
  static __host__ __device__ int bar_off(int DB) { return DB * kAtomBytes * 2 * kStages; }
  static __host__ __device__ int xch_off(int DB) { return bar_off(DB) + 256; }        // float [2][2][128] row maxima of a pair
  static __host__ __device__ int stats_off(int DB) { return xch_off(DB) + 2048; }     // float [2][3][128]: l (half 0), l (half 1), m
  static __host__ __device__ int items_off(int DB) { return stats_off(DB) + 3072; }
  static __host__ __device__ int total(int DB) { return items_off(DB) + kMaxItems * (int)sizeof(Item); }
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

// This CTA's work list.  Both decompositions keep the CTAs that run AT THE SAME TIME on neighbouring data, like the wave
// order of a plain grid does: the tower's q/k/v buffer (161 MB for 32 frames) and the KV windows of eight streams (134 MB
// per layer) are larger than the L2, and every K / V tile is wanted by all query tiles of its head.  (Contiguous group
// ranges per CTA had 148 heads from all over the buffer in flight: 712 MB of DRAM reads per launch instead of 179 MB.)
//   mode 0: groups c, c + G, c + 2 G, ... (G = grid size): the 148 groups in flight are ~25 neighbouring heads, the six query
//           tiles of a head run side by side and share its K / V through the L2.
//   mode 1: the CTAs form q_tiles teams (team = c mod q_tiles takes query tile `team` of every (segment, kv head)); inside a
//           team the (segment, kv head, key tile) units are dealt as equal contiguous ranges (stream-K).  Rank r of every team
//           walks the same keys at the same time, so a K / V tile comes from DRAM once; a (segment, kv head) cut by a range
//           boundary is shared with the neighbouring ranks and merged by its last part.
__device__ __forceinline__ int build_items(const AttnProblem& p, const AttnPersistPlan& pl, Item* items, int lane) {
  const int G = (int)gridDim.x, c = (int)blockIdx.x;
  int count = 0;
  if (pl.mode == 0) {
    count = c < pl.n_groups ? min((pl.n_groups - 1 - c) / G + 1, kMaxItems) : 0;
    for (int j = lane; j < count; j += 32) {
      const int g = c + j * G;
      const int seg = g / pl.gps, rem = g - seg * pl.gps;
      const int blk = rem / pl.q_tiles, qt = rem - blk * pl.q_tiles;
      Item it{seg, blk, qt, 0, (p.segs[seg].klen + kTile - 1) / kTile, 1, 0, g};
      items[j] = it;
    }
  } else {
    const int R = G / pl.q_tiles, team = c % pl.q_tiles, rank = c / pl.q_tiles;
    const long long U = pl.units;
    if (rank < R) {
      const int u0 = (int)((long long)rank * U / R), u1 = (int)((long long)(rank + 1) * U / R);
      if (u0 < u1) {
        int s0 = 0;
        while (s0 + 1 < p.n_segs && pl.unit_base[s0 + 1] <= u0) ++s0;
        const int T0 = (p.segs[s0].klen + kTile - 1) / kTile;
        const int sg_first = s0 * p.n_blocks + (u0 - pl.unit_base[s0]) / T0;       // (segment, kv head) pairs in unit order
        const int n_sg = p.n_segs * p.n_blocks;
        for (int r = 0; r < kMaxItems / 32; ++r) {
          const int j = r * 32 + lane;
          const int sg = sg_first + j;
          bool valid = sg < n_sg;
          Item it{0, 0, 0, 0, 0, 1, 0, 0};
          if (valid) {
            const int seg = sg / p.n_blocks, blk = sg - seg * p.n_blocks;
            const int T = (p.segs[seg].klen + kTile - 1) / kTile;
            const int gs = pl.unit_base[seg] + blk * T, ge = gs + T;
            valid = gs < u1;
            if (valid) {
              const int a = max(gs, u0), b = min(ge, u1);
              const int first_rank = (int)(((long long)(gs + 1) * R + U - 1) / U) - 1;
              const int last_rank = (int)(((long long)ge * R + U - 1) / U) - 1;
              it.seg = seg; it.blk = blk; it.qt = team;
              it.kt0 = a - gs; it.nkt = b - a; it.parts = last_rank - first_rank + 1; it.part = rank - first_rank;
              it.group_lin = sg * pl.q_tiles + team;
              items[j] = it;
            }
          }
          count += __popc(__ballot_sync(0xffffffffu, valid));
        }
      }
    }
  }
  return count;
}

__global__ void __launch_bounds__(kThreads, 1)
attention_persistent_kernel(const __grid_constant__ CUtensorMap tmap_k, const __grid_constant__ CUtensorMap tmap_v,
                            const __grid_constant__ AttnProblem p, const __grid_constant__ AttnPersistPlan pl) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem_raw) & 1023u) != 0) { if (threadIdx.x == 0) printf("aha_b200: attention shared memory is not 1024-byte aligned\n"); __trap(); }
  const int DB = (p.d_pad + 63) >> 6;
  uint8_t* sK = smem + PSmem::k_off(DB);
  uint8_t* sV = smem + PSmem::v_off(DB);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + PSmem::bar_off(DB));
  uint64_t* k_full = bars;                   // kStages
  uint64_t* k_empty = k_full + kStages;      // kStages
  uint64_t* v_full = k_empty + kStages;      // kStages
  uint64_t* v_empty = v_full + kStages;      // kStages
  uint64_t* s_full = v_empty + kStages;      // 2: scores of global tile n in buffer n & 1
  uint64_t* p_full = s_full + 2;             // 2 (256 arrivals)
  uint64_t* pv_done = p_full + 2;            // 2: P.V of global tile n commits to pv_done[n & 1]
  uint64_t* q_full = pv_done + 2;            // 2 (128 arrivals): Q of item i in tensor-memory buffer i & 1
  uint64_t* stats_full = q_full + 2;         // 2 (256 arrivals): (l, m) of item i in stats[i & 1]
  uint64_t* stats_free = stats_full + 2;     // 2 (128 arrivals)
  uint64_t* o_full = stats_free + 2;         // 1: last P.V of an item committed
  uint64_t* o_free = o_full + 1;             // 1 (128 arrivals): the epilogue warps have O in registers
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_free + 1);
  int* n_items_slot = reinterpret_cast<int*>(tmem_slot + 1);
  int* last_flag = n_items_slot + 1;
  static_assert((4 * kStages + 14) * 8 + 12 <= 256, "barrier block");
  float* xch = reinterpret_cast<float*>(smem + PSmem::xch_off(DB));
  float* stats = reinterpret_cast<float*>(smem + PSmem::stats_off(DB));
  Item* items = reinterpret_cast<Item*>(smem + PSmem::items_off(DB));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stage_bytes = DB * kAtomBytes;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmap_k); prefetch_tensormap(&tmap_v);
    if (p.segs[0].k_tmap) { prefetch_tensormap(p.segs[0].k_tmap); prefetch_tensormap(p.segs[0].v_tmap); }
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&k_full[s], 1); mbar_init(&k_empty[s], 1); mbar_init(&v_full[s], 1); mbar_init(&v_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&s_full[b], 1); mbar_init(&p_full[b], 256); mbar_init(&pv_done[b], 1);
      mbar_init(&q_full[b], 128); mbar_init(&stats_full[b], 256); mbar_init(&stats_free[b], 128);
    }
    mbar_init(o_full, 1); mbar_init(o_free, 128);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc<512>(tmem_slot);
    const int n = build_items(p, pl, items, lane);
    if (lane == 0) *n_items_slot = n;
  }
  // the weights the following GEMMs will stream (independent of every earlier kernel, so before griddepcontrol.wait)
  if (warp == 3 && lane == 0 && p.pf_early) l2_prefetch_slice(p.pf, blockIdx.x, gridDim.x);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_items = *n_items_slot;
  unsigned long long* ts = (p.dbg_ts && blockIdx.x == 0) ? p.dbg_ts : nullptr;     // bring-up: phase stamps of CTA 0
  if (threadIdx.x == 128) dbg_stamp(ts, 0);
  pdl_wait();                 // Q, K and V were all written by earlier kernels on the stream
  pdl_launch_dependents();
  if (threadIdx.x == 128) dbg_stamp(ts, 1);

  if (warp == 0 || warp == 3) {
    // ===================== TMA producers: warp 0 the K ring, warp 3 the V ring =====================
    if (elect_one()) {
      const bool is_k = warp == 0;
      uint64_t* full = is_k ? k_full : v_full;
      uint64_t* empty = is_k ? k_empty : v_empty;
      uint8_t* ring = is_k ? sK : sV;
      const int col0 = is_k ? p.k_col0 : p.v_col0, col_step = is_k ? p.k_col_step : p.v_col_step;
      // rolling L2 window kv_prefetch tiles ahead of the ring (it runs on into the next items)
      int pf_i = 0, pf_t = 0;
      auto prefetch_next = [&]() {
        if (pf_i >= n_items) return;
        const Item it = items[pf_i];
        const AttnSeg& seg = p.segs[it.seg];
        const void* tm = is_k ? (seg.k_tmap ? seg.k_tmap : static_cast<const void*>(&tmap_k))
                              : (seg.v_tmap ? seg.v_tmap : static_cast<const void*>(&tmap_v));
        const int row = (is_k ? seg.k_row0 : seg.v_row0) + it.blk * p.k_head_rows + (it.kt0 + pf_t) * kTile;
        for (int b = 0; b < DB; ++b) tma_prefetch_2d(tm, col0 + it.blk * col_step + b * 64, row);
        if (++pf_t == it.nkt) { pf_t = 0; ++pf_i; }
      };
      for (int i = 0; i < p.kv_prefetch; ++i) prefetch_next();
      int n = 0;
      for (int i = 0; i < n_items; ++i) {
        const Item it = items[i];
        const AttnSeg& seg = p.segs[it.seg];
        const void* tm = is_k ? (seg.k_tmap ? seg.k_tmap : static_cast<const void*>(&tmap_k))
                              : (seg.v_tmap ? seg.v_tmap : static_cast<const void*>(&tmap_v));
        const int row0 = (is_k ? seg.k_row0 : seg.v_row0) + it.blk * p.k_head_rows;
        const int col = col0 + it.blk * col_step;
        for (int t = 0; t < it.nkt; ++t, ++n) {
          const int s = n % kStages;
          const uint32_t ph = (n / kStages) & 1;
          if (p.kv_prefetch) prefetch_next();
          mbar_wait_relaxed(&empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&full[s], stage_bytes);
          for (int b = 0; b < DB; ++b)
            tma_load_2d(ring + s * stage_bytes + b * kAtomBytes, tm, &full[s], col + b * 64, row0 + (it.kt0 + t) * kTile);
        }
      }
      if (is_k && !p.pf_early) l2_prefetch_slice(p.pf, blockIdx.x, gridDim.x);
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one() && n_items > 0) {
      const uint32_t idesc_qk = make_idesc_bf16_f32(kTile, kTile);
      const uint32_t idesc_pv = make_idesc_bf16_f32_bmn(kTile, (uint32_t)p.d_pad);
      const int ksteps = p.d_pad >> 4;
      // The chain P(n) ready -> P.V(n) -> Q.K^T(n + 2) (its score buffer is the one P(n) sits in) -> S(n + 2) ready must be
      // shorter than the softmax of tile n + 1, or the softmax warps wait for scores (measured: 20 % of their time when
      // this thread found out about K(n + 2) / V(n) and built its descriptors AFTER P(n) had arrived).  So everything that
      // does not depend on P(n) happens before the wait for it; behind the wait there are only the MMA and commit issues.
      int qi = 0, qt_in = 0, nq = 0;          // the next Q.K^T: item, tile inside the item, global tile index
      uint32_t qk_ka = 0, qk_qa = 0; int qk_s = 0;
      auto qk_prepare = [&]() {
        qk_s = nq % kStages;
        mbar_wait(&k_full[qk_s], (nq / kStages) & 1);
        qk_ka = smem_u32(sK + qk_s * stage_bytes);
        qk_qa = tmem_base + kColQ + (uint32_t)(qi & 1) * 64u;
      };
      auto qk_fire = [&]() {
        // (Q of a new item: loaded an item ahead, so this wait normally falls through.  It cannot move in front of P.V(n): with
        // one-tile items Q(i + 2) is loaded after the epilogue of item i, which needs P.V(n) itself)
        if (qt_in == 0) { mbar_wait(&q_full[qi & 1], (qi >> 1) & 1); tc_fence_after(); }
        const uint64_t d0 = make_kmajor_sw128_desc(qk_ka);
        const uint32_t sd = tmem_base + ((nq & 1) ? kColS1 : kColS0);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          if (ks < ksteps) {
            // 64-column atom ks >> 2 (kAtomBytes apart), 32 bytes per K = 16 step inside it; the address field counts 16 B
            const uint64_t o = (uint64_t)((ks >> 2) * (kAtomBytes >> 4) + (ks & 3) * 2);
            umma_f16_ts(sd, qk_qa + (uint32_t)ks * 8u, d0 + o, idesc_qk, ks != 0);
          }
        }
        umma_commit(&k_empty[qk_s]);      // the K stage is free once these MMAs have read it
        umma_commit(&s_full[nq & 1]);
        ++nq;
        if (++qt_in == items[qi].nkt) { qt_in = 0; ++qi; }
      };
      qk_prepare(); tc_fence_after(); qk_fire();                              // S(0)
      if (qi < n_items) { qk_prepare(); tc_fence_after(); qk_fire(); }        // S(1)
      int np = 0;
      long long m_pre = 0, m_pwait = 0, m_fire = 0;   // bring-up: cycles in the early waits / waiting for P / issuing
      for (int pi = 0; pi < n_items; ++pi) {
        const int nkt = items[pi].nkt;
        for (int t = 0; t < nkt; ++t, ++np) {
          const int s = np % kStages;
          const long long c0 = ts ? clock64() : 0;
          mbar_wait(&v_full[s], (np / kStages) & 1);
          if (t == 0 && pi > 0) mbar_wait(o_free, (pi - 1) & 1);    // the previous item's O has left tensor memory
          const bool more = qi < n_items;
          if (more) qk_prepare();                                   // K(n + 2) (and Q of a new item) are there
          // V tile is MN-major: 64-column blocks kAtomBytes apart (LBO), 8-key groups 1024 B apart (SBO); one K = 16 step
          // (16 keys) is 2048 bytes further
          const uint64_t dv0 = make_mnmajor_sw128_desc(smem_u32(sV + s * stage_bytes), kAtomBytes, 1024);
          const uint32_t pt = tmem_base + ((np & 1) ? kColS1 : kColS0);     // P(n) sits in the first 64 columns of S(n)
          const long long c1 = ts ? clock64() : 0;
          mbar_wait(&p_full[np & 1], (np >> 1) & 1);
          const long long c2 = ts ? clock64() : 0;
          tc_fence_after();
#pragma unroll
          for (int ks = 0; ks < kTile / 16; ++ks)
            umma_f16_ts(tmem_base + kColO, pt + (uint32_t)ks * 8u, dv0 + (uint64_t)(ks * (2048 >> 4)), idesc_pv, (t | ks) != 0);
          umma_commit(&v_empty[s]);
          umma_commit(&pv_done[np & 1]);
          if (t == nkt - 1) umma_commit(o_full);
          if (more) qk_fire();                                      // S(n + 2): executes behind P.V(n) (issue order)
          if (ts) { const long long c3 = clock64(); m_pre += c1 - c0; m_pwait += c2 - c1; m_fire += c3 - c2; }
        }
      }
      if (ts) { ts[19] = (unsigned long long)m_pre; ts[20] = (unsigned long long)m_pwait; ts[21] = (unsigned long long)m_fire; }
    }
  } else if (warp >= 4 && warp < 12) {
    // ===================== online softmax =====================
    // A query row is shared by two threads, lane l of warps w and w + 4 (same TMEM lane quadrant): `half` picks the 64 of
    // the tile's 128 key columns the thread owns.
    const int ew = warp & 3;
    const int half = warp >= 8 ? 1 : 0;
    const int row = ew * 32 + lane;            // tile row == TMEM lane
    const uint32_t lane_addr = tmem_base + ((uint32_t)(ew * 32) << 16);
    const float sl2 = p.scale * 1.4426950408889634f;
    const int col0 = half * 64;
    unsigned long long* ts1 = threadIdx.x == 128 ? ts : nullptr;
    long long t_prev = 0, acc_wait = 0, acc_work = 0;       // bring-up: cycles waited for scores / spent on tiles (thread 128)
    int n = 0;
    for (int i = 0; i < n_items; ++i) {
      const Item item = items[i];
      const AttnSeg& seg = p.segs[item.seg];
      const int m = item.qt * kTile + row;     // query row inside the (segment, block)
      const int limit = seg.off + m / p.group; // logical key index must be <= limit
      const bool all_visible = seg.off >= seg.klen - 1;
      float m_used = -INFINITY, l_sum = 0.f;
      for (int t = 0; t < item.nkt; ++t, ++n) {
        const uint32_t s_base = lane_addr + ((n & 1) ? kColS1 : kColS0);
        const int kt = item.kt0 + t;
        const long long w0 = ts1 ? clock64() : 0;
        mbar_wait(&s_full[n & 1], (n >> 1) & 1);
        tc_fence_after();
        if (ts1) { const long long w1 = clock64(); acc_wait += w1 - w0; if (t_prev) acc_work += w0 - t_prev; t_prev = w1; }
        if (t == 0 && i < 6) dbg_stamp(ts1, 2 + i);
        uint32_t sv[64];
        tmem_ld_32x32b_x32p(s_base + col0, sv);
        tmem_ld_32x32b_x32p(s_base + col0 + 32, sv + 32);
        tmem_ld_wait();
        if (all_visible) {
          if (kt * kTile + kTile > seg.klen) {
            // ragged last tile without a causal limit (SigLIP's 729 keys, a full cache): only slots >= klen are masked
            const int nvalid = seg.klen - kt * kTile - col0;
#pragma unroll
            for (int c = 0; c < 64; ++c)
              if (c >= nvalid) sv[c] = 0xff800000u;   // -inf
          }
        } else {
          // slow path (growing cache: causal limit through the ring geometry): knock masked scores out before the statistics
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            const uint32_t vm = visible_mask32(seg, kt * kTile + col0 + c * 32, limit);
#pragma unroll
            for (int u = 0; u < 32; ++u)
              if (!((vm >> u) & 1u)) sv[c * 32 + u] = 0xff800000u;   // -inf
          }
        }
        // four independent maximum chains
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          mx0 = fmaxf(mx0, __uint_as_float(sv[c]));
          mx1 = fmaxf(mx1, __uint_as_float(sv[16 + c]));
          mx2 = fmaxf(mx2, __uint_as_float(sv[32 + c]));
          mx3 = fmaxf(mx3, __uint_as_float(sv[48 + c]));
        }
        float tmax = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
        {
          // row maximum over both halves: the two threads of a row then take identical rescale decisions.  The slots
          // alternate with the tile parity; the pair barrier of tile n + 1 orders their reuse at tile n + 2.
          float* x = xch + (n & 1) * 256;
          x[half * 128 + row] = tmax;
          asm volatile("bar.sync %0, 64;" ::"r"(2 + ew) : "memory");
          tmax = fmaxf(tmax, x[(half ^ 1) * 128 + row]);
        }
        tmax *= sl2;                                 // scale > 0: max commutes with the scaling
        // one compare decides (m_used = -inf: any finite tmax passes; tmax = -inf never does); the slow path sits behind
        // one warp vote.  The running maximum may lag the true one by 2^8 (lazy rescale), so after the first tiles it is rare.
        const bool rescale = tmax > m_used + kRescaleThreshold;
        if (__any_sync(0xffffffffu, rescale)) {
          float alpha = 1.f;
          if (rescale) {
            alpha = (m_used == -INFINITY) ? 0.f : ex2_approx(m_used - tmax);
            m_used = tmax;
          }
          if (t > 0) {
            // O may only be rescaled between two P.V products: wait for the previous one (the MMAs complete in order)
            mbar_wait(&pv_done[(n - 1) & 1], ((n - 1) >> 1) & 1);
            tc_fence_after();
#pragma unroll 1
            for (int cc = half * 16; cc < p.d_pad; cc += 32) {
              uint32_t r[16];
              tmem_ld_32x32b_x16(lane_addr + kColO + cc, r);
              tmem_ld_wait();
#pragma unroll
              for (int u = 0; u < 16; ++u) r[u] = __float_as_uint(__uint_as_float(r[u]) * alpha);
              tmem_st_32x32b_x16(lane_addr + kColO + cc, r);
            }
            tmem_st_wait();
          }
          l_sum *= alpha;
        }
        // probabilities -> packed bf16 pairs in the first 64 columns of this score buffer (both threads of the row have
        // their scores in registers - the pair barrier above)
        const float mref = (m_used == -INFINITY) ? 0.f : m_used;
        float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
        for (int cc = 0; cc < 64; cc += 32) {
          uint32_t pk[16];
#pragma unroll
          for (int u = 0; u < 32; u += 2) {
            const float x0 = ex2_approx(fmaf(__uint_as_float(sv[cc + u]), sl2, -mref));
            const float x1 = ex2_approx(fmaf(__uint_as_float(sv[cc + u + 1]), sl2, -mref));
            acc0 += x0; acc1 += x1;
            pk[u >> 1] = pack_bf16x2(x0, x1);
          }
          tmem_st_32x32b_x16(s_base + (uint32_t)(half * 32 + (cc >> 1)), pk);
        }
        l_sum += acc0 + acc1;
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&p_full[n & 1]);
      }
      // hand (l, m) of this item to the epilogue warps and go on with the next item's scores
      const int b = i & 1;
      if (i >= 2) mbar_wait(&stats_free[b], ((i >> 1) + 1) & 1);
      float* st = stats + b * 384;
      st[half * 128 + row] = l_sum;
      if (half == 0) st[256 + row] = m_used;
      mbar_arrive(&stats_full[b]);
    }
    if (ts1) { ts1[16] = (unsigned long long)acc_wait; ts1[17] = (unsigned long long)acc_work; ts1[18] = (unsigned long long)n; }
  } else if (warp >= 12) {
    // ===================== Q loader + epilogue: one thread per query row =====================
    const int ew = warp & 3;
    const int row = ew * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(ew * 32) << 16);
    unsigned long long* ts2 = threadIdx.x == 384 ? ts : nullptr;
    auto load_q = [&](int i) {
      // this row of Q goes from global memory straight into tensor memory (packed bf16 pairs, 64 columns per buffer);
      // elements beyond d_pad and rows beyond the matrix are 0
      const Item it = items[i];
      const AttnSeg& seg = p.segs[it.seg];
      const long long q_row = (long long)seg.q_row0 + (long long)it.blk * p.q_block_rows + (long long)it.qt * kTile + row;
      const bf16* qsrc = p.q + q_row * (long long)p.q_ld + (p.q_col0 + it.blk * p.q_col_step);
      const bool rv = q_row < (long long)p.q_rows_total;
      const uint32_t qaddr = lane_addr + kColQ + (uint32_t)(i & 1) * 64u;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t qa[16], qb[16];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint4 u = make_uint4(0u, 0u, 0u, 0u), w = make_uint4(0u, 0u, 0u, 0u);
          if (rv && h * 64 + c * 8 < p.d_pad) u = *reinterpret_cast<const uint4*>(qsrc + h * 64 + c * 8);
          if (rv && h * 64 + 32 + c * 8 < p.d_pad) w = *reinterpret_cast<const uint4*>(qsrc + h * 64 + 32 + c * 8);
          qa[4 * c] = u.x; qa[4 * c + 1] = u.y; qa[4 * c + 2] = u.z; qa[4 * c + 3] = u.w;
          qb[4 * c] = w.x; qb[4 * c + 1] = w.y; qb[4 * c + 2] = w.z; qb[4 * c + 3] = w.w;
        }
        tmem_st_32x32b_x16(qaddr + h * 32, qa);
        tmem_st_32x32b_x16(qaddr + h * 32 + 16, qb);
      }
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(&q_full[i & 1]);
    };
    if (n_items > 0) load_q(0);
    for (int i = 0; i < n_items; ++i) {
      // Q(i + 1) lands in the other buffer (its previous tenant, item i - 1, is complete: its epilogue ran above)
      if (i + 1 < n_items) load_q(i + 1);
      const Item item = items[i];
      const AttnSeg& seg = p.segs[item.seg];
      const int m = item.qt * kTile + row;
      const bool row_valid = m < p.nq_rows;
      const int tok = m / p.group;
      const int b = i & 1;
      mbar_wait_relaxed(&stats_full[b], (i >> 1) & 1, 400);
      if (i == 2) dbg_stamp(ts2, 15);
      const float* st = stats + b * 384;
      const float l_sum = st[row] + st[128 + row];      // half 0's partial + half 1's, in that order
      const float m_used = st[256 + row];
      mbar_arrive(&stats_free[b]);
      mbar_wait(o_full, i & 1);
      tc_fence_after();
      if (i == 2) dbg_stamp(ts2, 14);
      bf16* orow = p.out + (size_t)(seg.out_row0 + tok) * p.out_ld + (size_t)(item.blk * p.group + (m - tok * p.group)) * p.d;
      const bool vec_ok = (p.d & 7) == 0 && (p.out_ld & 7) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15) == 0;
      if (item.parts == 1) {
        const float inv = l_sum > 0.f ? 1.f / l_sum : 0.f;
        // 32 columns per tensor-memory round trip (next to the running MMAs, which read P and Q from tensor memory, one
        // tcgen05.ld round trip is ~300 cycles; O - and with it the next item's first P.V - stays locked until the last one)
#pragma unroll 1
        for (int c0 = 0; c0 < p.d_pad; c0 += 32) {
          uint32_t r[32];
          const int ncol = min(32, p.d_pad - c0);          // 16 or 32
          if (ncol == 32) tmem_ld_32x32b_x32p(lane_addr + kColO + c0, r);
          else tmem_ld_32x32b_x16p(lane_addr + kColO + c0, r);
          tmem_ld_wait();
          if (c0 + 32 >= p.d_pad) { tc_fence_before(); mbar_arrive(o_free); }   // O is in registers: the next item's P.V may start
          if (row_valid) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              const int c8 = c0 + g * 8;
              if (g * 8 < ncol) {
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = __uint_as_float(r[g * 8 + u]) * inv;
                if (vec_ok && c8 + 8 <= p.d) {
                  st_v4(orow + c8, make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                                              pack_bf16x2(v[6], v[7])));
                } else {
#pragma unroll
                  for (int u = 0; u < 8; ++u)
                    if (c8 + u < p.d) orow[c8 + u] = f2b(v[u]);
                }
              }
            }
          }
        }
      } else {
        // ---- this CTA holds only some key tiles of the group: park the unnormalised O (fp32, column-major [d_pad][128])
        // and (m, l) in the L2-resident workspace; the last part to arrive merges the group in part order ----
        const size_t cta_floats = (size_t)kTile * p.d_pad + 2 * kTile;
        float* grp = p.ws_partial + (size_t)item.group_lin * kMaxParts * cta_floats;
        float* mine = grp + (size_t)item.part * cta_floats;
#pragma unroll 1
        for (int c0 = 0; c0 < p.d_pad; c0 += 32) {
          uint32_t r[32];
          const int ncol = min(32, p.d_pad - c0);
          if (ncol == 32) tmem_ld_32x32b_x32p(lane_addr + kColO + c0, r);
          else tmem_ld_32x32b_x16p(lane_addr + kColO + c0, r);
          tmem_ld_wait();
          if (c0 + 32 >= p.d_pad) { tc_fence_before(); mbar_arrive(o_free); }
          float* dst = mine + (size_t)c0 * kTile + row;      // a warp's store of one column is one 128-byte segment
#pragma unroll
          for (int u = 0; u < 32; ++u)
            if (u < ncol) __stcg(dst + (size_t)u * kTile, __uint_as_float(r[u]));
        }
        __stcg(mine + (size_t)kTile * p.d_pad + row, m_used);
        __stcg(mine + (size_t)kTile * p.d_pad + kTile + row, l_sum);
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (threadIdx.x == 12 * 32) {
          int* cnt = p.ws_pcounters + item.group_lin;
          const int old = atomicAdd(cnt, 1);
          const int last = (old == item.parts - 1) ? 1 : 0;
          if (last) *cnt = 0;                       // self-resetting
          *last_flag = last;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const bool merge = *last_flag != 0;
        asm volatile("bar.sync 1, 128;" ::: "memory");   // everyone has read the flag before the next item may overwrite it
        if (merge) {
          __threadfence();
          float ms[kMaxParts], ls[kMaxParts];
#pragma unroll
          for (int s = 0; s < kMaxParts; ++s) {
            const float* pst = grp + (size_t)(s < item.parts ? s : 0) * cta_floats + (size_t)kTile * p.d_pad;
            ms[s] = __ldcg(pst + row);
            ls[s] = __ldcg(pst + kTile + row);
          }
          float mx = -INFINITY;
#pragma unroll
          for (int s = 0; s < kMaxParts; ++s) if (s < item.parts) mx = fmaxf(mx, ms[s]);
          float wgt[kMaxParts];
          float L = 0.f;
#pragma unroll
          for (int s = 0; s < kMaxParts; ++s) {
            wgt[s] = (s < item.parts && ms[s] != -INFINITY) ? ex2_approx(ms[s] - mx) : 0.f;
            L += wgt[s] * ls[s];
          }
          const float inv = L > 0.f ? 1.f / L : 0.f;
#pragma unroll 1
          for (int cc = 0; cc < p.d_pad; cc += 16) {
            float v[kMaxParts][16];
#pragma unroll
            for (int s = 0; s < kMaxParts; ++s) {
              const float* peer = grp + (size_t)(s < item.parts ? s : 0) * cta_floats + (size_t)cc * kTile + row;
              if (s < item.parts) {
#pragma unroll
                for (int u = 0; u < 16; ++u) v[s][u] = __ldcg(peer + (size_t)u * kTile);
              } else {
#pragma unroll
                for (int u = 0; u < 16; ++u) v[s][u] = 0.f;
              }
            }
            float acc[16];
#pragma unroll
            for (int u = 0; u < 16; ++u) acc[u] = 0.f;
#pragma unroll
            for (int s = 0; s < kMaxParts; ++s) {
#pragma unroll
              for (int u = 0; u < 16; ++u) acc[u] += wgt[s] * v[s][u];
            }
            if (row_valid) {
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                const int c8 = cc + h * 8;
                if (vec_ok && c8 + 8 <= p.d) {
                  st_v4(orow + c8, make_uint4(pack_bf16x2(acc[8 * h] * inv, acc[8 * h + 1] * inv),
                                              pack_bf16x2(acc[8 * h + 2] * inv, acc[8 * h + 3] * inv),
                                              pack_bf16x2(acc[8 * h + 4] * inv, acc[8 * h + 5] * inv),
                                              pack_bf16x2(acc[8 * h + 6] * inv, acc[8 * h + 7] * inv)));
                } else {
#pragma unroll 4
                  for (int u = 0; u < 8; ++u)
                    if (c8 + u < p.d) orow[c8 + u] = f2b(acc[8 * h + u] * inv);
                }
              }
            }
          }
        }
      }
      if (i < 6) dbg_stamp(ts2, 8 + i);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace

static unsigned long long g_pattn_launches = 0;
extern "C" unsigned long long aha_attention_persistent_launch_count(void) { return g_pattn_launches; }
static unsigned long long* g_pattn_ts = nullptr;
extern "C" int aha_debug_attention_persistent_timeline(unsigned long long* out32) {
  if (!g_pattn_ts) { aha_set_error("ValueError: run with AHA_TS=2"); return AHA_ERR_VALUE; }
  AHA_CUDA(cudaDeviceSynchronize());
  AHA_CUDA(cudaMemcpy(out32, g_pattn_ts, 32 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  return AHA_OK;
}

int op_attention_persistent(const AttnProblem& p_in, cudaStream_t st) {
  { const char* e = getenv("AHA_ATTN_PERSIST"); if (e && atoi(e) == 0) return 0; }   // (read per launch: A/B runs and tests)
  if (p_in.d_pad % 16 != 0 || p_in.d_pad < 16 || p_in.d_pad > 128 || p_in.d > p_in.d_pad) return 0;
  if (p_in.n_segs <= 0 || p_in.n_segs > 32 || p_in.n_blocks <= 0 || p_in.nq_rows <= 0 || p_in.group <= 0) return 0;
  static int attr_smem[16] = {0};          // function attributes and the SM count are per device
  static int sm_counts[16] = {0};
  int dev = 0;
  AHA_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 16) return 0;
  if (!sm_counts[dev]) AHA_CUDA(cudaDeviceGetAttribute(&sm_counts[dev], cudaDevAttrMultiProcessorCount, dev));
  const int sm_count = sm_counts[dev];

  AttnPersistPlan pl{};
  pl.q_tiles = (p_in.nq_rows + kTile - 1) / kTile;
  pl.gps = p_in.n_blocks * pl.q_tiles;
  pl.n_groups = p_in.n_segs * pl.gps;
  int max_tiles = 0, min_tiles = 1 << 30;
  long long units = 0;
  for (int i = 0; i < p_in.n_segs; ++i) {
    if (p_in.segs[i].klen <= 0) return 0;
    const int T = (p_in.segs[i].klen + kTile - 1) / kTile;
    max_tiles = T > max_tiles ? T : max_tiles;
    min_tiles = T < min_tiles ? T : min_tiles;
    pl.unit_base[i] = (int)units;
    units += (long long)p_in.n_blocks * T;      // per team (one query tile of every (segment, kv head))
  }
  pl.unit_base[p_in.n_segs] = (int)units;
  if (units >= (1ll << 30)) return 0;
  pl.units = (int)units;
  int grid = 0;
  const int force_mode = getenv("AHA_ATTN_PERSIST_MODE") ? atoi(getenv("AHA_ATTN_PERSIST_MODE")) : -1;
  if ((pl.n_groups >= 2 * sm_count && force_mode != 1) || force_mode == 0) {
    // many groups: whole groups, contiguous ranges
    grid = pl.n_groups < sm_count ? pl.n_groups : sm_count;
    if ((pl.n_groups + grid - 1) / grid > kMaxItems) return 0;
    pl.mode = 0;
  } else if (!p_in.no_split && (max_tiles >= 16 || force_mode == 1) && pl.q_tiles <= sm_count / 4 &&
             units * pl.q_tiles >= 2ll * sm_count) {
    // few groups with long key axes: teams of CTAs, equal unit ranges inside a team (stream-K)
    grid = sm_count;
    const int R = grid / pl.q_tiles;                      // ranks per team
    const int L = (int)(units / R);                       // shortest range
    const int Lmax = (int)((units + R - 1) / R);          // longest range
    if (L < 1) return 0;
    const int parts = (max_tiles - 1 + L - 1) / L + 1;
    if (parts > kMaxParts) return 0;
    if ((Lmax + min_tiles - 1) / min_tiles + 1 > kMaxItems) return 0;
    if (pl.n_groups > GemmWorkspace::kPCounters / 2) return 0;
    const size_t need = (size_t)pl.n_groups * kMaxParts * ((size_t)kTile * p_in.d_pad + 2 * kTile) * sizeof(float);
    if (need > GemmWorkspace::kPartialBytes) return 0;
    pl.mode = 1;
  } else {
    return 0;
  }

  AttnProblem p = p_in;
  const int DB = (p.d_pad + 63) / 64;
  CUtensorMap tk, tv;
  int rc = get_tmap(p.k, (int)p.k_rows_total, p.k_ld, p.k_ld, kTile, &tk); if (rc) return rc;
  rc = get_tmap(p.v, (int)p.v_rows_total, p.v_ld, p.v_ld, kTile, &tv); if (rc) return rc;
  const int smem = PSmem::total(DB);
  if (attr_smem[dev] < smem) {
    AHA_CUDA(cudaFuncSetAttribute(attention_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_smem[dev] = smem;
  }
  p.merge_last = pl.mode == 1;
  if (pl.mode == 1) {
    GemmSplitScratch sc;
    rc = aha_split_scratch(st, &sc, false); if (rc) return rc;
    p.ws_partial = sc.partial; p.ws_pcounters = sc.pcounters;
  }
  { static const int kvpf = getenv("AHA_KV_PF") ? atoi(getenv("AHA_KV_PF")) : 3; p.kv_prefetch = (pl.mode == 1) ? kvpf : 0; }
  static const bool want_ts = getenv("AHA_TS") && atoi(getenv("AHA_TS")) == 2;
  p.dbg_ts = nullptr;
  if (want_ts) {
    if (!g_pattn_ts) { AHA_CUDA(cudaMalloc(&g_pattn_ts, 32 * sizeof(unsigned long long))); }
    AHA_CUDA(cudaMemsetAsync(g_pattn_ts, 0, 256, st)); { }
    p.dbg_ts = g_pattn_ts;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = aha_pdl_enabled() ? 1 : 0;
  // per-launch timing for bench.py's roofline leg: class 3 = many-group batched attention (SigLIP), 4 = chunk versus cache
  ProfSlot* slot = prof_begin(pl.mode == 0 ? 3 : 4, st);
  AHA_CUDA(cudaLaunchKernelEx(&cfg, attention_persistent_kernel, tk, tv, p, pl));
  ++g_aha_launch_count;
  ++g_pattn_launches;
  if (slot) {
    double flops = 0, bytes = 0;
    for (int i = 0; i < p.n_segs; ++i) {
      flops += 4.0 * p.n_blocks * (double)p.nq_rows * p.segs[i].klen * p.d;
      bytes += 2.0 * p.n_blocks * (2.0 * p.nq_rows + 2.0 * p.segs[i].klen) * p.d;
    }
    prof_end(slot, st, bytes, flops);
  }
  AHA_CHECK_LAUNCH();
  return 1;
}
