// KV-cache policies of the reference on a device ring buffer: SinkCache (test/sink_cache.py:7-164),
// SlidingWindowCache (test/sliding_window_cache.py:5-53), TrulyStaticCache (test/static_cache.py:5-47).
// Integer semantics (lengths, positions, which token lives where) are decided on the host in cache_plan_update();
// the kernels below only move / rotate bytes: coalesced 128-bit accesses, one pass over every byte they touch.
#include "kvcache.h"
#include "gemm.h"
#include "ptx_sm100.cuh"
#include <cuda.h>
#include <algorithm>
#include <vector>
#include <cstdlib>

namespace {

// physical slot of chunk token t under plan geometry
__device__ __forceinline__ int store_slot(int t, int store_begin, int store_phys0, int wrap_end, int R) {
  int x = store_phys0 + (t - store_begin);
  if (x >= wrap_end) x -= R;
  return x;
}

// ---- store: copy already-rotated k / v chunk rows into their slots.  One thread per 16 bytes. ----
__global__ void cache_store_kernel(const bf16* __restrict__ k, const bf16* __restrict__ v, long long k_hs, long long k_ts,
                                   long long v_hs, long long v_ts, bf16* __restrict__ K, bf16* __restrict__ V, int kvh,
                                   int d, int cap, int store_begin, int store_end, int store_phys0, int wrap_end, int R) {
  const int dv = d >> 3;
  const int n = store_end - store_begin;
  const long long total = (long long)kvh * n * dv * 2;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(idx % dv);
    long long r = idx / dv;
    const int t = store_begin + (int)(r % n); r /= n;
    const int h = (int)(r % kvh);
    const int which = (int)(r / kvh);
    const int slot = store_slot(t, store_begin, store_phys0, wrap_end, R);
    const bf16* src = which ? (v + h * v_hs + t * v_ts) : (k + h * k_hs + t * k_ts);
    bf16* dst = (which ? V : K) + ((size_t)h * cap + slot) * d;
    // sources coming from a transposed view may be only 2-byte aligned per row; rows themselves are contiguous
    if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
      reinterpret_cast<uint4*>(dst)[c] = reinterpret_cast<const uint4*>(src)[c];
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) dst[c * 8 + e] = src[c * 8 + e];
    }
  }
}

__global__ void ids_store_kernel(int32_t* __restrict__ ids, long long fed, int store_begin, int store_end, int store_phys0,
                                 int wrap_end, int R) {
  const int t = store_begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (t < store_end) ids[store_slot(t, store_begin, store_phys0, wrap_end, R)] = (int32_t)(fed + t);
}

// ---- sink re-rotation table (sink_cache.py:35-55), each fp32 op rounded separately exactly as torch evaluates
//      original_cos * shifted_cos + original_sin * shifted_sin  and  -original_sin * shifted_cos + original_cos * shifted_sin
__global__ void rerot_table_kernel(const bf16* __restrict__ cos_t, const bf16* __restrict__ sin_t, bf16* __restrict__ out,
                                   int S, int q, int keep, int d) {
  const int total = keep * d;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int r = idx / d, c = idx % d;
    const float oc = b2f(cos_t[(size_t)(S + q + r) * d + c]), sc = b2f(cos_t[(size_t)(S + r) * d + c]);
    const float os = b2f(sin_t[(size_t)(S + q + r) * d + c]), ss = b2f(sin_t[(size_t)(S + r) * d + c]);
    const float rc = __fadd_rn(__fmul_rn(oc, sc), __fmul_rn(os, ss));
    const float rs = __fadd_rn(__fmul_rn(-os, sc), __fmul_rn(oc, ss));
    out[idx] = f2b(rc);
    out[(size_t)total + idx] = f2b(rs);
  }
}

// ---- in-place re-rotation of the kept keys: k' = (k * c) + (rotate_half(k) * s), every op rounded to bf16
//      (sink_cache.py:27-33).  Thread = 8 elements of the low half + the matching 8 of the high half of one key.
//      Algorithmic bytes per key: 2*d read + 2*d written (+ the shared table rows, which stay in L2).
struct RerotLayers { bf16* k[64]; };

// Two elements of the low half (l) and the matching two of the high half (h) of one key, with their table entries:
//   low'  = bf16(bf16(l * c_lo) + bf16(-h * s_lo))      (rotate_half = -high on the low half)
//   high' = bf16(bf16(h * c_hi) + bf16( l * s_hi))      (rotate_half =  low  on the high half)
// The product of two bf16 values is exact in fp32, so the packed bf16 multiply (one RNE rounding of the exact product)
// gives the bits torch's "fp32 multiply, round to bf16" gives; the sums stay fp32-then-round like torch's (a packed bf16
// add would round the exact sum once, which is not the same thing), and their rounding is the final pack.  The
// single-value F2F conversions this replaces ran on the XU pipe (16 lanes per clock) and bounded the kernel: ncu showed
// 83 % XU utilisation against 28 % DRAM.
__device__ __forceinline__ uint32_t mul_bf16x2(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ void rerot_pair(uint32_t l, uint32_t h, uint32_t c_lo, uint32_t c_hi, uint32_t s_lo,
                                           uint32_t s_hi, uint32_t& out_lo, uint32_t& out_hi) {
  const float2 lc = unpack_bf16x2(mul_bf16x2(l, c_lo)), hs = unpack_bf16x2(mul_bf16x2(h, s_lo));
  const float2 hc = unpack_bf16x2(mul_bf16x2(h, c_hi)), ls = unpack_bf16x2(mul_bf16x2(l, s_hi));
  out_lo = pack_bf16x2(lc.x - hs.x, lc.y - hs.y);
  out_hi = pack_bf16x2(hc.x + ls.x, hc.y + ls.y);
}

// Thread = (kept key r, 16-byte column chunk c) of one layer: its cos/sin entries are loaded once and applied to the key
// of every kv head, so a thread has 2 * kvh independent 16-byte HBM loads in flight (8 at 7B) before the first use.
__global__ void __launch_bounds__(256, 3) rerotate_kernel(RerotLayers ls, const bf16* __restrict__ rerot, int kvh, int d,
                                                       int cap, int S, int R, int base, int keep) {
  pdl_wait();
  pdl_launch_dependents();
  const int tpk = d >> 4;  // threads per key
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= keep * tpk) return;
  bf16* K = ls.k[blockIdx.y];
  const int half = d >> 1;
  const int c = idx % tpk, r = idx / tpk;
  int slot = base + r;
  if (slot >= R) slot -= R;
  slot += S;
  const bf16* rc = rerot + (size_t)r * d + c * 8;
  const bf16* rs = rc + (size_t)keep * d;
  const uint4 clo = *reinterpret_cast<const uint4*>(rc), chi = *reinterpret_cast<const uint4*>(rc + half);
  const uint4 slo = *reinterpret_cast<const uint4*>(rs), shi = *reinterpret_cast<const uint4*>(rs + half);
  const uint32_t cl[4] = {clo.x, clo.y, clo.z, clo.w}, ch[4] = {chi.x, chi.y, chi.z, chi.w};
  const uint32_t sl[4] = {slo.x, slo.y, slo.z, slo.w}, sh[4] = {shi.x, shi.y, shi.z, shi.w};
  constexpr int kHeadsPerThread = 4;   // blockIdx.z picks the head group (2 per thread measured slower: 33 vs 31 us at 7B)
#pragma unroll 1
  for (int h0 = blockIdx.z * kHeadsPerThread; h0 < min(kvh, (int)(blockIdx.z + 1) * kHeadsPerThread); h0 += 4) {
    uint4 lo[4], hi[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (u < kHeadsPerThread && h0 + u < kvh) {
        const bf16* row = K + ((size_t)(h0 + u) * cap + slot) * d + c * 8;
        lo[u] = *reinterpret_cast<const uint4*>(row);
        hi[u] = *reinterpret_cast<const uint4*>(row + half);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (u < kHeadsPerThread && h0 + u < kvh) {
        const uint32_t l[4] = {lo[u].x, lo[u].y, lo[u].z, lo[u].w}, hh[4] = {hi[u].x, hi[u].y, hi[u].z, hi[u].w};
        uint32_t ol[4], oh[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) rerot_pair(l[e], hh[e], cl[e], ch[e], sl[e], sh[e], ol[e], oh[e]);
        bf16* row = K + ((size_t)(h0 + u) * cap + slot) * d + c * 8;
        st_v4(row, make_uint4(ol[0], ol[1], ol[2], ol[3]));
        st_v4(row + half, make_uint4(oh[0], oh[1], oh[2], oh[3]));
      }
    }
  }
}

// The same re-rotation for every stream of a batched prefill call in ONE launch: blockIdx.y = stream * n_layers + layer,
// slab pointers come from each cache's device-resident layer table.
struct RerotStream { bf16* const* layer_k; const bf16* rerot; int base, keep, S, R, cap; };
struct RerotBatch { RerotStream s[kMaxStreams]; };

__global__ void __launch_bounds__(256, 3) rerotate_batched_kernel(RerotBatch b, int n_layers, int kvh, int d) {
  pdl_wait();
  pdl_launch_dependents();
  const RerotStream& st = b.s[blockIdx.y / n_layers];
  const int keep = st.keep;
  const int tpk = d >> 4;  // threads per key
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= keep * tpk) return;
  bf16* K = st.layer_k[blockIdx.y % n_layers];
  const int half = d >> 1, cap = st.cap;
  const int c = idx % tpk, r = idx / tpk;
  int slot = st.base + r;
  if (slot >= st.R) slot -= st.R;
  slot += st.S;
  const bf16* rc = st.rerot + (size_t)r * d + c * 8;
  const bf16* rs = rc + (size_t)keep * d;
  const uint4 clo = *reinterpret_cast<const uint4*>(rc), chi = *reinterpret_cast<const uint4*>(rc + half);
  const uint4 slo = *reinterpret_cast<const uint4*>(rs), shi = *reinterpret_cast<const uint4*>(rs + half);
  const uint32_t cl[4] = {clo.x, clo.y, clo.z, clo.w}, ch[4] = {chi.x, chi.y, chi.z, chi.w};
  const uint32_t sl[4] = {slo.x, slo.y, slo.z, slo.w}, sh[4] = {shi.x, shi.y, shi.z, shi.w};
#pragma unroll 1
  for (int h0 = blockIdx.z * 4; h0 < min(kvh, (int)(blockIdx.z + 1) * 4); h0 += 4) {
    uint4 lo[4], hi[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (h0 + u < kvh) {
        const bf16* row = K + ((size_t)(h0 + u) * cap + slot) * d + c * 8;
        lo[u] = *reinterpret_cast<const uint4*>(row);
        hi[u] = *reinterpret_cast<const uint4*>(row + half);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (h0 + u < kvh) {
        const uint32_t l[4] = {lo[u].x, lo[u].y, lo[u].z, lo[u].w}, hh[4] = {hi[u].x, hi[u].y, hi[u].z, hi[u].w};
        uint32_t ol[4], oh[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) rerot_pair(l[e], hh[e], cl[e], ch[e], sl[e], sh[e], ol[e], oh[e]);
        bf16* row = K + ((size_t)(h0 + u) * cap + slot) * d + c * 8;
        st_v4(row, make_uint4(ol[0], ol[1], ol[2], ol[3]));
        st_v4(row + half, make_uint4(oh[0], oh[1], oh[2], oh[3]));
      }
    }
  }
}

// TMA-staged variant of the same re-rotation (d = 128): a CTA owns a block of kRrKeys consecutive kept keys and walks over
// (layer, kv head) units of one layer group.  Each unit is one contiguous 8 KB span of the ring (two pieces when it
// wraps) pulled into shared memory with cp.async.bulk, rotated in place and pushed back with a bulk store; kRrStages
// units are in flight per CTA, so the bytes in flight live in shared memory (up to 5 CTAs x 40 KB per SM) instead of
// registers.  The cos/sin rows of the block are read once into registers and reused for every unit.
constexpr int kRrKeys = 32, kRrStages = 5, kRrThreads = 256, kRrLayersPerCta = 4;

__global__ void __launch_bounds__(kRrThreads, 4) rerotate_tma_kernel(RerotLayers ls, int n_layers_total, const bf16* __restrict__ rerot,
                                                                     int kvh, int cap, int S, int R, int base, int keep) {
  using namespace sm100;
  constexpr int d = 128, kUnitBytes = kRrKeys * d * 2;
  __shared__ __align__(128) uint8_t stage[kRrStages][kUnitBytes];
  __shared__ __align__(8) uint64_t full[kRrStages];
  const int tid = threadIdx.x;
  const int r0 = blockIdx.x * kRrKeys;
  const int nk = min(kRrKeys, keep - r0);
  const int layer0 = blockIdx.y * kRrLayersPerCta;
  const int units = min(kRrLayersPerCta, n_layers_total - layer0) * kvh;
  int slot0 = base + r0;
  if (slot0 >= R) slot0 -= R;
  const int len1 = min(nk, R - slot0);     // keys before the ring wraps
  if (tid == 0) {
    for (int s = 0; s < kRrStages; ++s) mbar_init(&full[s], 1);
    fence_barrier_init();
  }
  __syncthreads();
  pdl_wait();
  pdl_launch_dependents();
  auto key_ptr = [&](int u, int slot) { return ls.k[layer0 + u / kvh] + ((size_t)(u % kvh) * cap + S + slot) * d; };
  auto issue_load = [&](int u) {
    const int s = u % kRrStages;
    mbar_arrive_expect_tx(&full[s], (uint32_t)nk * d * 2);
    bulk_load(stage[s], key_ptr(u, slot0), (uint32_t)len1 * d * 2, &full[s]);
    if (len1 < nk) bulk_load(stage[s] + len1 * d * 2, key_ptr(u, 0), (uint32_t)(nk - len1) * d * 2, &full[s]);
  };
  if (tid == 0)
    for (int u = 0; u < min(units, kRrStages); ++u) issue_load(u);
  // this thread's (key, 16-byte chunk) of every unit, and its table entries
  const int k = tid >> 3, c = tid & 7;
  const bool active = k < nk;
  uint32_t cl[4] = {0, 0, 0, 0}, ch[4] = {0, 0, 0, 0}, sl[4] = {0, 0, 0, 0}, sh[4] = {0, 0, 0, 0};
  if (active) {
    const bf16* rc = rerot + (size_t)(r0 + k) * d + c * 8;
    const bf16* rs = rc + (size_t)keep * d;
    const uint4 a = *reinterpret_cast<const uint4*>(rc), b = *reinterpret_cast<const uint4*>(rc + 64);
    const uint4 e = *reinterpret_cast<const uint4*>(rs), f = *reinterpret_cast<const uint4*>(rs + 64);
    cl[0] = a.x; cl[1] = a.y; cl[2] = a.z; cl[3] = a.w; ch[0] = b.x; ch[1] = b.y; ch[2] = b.z; ch[3] = b.w;
    sl[0] = e.x; sl[1] = e.y; sl[2] = e.z; sl[3] = e.w; sh[0] = f.x; sh[1] = f.y; sh[2] = f.z; sh[3] = f.w;
  }
  for (int u = 0; u < units; ++u) {
    const int s = u % kRrStages;
    mbar_wait(&full[s], (u / kRrStages) & 1);
    if (active) {
      uint4* plo = reinterpret_cast<uint4*>(stage[s] + k * (d * 2) + c * 16);
      uint4* phi = reinterpret_cast<uint4*>(stage[s] + k * (d * 2) + 128 + c * 16);
      const uint4 lo = *plo, hi = *phi;
      const uint32_t l[4] = {lo.x, lo.y, lo.z, lo.w}, hh[4] = {hi.x, hi.y, hi.z, hi.w};
      uint32_t ol[4], oh[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) rerot_pair(l[e], hh[e], cl[e], ch[e], sl[e], sh[e], ol[e], oh[e]);
      *plo = make_uint4(ol[0], ol[1], ol[2], ol[3]);
      *phi = make_uint4(oh[0], oh[1], oh[2], oh[3]);
    }
    fence_proxy_async();      // generic-proxy writes -> visible to the bulk store
    __syncthreads();
    if (tid == 0) {
      bulk_store(key_ptr(u, slot0), stage[s], (uint32_t)len1 * d * 2);
      if (len1 < nk) bulk_store(key_ptr(u, 0), stage[s] + len1 * d * 2, (uint32_t)(nk - len1) * d * 2);
      bulk_commit();
      // refill the stage whose store was issued one unit ago (its shared-memory reads are done by now)
      if (u >= 1 && u - 1 + kRrStages < units) {
        bulk_wait_read<1>();
        issue_load(u - 1 + kRrStages);
      }
    }
  }
  if (tid == 0) bulk_wait_all<0>();   // the stores must have left shared memory before the CTA exits
}

// ---- logical-order export (what Cache.update() returns): out[h][j][:] = buf[h][phys(j)][:] ----
__global__ void export_kernel(const bf16* __restrict__ buf, bf16* __restrict__ out, int kvh, int d, int cap, int len,
                              int S, int R, int base) {
  const int dv = d >> 3;
  const long long total = (long long)kvh * len * dv;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(idx % dv);
    long long t = idx / dv;
    const int j = (int)(t % len);
    const int h = (int)(t / len);
    int p = j;
    if (R > 0 && j >= S) { p = base + (j - S); if (p >= R) p -= R; p += S; }
    reinterpret_cast<uint4*>(out + ((size_t)h * len + j) * d)[c] =
        reinterpret_cast<const uint4*>(buf + ((size_t)h * cap + p) * d)[c];
  }
}
__global__ void export_ids_kernel(const int32_t* __restrict__ ids, int32_t* __restrict__ out, int len, int S, int R,
                                  int base) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= len) return;
  int p = j;
  if (R > 0 && j >= S) { p = base + (j - S); if (p >= R) p -= R; p += S; }
  out[j] = ids[p];
}

// compaction helpers: gather rows [from, from+n) of every head into scratch, then scatter to [to, to+n)
__global__ void move_rows_kernel(const bf16* __restrict__ src, int src_cap, int src_row0, bf16* __restrict__ dst,
                                 int dst_cap, int dst_row0, int kvh, int d, int n) {
  const int dv = d >> 3;
  const long long total = (long long)kvh * n * dv;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(idx % dv);
    long long t = idx / dv;
    const int r = (int)(t % n);
    const int h = (int)(t / n);
    reinterpret_cast<uint4*>(dst + ((size_t)h * dst_cap + dst_row0 + r) * d)[c] =
        reinterpret_cast<const uint4*>(src + ((size_t)h * src_cap + src_row0 + r) * d)[c];
  }
}
__global__ void move_ids_kernel(const int32_t* __restrict__ src, int32_t* __restrict__ dst, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[i];
}

// ---- fused RoPE + append for the prefill engine ----
// Qwen2 apply_rotary_pos_emb (transformers 4.49): x' = (x * cos) + (rotate_half(x) * sin), each op rounded to bf16.
// Thread = (stream, token, head slot, 16-element pair of 8-chunks).  Head slots: [0, heads) = q heads,
// [heads, heads+kvh) = k heads, [heads+kvh, heads+2kvh) = v heads (the packed projection's column blocks).
__global__ void __launch_bounds__(256) rope_append_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ cos,
                                                          const bf16* __restrict__ sin, bf16* __restrict__ qbuf,
                                                          RopeAppendArgs a, int q, int heads, int kvh, int d, int cap) {
  pdl_wait();
  pdl_launch_dependents();
  const int tph = d >> 4;                 // threads per head row
  const int slots = heads + 2 * kvh;
  const int ld = slots * d;
  const int s = blockIdx.y;
  const int group = heads / kvh;
  const int half = d >> 1;
  const RopeAppendStream& st = a.s[s];
  const long long total = (long long)q * slots * tph;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(idx % tph);
    long long t2 = idx / tph;
    const int slot = (int)(t2 % slots);
    const int t = (int)(t2 / slots);
    const bf16* src = qkv + ((size_t)s * q + t) * ld + (size_t)slot * d;
    const uint4 lo = *reinterpret_cast<const uint4*>(src + c * 8);
    const uint4 hi = *reinterpret_cast<const uint4*>(src + half + c * 8);
    bf16* dst;
    bool rotate = true;
    if (slot < heads) {
      const int g = slot / group, r = slot % group;
      dst = qbuf + ((((size_t)s * kvh + g) * q + t) * group + r) * d;
    } else {
      if (t < st.store_begin || t >= st.store_end) continue;
      const int phys = store_slot(t, st.store_begin, st.store_phys0, st.wrap_end, st.R);
      if (slot < heads + kvh) {
        dst = st.k + ((size_t)(slot - heads) * cap + phys) * d;
      } else {
        dst = st.v + ((size_t)(slot - heads - kvh) * cap + phys) * d;
        rotate = false;
      }
    }
    if (!rotate) {
      *reinterpret_cast<uint4*>(dst + c * 8) = lo;
      *reinterpret_cast<uint4*>(dst + half + c * 8) = hi;
      continue;
    }
    const bf16* cr = cos + ((size_t)s * q + t) * d;
    const bf16* sr = sin + ((size_t)s * q + t) * d;
    const uint4 clo = *reinterpret_cast<const uint4*>(cr + c * 8), chi = *reinterpret_cast<const uint4*>(cr + half + c * 8);
    const uint4 slo = *reinterpret_cast<const uint4*>(sr + c * 8), shi = *reinterpret_cast<const uint4*>(sr + half + c * 8);
    const uint32_t l[4] = {lo.x, lo.y, lo.z, lo.w}, hh[4] = {hi.x, hi.y, hi.z, hi.w};
    const uint32_t cl[4] = {clo.x, clo.y, clo.z, clo.w}, ch[4] = {chi.x, chi.y, chi.z, chi.w};
    const uint32_t sl[4] = {slo.x, slo.y, slo.z, slo.w}, sh[4] = {shi.x, shi.y, shi.z, shi.w};
    uint32_t ol[4], oh[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) rerot_pair(l[e], hh[e], cl[e], ch[e], sl[e], sh[e], ol[e], oh[e]);   // same formula as re-rotation
    *reinterpret_cast<uint4*>(dst + c * 8) = make_uint4(ol[0], ol[1], ol[2], ol[3]);
    *reinterpret_cast<uint4*>(dst + half + c * 8) = make_uint4(oh[0], oh[1], oh[2], oh[3]);
  }
}

inline int grid_for(long long items, int threads, int cap = 148 * 8) {
  long long g = (items + threads - 1) / threads;
  return (int)std::max<long long>(1, std::min<long long>(g, cap));
}

}  // namespace

// =====================================================================================
// host side
// =====================================================================================
int cache_plan_update(const aha_cache_s* c, const CacheLayer& ly, int q, UpdatePlan* plan) {
  UpdatePlan p;
  AHA_REQUIRE(q > 0, "cache update with an empty chunk");
  p.q = q; p.S = c->S; p.R = c->R; p.L = ly.populated ? ly.len : 0;
  p.new_base = ly.ring_base;
  const int W = c->W, S = c->S, R = c->R;
  if (c->policy == AHA_POLICY_STATIC) {
    if (!ly.populated) {            // static_cache.py:26-33: first call stores k[:, :, :W]
      p.store_begin = 0; p.store_end = std::min(q, W); p.store_phys0 = 0; p.klen = p.store_end;
    } else {                        // :34-36: later calls are ignored, the frozen prefix is returned
      p.store_begin = p.store_end = 0; p.klen = ly.len;
    }
    *plan = p;
    return AHA_OK;
  }
  if (!ly.populated) {              // sink_cache.py:124-127 / sliding_window_cache.py:27-30: stored whole, even if q > W
    AHA_REQUIRE(q <= c->cap, "first chunk is larger than the cache capacity (max(window, max_chunk))");
    p.store_begin = 0; p.store_end = q; p.store_phys0 = 0; p.klen = q; p.new_base = 0;
    *plan = p;
    return AHA_OK;
  }
  const int len = ly.len;
  const bool oversize = len > W;    // linear layout left behind by an oversize first chunk
  if (c->policy == AHA_POLICY_SLIDING) {
    const long long total = (long long)len + q;
    if (total <= W) {               // sliding_window_cache.py:33-42 with nothing to drop
      p.store_begin = 0; p.store_end = q; p.store_phys0 = (ly.ring_base + len) % R; p.wrap_end = W; p.klen = (int)total;
    } else if (q >= W) {            // cat(old, new)[-W:] keeps only the last W new tokens
      p.store_begin = q - W; p.store_end = q; p.store_phys0 = 0; p.new_base = 0; p.klen = W; p.evicted = len;
    } else if (oversize) {
      p.keep_copy = W - q; p.compact_from = len - p.keep_copy; p.new_base = 0;
      p.store_begin = 0; p.store_end = q; p.store_phys0 = p.keep_copy; p.klen = W; p.evicted = len - p.keep_copy;
    } else {
      const int e = (int)(total - W);
      p.evicted = e;
      p.new_base = (ly.ring_base + e) % R;
      p.store_begin = 0; p.store_end = q; p.store_phys0 = (ly.ring_base + len) % R; p.wrap_end = W; p.klen = W;
    }
    *plan = p;
    return AHA_OK;
  }
  // ---- sink ----
  if ((long long)q + len < W) {     // sink_cache.py:129-132 growing (only ever reached with ring_base == 0)
    p.store_begin = 0; p.store_end = q; p.store_phys0 = len; p.klen = len + q;
    *plan = p;
    return AHA_OK;
  }
  const int keep = W - S - q;       // sink_cache.py:134-162 shifting
  AHA_REQUIRE(keep > 0, "chunk does not fit behind the sinks: q >= window_length - num_sink_tokens");
  p.keep = keep;
  p.evicted = len + q - W;
  p.klen = W;
  p.store_begin = 0; p.store_end = q;
  if (oversize) {
    p.keep_copy = keep; p.compact_from = len - keep; p.new_base = 0;
    p.store_phys0 = S + keep;
  } else {
    p.new_base = (ly.ring_base + p.evicted) % R;
    p.store_phys0 = S + (p.new_base + keep) % R;
    p.wrap_end = W;
  }
  *plan = p;
  return AHA_OK;
}

int cache_ensure_layer(aha_cache_s* c, int layer) {
  AHA_REQUIRE(layer >= 0 && layer < 4096, "layer index out of range");
  if ((int)c->layers.size() <= layer) c->layers.resize(layer + 1);
  CacheLayer& ly = c->layers[layer];
  if (!ly.k) {
    const size_t bytes = (size_t)c->kvh * c->cap * c->d * sizeof(bf16);
    AHA_CUDA(cudaMalloc(&ly.k, bytes));
    AHA_CUDA(cudaMalloc(&ly.v, bytes));
    // attention loads whole 128-row tiles by TMA and masks only the scores: rows past the valid length meet P = 0 in the
    // P.V product, so they must hold finite values (recycled device memory may hold NaN / Inf bit patterns)
    AHA_CUDA(cudaMemset(ly.k, 0, bytes));
    AHA_CUDA(cudaMemset(ly.v, 0, bytes));
    c->dev_views_layers = 0;      // the device-resident views are rebuilt on the next batched call
  }
  if (layer + 1 > c->n_layers) c->n_layers = layer + 1;
  return AHA_OK;
}

int cache_layer0_bookkeeping(aha_cache_s* c, const UpdatePlan& p, const bf16* cos, const bf16* sin, cudaStream_t st) {
  const int q = p.q;
  if (c->policy == AHA_POLICY_SINK) {
    c->seen_tokens += q;                                          // sink_cache.py:105-106
    c->have_rope = (cos != nullptr && sin != nullptr);
    if (c->have_rope) {
      if (c->cos_rows == 0 || c->cos_rows < c->W) {               // sink_cache.py:116-121
        if (c->cos_rows + q > c->cos_cap_rows) {
          // a later chunk may be longer than the first one (the reference accepts any q): grow the tables, stream-ordered
          const int rows = std::max(c->cos_rows + q, 2 * c->cos_cap_rows);
          bf16 *nc = nullptr, *ns = nullptr;
          AHA_CUDA(cudaMallocAsync(&nc, (size_t)rows * c->d * sizeof(bf16), st));
          AHA_CUDA(cudaMallocAsync(&ns, (size_t)rows * c->d * sizeof(bf16), st));
          if (c->cos_rows > 0) {
            AHA_CUDA(cudaMemcpyAsync(nc, c->cos_table, (size_t)c->cos_rows * c->d * sizeof(bf16), cudaMemcpyDeviceToDevice, st));
            AHA_CUDA(cudaMemcpyAsync(ns, c->sin_table, (size_t)c->cos_rows * c->d * sizeof(bf16), cudaMemcpyDeviceToDevice, st));
          }
          AHA_CUDA(cudaFreeAsync(c->cos_table, st));
          AHA_CUDA(cudaFreeAsync(c->sin_table, st));
          c->cos_table = nc; c->sin_table = ns; c->cos_cap_rows = rows;
        }
        const size_t bytes = (size_t)q * c->d * sizeof(bf16);
        AHA_CUDA(cudaMemcpyAsync(c->cos_table + (size_t)c->cos_rows * c->d, cos, bytes, cudaMemcpyDeviceToDevice, st));
        AHA_CUDA(cudaMemcpyAsync(c->sin_table + (size_t)c->cos_rows * c->d, sin, bytes, cudaMemcpyDeviceToDevice, st));
        c->cos_rows += q;
      }
      if (p.keep > 0 && !c->rerot.count(q)) {                     // sink_cache.py:38-55, once per distinct q
        AHA_REQUIRE(c->cos_rows >= c->W, "re-rotation requested before the cos/sin table covers the window");
        bf16* t = nullptr;
        AHA_CUDA(cudaMalloc(&t, (size_t)2 * p.keep * c->d * sizeof(bf16)));
        c->rerot[q] = t;
        rerot_table_kernel<<<grid_for((long long)p.keep * c->d, 256), 256, 0, st>>>(c->cos_table, c->sin_table, t, c->S, q,
                                                                                    p.keep, c->d);
        ++g_aha_launch_count;
        AHA_CHECK_LAUNCH();
      }
    }
  }
  // provenance ids
  if (p.compact_from >= 0 && p.keep_copy > 0) {
    int32_t* tmp = nullptr;
    AHA_CUDA(cudaMallocAsync(&tmp, (size_t)p.keep_copy * 4, st));
    move_ids_kernel<<<(p.keep_copy + 255) / 256, 256, 0, st>>>(c->ids + p.compact_from, tmp, p.keep_copy);
    move_ids_kernel<<<(p.keep_copy + 255) / 256, 256, 0, st>>>(tmp, c->ids + c->S, p.keep_copy);
    g_aha_launch_count += 2;
    AHA_CUDA(cudaFreeAsync(tmp, st));
  }
  if (p.store_end > p.store_begin) {
    const int n = p.store_end - p.store_begin;
    ids_store_kernel<<<(n + 127) / 128, 128, 0, st>>>(c->ids, c->fed, p.store_begin, p.store_end, p.store_phys0,
                                                      p.wrap_end, p.R);
    ++g_aha_launch_count;
    AHA_CHECK_LAUNCH();
  }
  c->fed += q;
  return AHA_OK;
}

int cache_compact(aha_cache_s* c, int layer, const UpdatePlan& p, cudaStream_t st) {
  if (p.compact_from < 0 || p.keep_copy <= 0) return AHA_OK;
  const size_t half = (size_t)c->kvh * c->W * c->d;
  if (!c->scratch) AHA_CUDA(cudaMalloc(&c->scratch, half * sizeof(bf16)));
  CacheLayer& ly = c->layers[layer];
  const int dst_row0 = c->policy == AHA_POLICY_SINK ? c->S : 0;
  bf16* bufs[2] = {ly.k, ly.v};
  for (int i = 0; i < 2; ++i) {
    const long long items = (long long)c->kvh * p.keep_copy * (c->d / 8);
    move_rows_kernel<<<grid_for(items, 256), 256, 0, st>>>(bufs[i], c->cap, p.compact_from, c->scratch, c->W, 0, c->kvh,
                                                          c->d, p.keep_copy);
    move_rows_kernel<<<grid_for(items, 256), 256, 0, st>>>(c->scratch, c->W, 0, bufs[i], c->cap, dst_row0, c->kvh, c->d,
                                                          p.keep_copy);
    g_aha_launch_count += 2;
  }
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}

int cache_rerotate(aha_cache_s* c, int layer0, int n, const UpdatePlan& p, cudaStream_t st) {
  if (p.keep <= 0 || !c->have_rope) return AHA_OK;
  auto it = c->rerot.find(p.q);
  AHA_REQUIRE(it != c->rerot.end(), "re-rotation table missing for this chunk length (layer 0 must be updated first)");
  // measured on B200 at 7B: 33.0 us per launch against 29.1 us for the register-staged kernel above -> off by default
  static const int use_tma = getenv("AHA_REROT_TMA") ? atoi(getenv("AHA_REROT_TMA")) : 0;
  if (use_tma && c->d == 128) {
    // 62 key blocks x 7 layer groups = 434 CTAs at 7B, one launch per 64 layers
    ProfSlot* slot = prof_begin(2, st);
    for (int l0 = 0; l0 < n; l0 += 64) {
      const int cnt = std::min(64, n - l0);
      RerotLayers ls;
      for (int i = 0; i < cnt; ++i) ls.k[i] = c->layers[layer0 + l0 + i].k;
      dim3 grid((p.keep + kRrKeys - 1) / kRrKeys, (cnt + kRrLayersPerCta - 1) / kRrLayersPerCta);
      int rc = launch_pdl(rerotate_tma_kernel, grid, dim3(kRrThreads), 0, st, ls, cnt, (const bf16*)it->second, c->kvh, c->cap,
                          c->S, c->R, p.new_base, p.keep);
      if (rc) return rc;
    }
    if (slot) prof_end(slot, st, 2.0 * n * (double)p.keep * c->kvh * c->d * sizeof(bf16), 0.0);
    AHA_CHECK_LAUNCH();
    return AHA_OK;
  }
  for (int l0 = 0; l0 < n; l0 += 64) {
    const int cnt = std::min(64, n - l0);
    RerotLayers ls;
    for (int i = 0; i < cnt; ++i) ls.k[i] = c->layers[layer0 + l0 + i].k;
    const long long items = (long long)p.keep * (c->d / 16);      // one thread per (kept key, 16-byte chunk), all kv heads
    dim3 grid((unsigned)((items + 255) / 256), cnt, (c->kvh + 3) / 4);
    ProfSlot* slot = prof_begin(2, st);
    int rc = launch_pdl(rerotate_kernel, grid, dim3(256), 0, st, ls, (const bf16*)it->second, c->kvh, c->d, c->cap, c->S, c->R, p.new_base, p.keep);
    if (slot) prof_end(slot, st, 2.0 * cnt * (double)p.keep * c->kvh * c->d * sizeof(bf16), 0.0);   // read + write of every kept key
    if (rc) return rc;
  }
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}

int cache_ensure_device_views(aha_cache_s* c, int n_layers) {
  if (c->dev_views_layers >= n_layers && c->tmaps_dev) return AHA_OK;
  AHA_REQUIRE((int)c->layers.size() >= n_layers, "device views requested for layers that are not allocated");
  std::vector<CUtensorMap> maps(2 * (size_t)n_layers);
  std::vector<bf16*> ks(n_layers);
  for (int l = 0; l < n_layers; ++l) {
    const CacheLayer& ly = c->layers[l];
    AHA_REQUIRE(ly.k && ly.v, "device views: layer slab missing");
    int rc = get_tmap(ly.k, c->kvh * c->cap, c->d, c->d, 128, &maps[2 * l]); if (rc) return rc;
    rc = get_tmap(ly.v, c->kvh * c->cap, c->d, c->d, 128, &maps[2 * l + 1]); if (rc) return rc;
    ks[l] = ly.k;
  }
  // (rare: once per cache) old tables may still be referenced by kernels in flight: free them only after a device sync
  if (c->tmaps_dev || c->layer_k_dev) { AHA_CUDA(cudaDeviceSynchronize()); cudaFree(c->tmaps_dev); cudaFree(c->layer_k_dev); c->tmaps_dev = nullptr; c->layer_k_dev = nullptr; }
  AHA_CUDA(cudaMalloc(&c->tmaps_dev, maps.size() * sizeof(CUtensorMap)));
  AHA_CUDA(cudaMalloc(&c->layer_k_dev, ks.size() * sizeof(bf16*)));
  AHA_CUDA(cudaMemcpy(c->tmaps_dev, maps.data(), maps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
  AHA_CUDA(cudaMemcpy(c->layer_k_dev, ks.data(), ks.size() * sizeof(bf16*), cudaMemcpyHostToDevice));
  c->dev_views_layers = n_layers;
  return AHA_OK;
}

int cache_rerotate_batched(aha_cache_s* const* caches, const UpdatePlan* plans, int n_streams, int n_layers, cudaStream_t st) {
  RerotBatch b{};
  int n = 0, max_keep = 0, kvh = 0, d = 0;
  double bytes = 0;
  for (int s = 0; s < n_streams; ++s) {
    aha_cache_s* c = caches[s];
    const UpdatePlan& p = plans[s];
    if (p.keep <= 0 || !c->have_rope) continue;
    auto it = c->rerot.find(p.q);
    AHA_REQUIRE(it != c->rerot.end(), "re-rotation table missing for this chunk length (layer 0 must be updated first)");
    AHA_REQUIRE(n == 0 || (c->kvh == kvh && c->d == d), "batched re-rotation: cache geometries differ");
    int rc = cache_ensure_device_views(c, n_layers); if (rc) return rc;
    kvh = c->kvh; d = c->d;
    b.s[n++] = RerotStream{c->layer_k_dev, it->second, p.new_base, p.keep, c->S, c->R, c->cap};
    max_keep = std::max(max_keep, p.keep);
    bytes += 2.0 * n_layers * (double)p.keep * c->kvh * c->d * sizeof(bf16);
  }
  if (n == 0) return AHA_OK;
  const long long items = (long long)max_keep * (d / 16);
  dim3 grid((unsigned)((items + 255) / 256), n * n_layers, (kvh + 3) / 4);
  ProfSlot* slot = prof_begin(2, st);
  int rc = launch_pdl(rerotate_batched_kernel, grid, dim3(256), 0, st, b, n_layers, kvh, d);
  if (slot) prof_end(slot, st, bytes, 0.0);
  if (rc) return rc;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}

int cache_store(aha_cache_s* c, int layer, const UpdatePlan& p, const bf16* k, const bf16* v, long long k_hs,
                long long k_ts, long long v_hs, long long v_ts, cudaStream_t st) {
  const int n = p.store_end - p.store_begin;
  if (n <= 0) return AHA_OK;
  CacheLayer& ly = c->layers[layer];
  const long long items = (long long)c->kvh * n * (c->d / 8) * 2;
  cache_store_kernel<<<grid_for(items, 256), 256, 0, st>>>(k, v, k_hs, k_ts, v_hs, v_ts, ly.k, ly.v, c->kvh, c->d, c->cap,
                                                           p.store_begin, p.store_end, p.store_phys0, p.wrap_end, p.R);
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}

void cache_commit(aha_cache_s* c, int layer, const UpdatePlan& p) {
  CacheLayer& ly = c->layers[layer];
  if (c->policy == AHA_POLICY_STATIC) {
    if (!ly.populated) { ly.populated = true; ly.len = p.klen; }
    return;
  }
  ly.populated = true;
  ly.len = p.klen;
  ly.ring_base = p.new_base;
}

int op_rope_append(const bf16* qkv, const bf16* cos, const bf16* sin, bf16* qbuf, const RopeAppendArgs& a, int n_streams,
                   int q, int heads, int kvh, int d, int cap, cudaStream_t st) {
  AHA_REQUIRE(n_streams > 0 && n_streams <= kMaxStreams && d % 16 == 0 && heads % kvh == 0, "rope_append: bad shape");
  const long long items = (long long)q * (heads + 2 * kvh) * (d / 16);
  dim3 grid(grid_for(items, 256), n_streams);
  return launch_pdl(rope_append_kernel, grid, dim3(256), 0, st, qkv, cos, sin, qbuf, a, q, heads, kvh, d, cap);
}

// =====================================================================================
// C-ABI
// =====================================================================================
extern "C" {

int aha_cache_create(int n_layers, int kv_heads, int head_dim, int policy, int window, int sinks, int max_chunk,
                     int device, aha_cache_t* out) {
  AHA_REQUIRE(out != nullptr, "aha_cache_create: out is NULL");
  AHA_REQUIRE(policy >= 0 && policy <= 2, "unknown cache policy");
  AHA_REQUIRE(kv_heads > 0 && head_dim > 0 && head_dim % 16 == 0, "head_dim must be a positive multiple of 16");
  AHA_REQUIRE(window > 0, "window_length must be positive");
  if (policy != AHA_POLICY_SINK) sinks = 0;
  AHA_REQUIRE(sinks >= 0 && sinks < window, "num_sink_tokens must be in [0, window_length)");
  AHA_REQUIRE(n_layers >= 0 && n_layers <= 4096, "bad layer count");
  AHA_CUDA(cudaSetDevice(device));
  aha_cache_s* c = new aha_cache_s();
  c->device = device; c->kvh = kv_heads; c->d = head_dim; c->policy = policy; c->W = window; c->S = sinks;
  c->R = window - sinks;
  c->max_chunk = std::max(max_chunk, 1);
  c->cap = std::max(window, c->max_chunk);
  cudaError_t e = cudaMalloc(&c->ids, (size_t)c->cap * sizeof(int32_t));
  if (e == cudaSuccess && policy == AHA_POLICY_SINK) {
    const size_t rows = (size_t)window + c->max_chunk;
    e = cudaMalloc(&c->cos_table, rows * head_dim * sizeof(bf16));
    if (e == cudaSuccess) e = cudaMalloc(&c->sin_table, rows * head_dim * sizeof(bf16));
    c->cos_cap_rows = (int)rows;
  }
  if (e != cudaSuccess) { aha_cache_destroy(c); return aha_cuda_fail(e, "cudaMalloc(cache)", __FILE__, __LINE__); }
  for (int l = 0; l < n_layers; ++l) {
    int rc = cache_ensure_layer(c, l);
    if (rc) { aha_cache_destroy(c); return rc; }
  }
  *out = c;
  return AHA_OK;
}

int aha_cache_reset(aha_cache_t c) {
  AHA_REQUIRE(c != nullptr, "aha_cache_reset: NULL handle");
  for (auto& ly : c->layers) { ly.populated = false; ly.len = 0; ly.ring_base = 0; }
  c->seen_tokens = 0; c->fed = 0; c->cos_rows = 0; c->have_rope = false;
  for (auto& kv : c->rerot) cudaFree(kv.second);
  c->rerot.clear();
  return AHA_OK;
}

int aha_cache_destroy(aha_cache_t c) {
  if (!c) return AHA_OK;
  for (auto& ly : c->layers) { cudaFree(ly.k); cudaFree(ly.v); }
  for (auto& kv : c->rerot) cudaFree(kv.second);
  cudaFree(c->ids); cudaFree(c->cos_table); cudaFree(c->sin_table); cudaFree(c->scratch);
  cudaFree(c->tmaps_dev); cudaFree(c->layer_k_dev);
  delete c;
  return AHA_OK;
}

int aha_cache_update(aha_cache_t c, int layer, const void* k, const void* v, long long k_hs, long long k_ts,
                     long long v_hs, long long v_ts, const void* cos, const void* sin, int q, void* stream) {
  AHA_REQUIRE(c != nullptr && k != nullptr && v != nullptr, "aha_cache_update: NULL argument");
  cudaStream_t st = (cudaStream_t)stream;
  int rc = cache_ensure_layer(c, layer); if (rc) return rc;
  UpdatePlan p;
  rc = cache_plan_update(c, c->layers[layer], q, &p); if (rc) return rc;
  if (layer == 0) { rc = cache_layer0_bookkeeping(c, p, (const bf16*)cos, (const bf16*)sin, st); if (rc) return rc; }
  else if (c->policy == AHA_POLICY_SINK) c->have_rope = (cos != nullptr && sin != nullptr);
  rc = cache_compact(c, layer, p, st); if (rc) return rc;
  rc = cache_rerotate(c, layer, 1, p, st); if (rc) return rc;
  rc = cache_store(c, layer, p, (const bf16*)k, (const bf16*)v, k_hs, k_ts, v_hs, v_ts, st); if (rc) return rc;
  cache_commit(c, layer, p);
  return AHA_OK;
}

int aha_cache_seq_length(aha_cache_t c, int layer, int* out) {
  AHA_REQUIRE(c != nullptr && out != nullptr, "aha_cache_seq_length: NULL argument");
  *out = (layer >= 0 && layer < (int)c->layers.size() && c->layers[layer].populated) ? c->layers[layer].len : 0;
  return AHA_OK;
}

int aha_cache_seen_tokens(aha_cache_t c, long long* out) {
  AHA_REQUIRE(c != nullptr && out != nullptr, "aha_cache_seen_tokens: NULL argument");
  *out = c->seen_tokens;
  return AHA_OK;
}

int aha_cache_export(aha_cache_t c, int layer, void* K_out, void* V_out, int32_t* prov, void* stream) {
  AHA_REQUIRE(c != nullptr, "aha_cache_export: NULL handle");
  AHA_REQUIRE(layer >= 0 && layer < (int)c->layers.size() && c->layers[layer].populated, "layer not populated");
  cudaStream_t st = (cudaStream_t)stream;
  const CacheLayer& ly = c->layers[layer];
  const int len = ly.len;
  if (len == 0) return AHA_OK;
  const bool ring = (c->policy != AHA_POLICY_STATIC) && len <= c->W;
  const int R = ring ? c->R : 0;
  const long long items = (long long)c->kvh * len * (c->d / 8);
  if (K_out) { export_kernel<<<grid_for(items, 256), 256, 0, st>>>(ly.k, (bf16*)K_out, c->kvh, c->d, c->cap, len, c->S, R, ly.ring_base); ++g_aha_launch_count; }
  if (V_out) { export_kernel<<<grid_for(items, 256), 256, 0, st>>>(ly.v, (bf16*)V_out, c->kvh, c->d, c->cap, len, c->S, R, ly.ring_base); ++g_aha_launch_count; }
  if (prov) { export_ids_kernel<<<(len + 255) / 256, 256, 0, st>>>(c->ids, prov, len, c->S, R, ly.ring_base); ++g_aha_launch_count; }
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}

}  // extern "C"
