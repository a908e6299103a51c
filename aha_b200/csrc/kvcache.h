// Device ring-buffer KV cache with the reference's three eviction policies.
//
// HBM layout (one slab per layer, allocated on first touch):
//     K : [kv_heads][cap][d] bf16      V : [kv_heads][cap][d] bf16        cap = max(window, max_chunk)
// Logical order (what Cache.update() returns in the reference) is mapped onto physical slots by
//     phys(j) = j                                   j <  S              (sinks never move)
//             = S + (ring_base + j - S) mod R       j >= S, R = W - S   (window ring; sliding policy has S = 0)
// so an eviction only advances ring_base: surviving keys/values are never copied.  The sink policy additionally
// re-rotates every kept key in place by -q positions (test/sink_cache.py:141-150), bit-exactly as the reference does.
// A chunk larger than the window stored into an empty cache (the reference keeps it whole, sink_cache.py:124-127)
// is kept in linear layout and compacted into the ring on the next update.
#pragma once
#include "common.cuh"
#include <map>
#include <vector>

struct CacheLayer {
  bf16* k = nullptr;   // [kv_heads][cap][d]
  bf16* v = nullptr;
  bool populated = false;
  int len = 0;         // logical length == get_seq_length(layer)
  int ring_base = 0;   // physical ring offset of logical window index 0
};

// What one update() does to one layer, computed on the host from integer state only.
struct UpdatePlan {
  int q = 0;
  int L = 0;            // length before the update (positions of the chunk start here)
  int klen = 0;         // keys returned / attended after the update
  int store_begin = 0;  // chunk tokens [store_begin, store_end) are written ...
  int store_end = 0;
  int store_phys0 = 0;  // ... to physical slots store_phys0 + (t - store_begin), wrapping from wrap_end back by R
  int wrap_end = 0x7fffffff;
  int R = 0;
  int S = 0;
  int new_base = 0;     // ring_base after the update
  int keep = 0;         // sink shift: number of kept window keys to re-rotate (0 = none)
  int compact_from = -1;  // >= 0: linear oversize layout, kept tokens live at physical [compact_from, compact_from+keep_copy)
  int keep_copy = 0;
  int evicted = 0;
};

struct aha_cache_s {
  int device = 0;
  int n_layers = 0, kvh = 0, d = 0;
  int policy = 0, W = 0, S = 0, R = 0, cap = 0, max_chunk = 0;
  std::vector<CacheLayer> layers;
  // layer-0 bookkeeping
  long long seen_tokens = 0;      // SinkCache._seen_tokens
  long long fed = 0;              // global token id generator (provenance)
  int32_t* ids = nullptr;         // [cap] global token id of every physical slot (layer-0 view)
  // SinkCache._cos_cache / _sin_cache (sink_cache.py:109-121): rows appended at layer 0 until there are >= W rows
  bf16* cos_table = nullptr;      // [W + max_chunk][d]
  bf16* sin_table = nullptr;
  int cos_rows = 0;
  bool have_rope = false;
  // SinkCache.cos_sin_rerotation_cache (sink_cache.py:35-55): q -> [2][W-S-q][d] bf16
  std::map<int, bf16*> rerot;
  bf16* scratch = nullptr;        // compaction scratch [kv_heads][W][d] x 2, allocated on demand
  // device-resident views of the layer slabs for launches that cover several streams at once (aha_prefill_chunk with
  // n_streams > 1): tmaps_dev[2 * layer + {0: K, 1: V}] are the TMA descriptors the attention kernel picks per segment,
  // layer_k_dev[layer] the K slab pointers the batched re-rotation walks.  Built by cache_ensure_device_views().
  void* tmaps_dev = nullptr;      // CUtensorMap [2 * dev_views_layers]
  bf16** layer_k_dev = nullptr;
  int dev_views_layers = 0;
  int cos_cap_rows = 0;           // rows allocated in cos_table / sin_table
};

// Integer semantics of one update (shared by the HF-compatible aha_cache_update and by aha_prefill_chunk).
int cache_plan_update(const aha_cache_s* c, const CacheLayer& ly, int q, UpdatePlan* plan);
int cache_ensure_layer(aha_cache_s* c, int layer);
// layer-0 side effects: seen_tokens, cos/sin table rows, re-rotation table for this q, provenance ids.
// cos/sin: [q, d] bf16 device pointers (NULL when the caller has no RoPE).
int cache_layer0_bookkeeping(aha_cache_s* c, const UpdatePlan& p, const bf16* cos, const bf16* sin, cudaStream_t st);
// moves the kept tokens of an oversize linear layout into the ring (rare path)
int cache_compact(aha_cache_s* c, int layer, const UpdatePlan& p, cudaStream_t st);
int cache_ensure_device_views(aha_cache_s* c, int n_layers);
// re-rotates the kept keys of all `n_layers` layers of every stream in one launch (streams whose plan keeps nothing are skipped)
int cache_rerotate_batched(aha_cache_s* const* caches, const UpdatePlan* plans, int n_streams, int n_layers, cudaStream_t st);
// re-rotates the kept keys of layers [layer0, layer0+n) in one launch (plan.keep > 0)
int cache_rerotate(aha_cache_s* c, int layer0, int n, const UpdatePlan& p, cudaStream_t st);
// stores already-rotated k / v of the chunk (strided source) into the planned slots
int cache_store(aha_cache_s* c, int layer, const UpdatePlan& p, const bf16* k, const bf16* v, long long k_hs,
                long long k_ts, long long v_hs, long long v_ts, cudaStream_t st);
void cache_commit(aha_cache_s* c, int layer, const UpdatePlan& p);

// Fused RoPE + append used by aha_prefill_chunk (kvcache.cu): for up to kMaxStreams streams, reads the packed projection output
// qkv [n_streams*q, (heads + 2*kv_heads)*d], rotates q and k at the chunk positions (cos/sin [n_streams][q][d]),
// writes q to qbuf [n_streams][kv_heads][q*group][d] and k / v into every stream's cache slots.
struct RopeAppendStream {
  bf16* k; bf16* v;                // this layer's slabs
  int store_begin, store_end, store_phys0, wrap_end, R;
};
constexpr int kMaxStreams = 32;   // streams per aha_prefill_chunk call (AttnProblem::segs, kernel parameter blocks)
struct RopeAppendArgs { RopeAppendStream s[kMaxStreams]; };
int op_rope_append(const bf16* qkv, const bf16* cos, const bf16* sin, bf16* qbuf, const RopeAppendArgs& a, int n_streams,
                   int q, int heads, int kvh, int d, int cap, cudaStream_t st);
