// Internal host-side launchers of the memory-bound kernels (elementwise.cu), the attention kernel (attention.cu)
// and the post-processing kernel (postproc.cu).  All enqueue on `stream` and return AHA_OK or a negative status.
#pragma once
#include "common.cuh"

// ---- elementwise.cu ----
int op_rmsnorm(const bf16* x, const bf16* w, bf16* out, int rows, int cols, float eps, cudaStream_t st);
// folded RMSNorm (gemm.h, GemmCall): out = x and sumsq[row] = sum of x[row, :]^2 ; W[:, c] *= v[c]
int op_copy_sumsq(const bf16* x, bf16* out, float* sumsq, int rows, int cols, cudaStream_t st);
int op_fold_columns(bf16* w, const bf16* v, long long rows, int cols, cudaStream_t st);
int op_layernorm(const bf16* x, const bf16* w, const bf16* b, bf16* out, int rows, int cols, float eps, cudaStream_t st);
int op_preprocess(const uint8_t* in, bf16* out, long long n, cudaStream_t st);
// frames [n,3,image,image] (u8 if frames_u8 else bf16 pixel values) -> patches [n*side*side, kpad] bf16, (c,ky,kx) order
int op_patchify(const uint8_t* frames_u8, const bf16* pixels, bf16* out, int n, int image, int patch, int kpad,
                cudaStream_t st);
int op_pool(const bf16* x, bf16* out, int frames, int side, int channels, int stride, int mode, cudaStream_t st);
int pool_out_side(int side, int stride, int mode);
// cos/sin [n_streams][q][d] bf16 at positions pos0[s] + i
struct RopeTableArgs { int pos0[32]; };
int op_rope_table(const RopeTableArgs& a, int n_streams, int q, int d, const float* inv_freq_dev, bf16* cos_out,
                  bf16* sin_out, cudaStream_t st);
int op_embed_assemble(const bf16* table, const int64_t* ids, int n_ids, const bf16* frame_embeds, int n_frame_tokens,
                      bf16* out, int hidden, int vocab, cudaStream_t st);
// three score heads on [rows, H] (video_head_live_llava_qwen.py:185-188) + per-stream last-token scores
int op_heads(const bf16* h, const bf16* w_inf /*[2,H]*/, const bf16* w_rel /*[1,H]*/, const bf16* w_unc /*[1,H]*/,
             int rows, int hidden, int q, float* informative, float* relevance, float* uncertainty, float* scores3,
             cudaStream_t st);
// weight packing helpers (model creation)
int op_copy_rows(const bf16* src, int src_ld, bf16* dst, int dst_ld, int rows, int cols, int dst_row0, int dst_row_step,
                 cudaStream_t st);
int op_fill_zero(void* dst, size_t bytes, cudaStream_t st);

// ---- attention.cu ----
struct AttnSeg {           // one (stream or frame) attention problem
  int q_row0;              // first row of this problem's queries in the Q tensor map
  int k_row0, v_row0;      // first key row in the K / V tensor maps (per kv head: + g * kv_head_rows)
  int klen;                // keys visible (physical slots 0 .. klen-1)
  int off;                 // allowed(i, j) <=> logical(j) <= off + i
  int ring_S, ring_R, ring_base;  // physical slot p -> logical j: p < S ? p : S + (p - S - base + R) % R   (R == 0: j = p)
  int out_row0;            // first output token row
  const void* k_tmap;      // device-resident CUtensorMap of THIS segment's K / V (streams with separate cache slabs in one
  const void* v_tmap;      // launch); null: the launch-wide maps built from AttnProblem::k / v
};
struct AttnProblem {
  const bf16* q; int q_rows_total; int q_ld;   // Q matrix [rows, q_ld]; head h / kv-group columns start at q_col0 + idx * q_col_step
  const bf16* k; long long k_rows_total; int k_ld;
  const bf16* v; long long v_rows_total; int v_ld;
  int q_col0, q_col_step;   // column of head block b: q_col0 + b * q_col_step
  int k_col0, k_col_step, v_col0, v_col_step;
  int q_block_rows;         // rows between consecutive head blocks in the Q matrix (0 when blocks are column blocks)
  int k_head_rows;          // rows between consecutive kv heads in the K/V matrices (0 when heads are column blocks)
  int n_blocks;             // independent (kv-)head blocks per segment
  int nq_rows;              // query rows per (segment, block) = q * group
  int group;                // query rows per token inside a block (GQA group size; 1 for the vision tower)
  int d;                    // true head dim
  int d_pad;                // head dim as laid out in Q/K/V columns (multiple of 16, pads are zero in Q and K)
  float scale;
  bf16* out; int out_ld;    // out[token, (block*group + r) * d + c]
  L2Prefetch pf;            // weights of later kernels to request into L2 (common.cuh)
  unsigned long long* dbg_ts;   // bring-up: phase stamps of CTA (0,0,0), or null
  float* ws_partial;        // key-split scratch (filled in by op_attention): per CTA [128][d_pad] fp32 O, then m[128], l[128]
  int* ws_counters;         // arrival counter per (segment, block, query tile) in this launch's bank (see gemm.h)
  int* ws_zero;             // the other counter bank, zeroed by this launch
  int kv_prefetch;          // key-split launches: K / V tiles requested into L2 ahead of the TMA ring (before griddepcontrol.wait, then rolling)
  int ws_zero_n;
  int no_split;             // 1: never split the key axis (the kernel's splits wait for each other; see gemm_bf16)
  int merge_last;           // key-split launches whose CTAs are not all co-resident: nobody waits, the LAST split to arrive
                            // at the group's (self-resetting) counter merges every chunk (filled in by op_attention)
  int* ws_pcounters;
  int pf_early;             // 1: request them at kernel entry, 0: after this CTA's own K/V loads are in flight
  int n_segs;
  AttnSeg segs[32];
};
int op_attention(const AttnProblem& p, cudaStream_t st);
// attention_persistent.cu: work decomposition of the persistent kernel (filled in by op_attention_persistent)
struct AttnPersistPlan {
  int mode;          // 0: every CTA takes a contiguous range of whole (segment, block, query tile) groups;
                     // 1: stream-K - one team of CTAs per query tile, equal contiguous ranges of (segment, kv head, key tile)
                     //    units inside a team, split groups merged by the last part
  int q_tiles;       // query tiles per (segment, block)
  int gps;           // groups per segment = n_blocks * q_tiles
  int n_groups;      // n_segs * gps
  int units;         // mode 1: key tiles summed over all (segment, kv head) pairs = the units of ONE team (query tile)
  int unit_base[33]; // mode 1: first unit of segment s
};
// returns 1 when the launch was taken (enqueued), 0 when this shape stays with attention.cu, < 0 on errors
int op_attention_persistent(const AttnProblem& p, cudaStream_t st);

// ---- ingest.cu ----
// decoded frames [n, h, w, 3] u8 -> [n, 3, out, out] u8 canvas; flavour 0 = cv2 bilinear + BGR->RGB, 1 = PIL bicubic
int op_ingest_frames(const uint8_t* src_hwc, int n, int h, int w, int flavour, uint8_t* out_chw, int out, cudaStream_t st);

// ---- postproc.cu ----
int op_find_ticks(const double* scores, int n_streams, int T, double prominence, int distance, const double* thresh,
                  int32_t* peaks, int max_peaks, int32_t* n_peaks, double* smoothed, double* thresh_out,
                  cudaStream_t st);
