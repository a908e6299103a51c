/* aha_b200.h - C-ABI of libaha_b200.so: the B200-native (sm_100a) implementation of Aha!'s per-frame streaming
 * ingest-and-score step.
 *
 * The reference (aiden200/Aha-) has no FFI: the path is reached through three Python surfaces.
 * Every entry point below states the reference interface it replaces (file:line relative to the reference repo);
 * aha_b200/*.py are thin ctypes wrappers that keep the reference's class / method names on top of these calls.
 *
 * Conventions
 *   - plain C types only: device pointers, sizes, CUDA stream as void* (cudaStream_t); no torch types.
 *   - every function returns 0 on success, AHA_ERR_VALUE (-1) where the reference would raise ValueError
 *     (bad geometry / shape), AHA_ERR_RUNTIME (-2) where it would raise RuntimeError (CUDA failure);
 *     aha_last_error() returns the thread-local message.
 *   - all work is enqueued on the caller's stream; no hidden synchronisation, no host callbacks.
 *   - handles are bound to one device and are not thread-safe; one process (or thread + handle set) per GPU.
 *   - bf16 tensors are passed as const void* (uint16 storage).
 */
#ifndef AHA_B200_H_
#define AHA_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AHA_OK 0
#define AHA_ERR_VALUE (-1)
#define AHA_ERR_RUNTIME (-2)

#define AHA_POLICY_SINK 0    /* test/sink_cache.py:7-164 */
#define AHA_POLICY_SLIDING 1 /* test/sliding_window_cache.py:5-53 */
#define AHA_POLICY_STATIC 2  /* test/static_cache.py:5-47 */

#define AHA_MASK_SDPA 0 /* allowed(i,j) <=> j <= L + i           (transformers 4.49 sdpa/eager mask sliced to Klen) */
#define AHA_MASK_FA2 1  /* allowed(i,j) <=> j <= i + (Klen - q)  (flash-attn bottom-right causal) */

#define AHA_LOGITS_NONE 0
#define AHA_LOGITS_LAST 1 /* lm_head on the last token of every stream's chunk only */
#define AHA_LOGITS_ALL 2  /* video_head_live_llava_qwen.py:175 (all q positions) */

#define AHA_POOL_BILINEAR 0 /* video_head_live_llava_qwen.py:128-131 */
#define AHA_POOL_AVERAGE 1  /* :124-125 */
#define AHA_POOL_MAX 2      /* :126-127 */
#define AHA_POOL_ADAPTIVE 3 /* models/vision_live.py:21-26 (adaptive_avg_pool2d) */

typedef struct aha_model_s* aha_model_t;
typedef struct aha_cache_s* aha_cache_t;

/* Shapes of VideoHeadLiveLlavaQwenConfig (models/live_llava/video_head_live_llava_qwen.py:43-47) + SigLipVisionConfig
 * (LLaVA_NeXT/llava/model/multimodal_encoder/siglip_encoder.py:70-115). */
typedef struct {
  int hidden, inter, layers, heads, kv_heads, head_dim, vocab;
  float rope_theta, rms_eps;
  int v_hidden, v_inter, v_layers /* layers actually run: 26 */, v_heads, v_patch, v_image;
  float v_eps;
  int pool_stride, pool_mode;
  int max_tokens; /* largest streams*q one aha_prefill_chunk call will see (workspace sizing) */
  int max_frames; /* largest n_frames one aha_vision_embed call will see */
} aha_model_desc;

/* One state-dict entry, named exactly as in the reference checkpoint ("model.layers.0.self_attn.q_proj.weight", ...). */
typedef struct {
  const char* name;
  const void* ptr; /* device pointer, bf16, contiguous */
} aha_weight_entry;

const char* aha_last_error(void);
/* number of kernels this library has launched in this process (bench.py's gpu_launches) */
unsigned long long aha_launch_count(void);

/* Scheduling hint.  shared != 0: another stream of this process keeps the SMs busy (the vision tower of later frames),
 * so the weight-streaming GEMMs use their two-CTAs-per-SM variant, which interleaves with foreign CTAs; shared == 0
 * (default): a <= one-CTA-per-SM grid takes the whole SM with the deep operand ring.  Results are identical. */
int aha_set_sm_sharing(int shared);

/* Bring-up aids (AHA_TS=1 in the environment): SM-cycle phase stamps of CTA (0,0,0) of the most recent attention launch /
 * of the most recent GEMM launch per (grid.x, grid.z) class.  Not part of the reference-facing surface. */
int aha_debug_attention_timeline(unsigned long long* out16);
int aha_debug_attention_persistent_timeline(unsigned long long* out32);
/* launches of the persistent attention kernel so far (tests check which kernel a shape runs on) */
unsigned long long aha_attention_persistent_launch_count(void);   /* AHA_TS=2: CTA 0 of the persistent kernel */
int aha_debug_gemm_timeline(int* n_out, int* grid_x /*[64]*/, int* grid_z /*[64]*/, unsigned long long* out /*[64*16]*/);
/* AHA_TS=1: (accumulator complete, exit) of every CTA (up to 512) of the most recent launch of timeline slot `slot`, ns */
int aha_debug_gemm_tails(int slot, unsigned long long* out /*[512*2]*/);

/* Optional CUDA-event timing of every GEMM launch (bench.py's roofline leg): enable, run, then read the summed launch
 * durations [ms], launch count, algorithmic bytes and FLOPs of class 0 (weight-streaming GEMM, tokens <= 256), 1 (tall
 * GEMM) or 2 (KV re-rotation on eviction). */
int aha_profile_gemm(int enable);
int aha_profile_gemm_read(int cls, double* ms, long long* launches, double* bytes, double* flops);

/* ---- model: replaces build_model_and_tokenizer()'s from_pretrained result as far as the hot path needs it
 *      (models/modeling_live.py:96-181).  Packs the weights it is given into its own device layouts (fused QKV,
 *      interleaved gate/up, padded vision heads); the caller may free its copies afterwards. ---- */
int aha_model_create(const aha_model_desc* desc, const aha_weight_entry* weights, int n_weights, int device,
                     aha_model_t* out);
int aha_model_destroy(aha_model_t m);

/* ---- KV caches: SinkCache(window_length, num_sink_tokens) test/sink_cache.py:10, SlidingWindowCache(window_length)
 *      test/sliding_window_cache.py:10, TrulyStaticCache(window_size) test/static_cache.py:11.
 *      Storage is a device ring buffer; eviction never moves surviving tokens. ---- */
int aha_cache_create(int n_layers, int kv_heads, int head_dim, int policy, int window, int sinks, int max_chunk,
                     int device, aha_cache_t* out);
int aha_cache_reset(aha_cache_t c);
int aha_cache_destroy(aha_cache_t c);
/* Cache.update(key_states, value_states, layer_idx, {"cos","sin"}): k/v are [1, kv_heads, q, d] bf16 with the given
 * element strides; k is already rotated at positions [L, L+q).  cos/sin are [q, d] bf16 (may be NULL: no re-rotation,
 * sink_cache.py:101).  test/sink_cache.py:74-164, test/sliding_window_cache.py:18-44, test/static_cache.py:18-38. */
int aha_cache_update(aha_cache_t c, int layer, const void* k, const void* v, long long k_head_stride,
                     long long k_token_stride, long long v_head_stride, long long v_token_stride, const void* cos,
                     const void* sin, int q, void* stream);
/* get_seq_length(layer_idx): sink_cache.py:57-62, sliding_window_cache.py:46-49, static_cache.py:43-47 */
int aha_cache_seq_length(aha_cache_t c, int layer, int* out);
/* SinkCache._seen_tokens (sink_cache.py:105-106) */
int aha_cache_seen_tokens(aha_cache_t c, long long* out);
/* Logical-order materialisation of what update() returns: K_out/V_out [kv_heads, len, d] bf16, provenance_out [len]
 * int32 global token ids (may each be NULL). */
int aha_cache_export(aha_cache_t c, int layer, void* K_out, void* V_out, int32_t* provenance_out, void* stream);

/* ---- vision: LiveMixin.visual_embed (models/modeling_live.py:31-37) = SigLipVisionTower.forward
 *      (siglip_encoder.py:576-589) -> connector (video_head_live_llava_qwen.py:107-108) -> post_projector_pooling
 *      (:117-136).  Give either uint8 frames [n,3,S,S] (preprocess of siglip_encoder.py:47-67 is fused in) or bf16
 *      pixel_values [n,3,S,S]; the other pointer NULL.  out: [n * frame_num_tokens, hidden] bf16. ---- */
int aha_vision_embed(aha_model_t m, const uint8_t* frames_u8, const void* pixel_values_bf16, int n_frames, void* out,
                     void* stream);

/* ---- chunk assembly: _encode_frame's torch.cat([embed_tokens(last_ids), frame_embeds]) (test/inference.py:204-214).
 *      ids: [n_ids] int64 device (may be NULL when n_ids == 0); frame_embeds [n_frame_tokens, hidden] bf16. ---- */
int aha_assemble_chunk(aha_model_t m, const int64_t* ids, int n_ids, const void* frame_embeds, int n_frame_tokens,
                       void* out, void* stream);

/* ---- the per-frame step: VideoHeadLiveLlavaQwenForCausalLM.forward(inputs_embeds=[1,q,H], past_key_values=cache,
 *      use_cache=True) (video_head_live_llava_qwen.py:138-188, inference branch) for n_streams independent streams at
 *      once: inputs_embeds [n_streams*q, H] bf16 (stream-major), one cache per stream, all with the same q.
 *      Outputs (each may be NULL): logits fp32 [n_streams*q, V] (ALL) or [n_streams, V] (LAST); informative fp32
 *      [n*q, 2]; relevance fp32 [n*q, 1] (post-sigmoid); uncertainty fp32 [n*q, 1] (log-variance); scores3 fp32
 *      [n_streams, 3] = (softmax(informative[-1])[1], relevance[-1], exp(uncertainty[-1])) (test/inference.py:222-227);
 *      hidden_out bf16 [n*q, H] = final-norm output. ---- */
int aha_prefill_chunk(aha_model_t m, const aha_cache_t* caches, int n_streams, const void* inputs_embeds, int q,
                      int mask_mode, int logits_mode, float* logits, float* informative, float* relevance,
                      float* uncertainty, float* scores3, void* hidden_out, void* stream);

/* ---- post-processing: find_ticks (test/live_infer_for_video.py:195-228, test/live_video/infer_live_video.py:70-91):
 *      savgol_filter(15,3,'interp') -> thresh = mean + 0.5*std (unless thresh[s] given, NaN = auto) -> find_peaks(height,
 *      prominence, distance).  scores: [n_streams, T] f64 device.  peaks: [n_streams, max_peaks] int32, n_peaks
 *      [n_streams] int32 (n_peaks[s] > max_peaks signals truncation); smoothed_out [n_streams, T] f64 and thresh_out
 *      [n_streams] f64 optional.  T >= 15 (scipy raises otherwise -> AHA_ERR_VALUE). ---- */
int aha_find_ticks(const double* scores, int n_streams, int T, double prominence, int distance,
                   const double* thresh_or_null, int32_t* peaks, int max_peaks, int32_t* n_peaks, double* smoothed_out,
                   double* thresh_out, void* stream);

/* ---- single operators, exported for the parity tests and for callers that run their own layer loop ---- */
/* ---- greedy next-token choice of the response decode loop (models/modeling_live.py:64-90) ----
 * logits [vocab] f32 (last position).  With n_generated > 0 and penalty != 1 the repetition penalty of transformers'
 * RepetitionPenaltyLogitsProcessor is applied to a copy in `scratch` [vocab] first.  out_id (device, int64) = argmax,
 * first maximum on ties. */
int aha_greedy_pick(const float* logits, int vocab, const long long* generated_ids, int n_generated,
                    float repetition_penalty, float* scratch, long long* out_id, void* stream);

/* ---- score fusion and highlight selection, float64, bit-identical to the reference's Python ----
 * aha_fuse_scores: test/grid_search.py:58-65  out = alpha*inf + beta*rel - (unc >= thr ? (unc - thr)*epsilon : 0), n values.
 * aha_knapsack_select: test/highlight_generator.py:8-37 (knapsack_selection), inf/rel/unc [n_streams, T]; every frame
 *   costs 1, value = inf*alpha + rel*beta + unc*epsilon, budget max_duration <= 1023; selected[n_streams, T] gets 0/1,
 *   scratch_bits [n_streams, T, aha_knapsack_words(max_duration)] uint32 is caller-allocated work space (one word per
 *   frame up to a budget of 31 - the reference's driver uses 20).
 * aha_highlight_mask: test/highlight_generator.py:76-86; mask [n_streams, last_frame + 1] = union over selected frames i
 *   of [gt_frames[i] - half_width, gt_frames[i] + half_width] clipped to [0, last_frame]. */
int aha_fuse_scores(const double* inf, const double* rel, const double* unc, long long n, double alpha, double beta,
                    double epsilon, double uncertainty_threshold, double* out, void* stream);
int aha_knapsack_words(int max_duration);
int aha_knapsack_select(const double* inf, const double* rel, const double* unc, int n_streams, int T, int max_duration,
                        double alpha, double beta, double epsilon, uint32_t* scratch_bits, uint8_t* selected,
                        void* stream);
int aha_highlight_mask(const uint8_t* selected, const int32_t* gt_frames, int n_streams, int T, int half_width,
                       int last_frame, uint8_t* mask, void* stream);

/* out[tokens, features] = X[tokens, K] . W[features, K]^T (+bias) (activation) (+residual); nn.Linear replacement. */
int aha_gemm_bf16(const void* x, int tokens, int ldx, const void* w, int features, int ldw, int K, void* out, int ldo,
                  int out_f32, int round_bf16, int act, const void* bias, const void* residual, int ldr, int res_mod,
                  int force_swap, int force_bn, int force_splits, void* stream);
/* Qwen2RMSNorm (transformers 4.49 modeling_qwen2.py) */
int aha_rmsnorm(const void* x, const void* weight, void* out, int rows, int cols, float eps, void* stream);
/* nn.LayerNorm (siglip_encoder.py:265,267) */
int aha_layernorm(const void* x, const void* weight, const void* bias, void* out, int rows, int cols, float eps,
                  void* stream);
/* SigLipImageProcessor.preprocess (siglip_encoder.py:47-67) + .to(bf16): uint8 [n] -> bf16 [n] */
int aha_preprocess(const uint8_t* frames_u8, void* out_bf16, long long n, void* stream);
/* Frame ingest: decoded frames [n, height, width, 3] u8 (device) -> the [n, 3, out_side, out_side] u8
 * canvas the reference's loaders hand to the image processor: aspect-preserving resize (longer side = out_side, shorter
 * int(short / long * out_side)), centred on black, CHW.  Bit-identical to
 *   flavour 0: load_video_for_testing, test/inference.py:540-563 - cv2.resize (8-bit INTER_LINEAR; the exact 2 x 2 case
 *              runs OpenCV's INTER_AREA fast path), cv2.copyMakeBorder, COLOR_BGR2RGB (input BGR, output RGB);
 *   flavour 1: load_one_frame, test/live_infer_for_video.py:107-119 - PIL Image.resize (BICUBIC), ImageOps.expand (RGB).
 * The coefficient tables of a geometry are built on the host and cached (first call of a geometry synchronises the
 * stream once).  ValueError for a flavour / geometry outside the supported range. */
int aha_ingest_frames(const uint8_t* frames_hwc, int n_frames, int height, int width, int flavour, uint8_t* out_chw,
                      int out_side, void* stream);
/* post_projector_pooling (video_head_live_llava_qwen.py:117-136): [frames, side*side, C] -> [frames, out*out, C] */
int aha_pool(const void* x, void* out, int frames, int side, int channels, int stride, int mode, void* stream);
/* Qwen2RotaryEmbedding: cos/sin [q, d] bf16 for positions pos0 .. pos0+q-1 */
int aha_rope_table(int pos0, int q, int head_dim, float theta, void* cos_out, void* sin_out, void* stream);
/* attention of a chunk against explicit K/V: q [kv_heads, nq, group, d], k/v [kv_heads, klen, d] bf16 contiguous, out
 * [nq, heads*d]; allowed(i,j) <=> j <= off + i (pass off >= klen for no mask).  Used by the parity tests. */
int aha_attention(const void* q, const void* k, const void* v, void* out, int heads, int kv_heads, int nq, int klen,
                  int head_dim, int off, float scale, void* stream);
/* aha_attention for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  Same arguments, same
 * kernel and results as aha_attention; the key-split scratch (fp32 partials + arrival counters) is a fixed device array
 * inside the library (8 MiB + 4 KiB, held by every GPU that loads the library until exit), so the call performs no
 * cudaMalloc / cudaFree / allocating library call.  Limit: (query tiles x kv_heads x key splits) x (128 x head_dim + 256)
 * x 4 bytes <= 8 MiB and at most 1024 (query tile, kv head) groups - ValueError beyond that (use aha_attention).  All
 * calls share the one array, which keeps the previous call's contents (the counters are cleared on `stream` before every
 * launch): calls that overlap on the GPU - other streams, other host threads - race on it and must be serialised. */
int aha_attention_noalloc(const void* q, const void* k, const void* v, void* out, int heads, int kv_heads, int nq, int klen,
                          int head_dim, int off, float scale, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* AHA_B200_H_ */
