// The ingest-and-score engine: weights packed into device layouts, the SigLIP -> projector -> pooling pass
// (LiveMixin.visual_embed, models/modeling_live.py:31-37) and the Qwen2 frame-chunk prefill against the KV caches
// with the three score heads (VideoHeadLiveLlavaQwenForCausalLM.forward, video_head_live_llava_qwen.py:138-188).
// Everything is a sequence of kernel launches on the caller's stream; no host synchronisation.
#include "ops.h"
#include "gemm.h"
#include "kvcache.h"
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

GemmWorkspace* aha_default_workspace(int* rc_out);

namespace {

struct VisionLayerW {
  bf16 *ln1_w, *ln1_b, *qkv_w, *qkv_b, *out_w, *out_b, *ln2_w, *ln2_b, *fc1_w, *fc1_b, *fc2_w, *fc2_b;
};
struct DecoderLayerW {
  bf16 *in_ln, *qkv_w, *qkv_b, *o_w, *post_ln, *gu_w, *down_w;
};

}  // namespace

// The decoder's weights in the order one prefill step streams them (qkv, o, gate|up, down per layer), as one circular
// byte stream.  `issued` is how far (relative to the start of the running step) L2 prefetch requests have been handed
// out; it never runs more than `cap` bytes ahead of the kernel being launched, so the prefetched lines fit the L2.
struct WeightStream {
  std::vector<const char*> ptr;
  std::vector<size_t> begin;      // begin[i] .. begin[i+1] = stream range of segment i
  size_t total = 0, cap = 0;
  long long issued = 0;
  void add(const void* p, size_t bytes) {
    if (begin.empty()) begin.push_back(0);
    ptr.push_back((const char*)p); total += bytes; begin.push_back(total);
  }
  size_t end_of(int seg) const { return begin[seg + 1]; }
  // ranges to request while a kernel runs that will have consumed the stream up to `consumed`
  L2Prefetch take(size_t consumed) {
    L2Prefetch pf{};
    if (cap == 0 || total == 0) return pf;
    if (issued < (long long)consumed) issued = (long long)consumed;
    long long room = (long long)(consumed + cap) - issued;
    for (int k = 0; k < 2 && room >= 16; ++k) {
      const size_t pos = (size_t)(issued % (long long)total);
      const int seg = (int)(std::upper_bound(begin.begin(), begin.end(), pos) - begin.begin()) - 1;
      size_t n = std::min<size_t>(begin[seg + 1] - pos, (size_t)room) & ~(size_t)15;
      if (n == 0) break;
      pf.p[k] = ptr[seg] + (pos - begin[seg]); pf.n[k] = n;
      issued += (long long)n; room -= (long long)n;
    }
    return pf;
  }
  void end_step() { issued = std::max<long long>(0, issued - (long long)total); }
};

struct aha_model_s {
  aha_model_desc d{};
  WeightStream wstream;
  int device = 0;
  // derived geometry
  int v_side = 0, v_patches = 0, v_hd = 0, v_hd_pad = 0, v_kpad = 0, v_qkv_cols = 0;
  int pooled_side = 0, frame_tokens = 0;
  int qkv_cols = 0;
  // one slab owns every packed weight and every workspace
  std::vector<void*> allocs;
  // vision weights
  bf16 *patch_w = nullptr, *patch_b = nullptr, *pos_emb = nullptr;
  std::vector<VisionLayerW> vl;
  bf16 *proj0_w = nullptr, *proj0_b = nullptr, *proj2_w = nullptr, *proj2_b = nullptr;
  // decoder weights
  bf16 *embed = nullptr, *lm_head = nullptr, *norm_w = nullptr, *w_inf = nullptr, *w_rel = nullptr, *w_unc = nullptr;
  std::vector<DecoderLayerW> dl;
  float* inv_freq = nullptr;
  // vision workspaces (max_frames)
  bf16 *ws_patches = nullptr, *ws_vx = nullptr, *ws_vh = nullptr, *ws_vqkv = nullptr, *ws_vattn = nullptr, *ws_vmlp = nullptr,
       *ws_p1 = nullptr, *ws_p2 = nullptr;
  // decoder workspaces (max_tokens)
  bf16 *ws_x = nullptr, *ws_h = nullptr, *ws_qkv = nullptr, *ws_qbuf = nullptr, *ws_attn = nullptr, *ws_act = nullptr,
       *ws_cos = nullptr, *ws_sin = nullptr, *ws_hn = nullptr;
  // RMSNorm folded around the decoder GEMMs (gemm.h, GemmCall): the norm weights are folded into the columns of the
  // qkv and gate|up matrices at load time; `ones` is the weight of the stand-alone norm kernel that the tall
  // orientation (more than 256 tokens per call) still runs, ws_ss_a / ws_ss_b hold the per-tile sums of squares of
  // the residual stream in front of the attention / the MLP block ([kSsTiles][max_tokens] floats each)
  bool fold_norm = false;
  bf16* ones = nullptr;
  float *ws_ss_a = nullptr, *ws_ss_b = nullptr;
  static constexpr int kSsTiles = 64;
};

namespace {

int dev_alloc(aha_model_s* m, void** out, size_t bytes) {
  void* p = nullptr;
  AHA_CUDA(cudaMalloc(&p, bytes));
  AHA_CUDA(cudaMemset(p, 0, bytes));
  m->allocs.push_back(p);
  *out = p;
  return AHA_OK;
}
template <typename T>
int dev_alloc_t(aha_model_s* m, T** out, size_t count) { return dev_alloc(m, reinterpret_cast<void**>(out), count * sizeof(T)); }

struct WeightTable {
  std::map<std::string, const bf16*> t;
  int get(const std::string& name, const bf16** out) const {
    auto it = t.find(name);
    if (it == t.end() || !it->second) { aha_set_error("ValueError: missing weight '" + name + "'"); return AHA_ERR_VALUE; }
    *out = it->second;
    return AHA_OK;
  }
};

// dst[rows, cols] (contiguous) <- src
int copy_w(aha_model_s* m, const WeightTable& wt, const std::string& name, size_t rows, size_t cols, bf16** out) {
  const bf16* src;
  int rc = wt.get(name, &src); if (rc) return rc;
  rc = dev_alloc_t(m, out, rows * cols); if (rc) return rc;
  AHA_CUDA(cudaMemcpy(*out, src, rows * cols * sizeof(bf16), cudaMemcpyDeviceToDevice));
  return AHA_OK;
}

#define TRY(expr) do { int _rc = (expr); if (_rc) return _rc; } while (0)

int build_model(aha_model_s* m, const WeightTable& wt) {
  const aha_model_desc& d = m->d;
  const std::string VT = "model.vision_tower.vision_tower.vision_model.";
  const int C = d.v_hidden;
  // ---- vision ----
  {
    const bf16* pw; TRY(wt.get(VT + "embeddings.patch_embedding.weight", &pw));
    const int kreal = 3 * d.v_patch * d.v_patch;
    TRY(dev_alloc_t(m, &m->patch_w, (size_t)C * m->v_kpad));
    TRY(op_copy_rows(pw, kreal, m->patch_w, m->v_kpad, C, kreal, 0, 1, 0));
    TRY(copy_w(m, wt, VT + "embeddings.patch_embedding.bias", 1, C, &m->patch_b));
    TRY(copy_w(m, wt, VT + "embeddings.position_embedding.weight", m->v_patches, C, &m->pos_emb));
  }
  m->vl.resize(d.v_layers);
  for (int l = 0; l < d.v_layers; ++l) {
    const std::string p = VT + "encoder.layers." + std::to_string(l) + ".";
    VisionLayerW& w = m->vl[l];
    TRY(copy_w(m, wt, p + "layer_norm1.weight", 1, C, &w.ln1_w));
    TRY(copy_w(m, wt, p + "layer_norm1.bias", 1, C, &w.ln1_b));
    // fused q|k|v projection with every head padded from v_hd to v_hd_pad rows (zero weight rows and zero bias), so
    // that the attention kernel finds zeros in the pad columns of Q and K
    TRY(dev_alloc_t(m, &w.qkv_w, (size_t)m->v_qkv_cols * C));
    TRY(dev_alloc_t(m, &w.qkv_b, (size_t)m->v_qkv_cols));
    const char* names[3] = {"q_proj", "k_proj", "v_proj"};
    for (int t = 0; t < 3; ++t) {
      const bf16 *sw, *sb;
      TRY(wt.get(p + "self_attn." + names[t] + ".weight", &sw));
      TRY(wt.get(p + "self_attn." + names[t] + ".bias", &sb));
      for (int h = 0; h < d.v_heads; ++h) {
        const int dst_row = (t * d.v_heads + h) * m->v_hd_pad;
        TRY(op_copy_rows(sw + (size_t)h * m->v_hd * C, C, w.qkv_w, C, m->v_hd, C, dst_row, 1, 0));
        TRY(op_copy_rows(sb + (size_t)h * m->v_hd, m->v_hd, w.qkv_b + dst_row, m->v_hd, 1, m->v_hd, 0, 1, 0));
      }
    }
    TRY(copy_w(m, wt, p + "self_attn.out_proj.weight", C, C, &w.out_w));
    TRY(copy_w(m, wt, p + "self_attn.out_proj.bias", 1, C, &w.out_b));
    TRY(copy_w(m, wt, p + "layer_norm2.weight", 1, C, &w.ln2_w));
    TRY(copy_w(m, wt, p + "layer_norm2.bias", 1, C, &w.ln2_b));
    TRY(copy_w(m, wt, p + "mlp.fc1.weight", d.v_inter, C, &w.fc1_w));
    TRY(copy_w(m, wt, p + "mlp.fc1.bias", 1, d.v_inter, &w.fc1_b));
    TRY(copy_w(m, wt, p + "mlp.fc2.weight", C, d.v_inter, &w.fc2_w));
    TRY(copy_w(m, wt, p + "mlp.fc2.bias", 1, C, &w.fc2_b));
  }
  const int H = d.hidden;
  TRY(copy_w(m, wt, "model.mm_projector.0.weight", H, C, &m->proj0_w));
  TRY(copy_w(m, wt, "model.mm_projector.0.bias", 1, H, &m->proj0_b));
  TRY(copy_w(m, wt, "model.mm_projector.2.weight", H, H, &m->proj2_w));
  TRY(copy_w(m, wt, "model.mm_projector.2.bias", 1, H, &m->proj2_b));
  // ---- decoder ----
  TRY(copy_w(m, wt, "model.embed_tokens.weight", d.vocab, H, &m->embed));
  {
    const bf16 *e, *lh;
    TRY(wt.get("model.embed_tokens.weight", &e));
    TRY(wt.get("lm_head.weight", &lh));
    if (e == lh) m->lm_head = m->embed;                       // tied embeddings share storage
    else TRY(copy_w(m, wt, "lm_head.weight", d.vocab, H, &m->lm_head));
  }
  TRY(copy_w(m, wt, "model.norm.weight", 1, H, &m->norm_w));
  TRY(copy_w(m, wt, "informative_head.weight", 2, H, &m->w_inf));
  TRY(copy_w(m, wt, "relevance_head.weight", 1, H, &m->w_rel));
  TRY(copy_w(m, wt, "uncertainty_head.weight", 1, H, &m->w_unc));
  const int qd = d.heads * d.head_dim, kd = d.kv_heads * d.head_dim;
  {
    static const int fold_env = getenv("AHA_FOLD_NORM") ? atoi(getenv("AHA_FOLD_NORM")) : 1;
    m->fold_norm = fold_env != 0 && H % 128 == 0 && H / 128 <= aha_model_s::kSsTiles;
    if (m->fold_norm) {
      std::vector<uint16_t> one(H, (uint16_t)0x3F80);    // bf16 1.0
      TRY(dev_alloc_t(m, &m->ones, (size_t)H));
      AHA_CUDA(cudaMemcpy(m->ones, one.data(), (size_t)H * 2, cudaMemcpyHostToDevice));
      TRY(dev_alloc_t(m, &m->ws_ss_a, (size_t)aha_model_s::kSsTiles * d.max_tokens));
      TRY(dev_alloc_t(m, &m->ws_ss_b, (size_t)aha_model_s::kSsTiles * d.max_tokens));
    }
  }
  m->dl.resize(d.layers);
  for (int l = 0; l < d.layers; ++l) {
    const std::string p = "model.layers." + std::to_string(l) + ".";
    DecoderLayerW& w = m->dl[l];
    TRY(copy_w(m, wt, p + "input_layernorm.weight", 1, H, &w.in_ln));
    TRY(dev_alloc_t(m, &w.qkv_w, (size_t)m->qkv_cols * H));
    TRY(dev_alloc_t(m, &w.qkv_b, (size_t)m->qkv_cols));
    const bf16 *qw, *kw, *vw, *qb, *kb, *vb;
    TRY(wt.get(p + "self_attn.q_proj.weight", &qw)); TRY(wt.get(p + "self_attn.k_proj.weight", &kw));
    TRY(wt.get(p + "self_attn.v_proj.weight", &vw)); TRY(wt.get(p + "self_attn.q_proj.bias", &qb));
    TRY(wt.get(p + "self_attn.k_proj.bias", &kb)); TRY(wt.get(p + "self_attn.v_proj.bias", &vb));
    AHA_CUDA(cudaMemcpy(w.qkv_w, qw, (size_t)qd * H * 2, cudaMemcpyDeviceToDevice));
    AHA_CUDA(cudaMemcpy(w.qkv_w + (size_t)qd * H, kw, (size_t)kd * H * 2, cudaMemcpyDeviceToDevice));
    AHA_CUDA(cudaMemcpy(w.qkv_w + (size_t)(qd + kd) * H, vw, (size_t)kd * H * 2, cudaMemcpyDeviceToDevice));
    AHA_CUDA(cudaMemcpy(w.qkv_b, qb, (size_t)qd * 2, cudaMemcpyDeviceToDevice));
    AHA_CUDA(cudaMemcpy(w.qkv_b + qd, kb, (size_t)kd * 2, cudaMemcpyDeviceToDevice));
    AHA_CUDA(cudaMemcpy(w.qkv_b + qd + kd, vb, (size_t)kd * 2, cudaMemcpyDeviceToDevice));
    if (m->fold_norm) TRY(op_fold_columns(w.qkv_w, w.in_ln, m->qkv_cols, H, 0));
    TRY(copy_w(m, wt, p + "self_attn.o_proj.weight", H, qd, &w.o_w));
    TRY(copy_w(m, wt, p + "post_attention_layernorm.weight", 1, H, &w.post_ln));
    // gate / up rows interleaved (gate f -> row 2f, up f -> row 2f+1): the SiLU*mul epilogue then finds both
    // halves of an output element inside one accumulator tile
    const bf16 *gw, *uw;
    TRY(wt.get(p + "mlp.gate_proj.weight", &gw)); TRY(wt.get(p + "mlp.up_proj.weight", &uw));
    TRY(dev_alloc_t(m, &w.gu_w, (size_t)2 * d.inter * H));
    TRY(op_copy_rows(gw, H, w.gu_w, H, d.inter, H, 0, 2, 0));
    TRY(op_copy_rows(uw, H, w.gu_w, H, d.inter, H, 1, 2, 0));
    if (m->fold_norm) TRY(op_fold_columns(w.gu_w, w.post_ln, 2LL * d.inter, H, 0));
    TRY(copy_w(m, wt, p + "mlp.down_proj.weight", H, d.inter, &w.down_w));
  }
  for (int l = 0; l < d.layers; ++l) {
    const DecoderLayerW& w = m->dl[l];
    m->wstream.add(w.qkv_w, (size_t)m->qkv_cols * H * 2);
    m->wstream.add(w.o_w, (size_t)H * qd * 2);
    m->wstream.add(w.gu_w, (size_t)2 * d.inter * H * 2);
    m->wstream.add(w.down_w, (size_t)H * d.inter * 2);
  }
  {
    // rolling L2 prefetch window in MB (AHA_PF_MB, 0 disables).  Measured on B200 at 7B with the current kernels
    // (profiles/r01_l2_prefetch_sweep.txt): GEMM-tail prefetch of 16-28 MB is worth 2.4 % of a decoder step (3.56 ->
    // 3.47 ms), 64 MB and more costs time, prefetching from the attention kernel brings nothing.
    const char* e = getenv("AHA_PF_MB");
    m->wstream.cap = (size_t)(e ? std::max(0, atoi(e)) : 20) << 20;
  }
  // Qwen2RotaryEmbedding: inv_freq = 1 / theta^(arange(0, d, 2) / d), float32 like torch
  {
    std::vector<float> inv(d.head_dim / 2);
    for (int j = 0; j < d.head_dim / 2; ++j)
      inv[j] = 1.0f / powf(d.rope_theta, (float)(2 * j) / (float)d.head_dim);
    TRY(dev_alloc_t(m, &m->inv_freq, inv.size()));
    AHA_CUDA(cudaMemcpy(m->inv_freq, inv.data(), inv.size() * sizeof(float), cudaMemcpyHostToDevice));
  }
  // ---- workspaces ----
  const size_t MV = (size_t)d.max_frames * m->v_patches;
  TRY(dev_alloc_t(m, &m->ws_patches, MV * m->v_kpad));
  TRY(dev_alloc_t(m, &m->ws_vx, MV * C));
  TRY(dev_alloc_t(m, &m->ws_vh, MV * C));
  TRY(dev_alloc_t(m, &m->ws_vqkv, MV * m->v_qkv_cols));
  TRY(dev_alloc_t(m, &m->ws_vattn, MV * C));
  TRY(dev_alloc_t(m, &m->ws_vmlp, MV * d.v_inter));
  TRY(dev_alloc_t(m, &m->ws_p1, MV * H));
  TRY(dev_alloc_t(m, &m->ws_p2, MV * H));
  const size_t MT = (size_t)d.max_tokens;
  TRY(dev_alloc_t(m, &m->ws_x, MT * H));
  TRY(dev_alloc_t(m, &m->ws_h, MT * H));
  TRY(dev_alloc_t(m, &m->ws_qkv, MT * m->qkv_cols));
  TRY(dev_alloc_t(m, &m->ws_qbuf, (MT + 128) * qd));
  TRY(dev_alloc_t(m, &m->ws_attn, MT * qd));
  TRY(dev_alloc_t(m, &m->ws_act, MT * d.inter));
  TRY(dev_alloc_t(m, &m->ws_cos, MT * d.head_dim));
  TRY(dev_alloc_t(m, &m->ws_sin, MT * d.head_dim));
  TRY(dev_alloc_t(m, &m->ws_hn, MT * H));
  AHA_CUDA(cudaDeviceSynchronize());
  return AHA_OK;
}

}  // namespace

namespace {

// folded RMSNorm arguments of one GEMM (gemm.h): consumer side (rs, rs_n) and / or producer side (ss_out, ss_tiles)
struct NormFold { const float* rs; int rs_n; float* ss_out; int* ss_tiles; int ld; float eps; };

int linear(const bf16* x, int tokens, int ldx, const bf16* w, int features, int K, void* out, int ldo, int act,
           const bf16* bias, const bf16* residual, int ldr, int res_mod, int out_f32, cudaStream_t st,
           const L2Prefetch* pf = nullptr, const NormFold* nf = nullptr) {
  int rc;
  GemmWorkspace* ws = aha_default_workspace(&rc);
  if (!ws) return rc;
  GemmCall c{};
  c.x = x; c.tokens = tokens; c.ldx = ldx; c.w = w; c.features = features; c.ldw = K; c.K = K;
  c.out = out; c.ldo = ldo; c.out_f32 = out_f32; c.round_bf16 = out_f32; c.act = act; c.bias = bias;
  c.residual = residual; c.ldr = ldr; c.res_mod = res_mod; c.force_swap = -1; c.force_bn = 0; c.force_splits = 0; c.force_deep = -1;
  if (pf) c.pf = *pf;
  if (nf) {
    c.rs_partials = nf->rs; c.rs_n = nf->rs_n; c.rs_ld = nf->ld; c.rs_eps = nf->eps;
    c.sumsq_out = nf->ss_out; c.sumsq_ld = nf->ld; c.sumsq_tiles = nf->ss_tiles;
  }
  return gemm_bf16(c, ws, st);
}

// SigLipVisionTower.forward -> connector -> post_projector_pooling for n frames (n <= max_frames)
int vision_forward(aha_model_s* m, const uint8_t* frames_u8, const bf16* pixels, int n, bf16* out, cudaStream_t st) {
  const aha_model_desc& d = m->d;
  const int C = d.v_hidden, P = m->v_patches, M = n * P, H = d.hidden;
  TRY(op_patchify(frames_u8, pixels, m->ws_patches, n, d.v_image, d.v_patch, m->v_kpad, st));
  // conv14/14 as a GEMM; bias in the epilogue, position embedding as a row-periodic residual (siglip_encoder.py:169-174)
  TRY(linear(m->ws_patches, M, m->v_kpad, m->patch_w, C, m->v_kpad, m->ws_vx, C, GEMM_ACT_NONE, m->patch_b, m->pos_emb, C, P, 0, st));
  AttnProblem ap{};
  ap.q = ap.k = ap.v = m->ws_vqkv;
  ap.q_rows_total = M; ap.k_rows_total = M; ap.v_rows_total = M;
  ap.q_ld = ap.k_ld = ap.v_ld = m->v_qkv_cols;
  ap.q_col0 = 0; ap.k_col0 = d.v_heads * m->v_hd_pad; ap.v_col0 = 2 * d.v_heads * m->v_hd_pad;
  ap.q_col_step = ap.k_col_step = ap.v_col_step = m->v_hd_pad;
  ap.q_block_rows = 0; ap.k_head_rows = 0;
  ap.n_blocks = d.v_heads; ap.nq_rows = P; ap.group = 1; ap.d = m->v_hd; ap.d_pad = m->v_hd_pad;
  ap.scale = 1.0f / sqrtf((float)m->v_hd);
  ap.out = m->ws_vattn; ap.out_ld = C;
  ap.no_split = aha_sm_shared() ? 1 : 0;   // the decoder stream owns the waiting-split kernels while the SMs are shared
  for (int l = 0; l < d.v_layers; ++l) {
    const VisionLayerW& w = m->vl[l];
    TRY(op_layernorm(m->ws_vx, w.ln1_w, w.ln1_b, m->ws_vh, M, C, d.v_eps, st));
    TRY(linear(m->ws_vh, M, C, w.qkv_w, m->v_qkv_cols, C, m->ws_vqkv, m->v_qkv_cols, GEMM_ACT_NONE, w.qkv_b, nullptr, 0, 0, 0, st));
    for (int f0 = 0; f0 < n; f0 += 32) {
      ap.n_segs = std::min(32, n - f0);
      for (int i = 0; i < ap.n_segs; ++i) {
        AttnSeg& s = ap.segs[i];
        s.q_row0 = s.k_row0 = s.v_row0 = s.out_row0 = (f0 + i) * P;
        s.klen = P; s.off = P; s.ring_S = 0; s.ring_R = 0; s.ring_base = 0;
      }
      TRY(op_attention(ap, st));
    }
    TRY(linear(m->ws_vattn, M, C, w.out_w, C, C, m->ws_vx, C, GEMM_ACT_NONE, w.out_b, m->ws_vx, C, 0, 0, st));
    TRY(op_layernorm(m->ws_vx, w.ln2_w, w.ln2_b, m->ws_vh, M, C, d.v_eps, st));
    TRY(linear(m->ws_vh, M, C, w.fc1_w, d.v_inter, C, m->ws_vmlp, d.v_inter, GEMM_ACT_GELU_TANH, w.fc1_b, nullptr, 0, 0, 0, st));
    TRY(linear(m->ws_vmlp, M, d.v_inter, w.fc2_w, C, d.v_inter, m->ws_vx, C, GEMM_ACT_NONE, w.fc2_b, m->ws_vx, C, 0, 0, st));
  }
  // mm_projector mlp2x_gelu on all patches (multimodal_projector/builder.py:42-48), then pooling
  TRY(linear(m->ws_vx, M, C, m->proj0_w, H, C, m->ws_p1, H, GEMM_ACT_GELU_ERF, m->proj0_b, nullptr, 0, 0, 0, st));
  TRY(linear(m->ws_p1, M, H, m->proj2_w, H, H, m->ws_p2, H, GEMM_ACT_NONE, m->proj2_b, nullptr, 0, 0, 0, st));
  TRY(op_pool(m->ws_p2, out, n, m->v_side, H, d.pool_stride, d.pool_mode, st));
  return AHA_OK;
}

int prefill(aha_model_s* m, const aha_cache_t* caches, int ns, const bf16* x_in, int q, int mask_mode, int logits_mode,
            float* logits, float* informative, float* relevance, float* uncertainty, float* scores3, bf16* hidden_out,
            cudaStream_t st) {
  const aha_model_desc& d = m->d;
  const int H = d.hidden, nh = d.heads, kvh = d.kv_heads, hd = d.head_dim, G = nh / kvh;
  const int M = ns * q, qd = nh * hd;
  AHA_REQUIRE(ns > 0 && ns <= kMaxStreams, "prefill: between 1 and 32 streams per call");
  AHA_REQUIRE(q > 0 && M <= d.max_tokens, "prefill: streams * q exceeds max_tokens of the model handle");
  AHA_REQUIRE(mask_mode == AHA_MASK_SDPA || mask_mode == AHA_MASK_FA2, "prefill: unknown mask mode");
  std::vector<UpdatePlan> plans(ns);
  RopeTableArgs ra{};
  // every check comes before the first cache mutation: a rejected call leaves all caches as they were
  for (int s = 0; s < ns; ++s) {
    aha_cache_s* c = caches[s];
    AHA_REQUIRE(c != nullptr, "prefill: NULL cache");
    AHA_REQUIRE(c->kvh == kvh && c->d == hd, "prefill: cache geometry does not match the model");
    AHA_REQUIRE(q <= c->max_chunk, "prefill: chunk longer than the cache's max_chunk");
    AHA_REQUIRE(c->cap == caches[0]->cap, "prefill: all caches of one call must share window / max_chunk");
    for (int t = 0; t < s; ++t) AHA_REQUIRE(caches[t] != c, "prefill: the same cache given for two streams");
  }
  for (int s = 0; s < ns; ++s) {
    aha_cache_s* c = caches[s];
    for (int l = 0; l < d.layers; ++l) TRY(cache_ensure_layer(c, l));      // (allocation only, no state change)
    for (int l = 1; l < d.layers; ++l)
      AHA_REQUIRE(c->layers[l].populated == c->layers[0].populated && c->layers[l].len == c->layers[0].len &&
                      c->layers[l].ring_base == c->layers[0].ring_base,
                  "prefill: cache layers are not in lockstep");
    TRY(cache_plan_update(c, c->layers[0], q, &plans[s]));
    ra.pos0[s] = plans[s].L;   // position_ids = arange(L, L+q), L = get_seq_length()
    if (ns > 1) TRY(cache_ensure_device_views(c, d.layers));
  }
  TRY(op_rope_table(ra, ns, q, hd, m->inv_freq, m->ws_cos, m->ws_sin, st));
  for (int s = 0; s < ns; ++s) {
    aha_cache_s* c = caches[s];
    TRY(cache_layer0_bookkeeping(c, plans[s], m->ws_cos + (size_t)s * q * hd, m->ws_sin + (size_t)s * q * hd, st));
    for (int l = 0; l < d.layers; ++l) TRY(cache_compact(c, l, plans[s], st));
    if (ns == 1) TRY(cache_rerotate(c, 0, d.layers, plans[s], st));   // every layer's kept keys in one launch
  }
  if (ns > 1) TRY(cache_rerotate_batched(caches, plans.data(), ns, d.layers, st));   // ... of every stream
  // RMSNorm folded around the GEMMs when the call runs in the skinny orientation (at most 832 tokens)
  const bool fold = m->fold_norm && M <= kGemmSkinnyMaxTokens;
  int ss_na = 1, ss_nb = 1;          // partial rows per token in ws_ss_a / ws_ss_b
  if (fold) TRY(op_copy_sumsq(x_in, m->ws_x, m->ws_ss_a, M, H, st));
  else AHA_CUDA(cudaMemcpyAsync(m->ws_x, x_in, (size_t)M * H * sizeof(bf16), cudaMemcpyDeviceToDevice, st));
  const int ss_ld = d.max_tokens;
  AttnProblem ap{};
  ap.q = m->ws_qbuf; ap.q_rows_total = M * nh; ap.q_ld = hd;
  ap.q_col0 = ap.k_col0 = ap.v_col0 = 0; ap.q_col_step = ap.k_col_step = ap.v_col_step = 0;
  ap.q_block_rows = q * G;
  ap.n_blocks = kvh; ap.nq_rows = q * G; ap.group = G; ap.d = hd; ap.d_pad = hd;
  ap.scale = 1.0f / sqrtf((float)hd);
  ap.out = m->ws_attn; ap.out_ld = qd;
  ap.n_segs = 1;
  // bring-up knob (timing breakdowns only, results are garbage): bit mask of per-layer kernels to skip
  // 1 norms, 2 rope+append, 4 attention, 8 qkv, 16 o_proj, 32 gate|up, 64 down
  static const int skip = getenv("AHA_SKIP") ? atoi(getenv("AHA_SKIP")) : 0;
  for (int l = 0; l < d.layers; ++l) {
    const DecoderLayerW& w = m->dl[l];
    if (!fold && !(skip & 1)) TRY(op_rmsnorm(m->ws_x, m->fold_norm ? m->ones : w.in_ln, m->ws_h, M, H, d.rms_eps, st));
    WeightStream& wsm = m->wstream;
    static const int pf_gemm = getenv("AHA_PF_GEMM") ? atoi(getenv("AHA_PF_GEMM")) : 1;
    auto take_g = [&](int seg) { return pf_gemm ? wsm.take(wsm.end_of(seg)) : L2Prefetch{}; };
    L2Prefetch pf = take_g(4 * l);
    const NormFold nf_qkv{m->ws_ss_a, ss_na, nullptr, nullptr, ss_ld, d.rms_eps};
    const NormFold nf_o{nullptr, 0, m->ws_ss_b, &ss_nb, ss_ld, d.rms_eps};
    const NormFold nf_down{nullptr, 0, m->ws_ss_a, &ss_na, ss_ld, d.rms_eps};
    if (!(skip & 8)) TRY(linear(fold ? m->ws_x : m->ws_h, M, H, w.qkv_w, m->qkv_cols, H, m->ws_qkv, m->qkv_cols, GEMM_ACT_NONE, w.qkv_b, nullptr, 0, 0, 0, st, &pf, fold ? &nf_qkv : nullptr));
    RopeAppendArgs aa{};
    const int cap = caches[0]->cap;
    for (int s = 0; s < ns; ++s) {
      aha_cache_s* c = caches[s];
      aa.s[s] = RopeAppendStream{c->layers[l].k, c->layers[l].v, plans[s].store_begin, plans[s].store_end,
                                 plans[s].store_phys0, plans[s].wrap_end, plans[s].R};
    }
    if (!(skip & 2)) TRY(op_rope_append(m->ws_qkv, m->ws_cos, m->ws_sin, m->ws_qbuf, aa, ns, q, nh, kvh, hd, cap, st));
    // ONE attention launch covers all streams of the call: every stream is a segment with its own mask geometry and
    // (their cache slabs being separate allocations) its own device-resident K / V tensor maps
    ap.k = caches[0]->layers[l].k; ap.v = caches[0]->layers[l].v;
    ap.k_rows_total = ap.v_rows_total = (long long)kvh * cap;
    ap.k_ld = ap.v_ld = hd; ap.k_head_rows = cap;
    ap.n_segs = ns;
    for (int s = 0; s < ns; ++s) {
      aha_cache_s* c = caches[s];
      const UpdatePlan& p = plans[s];
      AttnSeg& sg = ap.segs[s];
      sg.q_row0 = s * kvh * q * G;
      sg.k_row0 = sg.v_row0 = 0;
      sg.klen = p.klen;
      sg.off = (mask_mode == AHA_MASK_SDPA) ? p.L : p.klen - q;
      const bool ring = c->policy != AHA_POLICY_STATIC && p.klen <= c->W;
      sg.ring_S = c->S; sg.ring_R = ring ? c->R : 0; sg.ring_base = p.new_base;
      sg.out_row0 = s * q;
      sg.k_tmap = sg.v_tmap = nullptr;
      if (ns > 1) {
        sg.k_tmap = (const char*)c->tmaps_dev + (size_t)(2 * l) * 128;       // sizeof(CUtensorMap) == 128
        sg.v_tmap = (const char*)c->tmaps_dev + (size_t)(2 * l + 1) * 128;
      }
    }
    static const int pf_attn = getenv("AHA_PF_ATTN") ? atoi(getenv("AHA_PF_ATTN")) : 0;
    ap.pf = pf_attn ? wsm.take(wsm.end_of(4 * l)) : L2Prefetch{};
    ap.pf_early = pf_attn == 2;
    if (!(skip & 4)) TRY(op_attention(ap, st));
    pf = take_g(4 * l + 1);
    if (!(skip & 16)) TRY(linear(m->ws_attn, M, qd, w.o_w, H, qd, m->ws_x, H, GEMM_ACT_NONE, nullptr, m->ws_x, H, 0, 0, st, &pf, fold ? &nf_o : nullptr));
    if (!fold && !(skip & 1)) TRY(op_rmsnorm(m->ws_x, m->fold_norm ? m->ones : w.post_ln, m->ws_h, M, H, d.rms_eps, st));
    pf = take_g(4 * l + 2);
    const NormFold nf_gu2{m->ws_ss_b, ss_nb, nullptr, nullptr, ss_ld, d.rms_eps};
    if (!(skip & 32)) TRY(linear(fold ? m->ws_x : m->ws_h, M, H, w.gu_w, 2 * d.inter, H, m->ws_act, d.inter, GEMM_ACT_SILU_MUL, nullptr, nullptr, 0, 0, 0, st, &pf, fold ? &nf_gu2 : nullptr));
    pf = take_g(4 * l + 3);
    if (!(skip & 64)) TRY(linear(m->ws_act, M, d.inter, w.down_w, H, d.inter, m->ws_x, H, GEMM_ACT_NONE, nullptr, m->ws_x, H, 0, 0, st, &pf, fold ? &nf_down : nullptr));
  }
  m->wstream.end_step();
  bf16* hn = hidden_out ? hidden_out : m->ws_hn;
  TRY(op_rmsnorm(m->ws_x, m->norm_w, hn, M, H, d.rms_eps, st));
  if (informative || relevance || uncertainty || scores3)
    TRY(op_heads(hn, m->w_inf, m->w_rel, m->w_unc, M, H, q, informative, relevance, uncertainty, scores3, st));
  if (logits_mode == AHA_LOGITS_ALL && logits) {
    TRY(linear(hn, M, H, m->lm_head, d.vocab, H, logits, d.vocab, GEMM_ACT_NONE, nullptr, nullptr, 0, 0, 1, st));
  } else if (logits_mode == AHA_LOGITS_LAST && logits) {
    TRY(linear(hn + (size_t)(q - 1) * H, ns, q * H, m->lm_head, d.vocab, H, logits, d.vocab, GEMM_ACT_NONE, nullptr, nullptr, 0, 0, 1, st));
  }
  for (int s = 0; s < ns; ++s)
    for (int l = 0; l < d.layers; ++l) cache_commit(caches[s], l, plans[s]);
  return AHA_OK;
}

}  // namespace

extern "C" {

int aha_model_create(const aha_model_desc* desc, const aha_weight_entry* weights, int n_weights, int device,
                     aha_model_t* out) {
  AHA_REQUIRE(desc && weights && out && n_weights > 0, "aha_model_create: NULL argument");
  const aha_model_desc& d = *desc;
  AHA_REQUIRE(d.hidden > 0 && d.hidden % 8 == 0 && d.inter % 8 == 0 && d.layers > 0, "bad decoder shape");
  AHA_REQUIRE(d.heads > 0 && d.kv_heads > 0 && d.heads % d.kv_heads == 0, "heads must be a multiple of kv_heads");
  AHA_REQUIRE(d.head_dim == 64 || d.head_dim == 128, "decoder head_dim must be 64 or 128");
  AHA_REQUIRE(d.v_hidden % d.v_heads == 0 && d.v_hidden % 8 == 0 && d.v_inter % 8 == 0, "bad vision shape");
  AHA_REQUIRE((d.v_hidden / d.v_heads) % 8 == 0 && d.v_hidden / d.v_heads <= 128, "vision head_dim must be a multiple of 8, <= 128");
  AHA_REQUIRE(d.v_image >= d.v_patch && d.v_patch > 0, "image smaller than one patch");  // conv stride = patch: trailing pixels are ignored
  AHA_REQUIRE(d.max_tokens > 0 && d.max_frames > 0, "max_tokens / max_frames must be positive");
  AHA_REQUIRE(d.pool_mode >= 0 && d.pool_mode <= 3, "Unexpected mm_spatial_pool_mode");
  AHA_CUDA(cudaSetDevice(device));
  aha_model_s* m = new aha_model_s();
  m->d = d; m->device = device;
  m->v_side = d.v_image / d.v_patch; m->v_patches = m->v_side * m->v_side;
  m->v_hd = d.v_hidden / d.v_heads; m->v_hd_pad = (m->v_hd + 15) / 16 * 16;
  m->v_kpad = (3 * d.v_patch * d.v_patch + 7) / 8 * 8;
  m->v_qkv_cols = 3 * d.v_heads * m->v_hd_pad;
  m->pooled_side = pool_out_side(m->v_side, d.pool_stride, d.pool_mode);
  m->frame_tokens = m->pooled_side * m->pooled_side;
  m->qkv_cols = (d.heads + 2 * d.kv_heads) * d.head_dim;
  WeightTable wt;
  for (int i = 0; i < n_weights; ++i)
    if (weights[i].name) wt.t[weights[i].name] = (const bf16*)weights[i].ptr;
  int rc = build_model(m, wt);
  if (rc) { aha_model_destroy(m); return rc; }
  *out = m;
  return AHA_OK;
}

int aha_model_destroy(aha_model_t m) {
  if (!m) return AHA_OK;
  for (void* p : m->allocs) cudaFree(p);
  delete m;
  return AHA_OK;
}

int aha_vision_embed(aha_model_t m, const uint8_t* frames_u8, const void* pixel_values_bf16, int n_frames, void* out,
                     void* stream) {
  AHA_REQUIRE(m && out, "aha_vision_embed: NULL argument");
  AHA_REQUIRE(n_frames > 0, "aha_vision_embed: no frames");
  AHA_REQUIRE((frames_u8 != nullptr) != (pixel_values_bf16 != nullptr), "give exactly one of frames_u8 / pixel_values");
  const size_t in_stride = (size_t)3 * m->d.v_image * m->d.v_image;
  for (int f0 = 0; f0 < n_frames; f0 += m->d.max_frames) {
    const int n = std::min(m->d.max_frames, n_frames - f0);
    int rc = vision_forward(m, frames_u8 ? frames_u8 + f0 * in_stride : nullptr,
                            pixel_values_bf16 ? (const bf16*)pixel_values_bf16 + f0 * in_stride : nullptr, n,
                            (bf16*)out + (size_t)f0 * m->frame_tokens * m->d.hidden, (cudaStream_t)stream);
    if (rc) return rc;
  }
  return AHA_OK;
}

int aha_assemble_chunk(aha_model_t m, const int64_t* ids, int n_ids, const void* frame_embeds, int n_frame_tokens,
                       void* out, void* stream) {
  AHA_REQUIRE(m && out, "aha_assemble_chunk: NULL argument");
  AHA_REQUIRE(n_ids == 0 || ids, "aha_assemble_chunk: ids is NULL");
  AHA_REQUIRE(n_frame_tokens == 0 || frame_embeds, "aha_assemble_chunk: frame_embeds is NULL");
  return op_embed_assemble(m->embed, ids, n_ids, (const bf16*)frame_embeds, n_frame_tokens, (bf16*)out, m->d.hidden,
                           m->d.vocab, (cudaStream_t)stream);
}

int aha_prefill_chunk(aha_model_t m, const aha_cache_t* caches, int n_streams, const void* inputs_embeds, int q,
                      int mask_mode, int logits_mode, float* logits, float* informative, float* relevance,
                      float* uncertainty, float* scores3, void* hidden_out, void* stream) {
  AHA_REQUIRE(m && caches && inputs_embeds, "aha_prefill_chunk: NULL argument");
  return prefill(m, caches, n_streams, (const bf16*)inputs_embeds, q, mask_mode, logits_mode, logits, informative,
                 relevance, uncertainty, scores3, (bf16*)hidden_out, (cudaStream_t)stream);
}

}  // extern "C"
