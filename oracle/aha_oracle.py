"""CPU oracle for the Aha! per-frame ingest-and-score path.  TEST INFRASTRUCTURE ONLY.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may import
this module; the product path (`aha_b200/`) never does and fails loudly without its CUDA library.

It is a plain-PyTorch (CPU) restatement of the reference's algorithm, function by function:

  preprocess            LLaVA_NeXT/llava/model/multimodal_encoder/siglip_encoder.py:47-67
  siglip_tower          siglip_encoder.py:156-174 (embeddings), :197-239 (attention), :251-255 (MLP),
                        :269-305 (layer), :570-589 (tower surgery: last layer dropped, hidden_states[-1])
  connector             multimodal_projector/builder.py:42-48 (mlp2x_gelu), called models/modeling_live.py:34
  post_projector_pooling models/live_llava/video_head_live_llava_qwen.py:117-136
  visual_embed          models/modeling_live.py:31-37
  OracleKVCache         test/sink_cache.py:74-164, test/sliding_window_cache.py:18-49, test/static_cache.py:18-47
  qwen2_prefill         transformers==4.49.0 models/qwen2/modeling_qwen2.py (pinned by requirements.txt:57; NOT in
                        /root/reference) - restated from its published algorithm, anchored on
                        the reference call site video_head_live_llava_qwen.py:160-188
  heads / frame scores  video_head_live_llava_qwen.py:175-188, test/inference.py:222-229
  LiveInferOracle       test/inference.py:112-131, 196-229, 283-334
  find_ticks            test/live_infer_for_video.py:195-228, test/live_video/infer_live_video.py:70-91

Parity pin: the reference ships no golden vectors.  The pieces of the reference that import
in the authoring container (the three cache classes, the vendored SigLIP tower and image processor, scipy) were
executed there against this restatement by `oracle/validate_against_reference.py`, and the outputs are committed
as fixtures under `tests/golden/` by `tests/golden/make_golden.py`.  The Qwen2 decoder restatement is pinned by
execution too (tests/test_oracle_qwen2_pin.py, tests/golden/make_golden_qwen2.py): bit-exact in fp32 and bf16 against
the installed transformers `Qwen2Model` (eager attention, DynamicCache) chunk by chunk, and - for the 4.49-only
semantics (positions from get_seq_length(), mask sliced to the Klen the cache returns, cos / sin handed to update()) -
against HF's own Qwen2DecoderLayer / rotary / norm modules driven with the reference's three cache classes and the 4.49
additive mask, through eviction + re-rotation, the bidirectional regime of a full cache and the frozen static prefix.
What remains restated without an executable counterpart: mm_projector + pooling glue (two Linear calls and
F.interpolate), the three Linear heads and the _encode_frame loop (the modules that hold them import peft /
bitsandbytes / deepspeed, absent here).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F


# ----------------------------------------------------------------------------------------------
# Configuration
# ----------------------------------------------------------------------------------------------
@dataclass
class VisionDims:
    hidden: int = 1152
    inter: int = 4304
    layers: int = 26          # layers actually run (27 minus the one deleted at siglip_encoder.py:570)
    heads: int = 16
    patch: int = 14
    image: int = 384
    eps: float = 1e-6

    @property
    def side(self) -> int:
        return self.image // self.patch

    @property
    def patches(self) -> int:
        return self.side * self.side


@dataclass
class AhaDims:
    hidden: int
    inter: int
    layers: int
    heads: int
    kv_heads: int
    head_dim: int
    vocab: int
    rope_theta: float = 1e6
    rms_eps: float = 1e-6
    tie_embeddings: bool = False
    vision: VisionDims = field(default_factory=VisionDims)
    pool_stride: int = 4
    pool_mode: str = "bilinear"

    @property
    def pooled_side(self) -> int:
        s = self.vision.side
        return math.ceil(s / self.pool_stride) if self.pool_mode == "bilinear" else s // self.pool_stride

    @property
    def frame_num_tokens(self) -> int:
        return self.pooled_side ** 2


def dims_preset(name: str) -> AhaDims:
    """'7b' / 'tiny' are the lmms-lab llava-onevision-qwen2-{7b,0.5b}-ov shapes (models/arguments_live.py:12);
    'micro' is a shrunken shape for CPU-speed tests (same structure, 6x6 patches -> 2x2 pooled tokens)."""
    if name == "7b":
        return AhaDims(3584, 18944, 28, 28, 4, 128, 152064)
    if name == "tiny":
        return AhaDims(896, 4864, 24, 14, 2, 64, 151936, tie_embeddings=True)
    if name == "micro":
        return AhaDims(128, 320, 3, 4, 2, 32, 512,
                       vision=VisionDims(hidden=96, inter=200, layers=2, heads=4, patch=14, image=84))
    if name == "micro72":
        # head_dim 72 vision (the SigLIP-so400m head size) and head_dim 64 decoder on a small grid
        return AhaDims(256, 512, 2, 4, 2, 64, 768,
                       vision=VisionDims(hidden=144, inter=304, layers=2, heads=2, patch=14, image=126))
    raise ValueError(f"unknown dims preset {name}")


# ----------------------------------------------------------------------------------------------
# Random-init weights (HF-style normal(0, 0.02); shared by the oracle and the CUDA path)
# ----------------------------------------------------------------------------------------------
VT = "model.vision_tower.vision_tower.vision_model."


def init_weights(dims: AhaDims, seed: int = 0, dtype=torch.bfloat16, head_spread: float = 1.5,
                 device: str = "cpu") -> Dict[str, torch.Tensor]:
    g = torch.Generator(device=device)
    g.manual_seed(seed)

    def rn(*shape, std=0.02):
        return (torch.randn(*shape, generator=g, device=device, dtype=torch.float32) * std).to(dtype)

    def ones_noise(n):
        return (1.0 + 0.1 * torch.randn(n, generator=g, device=device, dtype=torch.float32)).to(dtype)

    w: Dict[str, torch.Tensor] = {}
    v = dims.vision
    w[VT + "embeddings.patch_embedding.weight"] = rn(v.hidden, 3, v.patch, v.patch)
    w[VT + "embeddings.patch_embedding.bias"] = rn(v.hidden)
    w[VT + "embeddings.position_embedding.weight"] = rn(v.patches, v.hidden)
    for i in range(v.layers):
        p = VT + f"encoder.layers.{i}."
        w[p + "layer_norm1.weight"] = ones_noise(v.hidden)
        w[p + "layer_norm1.bias"] = rn(v.hidden)
        for n in ("q_proj", "k_proj", "v_proj", "out_proj"):
            w[p + f"self_attn.{n}.weight"] = rn(v.hidden, v.hidden)
            w[p + f"self_attn.{n}.bias"] = rn(v.hidden)
        w[p + "layer_norm2.weight"] = ones_noise(v.hidden)
        w[p + "layer_norm2.bias"] = rn(v.hidden)
        w[p + "mlp.fc1.weight"] = rn(v.inter, v.hidden)
        w[p + "mlp.fc1.bias"] = rn(v.inter)
        w[p + "mlp.fc2.weight"] = rn(v.hidden, v.inter)
        w[p + "mlp.fc2.bias"] = rn(v.hidden)
    H = dims.hidden
    w["model.mm_projector.0.weight"] = rn(H, v.hidden)
    w["model.mm_projector.0.bias"] = rn(H)
    w["model.mm_projector.2.weight"] = rn(H, H)
    w["model.mm_projector.2.bias"] = rn(H)
    w["model.embed_tokens.weight"] = rn(dims.vocab, H)
    qd, kd = dims.heads * dims.head_dim, dims.kv_heads * dims.head_dim
    for i in range(dims.layers):
        p = f"model.layers.{i}."
        w[p + "input_layernorm.weight"] = ones_noise(H)
        w[p + "self_attn.q_proj.weight"] = rn(qd, H)
        w[p + "self_attn.q_proj.bias"] = rn(qd)
        w[p + "self_attn.k_proj.weight"] = rn(kd, H)
        w[p + "self_attn.k_proj.bias"] = rn(kd)
        w[p + "self_attn.v_proj.weight"] = rn(kd, H)
        w[p + "self_attn.v_proj.bias"] = rn(kd)
        w[p + "self_attn.o_proj.weight"] = rn(H, qd)
        w[p + "post_attention_layernorm.weight"] = ones_noise(H)
        w[p + "mlp.gate_proj.weight"] = rn(dims.inter, H)
        w[p + "mlp.up_proj.weight"] = rn(dims.inter, H)
        w[p + "mlp.down_proj.weight"] = rn(H, dims.inter)
    w["model.norm.weight"] = ones_noise(H)
    w["lm_head.weight"] = w["model.embed_tokens.weight"] if dims.tie_embeddings else rn(dims.vocab, H)
    # score heads: std chosen so the logits have O(head_spread) spread (plain 0.02 init pins every score at 0.5)
    hs = head_spread / math.sqrt(H)
    w["informative_head.weight"] = rn(2, H, std=hs)
    w["relevance_head.weight"] = rn(1, H, std=hs)
    w["uncertainty_head.weight"] = rn(1, H, std=hs)
    return w


def synthetic_frames(n: int, image: int = 384, seed: int = 1, smooth: int = 4) -> torch.Tensor:
    """uint8 [n,3,image,image] frames; low-pass filtered over time so the score stream has structure
    (SURVEY.md 8(d)).  A coarse random field is upsampled so the picture is not pure white noise."""
    g = torch.Generator().manual_seed(seed)
    coarse = torch.rand(n + smooth, 3, 24, 24, generator=g)
    if smooth > 1:
        k = torch.ones(smooth) / smooth
        coarse = torch.stack([(coarse[i:i + smooth] * k[:, None, None, None]).sum(0) for i in range(n)])
    else:
        coarse = coarse[:n]
    up = F.interpolate(coarse, size=(image, image), mode="bilinear", align_corners=False)
    noise = torch.rand(n, 3, image, image, generator=g) * 0.2
    return ((up * 0.8 + noise).clamp(0, 1) * 255).to(torch.uint8)


def synthetic_frames_varied(n: int, image: int = 384, seed: int = 5) -> torch.Tensor:
    """uint8 frames whose brightness and coarse pattern change from frame to frame, so that the per-frame scores
    spread far more than one bf16 step (rank correlation is meaningless on a flat score stream)."""
    g = torch.Generator().manual_seed(seed)
    a = torch.rand(n, 1, 1, 1, generator=g)
    base = torch.rand(n, 3, image, image, generator=g)
    pat = torch.rand(n, 3, 6, 6, generator=g)
    up = F.interpolate(pat, size=(image, image), mode="nearest")
    return ((0.15 + 0.85 * a) * (0.6 * up + 0.4 * base) * 255).clamp(0, 255).to(torch.uint8)


# ----------------------------------------------------------------------------------------------
# A1 preprocess (siglip_encoder.py:47-67; frames are already 384x384 so the bicubic resize is the identity)
# ----------------------------------------------------------------------------------------------
def preprocess_lut() -> np.ndarray:
    """The image processor is a pure function of the uint8 value: f32(f64(u8) * (1/255)), then (x - 0.5) / 0.5
    in float32 (transformers image_transforms.rescale / normalize).  Returns the 256-entry float32 table."""
    u = np.arange(256, dtype=np.uint8)
    x = (u.astype(np.float64) * (1 / 255)).astype(np.float32)
    mean = np.array(0.5, dtype=np.float32)
    std = np.array(0.5, dtype=np.float32)
    return ((x - mean) / std).astype(np.float32)


def preprocess(frames_u8: torch.Tensor) -> torch.Tensor:
    """uint8 [T,3,S,S] -> float32 [T,3,S,S] in [-1, 1]."""
    lut = torch.from_numpy(preprocess_lut()).to(frames_u8.device)
    return lut[frames_u8.long()]


# ----------------------------------------------------------------------------------------------
# A2 SigLIP tower (26 pre-LN blocks, eager attention with fp32 softmax, GELU-tanh MLP, no post_layernorm)
# ----------------------------------------------------------------------------------------------
def siglip_tower(w: Dict[str, torch.Tensor], dims: AhaDims, pixel_values: torch.Tensor) -> torch.Tensor:
    v = dims.vision
    B = pixel_values.shape[0]
    x = F.conv2d(pixel_values, w[VT + "embeddings.patch_embedding.weight"], w[VT + "embeddings.patch_embedding.bias"],
                 stride=v.patch)
    x = x.flatten(2).transpose(1, 2)                                   # [B, patches, C]
    x = x + w[VT + "embeddings.position_embedding.weight"][None]
    hd = v.hidden // v.heads
    scale = hd ** -0.5
    for i in range(v.layers):
        p = VT + f"encoder.layers.{i}."
        res = x
        h = F.layer_norm(x, (v.hidden,), w[p + "layer_norm1.weight"], w[p + "layer_norm1.bias"], v.eps)
        q = F.linear(h, w[p + "self_attn.q_proj.weight"], w[p + "self_attn.q_proj.bias"])
        k = F.linear(h, w[p + "self_attn.k_proj.weight"], w[p + "self_attn.k_proj.bias"])
        vv = F.linear(h, w[p + "self_attn.v_proj.weight"], w[p + "self_attn.v_proj.bias"])
        q = q.view(B, -1, v.heads, hd).transpose(1, 2)
        k = k.view(B, -1, v.heads, hd).transpose(1, 2)
        vv = vv.view(B, -1, v.heads, hd).transpose(1, 2)
        a = torch.matmul(q, k.transpose(2, 3)) * scale
        a = F.softmax(a, dim=-1, dtype=torch.float32).to(q.dtype)
        o = torch.matmul(a, vv).transpose(1, 2).contiguous().reshape(B, -1, v.hidden)
        o = F.linear(o, w[p + "self_attn.out_proj.weight"], w[p + "self_attn.out_proj.bias"])
        x = res + o
        res = x
        h = F.layer_norm(x, (v.hidden,), w[p + "layer_norm2.weight"], w[p + "layer_norm2.bias"], v.eps)
        h = F.linear(h, w[p + "mlp.fc1.weight"], w[p + "mlp.fc1.bias"])
        h = F.gelu(h, approximate="tanh")
        h = F.linear(h, w[p + "mlp.fc2.weight"], w[p + "mlp.fc2.bias"])
        x = res + h
    return x


# ----------------------------------------------------------------------------------------------
# A3 connector, A4 pooling, visual_embed
# ----------------------------------------------------------------------------------------------
def connector(w: Dict[str, torch.Tensor], x: torch.Tensor) -> torch.Tensor:
    h = F.linear(x, w["model.mm_projector.0.weight"], w["model.mm_projector.0.bias"])
    h = F.gelu(h)
    return F.linear(h, w["model.mm_projector.2.weight"], w["model.mm_projector.2.bias"])


def post_projector_pooling(dims: AhaDims, feat: torch.Tensor) -> torch.Tensor:
    side = dims.vision.side
    nf, _, nd = feat.shape
    x = feat.view(nf, side, side, -1).permute(0, 3, 1, 2).contiguous()
    if dims.pool_mode == "average":
        x = F.avg_pool2d(x, dims.pool_stride)
    elif dims.pool_mode == "max":
        x = F.max_pool2d(x, dims.pool_stride)
    elif dims.pool_mode == "bilinear":
        out = [math.ceil(side / dims.pool_stride)] * 2
        x = F.interpolate(x, size=out, mode="bilinear")
    else:
        raise ValueError(f"Unexpected mm_spatial_pool_mode: {dims.pool_mode}")
    return x.permute(0, 2, 3, 1).reshape(nf, -1, nd).contiguous()


def adaptive_pool_alt(feat: torch.Tensor, side: int, pooled: Tuple[int, int] = (7, 7)) -> torch.Tensor:
    """A4b: models/vision_live.py:21-26 (adaptive average pooling of the tower output; dead code in the reference)."""
    nf, _, nd = feat.shape
    x = feat.view(nf, side, side, nd).permute(0, 3, 1, 2)
    x = F.adaptive_avg_pool2d(x, pooled)
    return x.flatten(2).permute(0, 2, 1).contiguous()


def visual_embed(w: Dict[str, torch.Tensor], dims: AhaDims, pixel_values: torch.Tensor) -> torch.Tensor:
    """[F,3,S,S] (model dtype) -> [F * frame_num_tokens, H] (models/modeling_live.py:31-37)."""
    x = siglip_tower(w, dims, pixel_values)
    x = connector(w, x)
    x = post_projector_pooling(dims, x)
    return x.view(-1, x.shape[-1])


# ----------------------------------------------------------------------------------------------
# A6-A8 KV-cache policies.  One class, three policies, written index-first: every cached token carries
# its global token id (`ids`), so tests can check eviction provenance bit-exactly.
# ----------------------------------------------------------------------------------------------
def rotate_half(x: torch.Tensor) -> torch.Tensor:
    h = x.shape[-1] // 2
    return torch.cat((-x[..., h:], x[..., :h]), dim=-1)


class OracleKVCache:
    """policy: 'sink' (test/sink_cache.py), 'sliding' (test/sliding_window_cache.py), 'static' (test/static_cache.py).

    update(k, v, layer, cos, sin) takes k/v of shape [1, kvH, q, d] (k already rotated at positions [L, L+q)) and
    cos/sin of shape [1, q, d]; returns the keys/values the attention sees, like the HF Cache API.
    """

    def __init__(self, policy: str, window: int, sinks: int = 0):
        assert policy in ("sink", "sliding", "static")
        self.policy, self.W, self.S = policy, int(window), int(sinks)
        self.keys: List[torch.Tensor] = []
        self.vals: List[torch.Tensor] = []
        self.ids: List[torch.Tensor] = []          # provenance: global token id of every cached slot, per layer
        self.fed = 0                                # tokens offered to layer 0 so far (id generator)
        self.seen_tokens = 0                        # SinkCache._seen_tokens (sink_cache.py:105-106); others never bump it
        self.cos_table: Optional[torch.Tensor] = None   # SinkCache._cos_cache (sink_cache.py:109-121)
        self.sin_table: Optional[torch.Tensor] = None
        self.rerot: Dict[int, Tuple[torch.Tensor, torch.Tensor]] = {}

    def get_seq_length(self, layer: int = 0) -> int:
        return 0 if len(self.keys) <= layer else self.keys[layer].shape[-2]

    def get_max_length(self) -> int:
        return self.W

    def _rerotation(self, q: int, dtype) -> Tuple[torch.Tensor, torch.Tensor]:
        # sink_cache.py:35-55: rotate a key forward-encoded at position p+q back to position p, for p = S..W-q-1
        if q not in self.rerot:
            c = self.cos_table[: self.W].float()
            s = self.sin_table[: self.W].float()
            oc, sc = c[self.S + q:], c[self.S: -q]
            os_, ss = s[self.S + q:], s[self.S: -q]
            rc = oc * sc + os_ * ss
            rs = -os_ * sc + oc * ss
            self.rerot[q] = (rc.to(dtype)[None], rs.to(dtype)[None])
        return self.rerot[q]

    def update(self, k: torch.Tensor, v: torch.Tensor, layer: int, cos: Optional[torch.Tensor] = None,
               sin: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        q = k.shape[-2]
        if layer == 0:
            self._new_ids = torch.arange(self.fed, self.fed + q, dtype=torch.int64)
            self.fed += q
        new_ids = self._new_ids
        if self.policy == "static":
            if len(self.keys) <= layer:
                self.keys.append(k[:, :, : self.W]); self.vals.append(v[:, :, : self.W]); self.ids.append(new_ids[: self.W])
            return self.keys[layer], self.vals[layer]
        if self.policy == "sliding":
            if len(self.keys) <= layer:
                self.keys.append(k); self.vals.append(v); self.ids.append(new_ids)
                return k, v
            self.keys[layer] = torch.cat([self.keys[layer], k], dim=-2)[:, :, -self.W:]
            self.vals[layer] = torch.cat([self.vals[layer], v], dim=-2)[:, :, -self.W:]
            self.ids[layer] = torch.cat([self.ids[layer], new_ids])[-self.W:]
            return self.keys[layer], self.vals[layer]
        # ---- sink ----
        rope = cos is not None and sin is not None
        if layer == 0:
            self.seen_tokens += q
            if rope:
                if self.cos_table is None:
                    self.cos_table, self.sin_table = cos[0], sin[0]
                elif self.cos_table.shape[0] < self.W:
                    self.cos_table = torch.cat([self.cos_table, cos[0]], dim=0)
                    self.sin_table = torch.cat([self.sin_table, sin[0]], dim=0)
        if len(self.keys) <= layer:
            self.keys.append(k); self.vals.append(v); self.ids.append(new_ids)
        elif q + self.get_seq_length(layer) < self.W:
            self.keys[layer] = torch.cat([self.keys[layer], k], dim=-2)
            self.vals[layer] = torch.cat([self.vals[layer], v], dim=-2)
            self.ids[layer] = torch.cat([self.ids[layer], new_ids])
        else:
            keep = self.W - self.S - q
            if keep <= 0:
                raise ValueError("chunk does not fit behind the sinks: q >= window_length - num_sink_tokens")
            old_k, old_v, old_i = self.keys[layer], self.vals[layer], self.ids[layer]
            kept = old_k[:, :, -keep:]
            if rope:
                rc, rs = self._rerotation(q, k.dtype)
                kept = (kept * rc) + (rotate_half(kept) * rs)
            self.keys[layer] = torch.cat([old_k[:, :, : self.S], kept, k], dim=-2)
            self.vals[layer] = torch.cat([old_v[:, :, : self.S], old_v[:, :, -keep:], v], dim=-2)
            self.ids[layer] = torch.cat([old_i[: self.S], old_i[-keep:], new_ids])
        return self.keys[layer], self.vals[layer]


# ----------------------------------------------------------------------------------------------
# A9 Qwen2 decoder prefill (transformers 4.49 semantics) + A10 lm_head + A11 heads
# ----------------------------------------------------------------------------------------------
def rope_cos_sin(positions: torch.Tensor, head_dim: int, theta: float, dtype) -> Tuple[torch.Tensor, torch.Tensor]:
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.int64).float() / head_dim))
    inv_freq = inv_freq.to(positions.device)         # (computed on the host either way: same fp32 bits on any device)
    freqs = positions.float()[:, None] * inv_freq[None, :]
    emb = torch.cat((freqs, freqs), dim=-1)
    return emb.cos().to(dtype)[None], emb.sin().to(dtype)[None]      # [1, q, d]


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    dt = x.dtype
    h = x.float()
    h = h * torch.rsqrt(h.pow(2).mean(-1, keepdim=True) + eps)
    return weight * h.to(dt)


@dataclass
class PrefillOut:
    hidden: torch.Tensor                 # [1, q, H] after the final norm
    logits: Optional[torch.Tensor]       # [1, q, V] fp32 (None when skipped)
    informative_logits: torch.Tensor     # [1, q, 2] fp32
    relevance_logits: torch.Tensor       # [1, q, 1] fp32, post-sigmoid
    uncertainty: torch.Tensor            # [1, q, 1] fp32 (log-variance)
    positions: torch.Tensor              # [q] int64 positions the chunk was rotated at
    klen: int                            # keys seen by layer-0 attention


def qwen2_prefill(w: Dict[str, torch.Tensor], dims: AhaDims, cache: OracleKVCache, inputs_embeds: torch.Tensor,
                  mask_mode: str = "sdpa", want_logits: bool = True, attn_impl: str = "eager",
                  lora: Optional[Dict[str, Tuple[torch.Tensor, torch.Tensor, float]]] = None) -> PrefillOut:
    """One chunk [1,q,H] against the cache.  mask_mode 'sdpa': allowed(i,j) <=> j <= L+i (mask sliced to Klen
    columns); 'fa2': bottom-right aligned causal j <= i + (Klen - q).
    `lora`: {module name: (A [r, in], B [out, r], alpha / r)} - UNMERGED adapters as the reference runs them through peft
    (models/modeling_live.py:153-173): y = base(x) + lora_B(lora_A(x)) * scaling, every op in the model dtype."""
    if lora:
        _linear = F.linear

        def lin(x_, weight, bias=None):
            y = _linear(x_, weight, bias)
            for name, (A_, B_, sc) in lora.items():
                if w.get(name + ".weight") is weight:
                    return y + _linear(_linear(x_, A_), B_) * sc
            return y
    else:
        lin = F.linear
    x = inputs_embeds
    dt = x.dtype
    q = x.shape[1]
    L = cache.get_seq_length()
    pos = torch.arange(L, L + q, dtype=torch.int64, device=x.device)
    cos, sin = rope_cos_sin(pos, dims.head_dim, dims.rope_theta, dt)
    nh, nkv, d = dims.heads, dims.kv_heads, dims.head_dim
    klen0 = None
    for li in range(dims.layers):
        p = f"model.layers.{li}."
        res = x
        h = rms_norm(x, w[p + "input_layernorm.weight"], dims.rms_eps)
        qs = lin(h, w[p + "self_attn.q_proj.weight"], w[p + "self_attn.q_proj.bias"]).view(1, q, nh, d).transpose(1, 2)
        ks = lin(h, w[p + "self_attn.k_proj.weight"], w[p + "self_attn.k_proj.bias"]).view(1, q, nkv, d).transpose(1, 2)
        vs = lin(h, w[p + "self_attn.v_proj.weight"], w[p + "self_attn.v_proj.bias"]).view(1, q, nkv, d).transpose(1, 2)
        c, s = cos[:, None], sin[:, None]
        qs = (qs * c) + (rotate_half(qs) * s)
        ks = (ks * c) + (rotate_half(ks) * s)
        K, V = cache.update(ks, vs, li, cos, sin)
        klen = K.shape[-2]
        if klen0 is None:
            klen0 = klen
        off = L if mask_mode == "sdpa" else klen - q
        jj = torch.arange(klen, device=x.device)[None, :]
        ii = torch.arange(q, device=x.device)[:, None]
        allowed = jj <= (off + ii)
        Kr = K.repeat_interleave(nh // nkv, dim=1)
        Vr = V.repeat_interleave(nh // nkv, dim=1)
        if attn_impl == "sdpa":
            o = F.scaled_dot_product_attention(qs, Kr, Vr, attn_mask=allowed[None, None])
        else:
            a = torch.matmul(qs, Kr.transpose(2, 3)) * (d ** -0.5)
            a = a.masked_fill(~allowed[None, None], torch.finfo(a.dtype).min)
            a = F.softmax(a, dim=-1, dtype=torch.float32).to(dt)
            o = torch.matmul(a, Vr)
        o = o.transpose(1, 2).reshape(1, q, nh * d)
        x = res + lin(o, w[p + "self_attn.o_proj.weight"])
        res = x
        h = rms_norm(x, w[p + "post_attention_layernorm.weight"], dims.rms_eps)
        h = lin(F.silu(lin(h, w[p + "mlp.gate_proj.weight"])) * lin(h, w[p + "mlp.up_proj.weight"]),
                     w[p + "mlp.down_proj.weight"])
        x = res + h
    x = rms_norm(x, w["model.norm.weight"], dims.rms_eps)
    logits = F.linear(x, w["lm_head.weight"]).float() if want_logits else None
    inf = F.linear(x, w["informative_head.weight"]).float()
    rel = torch.sigmoid(F.linear(x, w["relevance_head.weight"]).float())
    unc = F.linear(x, w["uncertainty_head.weight"]).float()
    return PrefillOut(x, logits, inf, rel, unc, pos, klen0)


def frame_scores(out: PrefillOut) -> Tuple[float, float, float]:
    """test/inference.py:222-229: last token of the chunk only."""
    informative = out.informative_logits[0, -1].softmax(dim=-1)[1].item()
    relevance = out.relevance_logits[0, -1].item()
    uncertainty = torch.exp(out.uncertainty[0, -1]).item()
    return informative, relevance, uncertainty


# ----------------------------------------------------------------------------------------------
# A5 + A12: the stream loop (test/inference.py:112-131, 133-155, 169-229, 283-334), ingest-and-score only
# ----------------------------------------------------------------------------------------------
def make_cache(alt_cache: str = "default_sink", window_length: int = 2048, num_sink_tokens: int = 32,
               sink_cache: bool = False, n_sys: int = 0) -> OracleKVCache:
    """test/inference.py:133-155 (_init_cache)."""
    if sink_cache:
        return OracleKVCache("sink", window_length + num_sink_tokens - n_sys, n_sys)
    if alt_cache == "default_sink":
        return OracleKVCache("sink", window_length, num_sink_tokens)
    if alt_cache == "sliding_window":
        return OracleKVCache("sliding", window_length)
    if alt_cache == "static":
        return OracleKVCache("static", window_length)
    raise ValueError(alt_cache)


def synthetic_prompt_ids(n: int, vocab: int, seed: int) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, vocab, (1, n), generator=g, dtype=torch.int64)


class LiveInferOracle:
    """Mirror of LiveInferForBenchmark restricted to ingest-and-score (response generation never triggers:
    stream_end_prob_threshold=None, stream_end_score_sum_threshold=inf as in SURVEY.md 8(d))."""

    def __init__(self, w, dims: AhaDims, dtype=torch.bfloat16, alt_cache="default_sink", window_length=2048,
                 num_sink_tokens=32, sink_cache=False, n_sys=36, n_query=24, frame_fps=1.0, mask_mode="sdpa",
                 score_heads=("relevance_score", "informative_score"), want_logits=False, prompt_seed=11):
        self.w, self.dims, self.dtype = w, dims, dtype
        self.cache_args = dict(alt_cache=alt_cache, window_length=window_length, num_sink_tokens=num_sink_tokens,
                               sink_cache=sink_cache, n_sys=n_sys)
        self.device = w["model.embed_tokens.weight"].device      # the oracle runs where its weights live
        self.start_ids = synthetic_prompt_ids(n_sys, dims.vocab, prompt_seed).to(self.device)
        self.query_ids = synthetic_prompt_ids(n_query, dims.vocab, prompt_seed + 1).to(self.device) if n_query > 0 else None
        self.frame_fps = frame_fps
        self.mask_mode = mask_mode
        self.score_heads = list(score_heads)
        self.want_logits = want_logits
        self.reset()

    def reset(self):
        self.cache = make_cache(**self.cache_args)
        self.frame_embeds_queue: List[Tuple[float, torch.Tensor]] = []
        self.query_queue: List[Tuple[float, torch.Tensor]] = []
        self.video_time = 0.0
        self.init_vision_time = False
        self.debug_data_list: List[dict] = []
        self.trace: List[dict] = []           # per-chunk integer trace (positions, Klen, cache ids)
        self.hidden_last: List[torch.Tensor] = []   # final-norm hidden state of the last token of every frame chunk (fp32)
        self.stream_end_score_sum = 0.0

    def input_video_stream(self, frames_u8: torch.Tensor, batch_size: int = 32):
        pv = preprocess(frames_u8.to(self.device)).to(self.dtype)
        nt = self.dims.frame_num_tokens
        for b in range(0, len(pv), batch_size):
            emb = visual_embed(self.w, self.dims, pv[b:b + batch_size]).split(nt)
            for r, f in enumerate(emb):
                self.frame_embeds_queue.append(((r + b) / self.frame_fps, f))

    def input_query_stream(self, time: float = 0.0):
        if self.query_ids is not None:
            self.query_queue.append((time, self.query_ids))

    def _embed(self, ids: torch.Tensor) -> torch.Tensor:
        return F.embedding(ids, self.w["model.embed_tokens.weight"])

    def _record(self, kind: str, out: PrefillOut):
        self.trace.append(dict(kind=kind, q=int(out.positions.numel()), pos0=int(out.positions[0]), klen=int(out.klen),
                               seq_len=self.cache.get_seq_length(), ids=self.cache.ids[0].clone()))

    def _encode_query(self):
        _, ids = self.query_queue.pop(0)
        out = qwen2_prefill(self.w, self.dims, self.cache, self._embed(ids), self.mask_mode, want_logits=True)
        self._record("query", out)
        self.last_query_logits = out.logits[:, -1]

    def _encode_frame(self):
        _, emb = self.frame_embeds_queue.pop(0)
        if not self.init_vision_time:
            last_ids = self.start_ids
            self.init_vision_time = True
        else:
            last_ids = torch.zeros(1, 0, dtype=torch.int64, device=self.device)
        x = torch.cat([self._embed(last_ids).view(1, -1, self.dims.hidden), emb.view(1, -1, self.dims.hidden)], dim=1)
        out = qwen2_prefill(self.w, self.dims, self.cache, x, self.mask_mode, want_logits=self.want_logits)
        self._record("frame", out)
        self.last_out = out
        self.hidden_last.append(out.hidden[0, -1].float().cpu())
        return frame_scores(out)

    def inference(self):
        while self.frame_embeds_queue:
            if self.query_queue and self.video_time >= self.query_queue[0][0]:
                self._encode_query()
            inf, rel, unc = self._encode_frame()
            self.debug_data_list.append(dict(time=self.video_time, informative_score=inf, relevance_score=rel,
                                             uncertainty_score=unc))
            scores = dict(informative_score=inf, relevance_score=rel)
            self.stream_end_score_sum += sum(v for k, v in scores.items() if k in self.score_heads)
            self.video_time += 1 / self.frame_fps
        return self.debug_data_list


def aligned_heads(w: Dict[str, torch.Tensor], hidden_last: List[torch.Tensor], logit_std=(1.0, 1.0, 0.3),
                  skip: int = 2) -> Dict[str, torch.Tensor]:
    """TEST CONSTRUCTION, not part of the reference: score heads with a trained model's spread on random-init weights.
    With N(0, 0.02) heads the per-frame scores of a random-init model are nearly flat: the head direction is random, so it
    picks up 1/H of the frame-to-frame variance of the final hidden state while bf16 rounding noise is isotropic, and
    rank correlation measures the noise.  A trained head points where the frames differ; this returns a copy of `w` whose
    three heads lie along the first three principal directions of the given final hidden states (one per frame, e.g. of
    an fp32 oracle run), scaled so the head outputs have the given standard deviations over those frames.  Everything
    else - and the arithmetic every implementation runs - is unchanged; all implementations under test get the SAME
    weights.  The first `skip` frames (the chunk that carries the system prompt is an outlier) are left out of the fit."""
    Hm = torch.stack([h.float().cpu() for h in hidden_last[skip:]])
    Hc = Hm - Hm.mean(0, keepdim=True)
    _, _, Vt = torch.linalg.svd(Hc, full_matrices=False)
    out = dict(w)
    dt, dev = w["relevance_head.weight"].dtype, w["relevance_head.weight"].device
    dirs = []
    for k in range(3):
        v = Vt[k]
        dirs.append(v * (logit_std[k] / max(float((Hc @ v).std()), 1e-12)))
    # informative = softmax over two logits: +-v/2 gives logit difference v.h
    out["informative_head.weight"] = torch.stack([-0.5 * dirs[0], 0.5 * dirs[0]]).to(dt).to(dev)
    out["relevance_head.weight"] = dirs[1][None].to(dt).to(dev)
    out["uncertainty_head.weight"] = dirs[2][None].to(dt).to(dev)
    return out


# ----------------------------------------------------------------------------------------------
# A13 smoothing + peak picking
# ----------------------------------------------------------------------------------------------
def find_ticks(scores, fps: float = 1.0, prominence: float = 0.02, thresh=False) -> Tuple[List[float], np.ndarray, float]:
    """test/live_infer_for_video.py:195-228 (min_separation is hard-reset to 10 there, :213-214)."""
    from scipy.signal import find_peaks, savgol_filter
    arr = np.array(scores)
    smoothed = savgol_filter(arr, 15, 3)
    if not thresh:
        thresh = smoothed.mean() + 0.5 * smoothed.std()
    distance = int(10 * fps)
    peaks, _ = find_peaks(smoothed, height=thresh, prominence=prominence, distance=distance)
    return [float(p / fps) for p in peaks], smoothed, float(thresh)


# scipy.signal.savgol_coeffs(15, 3) as exact doubles (scipy 1.18.1 on this image; checked in tests/test_oracle.py)
SAVGOL_15_3 = np.array([float.fromhex(h) for h in (
    '-0x1.2121212121230p-4', '-0x1.818181818180fp-7', '0x1.375ec1375ec3cp-5', '0x1.427d91427d939p-4',
    '0x1.c43a61c43a64ep-4', '0x1.1072e91072eafp-3', '0x1.2c3ff12c3ff34p-3', '0x1.35844935844b6p-3',
    '0x1.2c3ff12c3ff36p-3', '0x1.1072e91072eb1p-3', '0x1.c43a61c43a653p-4', '0x1.427d91427d93cp-4',
    '0x1.375ec1375ec3cp-5', '-0x1.818181818183bp-7', '-0x1.2121212121233p-4')])


def find_ticks_restated(scores, fps: float = 1.0, prominence: float = 0.02, thresh=None):
    """scipy-free restatement of the same pass; used to pin the CUDA kernel's
    algorithm against scipy itself.  Returns (peak indices, smoothed, thresh)."""
    x = np.asarray(scores, dtype=np.float64)
    n = x.shape[0]
    if n < 15:
        raise ValueError("If mode is 'interp', window_length must be less than or equal to the size of x.")
    half = 7
    # interior: scipy.ndimage.correlate1d's symmetric-kernel order, tmp = x[i]*f[7] + sum_{l=-7..-1} (x[i+l]+x[i-l])*f[l+7]
    # with f = savgol_coeffs(15,3)[::-1]; reproducing that order makes the interior bit-identical to scipy.
    f = SAVGOL_15_3[::-1]
    y = np.empty(n)
    for i in range(half, n - half):
        tmp = x[i] * f[half]
        for l in range(-half, 0):
            tmp += (x[i + l] + x[i - l]) * f[l + half]
        y[i] = tmp
    V15 = np.vander(np.arange(15, dtype=np.float64), 4, increasing=True)
    P = np.linalg.pinv(V15)
    E_head = V15[:half] @ P
    E_tail = V15[15 - half:] @ P
    y[:half] = E_head @ x[:15]
    y[n - half:] = E_tail @ x[n - 15:]
    if thresh is None:
        thresh = y.mean() + 0.5 * y.std()
    # local maxima with plateau midpoints
    peaks = []
    i = 1
    while i < n - 1:
        if y[i - 1] < y[i]:
            a = i + 1
            while a < n - 1 and y[a] == y[i]:
                a += 1
            if y[a] < y[i]:
                peaks.append((i + a - 1) // 2)
                i = a
        i += 1
    peaks = [p for p in peaks if y[p] >= thresh]
    # distance: greedy by descending height
    D = max(1, math.ceil(int(10 * fps)))
    keep = [True] * len(peaks)
    order = np.argsort(np.array([y[p] for p in peaks]), kind="stable") if peaks else []
    for j in reversed(list(order)):
        if not keep[j]:
            continue
        k = j - 1
        while k >= 0 and peaks[j] - peaks[k] < D:
            keep[k] = False
            k -= 1
        k = j + 1
        while k < len(peaks) and peaks[k] - peaks[j] < D:
            keep[k] = False
            k += 1
    peaks = [p for p, kp in zip(peaks, keep) if kp]
    out = []
    for p in peaks:
        lm = y[p]
        i = p
        while i >= 0 and y[i] <= y[p]:
            lm = min(lm, y[i]); i -= 1
        rm = y[p]
        i = p
        while i < n and y[i] <= y[p]:
            rm = min(rm, y[i]); i += 1
        if y[p] - max(lm, rm) >= prominence:
            out.append(p)
    return np.array(out, dtype=np.int64), y, float(thresh)


# ----------------------------------------------------------------------------------------------
# N3: score fusion and highlight selection (pure-Python restatement; test infrastructure only)
# ----------------------------------------------------------------------------------------------
def fuse_scores(debug_data, alpha, beta, epsilon, uncertainty_threshold):
    """test/grid_search.py:58-65 (hisum) and :95-104 (tvsum): one fused score per frame."""
    out = []
    for e in debug_data:
        s = alpha * e["informative_score"] + beta * e["relevance_score"]
        if e["uncertainty_score"] >= uncertainty_threshold:
            s -= (e["uncertainty_score"] - uncertainty_threshold) * epsilon
        out.append(s)
    return out


def knapsack_selection(frames_with_index, max_duration, alpha, beta, epsilon):
    """test/highlight_generator.py:8-37: 0/1 knapsack with unit costs; keeps the reference's DP recurrence, its
    float comparisons and its backtracking rule, so ties resolve the same way."""
    n = len(frames_with_index)
    prev = [0] * (max_duration + 1)
    changed = []                                   # changed[i-1][j] <=> dp[i][j] != dp[i-1][j]
    for i in range(1, n + 1):
        f = frames_with_index[i - 1]
        value = f["informative_score"] * alpha + f["relevance_score"] * beta + f["uncertainty_score"] * epsilon
        cur = list(prev)
        for j in range(1, max_duration + 1):
            cur[j] = max(prev[j], prev[j - 1] + value)
        changed.append([cur[j] != prev[j] for j in range(max_duration + 1)])
        prev = cur
    selected, rc = [], max_duration
    for i in range(n, 0, -1):
        if changed[i - 1][rc]:
            selected.append(frames_with_index[i - 1]["idx"])
            rc -= 1
    return set(selected)


def highlight_indices(selected, ground_truth_frames, fps):
    """test/highlight_generator.py:76-86."""
    last_frame = ground_truth_frames[-1]
    half_width = int(fps // 2)
    out = set()
    for idx in selected:
        g = ground_truth_frames[idx]
        for i in range(max(0, g - half_width), min(last_frame + 1, g + half_width + 1)):
            out.add(i)
    return sorted(out)


# ----------------------------------------------------------------------------------------------
# N1: greedy response decode (restatement; test infrastructure only)
# ----------------------------------------------------------------------------------------------
def repetition_penalty_logits(scores: torch.Tensor, input_ids: torch.Tensor, penalty: float) -> torch.Tensor:
    """transformers RepetitionPenaltyLogitsProcessor.__call__ (logits_process.py, 4.49): scores [1,V], input_ids [1,n]."""
    score = torch.gather(scores, 1, input_ids)
    score = torch.where(score < 0, score * penalty, score / penalty)
    return scores.scatter(1, input_ids, score)


def greedy_step_logits(w, dims, cache, inputs_embeds, mask_mode="sdpa") -> torch.Tensor:
    """One decode step of fast_greedy_generate (models/modeling_live.py:72-73): last-position logits [1, V] float32."""
    out = qwen2_prefill(w, dims, cache, inputs_embeds, mask_mode, want_logits=True)
    return out.logits[:, -1, :]


def fast_greedy_generate(w, dims, cache, inputs_embeds, eos_token_id, max_new_tokens, repetition_penalty=None,
                         generated_token_ids=None, mask_mode="sdpa", forced_ids=None):
    """models/modeling_live.py:64-90.  Returns (output ids, generated_token_ids, per-step processed logits).  With
    `forced_ids` the chosen token of every step is overridden (teacher forcing, for tolerance-aware parity checks)."""
    generated_token_ids = [] if generated_token_ids is None else generated_token_ids
    out_ids, step_logits = [], []
    for i in range(max_new_tokens):
        logits = greedy_step_logits(w, dims, cache, inputs_embeds, mask_mode)
        if repetition_penalty is not None and len(generated_token_ids) > 0:
            logits = repetition_penalty_logits(logits, torch.tensor(generated_token_ids).unsqueeze(0), repetition_penalty)
        step_logits.append(logits[0].clone())
        tok = int(logits.argmax(dim=-1)) if forced_ids is None else int(forced_ids[i])
        if repetition_penalty is not None and tok != eos_token_id:
            generated_token_ids.append(tok)
        out_ids.append(tok)
        if tok == eos_token_id or (forced_ids is not None and i + 1 >= len(forced_ids)):
            break
        inputs_embeds = F.embedding(torch.tensor([[tok]]), w["model.embed_tokens.weight"])
    return out_ids, generated_token_ids, step_logits
