"""Host-side mirror of the reference model surface for the ingest-and-score path.

    VideoHeadLiveLlavaQwenForCausalLM.forward      models/live_llava/video_head_live_llava_qwen.py:138-188, 317-330
    LiveMixin.visual_embed                         models/modeling_live.py:31-37
    VideoHeadCausalLMOutputWithPast                models/live_llava/video_head_live_llava_qwen.py:50-61

All arithmetic runs in libaha_b200.so (aha_vision_embed / aha_prefill_chunk); PyTorch only owns the device buffers.
Training (labels / losses, :190-308) is out of scope and raises NotImplementedError.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Union

import torch

from . import _lib
from .caches import _NativeCache
from .config import AhaConfig


@dataclass
class VideoHeadCausalLMOutputWithPast:
    loss: Optional[torch.Tensor] = None
    lm_loss: Optional[torch.Tensor] = None
    video_loss: Optional[torch.Tensor] = None
    informative_logits: Optional[torch.Tensor] = None   # [B, q, 2] fp32
    relevance_logits: Optional[torch.Tensor] = None     # [B, q, 1] fp32, post-sigmoid
    uncertainty: Optional[torch.Tensor] = None          # [B, q, 1] fp32, log-variance
    logits: Optional[torch.Tensor] = None               # [B, q, V] fp32 (or [B, 1, V] in last-token mode)
    past_key_values: Optional[object] = None
    hidden_states: Optional[torch.Tensor] = None
    attentions: Optional[torch.Tensor] = None
    frame_scores: Optional[torch.Tensor] = None         # [B, 3] fp32 device: (informative, relevance, uncertainty) of the last token

    def __getitem__(self, k):
        return getattr(self, k)


class _InputEmbeddings:
    """`model.get_input_embeddings()(ids)` (test/inference.py:211): gathers rows of embed_tokens on the device."""

    def __init__(self, model: "VideoHeadLiveLlavaQwenForCausalLM"):
        self._m = model

    def __call__(self, ids: torch.Tensor) -> torch.Tensor:
        shape = tuple(ids.shape)
        flat = ids.reshape(-1).to(device=self._m.device, dtype=torch.int64).contiguous()
        if flat.numel() and (int(flat.min()) < 0 or int(flat.max()) >= self._m.config.vocab_size):
            raise IndexError("index out of range in self")
        out = torch.empty(flat.numel(), self._m.config.hidden_size, dtype=torch.bfloat16, device=self._m.device)
        if flat.numel():
            _lib.check(_lib.load().aha_assemble_chunk(self._m._h, _lib.ptr(flat), flat.numel(), None, 0, _lib.ptr(out),
                                                      _lib.stream_ptr()))
        return out.view(*shape, self._m.config.hidden_size)


class VideoHeadLiveLlavaQwenForCausalLM:
    """Inference-only, B200-native stand-in for the reference class of the same name."""

    def __init__(self, config: AhaConfig, state_dict: Dict[str, torch.Tensor], device: Union[int, str, torch.device] = "cuda",
                 max_tokens: int = 256, max_frames: int = 32, logits_mode: str = "all", mask_mode: str = "sdpa"):
        self.config = config
        self.device = torch.device(device if not isinstance(device, int) else f"cuda:{device}")
        if self.device.type != "cuda":
            raise RuntimeError("aha_b200 has no CPU path: a CUDA device is required")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.logits_mode = {"none": _lib.LOGITS_NONE, "last": _lib.LOGITS_LAST, "all": _lib.LOGITS_ALL}[logits_mode]
        self.mask_mode = {"sdpa": _lib.MASK_SDPA, "eager": _lib.MASK_SDPA, "flash_attention_2": _lib.MASK_FA2,
                          "fa2": _lib.MASK_FA2}[mask_mode]
        self.currently_training = False
        self.max_tokens, self.max_frames = int(max_tokens), int(max_frames)
        v = config.vision
        desc = _lib.ModelDesc(
            hidden=config.hidden_size, inter=config.intermediate_size, layers=config.num_hidden_layers,
            heads=config.num_attention_heads, kv_heads=config.num_key_value_heads, head_dim=config.head_dim,
            vocab=config.vocab_size, rope_theta=config.rope_theta, rms_eps=config.rms_norm_eps,
            v_hidden=v.hidden_size, v_inter=v.intermediate_size, v_layers=v.num_hidden_layers, v_heads=v.num_attention_heads,
            v_patch=v.patch_size, v_image=v.image_size, v_eps=v.layer_norm_eps,
            pool_stride=config.video_pooling_stride, pool_mode=_lib.POOL_MODES[config.mm_spatial_pool_mode],
            max_tokens=self.max_tokens, max_frames=self.max_frames)
        keep = []
        entries = (_lib.WeightEntry * len(state_dict))()
        for i, (name, t) in enumerate(state_dict.items()):
            t = t.to(device=self.device, dtype=torch.bfloat16).contiguous()
            keep.append(t)
            entries[i].name = name.encode()
            entries[i].ptr = t.data_ptr()
        # tied lm_head: give it the embed_tokens pointer so the handle shares storage
        if config.tie_word_embeddings and "lm_head.weight" in state_dict and "model.embed_tokens.weight" in state_dict:
            names = list(state_dict.keys())
            entries[names.index("lm_head.weight")].ptr = entries[names.index("model.embed_tokens.weight")].ptr
        h = ctypes.c_void_p()
        torch.cuda.synchronize(self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().aha_model_create(ctypes.byref(desc), entries, len(state_dict), self.device.index,
                                                    ctypes.byref(h)))
        self._h = h
        del keep                                   # the handle owns packed copies

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h is not None:
            try:
                _lib.load().aha_model_destroy(h)
            except Exception:
                pass

    # ---- small pieces of the nn.Module surface the drivers touch ----
    def eval(self):
        return self

    def to(self, *a, **k):
        return self

    def get_input_embeddings(self):
        return _InputEmbeddings(self)

    def get_vision_tower(self):
        """`model.get_vision_tower().image_processor` (test/inference.py:51): the SigLIP image processor of the reference,
        for callers that run it themselves; `visual_embed` also takes the raw uint8 frames and fuses it in."""
        if getattr(self, "_vision_tower", None) is None:
            from .compat import _VisionTower
            self._vision_tower = _VisionTower(self.config.vision.image_size)
        return self._vision_tower

    dtype = torch.bfloat16

    def new_cache(self, alt_cache="default_sink", window_length=2048, num_sink_tokens=32, sink_cache=False, n_sys=0,
                  max_chunk=None):
        """_init_cache (test/inference.py:133-155) with the slabs allocated up front for this model's geometry."""
        from .caches import make_cache
        c = make_cache(alt_cache, window_length, num_sink_tokens, sink_cache, n_sys)
        cfg = self.config
        return c.allocate(cfg.num_key_value_heads, cfg.head_dim, cfg.num_hidden_layers,
                          max_chunk or min(self.max_tokens, 4096), self.device)

    # ---- vision ----
    def vision_stream(self) -> "torch.cuda.Stream":
        """Second CUDA stream for the SigLIP tower: tensor-bound vision batches of later frames run underneath the
        HBM-bound decoder steps of earlier ones.  (The vision and decoder workspaces of the handle are disjoint and the
        library's split-K scratch is per stream, so the two streams never share mutable state.)"""
        if getattr(self, "_vision_stream", None) is None:
            self._vision_stream = torch.cuda.Stream(device=self.device)
        return self._vision_stream

    def decoder_stream(self) -> "torch.cuda.Stream":
        """High-priority stream for the decoder steps while vision batches are in flight on `vision_stream()`: without a
        priority the block scheduler drains a 5000-CTA vision GEMM before it dispatches the decoder's next (small) grid,
        which serialises the two streams again."""
        if getattr(self, "_decoder_stream", None) is None:
            self._decoder_stream = torch.cuda.Stream(device=self.device, priority=-1)
        return self._decoder_stream

    def visual_embed(self, frames: torch.Tensor) -> torch.Tensor:
        """[F, 3, S, S] uint8 (raw frames; the image processor's rescale/normalise is fused in) or bf16 pixel_values
        -> [F * frame_num_tokens, hidden] bf16 (models/modeling_live.py:31-37)."""
        if frames.dim() != 4 or frames.shape[1] != 3 or frames.shape[2] != self.config.vision.image_size \
                or frames.shape[3] != self.config.vision.image_size:
            raise ValueError(f"frames must be [F, 3, {self.config.vision.image_size}, {self.config.vision.image_size}]")
        frames = frames.to(self.device).contiguous()
        n = frames.shape[0]
        out = torch.empty(n * self.config.frame_num_tokens, self.config.hidden_size, dtype=torch.bfloat16, device=self.device)
        if frames.dtype == torch.uint8:
            u8, px = _lib.ptr(frames), None
        elif frames.dtype == torch.bfloat16:
            u8, px = None, _lib.ptr(frames)
        else:
            raise TypeError("frames must be uint8 (raw) or bfloat16 (pixel_values)")
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().aha_vision_embed(self._h, u8, px, n, _lib.ptr(out), _lib.stream_ptr()))
        return out

    def assemble_chunk(self, ids: Optional[torch.Tensor], frame_embeds: Optional[torch.Tensor]) -> torch.Tensor:
        """torch.cat([embed_tokens(ids), frame_embeds]) of _encode_frame (test/inference.py:204-214) -> [1, q, H]."""
        n_ids = 0 if ids is None else int(ids.numel())
        n_f = 0 if frame_embeds is None else int(frame_embeds.shape[-2])
        out = torch.empty(1, n_ids + n_f, self.config.hidden_size, dtype=torch.bfloat16, device=self.device)
        idt = None if n_ids == 0 else ids.reshape(-1).to(device=self.device, dtype=torch.int64).contiguous()
        if n_ids and (int(idt.min()) < 0 or int(idt.max()) >= self.config.vocab_size):
            raise IndexError("index out of range in self")      # what torch's embedding raises (the kernel does not check)
        fe = None if n_f == 0 else frame_embeds.to(self.device).contiguous()
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().aha_assemble_chunk(self._h, _lib.ptr(idt), n_ids, _lib.ptr(fe), n_f, _lib.ptr(out),
                                                      _lib.stream_ptr()))
        return out

    # ---- the per-frame step ----
    def forward(self, input_ids=None, attention_mask=None, position_ids=None, past_key_values=None, inputs_embeds=None,
                labels=None, informative_labels=None, relevance_labels=None, use_cache=None, output_attentions=None,
                output_hidden_states=None, frames=None, return_dict=None, scores_out: Optional[torch.Tensor] = None,
                **kwargs) -> VideoHeadCausalLMOutputWithPast:
        logits_mode = kwargs.pop("logits_mode", None)       # per-call override: "none" | "last" | "all"
        logits_mode = self.logits_mode if logits_mode is None else \
            {"none": _lib.LOGITS_NONE, "last": _lib.LOGITS_LAST, "all": _lib.LOGITS_ALL}[logits_mode]
        if labels is not None or informative_labels is not None or relevance_labels is not None:
            raise NotImplementedError("training losses are out of scope of aha_b200 (inference hot path only)")
        if inputs_embeds is None:
            if input_ids is None or frames is not None:
                raise NotImplementedError("pass inputs_embeds (the live drivers always do, test/inference.py:217)")
            inputs_embeds = self.get_input_embeddings()(input_ids)
        if past_key_values is None:
            raise ValueError("aha_b200 prefill needs past_key_values (a SinkCache / SlidingWindowCache / TrulyStaticCache)")
        caches: Sequence[_NativeCache] = past_key_values if isinstance(past_key_values, (list, tuple)) else [past_key_values]
        x = inputs_embeds
        if x.dim() != 3 or x.shape[0] != len(caches) or x.shape[2] != self.config.hidden_size:
            raise ValueError(f"inputs_embeds must be [{len(caches)}, q, {self.config.hidden_size}], got {tuple(x.shape)}")
        if x.dtype != torch.bfloat16:
            raise TypeError("inputs_embeds must be bfloat16")
        x = x.to(self.device).contiguous()
        ns, q, H = x.shape
        cfg = self.config
        for c in caches:
            if not isinstance(c, _NativeCache):
                raise TypeError("past_key_values must be aha_b200 caches")
            if c.handle is None:
                c.allocate(cfg.num_key_value_heads, cfg.head_dim, cfg.num_hidden_layers, max(q, min(self.max_tokens, 4096)), self.device)
        dev = self.device
        f32 = dict(dtype=torch.float32, device=dev)
        inf = torch.empty(ns, q, 2, **f32)
        rel = torch.empty(ns, q, 1, **f32)
        unc = torch.empty(ns, q, 1, **f32)
        scores = scores_out if scores_out is not None else torch.empty(ns, 3, **f32)
        if scores.dtype != torch.float32 or scores.numel() != ns * 3 or not scores.is_contiguous():
            raise ValueError("scores_out must be a contiguous float32 [n_streams, 3] tensor")
        logits = None
        if logits_mode == _lib.LOGITS_ALL:
            logits = torch.empty(ns, q, cfg.vocab_size, **f32)
        elif logits_mode == _lib.LOGITS_LAST:
            logits = torch.empty(ns, 1, cfg.vocab_size, **f32)
        hidden = torch.empty(ns, q, H, dtype=torch.bfloat16, device=dev) if output_hidden_states else None
        handles = (ctypes.c_void_p * ns)(*[c.handle for c in caches])
        with torch.cuda.device(dev):
            _lib.check(_lib.load().aha_prefill_chunk(
                self._h, handles, ns, _lib.ptr(x), q, self.mask_mode, logits_mode, _lib.ptr(logits), _lib.ptr(inf),
                _lib.ptr(rel), _lib.ptr(unc), _lib.ptr(scores), _lib.ptr(hidden), _lib.stream_ptr()))
        return VideoHeadCausalLMOutputWithPast(
            informative_logits=inf, relevance_logits=rel, uncertainty=unc, logits=logits, past_key_values=past_key_values,
            hidden_states=hidden, frame_scores=scores)

    __call__ = forward


def build_model_from_state_dict(config: AhaConfig, state_dict, **kw) -> VideoHeadLiveLlavaQwenForCausalLM:
    return VideoHeadLiveLlavaQwenForCausalLM(config, state_dict, **kw)
