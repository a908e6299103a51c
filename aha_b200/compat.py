"""The pieces of the reference's object graph that its drivers touch around the hot path, so that the driver code of
test/inference.py and test/live_infer_for_video.py runs against the native engine as it is written:

    model.get_vision_tower().image_processor.preprocess(frames, return_tensors='pt')['pixel_values']
                                                     test/inference.py:51, :176; siglip_encoder.py:47-67
    build_model_and_tokenizer(is_training=False, set_vision_inside=True, torch_dtype=..., **asdict(args))
                                                     models/__init__.py:8-13, test/inference.py:45
    tokenizer.apply_chat_template([...], add_stream_prompt= / add_stream_generation_prompt= / add_stream_query_prompt=,
                                  return_tensors='pt')     models/tokenization_live.py, test/inference.py:92-94, :250

Weights: a local checkpoint directory (`llm_pretrained`, *.safetensors + config.json; LoRA adapters from
`lora_pretrained` are folded in with merge_lora_state_dict) or an explicit `state_dict=` / `config=`.  There is no network
in the build environment, so the tokenizer falls back to `SyntheticTokenizer` (fixed id chunks of the right lengths) when no local tokenizer files are given.
"""
from __future__ import annotations

import glob
import json
import os
from typing import Dict, List, Optional

import numpy as np
import torch

from .config import AhaConfig, VisionConfig, config_preset


class SigLipImageProcessor:
    """LLaVA_NeXT/llava/model/multimodal_encoder/siglip_encoder.py:34-67 for frames that are already `size` x `size`
    (every loader of the reference guarantees that, test/inference.py:505-558): rescale by 1/255, normalise with
    mean = std = 0.5, channels first.  Returns float32 like the reference; the 256 possible values are computed with the
    reference's numpy float32 arithmetic, so `.to(torch.bfloat16)` afterwards gives the bits the reference feeds its tower."""

    def __init__(self, image_mean=(0.5, 0.5, 0.5), image_std=(0.5, 0.5, 0.5), size=(384, 384), rescale_factor=1 / 255):
        self.image_mean, self.image_std, self.size, self.rescale_factor = image_mean, image_std, size, rescale_factor
        self.crop_size = {"height": size[0], "width": size[1]}
        # transformers image_transforms.rescale multiplies in float64 and casts to float32; normalize runs in float32
        lut = (np.arange(256, dtype=np.uint8).astype(np.float64) * rescale_factor).astype(np.float32)
        lut = (lut - np.array(image_mean[0], dtype=np.float32)) / np.array(image_std[0], dtype=np.float32)
        self._lut = torch.from_numpy(lut.astype(np.float32))

    def preprocess(self, images, return_tensors="pt"):
        if isinstance(images, (list, tuple)):
            images = torch.stack([torch.as_tensor(np.asarray(im)) for im in images])
        images = torch.as_tensor(images)
        if images.dim() == 3:
            images = images[None]
        if images.dtype != torch.uint8:
            raise TypeError("image_processor.preprocess expects uint8 frames [T, 3, S, S]")
        if tuple(images.shape[-2:]) != tuple(self.size):
            raise ValueError(f"frames must already be {self.size[0]} x {self.size[1]} (use aha_b200.load_video_for_testing / "
                             f"load_one_frame for the resize + border step)")
        lut = self._lut.to(images.device)
        return {"pixel_values": lut[images.long()]}


class _VisionTower:
    def __init__(self, size):
        self.image_processor = SigLipImageProcessor(size=(size, size))
        self.is_loaded = True


class SyntheticTokenizer:
    """Stand-in for the reference's chat-template tokenizer when no tokenizer files are available: every template call
    returns a fixed id chunk of the configured length (system prompt, stream prompt, generation prompt, query)."""

    def __init__(self, vocab_size: int, n_sys: int = 36, n_stream_prompt: int = 4, n_generation: int = 4, n_query: int = 24,
                 seed: int = 11, eos_token_id: int = 0):
        def ids(n, s):
            g = torch.Generator().manual_seed(s)
            return torch.randint(0, vocab_size, (1, n), generator=g, dtype=torch.int64)
        self._sys, self._query = ids(n_sys, seed), ids(n_query, seed + 1)
        self._stream, self._gen = ids(n_stream_prompt, seed + 2), ids(n_generation, seed + 3)
        self.eos_token_id = eos_token_id
        self.vocab_size = vocab_size

    def apply_chat_template(self, conversation, add_stream_prompt=False, add_stream_generation_prompt=False,
                            add_stream_query_prompt=False, return_tensors="pt", **kw):
        turn = conversation[0] if conversation else {}
        if turn.get("role") == "system":
            return self._sys.clone()
        if turn.get("role") == "user":
            return self._query.clone()
        if add_stream_generation_prompt:
            return self._gen.clone()
        return self._stream.clone()

    def decode(self, ids, **kw):
        ids = ids.tolist() if hasattr(ids, "tolist") else list(ids)
        return " ".join(f"<{int(i)}>" for i in (ids[0] if ids and isinstance(ids[0], list) else ids))


def _config_from_hf_json(path: str) -> AhaConfig:
    with open(path) as f:
        c = json.load(f)
    heads = c["num_attention_heads"]
    return AhaConfig(c["hidden_size"], c["intermediate_size"], c["num_hidden_layers"], heads, c["num_key_value_heads"],
                     c.get("head_dim", c["hidden_size"] // heads), c["vocab_size"], rope_theta=c.get("rope_theta", 1e6),
                     rms_norm_eps=c.get("rms_norm_eps", 1e-6), tie_word_embeddings=c.get("tie_word_embeddings", False),
                     video_pooling_stride=c.get("video_pooling_stride", 4), mm_spatial_pool_mode=c.get("mm_spatial_pool_mode", "bilinear"),
                     eos_token_id=c.get("eos_token_id", 151645) if isinstance(c.get("eos_token_id", 151645), int) else 151645)


def _load_safetensors_dir(path: str) -> Dict[str, torch.Tensor]:
    from safetensors.torch import load_file
    files = sorted(glob.glob(os.path.join(path, "*.safetensors")))
    if not files:
        raise FileNotFoundError(f"no *.safetensors under {path}")
    sd: Dict[str, torch.Tensor] = {}
    for f in files:
        sd.update(load_file(f))
    return sd


def build_model_and_tokenizer(is_training: bool = False, set_vision_inside: bool = True, torch_dtype=torch.bfloat16,
                              state_dict: Optional[Dict[str, torch.Tensor]] = None, config: Optional[AhaConfig] = None,
                              tokenizer=None, llm_pretrained: Optional[str] = None, lora_pretrained: Optional[str] = None,
                              lora_alpha: float = 32, lora_r: int = 16, attn_implementation: str = "sdpa",
                              device="cuda", max_tokens: int = 256, max_frames: int = 32, logits_mode: str = "all", **kwargs):
    """models/__init__.py:8-13 -> build_live (models/modeling_live.py:96-181), inference only.  Extra reference arguments
    (the whole `asdict(args)`) are accepted and ignored except the ones that shape the hot path."""
    from .model import VideoHeadLiveLlavaQwenForCausalLM
    from .weights import merge_lora_state_dict
    if is_training:
        raise NotImplementedError("aha_b200 is inference only (training is out of scope)")
    if torch_dtype not in (torch.bfloat16, None):
        raise TypeError("aha_b200 computes in bfloat16 (pass --bf16 true)")
    if state_dict is None:
        if not llm_pretrained or not os.path.isdir(llm_pretrained):
            raise FileNotFoundError(f"llm_pretrained must be a local checkpoint directory (no network here): {llm_pretrained!r}")
        state_dict = _load_safetensors_dir(llm_pretrained)
        if config is None:
            config = _config_from_hf_json(os.path.join(llm_pretrained, "config.json"))
    if config is None:
        raise ValueError("config is required with an explicit state_dict")
    for k in ("video_pooling_stride", "frame_resolution"):
        if kwargs.get(k) is not None and hasattr(config, k):
            setattr(config, k, kwargs[k])
    if lora_pretrained:
        adapter = _load_safetensors_dir(lora_pretrained)
        cfg_path = os.path.join(lora_pretrained, "adapter_config.json")
        if os.path.exists(cfg_path):
            with open(cfg_path) as f:
                ac = json.load(f)
            lora_alpha, lora_r = ac.get("lora_alpha", lora_alpha), ac.get("r", lora_r)
        state_dict = merge_lora_state_dict(state_dict, adapter, lora_alpha=lora_alpha, r=lora_r)
    mask_mode = {"flash_attention_2": "fa2", "sdpa": "sdpa", "eager": "sdpa"}.get(attn_implementation, "sdpa")
    model = VideoHeadLiveLlavaQwenForCausalLM(config, state_dict, device=device, max_tokens=max_tokens, max_frames=max_frames,
                                              logits_mode=logits_mode, mask_mode=mask_mode)
    if tokenizer is None:
        tokenizer = SyntheticTokenizer(config.vocab_size, eos_token_id=config.eos_token_id)
    return model, tokenizer
