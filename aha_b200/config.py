"""Shapes of the Aha! model the hot path needs: VideoHeadLiveLlavaQwenConfig (models/live_llava/
video_head_live_llava_qwen.py:43-47, models/configuration_live.py:22-36) + SigLipVisionConfig
(LLaVA_NeXT/llava/model/multimodal_encoder/siglip_encoder.py:70-115).  The lmms-lab llava-onevision-qwen2 config files
are not part of the reference tree (models/arguments_live.py:12 names them), so the shapes are spelled out here."""
from __future__ import annotations

import math
from dataclasses import dataclass, field


@dataclass
class VisionConfig:
    hidden_size: int = 1152
    intermediate_size: int = 4304
    num_hidden_layers: int = 26          # 27 minus the layer SigLipVisionTower deletes (siglip_encoder.py:570)
    num_attention_heads: int = 16
    patch_size: int = 14
    image_size: int = 384
    layer_norm_eps: float = 1e-6

    @property
    def side(self) -> int:
        return self.image_size // self.patch_size

    @property
    def num_patches(self) -> int:
        return self.side * self.side


@dataclass
class AhaConfig:
    hidden_size: int
    intermediate_size: int
    num_hidden_layers: int
    num_attention_heads: int
    num_key_value_heads: int
    head_dim: int
    vocab_size: int
    rope_theta: float = 1e6
    rms_norm_eps: float = 1e-6
    tie_word_embeddings: bool = False
    vision: VisionConfig = field(default_factory=VisionConfig)
    video_pooling_stride: int = 4                 # models/arguments_live.py:23
    mm_spatial_pool_mode: str = "bilinear"
    frame_resolution: int = 384                   # models/arguments_live.py:20
    v_placeholder: str = "<image>"
    eos_token_id: int = 151645

    @property
    def pooled_side(self) -> int:
        s = self.vision.side
        if self.mm_spatial_pool_mode == "bilinear":
            return math.ceil(s / self.video_pooling_stride)
        return s // self.video_pooling_stride

    @property
    def frame_num_tokens(self) -> int:            # models/arguments_live.py:22 (49 for 27x27 patches, stride 4, bilinear)
        return self.pooled_side ** 2


def config_preset(name: str) -> AhaConfig:
    if name == "7b":      # lmms-lab/llava-onevision-qwen2-7b-ov
        return AhaConfig(3584, 18944, 28, 28, 4, 128, 152064)
    if name == "tiny":    # lmms-lab/llava-onevision-qwen2-0.5b-ov
        return AhaConfig(896, 4864, 24, 14, 2, 64, 151936, tie_word_embeddings=True)
    if name == "micro72":  # small test shape: SigLIP head_dim 72, decoder head_dim 64, 9x9 patches -> 3x3 pooled tokens
        return AhaConfig(256, 512, 2, 4, 2, 64, 768,
                         vision=VisionConfig(hidden_size=144, intermediate_size=304, num_hidden_layers=2,
                                             num_attention_heads=2, patch_size=14, image_size=126),
                         frame_resolution=126)
    raise ValueError(f"unknown config preset {name}")
