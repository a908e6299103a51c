"""Random-init weights in the reference checkpoint's state-dict naming (there is no network for real checkpoints):
HF-style normal(0, 0.02) for Linear / Embedding, ~1 for norm scales, wider score heads so the scores spread."""
from __future__ import annotations

import math
from typing import Dict

import torch

from .config import AhaConfig

VT = "model.vision_tower.vision_tower.vision_model."


def random_init_state_dict(cfg: AhaConfig, seed: int = 0, device="cuda", head_spread: float = 1.5) -> Dict[str, torch.Tensor]:
    g = torch.Generator(device=device)
    g.manual_seed(seed)

    def rn(*shape, std=0.02):
        return (torch.randn(*shape, generator=g, device=device, dtype=torch.float32) * std).to(torch.bfloat16)

    def ones_noise(n):
        return (1.0 + 0.1 * torch.randn(n, generator=g, device=device, dtype=torch.float32)).to(torch.bfloat16)

    w: Dict[str, torch.Tensor] = {}
    v = cfg.vision
    w[VT + "embeddings.patch_embedding.weight"] = rn(v.hidden_size, 3, v.patch_size, v.patch_size)
    w[VT + "embeddings.patch_embedding.bias"] = rn(v.hidden_size)
    w[VT + "embeddings.position_embedding.weight"] = rn(v.num_patches, v.hidden_size)
    for i in range(v.num_hidden_layers):
        p = VT + f"encoder.layers.{i}."
        w[p + "layer_norm1.weight"] = ones_noise(v.hidden_size)
        w[p + "layer_norm1.bias"] = rn(v.hidden_size)
        for n in ("q_proj", "k_proj", "v_proj", "out_proj"):
            w[p + f"self_attn.{n}.weight"] = rn(v.hidden_size, v.hidden_size)
            w[p + f"self_attn.{n}.bias"] = rn(v.hidden_size)
        w[p + "layer_norm2.weight"] = ones_noise(v.hidden_size)
        w[p + "layer_norm2.bias"] = rn(v.hidden_size)
        w[p + "mlp.fc1.weight"] = rn(v.intermediate_size, v.hidden_size)
        w[p + "mlp.fc1.bias"] = rn(v.intermediate_size)
        w[p + "mlp.fc2.weight"] = rn(v.hidden_size, v.intermediate_size)
        w[p + "mlp.fc2.bias"] = rn(v.hidden_size)
    H = cfg.hidden_size
    w["model.mm_projector.0.weight"] = rn(H, v.hidden_size)
    w["model.mm_projector.0.bias"] = rn(H)
    w["model.mm_projector.2.weight"] = rn(H, H)
    w["model.mm_projector.2.bias"] = rn(H)
    w["model.embed_tokens.weight"] = rn(cfg.vocab_size, H)
    qd, kd = cfg.num_attention_heads * cfg.head_dim, cfg.num_key_value_heads * cfg.head_dim
    for i in range(cfg.num_hidden_layers):
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
        w[p + "mlp.gate_proj.weight"] = rn(cfg.intermediate_size, H)
        w[p + "mlp.up_proj.weight"] = rn(cfg.intermediate_size, H)
        w[p + "mlp.down_proj.weight"] = rn(H, cfg.intermediate_size)
    w["model.norm.weight"] = ones_noise(H)
    w["lm_head.weight"] = w["model.embed_tokens.weight"] if cfg.tie_word_embeddings else rn(cfg.vocab_size, H)
    hs = head_spread / math.sqrt(H)
    w["informative_head.weight"] = rn(2, H, std=hs)
    w["relevance_head.weight"] = rn(1, H, std=hs)
    w["uncertainty_head.weight"] = rn(1, H, std=hs)
    return w


def synthetic_frames(n: int, image: int = 384, seed: int = 1, device="cpu") -> torch.Tensor:
    """uint8 [n, 3, image, image] frames: a slowly drifting coarse field plus pixel noise."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    coarse = torch.rand(n + 4, 3, 24, 24, generator=g, device=device)
    coarse = torch.stack([coarse[i:i + 4].mean(0) for i in range(n)])
    up = torch.nn.functional.interpolate(coarse, size=(image, image), mode="bilinear", align_corners=False)
    noise = torch.rand(n, 3, image, image, generator=g, device=device) * 0.2
    return ((up * 0.8 + noise).clamp(0, 1) * 255).to(torch.uint8)


# ----------------------------------------------------------------------------------------------------------------------
# N2: LoRA adapters at load time.  The reference runs the adapters UNMERGED through peft (PeftModel.from_pretrained,
# models/modeling_live.py:170-173; r = 16, alpha = 32 on q/k/v/o/gate/up/down of every decoder layer, plus full copies of
# the `modules_to_save`, arguments_live.py:14-18): y = W x + (alpha / r) * B (A x).  The native engine streams ONE weight
# matrix per projection, so the adapters are folded in once: W' = W + (alpha / r) * B A, accumulated in fp32 and rounded
# to bf16 (14 skinny GEMMs per layer less on the hot path).
# ----------------------------------------------------------------------------------------------------------------------
def merge_lora_state_dict(base: Dict[str, torch.Tensor], adapter: Dict[str, torch.Tensor], lora_alpha: float = 32.0,
                          r: int = None, adapter_name: str = "default") -> Dict[str, torch.Tensor]:
    """`base`: the LLM checkpoint's state dict; `adapter`: a peft adapter state dict (adapter_model.safetensors keys such as
    `base_model.model.model.layers.0.self_attn.q_proj.lora_A.weight`, with or without the `.default` adapter infix, and
    `base_model.model.informative_head.modules_to_save.default.weight` / `...informative_head.weight` for saved modules).
    Returns a new state dict in the base naming with every adapter folded in."""
    out = dict(base)

    def strip(k: str) -> str:
        for pre in ("base_model.model.",):
            if k.startswith(pre):
                k = k[len(pre):]
        return k

    a_mats, b_mats, saved = {}, {}, {}
    for k, t in adapter.items():
        k = strip(k)
        k = k.replace(f".lora_A.{adapter_name}.", ".lora_A.").replace(f".lora_B.{adapter_name}.", ".lora_B.")
        k = k.replace(f".modules_to_save.{adapter_name}.", ".").replace(".original_module.", ".")
        if ".lora_A." in k:
            a_mats[k.replace(".lora_A.", ".")] = t
        elif ".lora_B." in k:
            b_mats[k.replace(".lora_B.", ".")] = t
        else:
            saved[k] = t
    if set(a_mats) != set(b_mats):
        raise ValueError(f"unpaired LoRA matrices: {sorted(set(a_mats) ^ set(b_mats))[:4]}")
    for name, A in a_mats.items():
        if name not in out:
            raise KeyError(f"LoRA target '{name}' is not in the base state dict")
        B = b_mats[name]
        rank = A.shape[0]
        if B.shape[1] != rank or (r is not None and rank != r):
            raise ValueError(f"LoRA rank mismatch for '{name}': A {tuple(A.shape)}, B {tuple(B.shape)}")
        W = out[name]
        scale = float(lora_alpha) / rank
        delta = (B.to(device=W.device, dtype=torch.float32) @ A.to(device=W.device, dtype=torch.float32)) * scale
        out[name] = (W.float() + delta).to(W.dtype)
    for name, t in saved.items():          # modules_to_save: full replacements (heads, projector, ...)
        out[name] = t.to(dtype=out[name].dtype, device=out[name].device) if name in out else t
    return out
