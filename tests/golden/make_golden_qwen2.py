"""Pins oracle.qwen2_prefill (the restatement of transformers 4.49's Qwen2 forward) to EXECUTED code.
Authoring container only for part (b) (needs /root/reference).  Run:  python tests/golden/make_golden_qwen2.py

  (a) hf_growing_*   the installed transformers `Qwen2Model(attn_implementation="eager")` with its own DynamicCache, the
                     oracle's weights, chunks 17 / 9 / 9 / 1 / 9: the growing (plain causal) regime end to end - position
                     ids from the cache length, rotary embedding, RMSNorm, biased q/k/v, GQA, fp32 softmax, SwiGLU MLP.
  (b) ref_<policy>_* the 4.49-only semantics: HF's own Qwen2DecoderLayer / rotary / norm modules driven layer by layer
                     with THE REFERENCE'S OWN cache classes (test/sink_cache.py, test/sliding_window_cache.py,
                     test/static_cache.py; 4.49 calling convention update(k, v, layer, {"cos", "sin"})), positions
                     arange(L, L + q) with L = cache.get_seq_length(), and the 4.49 additive mask
                     [q, L + q + 1] (finfo.min where j > L + i) sliced to the Klen columns the cache returns.  Windows
                     small enough that eviction + re-rotation, the bidirectional in-chunk regime (cache full) and the
                     frozen static prefix all occur.
Stored: bit patterns of the final-norm hidden states of every chunk (fp32 and bf16).  The test regenerates weights and
inputs from the seeds below and requires oracle.qwen2_prefill to reproduce the bits.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
from oracle import aha_oracle as o                      # noqa: E402

W_SEED, X_SEED = 5, 17
GROWING_CHUNKS = [17, 9, 9, 1, 9]
REF_CASES = [
    ("sink", dict(window_length=40, num_sink_tokens=4), [13, 9, 9, 9, 9, 9, 5, 9]),
    ("sink", dict(window_length=48, num_sink_tokens=0), [20, 9, 9, 9, 9, 9]),
    ("sliding", dict(window_length=30), [13, 9, 9, 9, 9, 4]),
    ("static", dict(window_size=64), [11, 9, 9, 9]),
    ("static", dict(window_size=6), [11, 9, 9]),
]


def qwen2_dims():
    return o.dims_preset("micro72")


def chunk_inputs(dims, case, step, q, dtype):
    g = torch.Generator().manual_seed(X_SEED + 1000 * case + step)
    return (torch.randn(1, q, dims.hidden, generator=g) * 0.5).to(dtype)


def hf_model(dims, w, dtype):
    from transformers import Qwen2Config, Qwen2Model
    cfg = Qwen2Config(hidden_size=dims.hidden, intermediate_size=dims.inter, num_hidden_layers=dims.layers,
                      num_attention_heads=dims.heads, num_key_value_heads=dims.kv_heads, vocab_size=dims.vocab,
                      rms_norm_eps=dims.rms_eps, max_position_embeddings=32768, rope_theta=dims.rope_theta,
                      use_sliding_window=False, attn_implementation="eager")
    m = Qwen2Model(cfg).eval()
    sd = {k[len("model."):]: v.float() for k, v in w.items()
          if k.startswith("model.layers.") or k in ("model.embed_tokens.weight", "model.norm.weight")}
    m.load_state_dict(sd, strict=True)
    inv = m.rotary_emb.inv_freq.clone()
    m = m.to(dtype)
    m.rotary_emb.inv_freq = inv            # the rotary frequencies stay fp32 (a non-persistent buffer in 4.49 as well)
    return m


def bits(t):
    t = t.contiguous()
    return t.view(torch.int32 if t.dtype == torch.float32 else torch.int16).numpy()


def run_hf_growing(dims, w, dtype):
    from transformers import DynamicCache
    m = hf_model(dims, w, dtype)
    cache = DynamicCache(config=m.config)
    outs = []
    with torch.no_grad():
        for step, q in enumerate(GROWING_CHUNKS):
            x = chunk_inputs(dims, 99, step, q, dtype)
            outs.append(m(inputs_embeds=x, past_key_values=cache, use_cache=True).last_hidden_state)
    return outs


class _RefCacheAdapter:
    """What transformers 4.49's Qwen2Attention does with a Cache: update(k, v, layer_idx, {"sin", "cos", ...})."""

    def __init__(self, cache):
        self.cache, self.cos, self.sin = cache, None, None

    def update(self, k, v, layer_idx, cache_kwargs=None):
        return self.cache.update(k, v, layer_idx, {"cos": self.cos, "sin": self.sin})


def run_ref_case(dims, w, dtype, case, policy, kw, chunks, classes):
    import copy
    m = hf_model(dims, w, dtype)
    SinkCache, SlidingWindowCache, TrulyStaticCache = classes
    cache = {"sink": SinkCache, "sliding": SlidingWindowCache, "static": TrulyStaticCache}[policy](**kw)
    ad = _RefCacheAdapter(cache)
    outs, lens = [], []
    with torch.no_grad():
        for step, q in enumerate(chunks):
            x = chunk_inputs(dims, case, step, q, dtype)
            L = cache.get_seq_length()
            pos = torch.arange(L, L + q)[None]
            cos, sin = m.rotary_emb(x, pos)
            ad.cos, ad.sin = cos, sin
            # Klen = what the reference cache returns for this chunk (probed on a copy with dummy keys)
            probe = copy.deepcopy(cache)
            dk = torch.zeros(1, dims.kv_heads, q, dims.head_dim, dtype=dtype)
            klen = probe.update(dk, dk, 0, {"cos": cos, "sin": sin})[0].shape[-2]
            full = torch.full((q, L + q + 1), torch.finfo(dtype).min, dtype=dtype)
            full = full * (torch.arange(L + q + 1)[None, :] > (L + torch.arange(q))[:, None])     # 4.49 mask builder
            mask = full[None, None, :, :klen]
            h = x
            for layer in m.layers:
                h = layer(h, attention_mask=mask, position_ids=pos, past_key_values=ad, use_cache=True,
                          position_embeddings=(cos, sin))
                if isinstance(h, tuple):
                    h = h[0]
            outs.append(m.norm(h))
            lens.append([L, klen, cache.get_seq_length()])
    return outs, lens


def main():
    import validate_against_reference as ref_loader
    dims = qwen2_dims()
    out = {}
    for dtype, tag in ((torch.float32, "f32"), (torch.bfloat16, "bf16")):      # (a): before the Cache.__init__ shim below
        w = o.init_weights(dims, seed=W_SEED, dtype=dtype)
        for i, h in enumerate(run_hf_growing(dims, w, dtype)):
            out[f"hf_growing_{tag}_{i}"] = bits(h)
    classes = ref_loader.load_reference_caches()      # patches transformers.Cache.__init__ (4.49-style no-arg ctor)
    for dtype, tag in ((torch.float32, "f32"), (torch.bfloat16, "bf16")):
        w = o.init_weights(dims, seed=W_SEED, dtype=dtype)
        for ci, (policy, kw, chunks) in enumerate(REF_CASES):
            hs, lens = run_ref_case(dims, w, dtype, ci, policy, kw, chunks, classes)
            out[f"ref{ci}_{tag}_lens"] = np.array(lens, dtype=np.int32)
            for i, h in enumerate(hs):
                out[f"ref{ci}_{tag}_{i}"] = bits(h)
    np.savez_compressed(os.path.join(HERE, "qwen2_hf.npz"), **out)
    print("wrote qwen2_hf.npz:", len(out), "arrays")


if __name__ == "__main__":
    main()
