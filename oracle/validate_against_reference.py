"""Pins the oracle restatement against the pieces of the reference that import in the authoring container.
Runs ONLY where /root/reference exists (never on the GPU box, never from tests).  Usage:

    python oracle/validate_against_reference.py

Checks (all must be bit-exact on CPU):
  1. OracleKVCache('sink'|'sliding'|'static') vs test/sink_cache.py, test/sliding_window_cache.py,
     test/static_cache.py driven with the transformers-4.49 calling convention update(k, v, layer, {"cos","sin"}).
  2. siglip_tower vs the vendored SigLipVisionModel (tower surgery of siglip_encoder.py:570-571 applied).
  3. preprocess vs SigLipImageProcessor.preprocess.
  4. find_ticks_restated vs scipy (identical peaks).
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import aha_oracle as o  # noqa: E402


def load_reference_caches():
    import transformers
    transformers.Cache.__init__ = lambda self, *a, **k: None      # 4.49-style no-arg base ctor
    sys.path.insert(0, REF)
    from test.sink_cache import SinkCache
    from test.sliding_window_cache import SlidingWindowCache
    from test.static_cache import TrulyStaticCache
    return SinkCache, SlidingWindowCache, TrulyStaticCache


def load_reference_siglip():
    llava = types.ModuleType("llava"); llava.__path__ = []
    lu = types.ModuleType("llava.utils"); lu.rank0_print = print
    sys.modules["llava"] = llava; sys.modules["llava.utils"] = lu
    spec = importlib.util.spec_from_file_location(
        "siglip_encoder", f"{REF}/LLaVA_NeXT/llava/model/multimodal_encoder/siglip_encoder.py")
    s = importlib.util.module_from_spec(spec); sys.modules["siglip_encoder"] = s; spec.loader.exec_module(s)
    return s


def check_caches():
    SinkCache, SlidingWindowCache, TrulyStaticCache = load_reference_caches()
    torch.manual_seed(0)
    d, kvh, layers = 16, 2, 2
    cases = [
        ("sink", dict(window_length=40, num_sink_tokens=4), [7, 9, 9, 9, 9, 9, 9, 9, 5, 9, 3]),
        ("sink", dict(window_length=64, num_sink_tokens=0), [10, 20, 20, 20, 20, 7]),
        ("sink", dict(window_length=30, num_sink_tokens=6), [40, 5, 5, 8]),          # first chunk longer than the window
        ("sliding", dict(window_length=20), [7, 9, 9, 30, 4]),
        ("static", dict(window_size=5), [7, 9, 4]),
        ("static", dict(window_size=50), [7, 9, 4]),
    ]
    for policy, kw, chunks in cases:
        for dtype in (torch.float32, torch.bfloat16):
            if policy == "sink":
                ref = SinkCache(**kw); mine = o.OracleKVCache("sink", kw["window_length"], kw["num_sink_tokens"])
            elif policy == "sliding":
                ref = SlidingWindowCache(**kw); mine = o.OracleKVCache("sliding", kw["window_length"])
            else:
                ref = TrulyStaticCache(**kw); mine = o.OracleKVCache("static", kw["window_size"])
            for q in chunks:
                L = ref.get_seq_length()
                assert L == mine.get_seq_length(), (policy, kw, q)
                pos = torch.arange(L, L + q)
                cos, sin = o.rope_cos_sin(pos, d, 1e6, dtype)
                for layer in range(layers):
                    k = torch.randn(1, kvh, q, d).to(dtype); v = torch.randn(1, kvh, q, d).to(dtype)
                    Kr, Vr = ref.update(k, v, layer, {"cos": cos, "sin": sin})
                    Km, Vm = mine.update(k, v, layer, cos, sin)
                    assert Kr.shape == Km.shape, (policy, kw, q, Kr.shape, Km.shape)
                    assert torch.equal(Kr, Km) and torch.equal(Vr, Vm), (policy, kw, q, layer)
                assert ref.get_seq_length() == mine.get_seq_length()
                if policy == "sink":
                    assert ref._seen_tokens == mine.seen_tokens
        print(f"cache {policy} {kw}: bit-exact over chunks {chunks}")


def check_siglip():
    s = load_reference_siglip()
    dims = o.dims_preset("micro72")
    v = dims.vision
    cfg = s.SigLipVisionConfig(hidden_size=v.hidden, intermediate_size=v.inter, num_hidden_layers=v.layers + 1,
                               num_attention_heads=v.heads, image_size=v.image, patch_size=v.patch)
    m = s.SigLipVisionModel(cfg)
    del m.vision_model.encoder.layers[-1:]
    m.vision_model.head = torch.nn.Identity()
    w = o.init_weights(dims, 3, dtype=torch.float32)
    sd = {k[len("model.vision_tower.vision_tower."):]: t for k, t in w.items() if k.startswith(o.VT)}
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all("post_layernorm" in k or "head" in k for k in missing), missing
    frames = o.synthetic_frames(2, v.image)
    proc = s.SigLipImageProcessor(size=(v.image, v.image))
    pv_ref = proc.preprocess(frames, return_tensors="pt")["pixel_values"]
    pv = o.preprocess(frames)
    assert pv_ref.dtype == torch.float32 and torch.equal(pv_ref, pv), "preprocess mismatch"
    print("preprocess: bit-exact vs SigLipImageProcessor")
    for dtype in (torch.float32, torch.bfloat16):
        mm = m.to(dtype).eval()
        with torch.no_grad():
            ref = mm(pv.to(dtype), output_hidden_states=True).hidden_states[-1]
            mine = o.siglip_tower({k: t.to(dtype) for k, t in w.items()}, dims, pv.to(dtype))
        assert torch.equal(ref, mine), (dtype, (ref.float() - mine.float()).abs().max())
        print(f"siglip tower {dtype}: bit-exact vs vendored SigLipVisionModel")


def check_ticks():
    rng = np.random.default_rng(7)
    n_ok = 0
    for k in range(64):
        T = int(rng.integers(15, 3000))
        t = np.arange(T)
        ar = np.zeros(T)
        e = rng.standard_normal(T)
        for i in range(1, T):
            ar[i] = 0.9 * ar[i - 1] + e[i]
        x = np.clip(0.5 + 0.25 * np.sin(2 * np.pi * t / rng.integers(50, 500)) + 0.1 * ar * 0.3 + 0.02 * rng.standard_normal(T), 0, 1)
        if k % 8 == 0:
            x = np.round(x, 1)      # plateaus
        ticks, sm, th = o.find_ticks(x, fps=1.0)
        pk, sm2, th2 = o.find_ticks_restated(x, fps=1.0)
        assert [int(p) for p in ticks] == pk.tolist(), (k, ticks, pk)
        assert np.abs(sm - sm2).max() < 1e-12
        n_ok += 1
    print(f"find_ticks restatement: identical peaks on {n_ok} streams")


if __name__ == "__main__":
    check_ticks()
    check_caches()
    check_siglip()
    print("ORACLE PINNED AGAINST REFERENCE: OK")
