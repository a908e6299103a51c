"""Generates the committed golden fixtures by EXECUTING THE REFERENCE where it imports (authoring container only:
needs /root/reference).  Run:  python tests/golden/make_golden.py

  caches.npz    test/sink_cache.py, test/sliding_window_cache.py, test/static_cache.py driven with the transformers-4.49
                calling convention update(k, v, layer, {"cos","sin"}); inputs are regenerated from seeds by the tests,
                outputs (lengths per step, final K/V bit patterns) are stored.
  siglip.npz    vendored SigLipVisionModel (tower surgery of siglip_encoder.py:570-571) + SigLipImageProcessor on two
                synthetic frames at the micro72 shape, fp32.
  ticks.npz     scipy.signal.savgol_filter / find_peaks with the arguments of test/live_infer_for_video.py:209-221.
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
import validate_against_reference as ref_loader          # noqa: E402

CACHE_CASES = [
    ("sink", dict(window_length=40, num_sink_tokens=4), [7, 9, 9, 9, 9, 9, 9, 9, 5, 9, 3]),
    ("sink", dict(window_length=64, num_sink_tokens=0), [10, 20, 20, 20, 20, 7]),
    ("sink", dict(window_length=30, num_sink_tokens=6), [40, 5, 5, 8]),
    ("sliding", dict(window_length=20), [7, 9, 9, 30, 4]),
    ("static", dict(window_size=5), [7, 9, 4]),
    ("static", dict(window_size=50), [7, 9, 4]),
]
CACHE_D, CACHE_KVH = 16, 2


def cache_inputs(case_idx, step, q, L):
    """Deterministic inputs shared by the generator and the tests."""
    g = torch.Generator().manual_seed(1000 * case_idx + step)
    k = torch.randn(1, CACHE_KVH, q, CACHE_D, generator=g).bfloat16()
    v = torch.randn(1, CACHE_KVH, q, CACHE_D, generator=g).bfloat16()
    cos, sin = o.rope_cos_sin(torch.arange(L, L + q), CACHE_D, 1e6, torch.bfloat16)
    return k, v, cos, sin


def bits(t):
    return t.contiguous().view(torch.int16).numpy()


def make_caches():
    SinkCache, SlidingWindowCache, TrulyStaticCache = ref_loader.load_reference_caches()
    out = {}
    for ci, (policy, kw, chunks) in enumerate(CACHE_CASES):
        c = {"sink": SinkCache, "sliding": SlidingWindowCache, "static": TrulyStaticCache}[policy](**kw)
        lens = []
        for step, q in enumerate(chunks):
            L = c.get_seq_length()
            k, v, cos, sin = cache_inputs(ci, step, q, L)
            K, V = c.update(k, v, 0, {"cos": cos, "sin": sin})
            lens.append([L, K.shape[-2], c.get_seq_length()])
        out[f"case{ci}_lens"] = np.array(lens, dtype=np.int32)
        out[f"case{ci}_K"] = bits(K)
        out[f"case{ci}_V"] = bits(V)
    np.savez_compressed(os.path.join(HERE, "caches.npz"), **out)
    print("caches.npz:", len(CACHE_CASES), "cases")


def make_siglip():
    s = ref_loader.load_reference_siglip()
    dims = o.dims_preset("micro72")
    v = dims.vision
    cfg = s.SigLipVisionConfig(hidden_size=v.hidden, intermediate_size=v.inter, num_hidden_layers=v.layers + 1,
                               num_attention_heads=v.heads, image_size=v.image, patch_size=v.patch)
    m = s.SigLipVisionModel(cfg)
    del m.vision_model.encoder.layers[-1:]
    m.vision_model.head = torch.nn.Identity()
    w = o.init_weights(dims, 3, dtype=torch.float32)
    sd = {k[len("model.vision_tower.vision_tower."):]: t for k, t in w.items() if k.startswith(o.VT)}
    m.load_state_dict(sd, strict=False)
    frames = o.synthetic_frames(2, v.image)
    pv = s.SigLipImageProcessor(size=(v.image, v.image)).preprocess(frames, return_tensors="pt")["pixel_values"]
    with torch.no_grad():
        hid = m.eval()(pv, output_hidden_states=True).hidden_states[-1]
    lut = s.SigLipImageProcessor(size=(16, 16)).preprocess(
        torch.arange(256, dtype=torch.uint8).repeat(3).view(1, 3, 16, 16), return_tensors="pt")["pixel_values"][0, 0].reshape(-1)
    np.savez_compressed(os.path.join(HERE, "siglip.npz"), frames=frames.numpy(), pixel_values_sample=pv[0, :, :4, :4].numpy(),
                        lut=lut.numpy(), hidden=hid.numpy().astype(np.float32))
    print("siglip.npz:", hid.shape)


def tick_stream(k, T):
    rng = np.random.default_rng(100 + k)
    t = np.arange(T)
    e = rng.standard_normal(T)
    ar = np.zeros(T)
    for i in range(1, T):
        ar[i] = 0.9 * ar[i - 1] + e[i]
    x = np.clip(0.5 + 0.25 * np.sin(2 * np.pi * t / rng.integers(50, 500)) + 0.03 * ar + 0.02 * rng.standard_normal(T), 0, 1)
    return np.round(x, 1) if k % 4 == 0 else x


def make_ticks():
    from scipy.signal import find_peaks, savgol_filter
    out = {}
    for k, (T, fps) in enumerate([(64, 1.0), (300, 1.0), (300, 2.0), (1000, 1.0), (15, 1.0), (777, 0.5), (3300, 1.0), (500, 1.0)]):
        x = tick_stream(k, T)
        sm = savgol_filter(np.array(list(x)), 15, 3)                     # test/live_infer_for_video.py:209
        thresh = sm.mean() + 0.5 * sm.std()                              # :216-217
        peaks, _ = find_peaks(sm, height=thresh, prominence=0.02, distance=int(10 * fps))   # :213-221
        out[f"s{k}_x"] = x
        out[f"s{k}_fps"] = np.array(fps)
        out[f"s{k}_peaks"] = peaks.astype(np.int32)
        out[f"s{k}_smoothed"] = sm
        out[f"s{k}_thresh"] = np.array(thresh)
    np.savez_compressed(os.path.join(HERE, "ticks.npz"), **out)
    print("ticks.npz: 8 streams")


if __name__ == "__main__":
    make_ticks()
    make_caches()
    make_siglip()
