"""The CUDA path (through the C-ABI wrappers) against the golden fixtures produced by executing the reference itself
(tests/golden/make_golden.py): cache policies bit-exact, peak frames identical, SigLIP tower within bf16 tolerance."""
import os

import numpy as np
import pytest
import torch

import aha_b200 as A
from oracle import aha_oracle as o
from tests.golden import make_golden as G

pytestmark = pytest.mark.gpu
GOLD = os.path.dirname(os.path.abspath(G.__file__))


@pytest.mark.parametrize("ci", range(len(G.CACHE_CASES)))
def test_cuda_cache_matches_reference_golden(cuda, ci):
    gold = np.load(os.path.join(GOLD, "caches.npz"))
    policy, kw, chunks = G.CACHE_CASES[ci]
    c = {"sink": A.SinkCache, "sliding": A.SlidingWindowCache, "static": A.TrulyStaticCache}[policy](**kw)
    lens = []
    for step, q in enumerate(chunks):
        L = c.get_seq_length()
        k, v, cos, sin = G.cache_inputs(ci, step, q, L)
        K, V = c.update(k.to(cuda), v.to(cuda), 0, {"cos": cos.to(cuda), "sin": sin.to(cuda)})
        lens.append([L, K.shape[-2], c.get_seq_length()])
    assert np.array_equal(np.array(lens, dtype=np.int32), gold[f"case{ci}_lens"])
    assert np.array_equal(K.cpu().contiguous().view(torch.int16).numpy(), gold[f"case{ci}_K"])
    assert np.array_equal(V.cpu().contiguous().view(torch.int16).numpy(), gold[f"case{ci}_V"])


def test_cuda_find_ticks_matches_scipy_golden(cuda):
    gold = np.load(os.path.join(GOLD, "ticks.npz"))
    for k in range(8):
        x, fps = gold[f"s{k}_x"], float(gold[f"s{k}_fps"])
        peaks, cnt, sm, th = A.find_ticks_batched(torch.from_numpy(x).to(cuda).view(1, -1), fps=fps, return_smoothed=True)
        assert peaks[0, :int(cnt[0])].cpu().tolist() == gold[f"s{k}_peaks"].tolist(), k
        assert np.abs(sm[0].cpu().numpy() - gold[f"s{k}_smoothed"]).max() < 1e-12
        assert abs(float(th[0]) - float(gold[f"s{k}_thresh"])) < 1e-12


def test_cuda_siglip_tower_matches_reference_golden(cuda):
    """visual_embed = tower -> projector -> pooling; the tower is pinned by the reference's own SigLipVisionModel output:
    the oracle projector/pooling applied to the golden tower output must equal the CUDA visual_embed within tolerance."""
    gold = np.load(os.path.join(GOLD, "siglip.npz"))
    dims, cfg = o.dims_preset("micro72"), A.config_preset("micro72")
    w = o.init_weights(dims, 3, dtype=torch.float32)
    frames = torch.from_numpy(gold["frames"])
    ref = o.post_projector_pooling(dims, o.connector(w, torch.from_numpy(gold["hidden"]))).reshape(-1, cfg.hidden_size)
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, {k: v.bfloat16() for k, v in w.items()}, device=cuda, max_tokens=32, max_frames=2)
    mine = model.visual_embed(frames.to(cuda)).float().cpu()
    assert (mine - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())
