"""find_ticks kernel (through the C-ABI) against scipy itself: identical peak frames."""
import numpy as np
import pytest
import torch

from aha_b200 import find_ticks, find_ticks_batched
from oracle import aha_oracle as o

pytestmark = pytest.mark.gpu


def _streams(n, T, seed=7, plateaus=False):
    rng = np.random.default_rng(seed)
    t = np.arange(T)
    out = np.empty((n, T))
    for k in range(n):
        e = rng.standard_normal(T)
        ar = np.zeros(T)
        for i in range(1, T):
            ar[i] = 0.9 * ar[i - 1] + e[i]
        x = 0.5 + 0.25 * np.sin(2 * np.pi * t / rng.integers(50, 500)) + 0.1 * ar * 0.3 + 0.02 * rng.standard_normal(T)
        x = np.clip(x, 0, 1)
        if plateaus and k % 2 == 0:
            x = np.round(x, 1)
        out[k] = x
    return out


@pytest.mark.parametrize("n,T,fps,plateaus", [(8, 64, 1.0, False), (16, 3300, 1.0, False), (8, 1000, 2.0, True),
                                               (4, 15, 1.0, False), (6, 777, 0.5, True), (4, 10000, 1.0, False),
                                               (3, 129, 1.0, False), (2, 30011, 1.0, False)])
def test_find_ticks_identical_to_scipy(cuda, n, T, fps, plateaus):
    x = _streams(n, T, plateaus=plateaus)
    peaks, cnt, sm, th = find_ticks_batched(torch.from_numpy(x).to(cuda), fps=fps, return_smoothed=True)
    peaks, cnt, sm, th = peaks.cpu().numpy(), cnt.cpu().numpy(), sm.cpu().numpy(), th.cpu().numpy()
    for k in range(n):
        ticks, smoothed, thresh = o.find_ticks(x[k], fps=fps)
        ref = [int(round(tk * fps)) for tk in ticks]
        assert np.abs(sm[k] - smoothed).max() < 1e-12
        # mean and std are added up in numpy's pairwise order: on the kernel's own smoothed signal the threshold is the
        # very double numpy computes (and within rounding of scipy's, whose filter differs in the last bit at most)
        assert th[k] == sm[k].mean() + 0.5 * sm[k].std(), (k, th[k], sm[k].mean() + 0.5 * sm[k].std())
        assert abs(th[k] - thresh) < 1e-12
        assert cnt[k] == len(ref), (k, cnt[k], len(ref))
        assert peaks[k, :cnt[k]].tolist() == ref, k


def test_find_ticks_dropin(cuda):
    x = _streams(1, 400, seed=3)[0]
    assert find_ticks(list(x), fps=1.0) == o.find_ticks(x, fps=1.0)[0]
    assert find_ticks(list(x), fps=1.0, thresh=0.55) == o.find_ticks(x, fps=1.0, thresh=0.55)[0]
    with pytest.raises(ValueError):
        find_ticks(list(x[:10]), fps=1.0)


def test_find_ticks_c5_property(cuda):
    """BASELINE config 5 size (4096 x 10000): scipy on a sample of streams + global invariants on all of them."""
    g = torch.Generator(device=cuda).manual_seed(7)
    T, n = 10000, 4096
    t = torch.arange(T, device=cuda, dtype=torch.float64)
    per = torch.randint(50, 500, (n, 1), generator=g, device=cuda).double()
    x = (0.5 + 0.25 * torch.sin(2 * torch.pi * t / per) + 0.05 * torch.randn(n, T, generator=g, device=cuda, dtype=torch.float64)).clamp(0, 1)
    peaks, cnt, sm, th = find_ticks_batched(x, fps=1.0, return_smoothed=True)
    cnt_h = cnt.cpu().numpy(); peaks_h = peaks.cpu().numpy()
    assert (cnt_h <= peaks.shape[1]).all()
    for k in range(0, n, 512):
        ref = [int(p) for p in o.find_ticks(x[k].cpu().numpy(), fps=1.0)[0]]
        assert peaks_h[k, :cnt_h[k]].tolist() == ref
    # invariants: sorted, >= 10 apart, above the threshold
    for k in range(0, n, 64):
        p = peaks_h[k, :cnt_h[k]]
        assert (np.diff(p) >= 10).all()
        assert (sm[k].cpu().numpy()[p] >= th[k].item()).all()
