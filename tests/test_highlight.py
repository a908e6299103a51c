"""N3 (score fusion + knapsack highlight selection): the oracle restatement against goldens produced by the reference's own
knapsack_selection (CPU), and the CUDA path against both (GPU, bit-exact)."""
import os

import numpy as np
import pytest
import torch

from oracle import aha_oracle as o

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "highlight.npz"))
NAMES = [str(n) for n in G["names"]]


def _frames(name, s):
    inf, rel, unc = G[f"{name}_inf"][s], G[f"{name}_rel"][s], G[f"{name}_unc"][s]
    return [dict(idx=i, informative_score=float(inf[i]), relevance_score=float(rel[i]), uncertainty_score=float(unc[i]))
            for i in range(len(inf))]


@pytest.mark.parametrize("name", NAMES)
def test_oracle_matches_reference_golden(name):
    budget, alpha, beta, eps, thr = G[f"{name}_par"]
    for s in range(G[f"{name}_inf"].shape[0]):
        fr = _frames(name, s)
        sel = o.knapsack_selection(fr, int(budget), alpha, beta, eps)
        assert sel == set(np.nonzero(G[f"{name}_sel"][s])[0].tolist())
        assert o.fuse_scores(fr, alpha, beta, eps, thr) == G[f"{name}_fused"][s].tolist()


def test_oracle_highlight_indices():
    gt = list(range(0, 300, 30))
    assert o.highlight_indices({0, 3, 9}, gt, 30.0) == sorted(set(range(0, 16)) | set(range(75, 106)) | set(range(255, 271)))
    assert o.highlight_indices(set(), gt, 30.0) == []


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_cuda_matches_reference_golden(cuda, name):
    import aha_b200 as A
    budget, alpha, beta, eps, thr = G[f"{name}_par"]
    inf, rel, unc = (torch.from_numpy(G[f"{name}_{k}"]).to(cuda) for k in ("inf", "rel", "unc"))
    sel = A.knapsack_select_batched(inf, rel, unc, int(budget), alpha, beta, eps)
    assert np.array_equal(sel.cpu().numpy(), G[f"{name}_sel"])
    fused = A.fuse_scores_batched(inf, rel, unc, alpha, beta, eps, thr)
    assert np.array_equal(fused.cpu().numpy(), G[f"{name}_fused"])          # bit-exact float64
    fr = _frames(name, 0)
    assert A.knapsack_selection(fr, int(budget), "relevance_score", alpha, beta, eps) == \
        set(np.nonzero(G[f"{name}_sel"][0])[0].tolist())
    assert A.fuse_scores(fr, alpha, beta, eps, thr) == G[f"{name}_fused"][0].tolist()


@pytest.mark.gpu
def test_cuda_highlight_pipeline_vs_oracle(cuda):
    import aha_b200 as A
    g = np.random.default_rng(11)
    n, T, fps = 64, 400, 30.0
    inf = g.random((n, T)).astype(np.float32).astype(np.float64)
    rel = g.random((n, T)).astype(np.float32).astype(np.float64)
    unc = (g.random((n, T)) * 0.2).astype(np.float32).astype(np.float64)
    gt = (np.arange(T) * 29 + 7).astype(np.int32)
    ti, tr, tu = (torch.from_numpy(x).to(cuda) for x in (inf, rel, unc))
    sel = A.knapsack_select_batched(ti, tr, tu, 20, 0.0, -1.0, -5.0)
    mask = A.highlight_mask_batched(sel, torch.from_numpy(gt), fps).cpu().numpy()
    for s in range(0, n, 7):
        fr = [dict(idx=i, informative_score=float(inf[s, i]), relevance_score=float(rel[s, i]),
                   uncertainty_score=float(unc[s, i])) for i in range(T)]
        ref_sel = o.knapsack_selection(fr, 20, 0.0, -1.0, -5.0)
        assert set(np.nonzero(sel[s].cpu().numpy())[0].tolist()) == ref_sel
        assert np.nonzero(mask[s])[0].tolist() == o.highlight_indices(ref_sel, gt.tolist(), fps)
    assert A.highlight_indices({0, 3}, gt.tolist(), fps) == o.highlight_indices({0, 3}, gt.tolist(), fps)


@pytest.mark.gpu
@pytest.mark.parametrize("budget", [31, 32, 33, 40, 64, 100, 250])
def test_cuda_knapsack_large_budgets(cuda, budget):
    """Budgets beyond one decision word per frame (the reference's knapsack_selection takes any budget; its driver uses
    20): several 32-bit words per frame, dp[b - 1] of lane 0 carried over from lane 31 of the word below.  Values with
    mixed signs, so that fewer than `budget` frames may be worth taking; T both above and below the budget."""
    import aha_b200 as A
    g = np.random.default_rng(100 + budget)
    for T in (budget + 57, max(8, budget - 9)):
        n = 5
        inf = (g.random((n, T)) - 0.4).astype(np.float32).astype(np.float64)
        rel = (g.random((n, T)) - 0.5).astype(np.float32).astype(np.float64)
        unc = (g.random((n, T)) * 0.3).astype(np.float32).astype(np.float64)
        sel = A.knapsack_select_batched(*(torch.from_numpy(x).to(cuda) for x in (inf, rel, unc)), budget, 0.7, 0.6, -1.5)
        for s_ in range(n):
            fr = [dict(idx=i, informative_score=float(inf[s_, i]), relevance_score=float(rel[s_, i]),
                       uncertainty_score=float(unc[s_, i])) for i in range(T)]
            assert set(np.nonzero(sel[s_].cpu().numpy())[0].tolist()) == o.knapsack_selection(fr, budget, 0.7, 0.6, -1.5), (T, s_)


@pytest.mark.gpu
def test_cuda_knapsack_errors(cuda):
    import aha_b200 as A
    x = torch.zeros(1, 8, dtype=torch.float64, device=cuda)
    with pytest.raises(ValueError, match="max_duration"):
        A.knapsack_select_batched(x, x, x, 1024, 1.0, 1.0, 1.0)
    with pytest.raises(TypeError):
        A.knapsack_select_batched(x.float(), x, x, 4, 1.0, 1.0, 1.0)
