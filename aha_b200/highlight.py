"""Score fusion and highlight selection on the GPU, same names and argument meaning as the
reference:

    fuse_scores                test/grid_search.py:58-65 / :95-104  (alpha*inf + beta*rel - uncertainty penalty)
    knapsack_selection         test/highlight_generator.py:8-37     (0/1 knapsack over frames, cost 1 each)
    highlight_indices          test/highlight_generator.py:76-86    (selected frames widened to half-second windows)

Everything is float64 on the device with the reference's evaluation order, so results are bit-identical to the Python
code.  The batched entry points take [N, T] tensors (N independent score streams)."""
from __future__ import annotations

from typing import Dict, List, Sequence, Set

import torch

from . import _lib


def _f64(x, device):
    t = torch.as_tensor(x, dtype=torch.float64, device=device)
    return t.contiguous()


def fuse_scores_batched(informative: torch.Tensor, relevance: torch.Tensor, uncertainty: torch.Tensor, alpha: float,
                        beta: float, epsilon: float, uncertainty_threshold: float) -> torch.Tensor:
    """[N, T] float64 CUDA tensors -> fused scores [N, T] float64."""
    for t in (informative, relevance, uncertainty):
        if t.dtype != torch.float64 or not t.is_cuda:
            raise TypeError("scores must be float64 CUDA tensors")
    inf, rel, unc = informative.contiguous(), relevance.contiguous(), uncertainty.contiguous()
    if not (inf.shape == rel.shape == unc.shape):
        raise ValueError("score tensors must have the same shape")
    out = torch.empty_like(inf)
    _lib.check(_lib.load().aha_fuse_scores(_lib.ptr(inf), _lib.ptr(rel), _lib.ptr(unc), inf.numel(), float(alpha),
                                           float(beta), float(epsilon), float(uncertainty_threshold), _lib.ptr(out),
                                           _lib.stream_ptr()))
    return out


def fuse_scores(debug_data: Sequence[Dict[str, float]], alpha: float, beta: float, epsilon: float,
                uncertainty_threshold: float, device="cuda") -> List[float]:
    """The per-frame loop of hisum/tvsum_score_calculation on one prediction's `debug_data`."""
    inf = _f64([e["informative_score"] for e in debug_data], device).reshape(1, -1)
    rel = _f64([e["relevance_score"] for e in debug_data], device).reshape(1, -1)
    unc = _f64([e["uncertainty_score"] for e in debug_data], device).reshape(1, -1)
    return fuse_scores_batched(inf, rel, unc, alpha, beta, epsilon, uncertainty_threshold)[0].tolist()


def knapsack_select_batched(informative: torch.Tensor, relevance: torch.Tensor, uncertainty: torch.Tensor,
                            max_duration: int, alpha: float, beta: float, epsilon: float) -> torch.Tensor:
    """[N, T] float64 CUDA tensors -> uint8 mask [N, T] of the frames knapsack_selection would return."""
    for t in (informative, relevance, uncertainty):
        if t.dtype != torch.float64 or not t.is_cuda or t.dim() != 2:
            raise TypeError("scores must be [N, T] float64 CUDA tensors")
    inf, rel, unc = informative.contiguous(), relevance.contiguous(), uncertainty.contiguous()
    n, T = inf.shape
    bits = torch.empty(n, T, _lib.load().aha_knapsack_words(int(max_duration)), dtype=torch.int32, device=inf.device)
    sel = torch.empty(n, T, dtype=torch.uint8, device=inf.device)
    _lib.check(_lib.load().aha_knapsack_select(_lib.ptr(inf), _lib.ptr(rel), _lib.ptr(unc), n, T, int(max_duration),
                                               float(alpha), float(beta), float(epsilon), _lib.ptr(bits), _lib.ptr(sel),
                                               _lib.stream_ptr()))
    return sel


def knapsack_selection(frames_with_index: Sequence[Dict], max_duration: int, weight, alpha: float, beta: float,
                       epsilon: float, device="cuda") -> Set[int]:
    """Drop-in for test/highlight_generator.py:8 (the `weight` argument is unused there as well)."""
    inf = _f64([f["informative_score"] for f in frames_with_index], device).reshape(1, -1)
    rel = _f64([f["relevance_score"] for f in frames_with_index], device).reshape(1, -1)
    unc = _f64([f["uncertainty_score"] for f in frames_with_index], device).reshape(1, -1)
    sel = knapsack_select_batched(inf, rel, unc, max_duration, alpha, beta, epsilon)[0].cpu().tolist()
    return set(frames_with_index[i]["idx"] for i, s in enumerate(sel) if s)


def highlight_mask_batched(selected: torch.Tensor, ground_truth_frames: torch.Tensor, fps: float) -> torch.Tensor:
    """selected [N, T] uint8, ground_truth_frames [T] int32 (source-video frame of every scored frame) -> uint8 mask
    [N, last_frame + 1] of the source frames that belong to the highlight."""
    if selected.dtype != torch.uint8 or not selected.is_cuda or selected.dim() != 2:
        raise TypeError("selected must be a [N, T] uint8 CUDA tensor")
    gt = ground_truth_frames.to(device=selected.device, dtype=torch.int32).contiguous()
    n, T = selected.shape
    if gt.numel() != T:
        raise ValueError("ground_truth_frames must have one entry per scored frame")
    last = int(gt[-1].item())
    mask = torch.empty(n, last + 1, dtype=torch.uint8, device=selected.device)
    _lib.check(_lib.load().aha_highlight_mask(_lib.ptr(selected.contiguous()), _lib.ptr(gt), n, T, int(fps // 2), last,
                                              _lib.ptr(mask), _lib.stream_ptr()))
    return mask


def highlight_indices(selected: Set[int], ground_truth_frames: Sequence[int], fps: float, device="cuda") -> List[int]:
    """Sorted source-frame indices of the highlight (test/highlight_generator.py:76-86)."""
    T = len(ground_truth_frames)
    sel = torch.zeros(1, T, dtype=torch.uint8, device=device)
    if selected:
        sel[0, torch.as_tensor(sorted(selected), device=device)] = 1
    mask = highlight_mask_batched(sel, torch.as_tensor(list(ground_truth_frames)), fps)
    return torch.nonzero(mask[0]).flatten().cpu().tolist()
