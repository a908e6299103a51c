"""find_ticks (test/live_infer_for_video.py:195-228, test/live_video/infer_live_video.py:70-91) on the GPU:
Savitzky-Golay(15, 3) smoothing + scipy.signal.find_peaks semantics in one kernel launch for any number of streams."""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from . import _lib


def find_ticks_batched(scores: torch.Tensor, fps: float = 1.0, prominence: float = 0.02, thresh=None,
                       max_peaks: Optional[int] = None, return_smoothed: bool = False):
    """scores: [N, T] float64 CUDA tensor.  Returns (peaks [N, max_peaks] int32, n_peaks [N] int32) (+ smoothed, thresh).
    `distance = int(10 * fps)` exactly as the reference (its min_separation argument is overwritten with 10)."""
    if scores.dim() != 2 or scores.dtype != torch.float64 or not scores.is_cuda:
        raise TypeError("scores must be a [N, T] float64 CUDA tensor")
    scores = scores.contiguous()
    n, T = scores.shape
    distance = int(10 * fps)
    if max_peaks is None:
        max_peaks = T // max(distance, 2) + 2
    peaks = torch.full((n, max_peaks), -1, dtype=torch.int32, device=scores.device)
    n_peaks = torch.zeros(n, dtype=torch.int32, device=scores.device)
    smoothed = torch.empty_like(scores) if return_smoothed else None
    th_out = torch.empty(n, dtype=torch.float64, device=scores.device) if return_smoothed else None
    th_in = None
    if thresh is not None and thresh is not False:
        th_in = torch.as_tensor(thresh, dtype=torch.float64, device=scores.device).expand(n).contiguous()
    _lib.check(_lib.load().aha_find_ticks(_lib.ptr(scores), n, T, float(prominence), distance, _lib.ptr(th_in),
                                          _lib.ptr(peaks), max_peaks, _lib.ptr(n_peaks), _lib.ptr(smoothed),
                                          _lib.ptr(th_out), _lib.stream_ptr()))
    if return_smoothed:
        return peaks, n_peaks, smoothed, th_out
    return peaks, n_peaks


def find_ticks(scores: Sequence[float], fps: float = 1.0, min_separation=10, prominence: float = 0.02, thresh=False,
               verbose: bool = False, device="cuda") -> List[float]:
    """Drop-in for LiveInferForDemo.find_ticks: list of peak times in seconds."""
    x = torch.as_tensor(list(scores) if not isinstance(scores, torch.Tensor) else scores, dtype=torch.float64,
                        device=device).reshape(1, -1)
    peaks, n = find_ticks_batched(x, fps=fps, prominence=prominence, thresh=(thresh if thresh else None))
    k = int(n.item())
    if k > peaks.shape[1]:
        raise RuntimeError("peak buffer overflow")
    ticks = [float(p / fps) for p in peaks[0, :k].tolist()]
    if verbose:
        print(f"Detected ticks {ticks}")
    return ticks
