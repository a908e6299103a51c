"""TEST INFRASTRUCTURE ONLY (imported by tests/, tests/golden/make_golden_ingest.py and nothing under aha_b200/).

CPU restatement (numpy, integer arithmetic) of the reference's frame ingest, SURVEY.md 8(f) row N4:

    load_video_for_testing   /root/reference/test/inference.py:497-582   cv2.resize (INTER_LINEAR) keeping the aspect
                             ratio -> cv2.copyMakeBorder to 384 x 384 with black -> BGR2RGB -> CHW; fps resampling by
                             the `cur_time >= frame_sec[frame_index]` rule (:529-570)
    load_one_frame           /root/reference/test/live_infer_for_video.py:98-118   PIL Image.resize (default filter:
                             BICUBIC) -> ImageOps.expand with black -> CHW

The resampling arithmetic lives in third-party libraries that are not under /root/reference: OpenCV (opencv-python,
requirements.txt; 4.13.0 is installed here) and Pillow (12.2.0 installed here).  Their published algorithms are
restated below:
  * OpenCV resize.cpp, 8-bit INTER_LINEAR: coordinates `fx = (float)((dx + 0.5) * scale - 0.5)`, 11-bit fixed-point
    coefficients `saturate_cast<short>(w * 2048)`, horizontal pass exact in int32, vertical pass
    `(((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2`; when both scale factors are exactly 2 the call is
    redirected to INTER_AREA (2 x 2 box mean, `(sum + 2) >> 2`).
  * Pillow Resample.c, 8 bits per channel: bicubic (a = -0.5) with the support stretched by the down-scale factor,
    coefficients normalised in double and rounded to 22-bit fixed point, horizontal pass into an 8-bit image, then the
    vertical pass, each `clip8((sum + 2^21) >> 22)`.
Versions: the reference pins opencv-python 4.10.0.84 and pillow 10.4.0 (requirements.txt:33,36); 4.13.0 / 12.2.0 are what
this image has.  The one known difference on this path: Pillow 12.2 resamples pictures more than 100 times taller than
wide vertical pass first (PIL/Image.py, resize); that geometry is refused (ValueError) instead of guessed.
PINNED: both restatements are executed against cv2 / PIL themselves (tests/golden/make_golden_ingest.py, run in the
build container; the goldens in tests/golden/ingest.npz are the libraries' outputs, not this file's)."""
from __future__ import annotations

import math
from typing import List, Optional, Tuple

import numpy as np

OUT = 384


def fit_geometry(width: int, height: int, out: int = OUT) -> Tuple[int, int, int, int]:
    """(new_width, new_height, left, top) of test/inference.py:540-556 / live_infer_for_video.py:107-117."""
    if width > height:
        new_w, new_h = out, int((height / width) * out)
    else:
        new_h, new_w = out, int((width / height) * out)
    return new_w, new_h, (out - new_w) // 2, (out - new_h) // 2


# ---------------------------------------------------------------- OpenCV INTER_LINEAR, 8-bit
def cv2_linear_table(src: int, dst: int) -> Tuple[np.ndarray, np.ndarray]:
    """index [dst] (left tap, already clamped) and int16 coefficient pairs [dst, 2] of resize.cpp."""
    scale = 1.0 / (dst / src)                       # cv::resize: inv_scale = dsize / ssize; scale = 1. / inv_scale
    idx = np.zeros(dst, np.int32)
    coef = np.zeros((dst, 2), np.int32)
    for d in range(dst):
        f = np.float32((d + 0.5) * scale - 0.5)
        s = int(math.floor(f))
        f = np.float32(f - np.float32(s))
        if s < 0:
            s, f = 0, np.float32(0.0)
        if s >= src - 1:
            s, f = src - 1, np.float32(0.0)
        idx[d] = s
        # saturate_cast<short>(float) = cvRound: round half to even
        coef[d, 0] = int(np.rint(np.float32(np.float32(1.0) - f) * np.float32(2048.0)))
        coef[d, 1] = int(np.rint(f * np.float32(2048.0)))
    return idx, coef


def cv2_linear_rows(src: int, dst: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """vertical taps: rows are clipped, the fraction is NOT reset at the borders (resize.cpp, resizeGeneric_)."""
    scale = 1.0 / (dst / src)
    r0 = np.zeros(dst, np.int32); r1 = np.zeros(dst, np.int32); coef = np.zeros((dst, 2), np.int32)
    for d in range(dst):
        f = np.float32((d + 0.5) * scale - 0.5)
        s = int(math.floor(f))
        f = np.float32(f - np.float32(s))
        r0[d] = min(max(s, 0), src - 1)
        r1[d] = min(max(s + 1, 0), src - 1)
        coef[d, 0] = int(np.rint(np.float32(np.float32(1.0) - f) * np.float32(2048.0)))
        coef[d, 1] = int(np.rint(f * np.float32(2048.0)))
    return r0, r1, coef


def cv2_resize_linear(img: np.ndarray, new_w: int, new_h: int) -> np.ndarray:
    """img [h, w, c] uint8 -> [new_h, new_w, c] uint8, bit-identical to cv2.resize(img, (new_w, new_h))."""
    h, w, _ = img.shape
    if (new_w, new_h) == (w, h):
        return img.copy()
    if w == 2 * new_w and h == 2 * new_h:           # INTER_LINEAR with both factors 2 runs the INTER_AREA fast path
        s = img.astype(np.int32)
        return ((s[0::2, 0::2] + s[0::2, 1::2] + s[1::2, 0::2] + s[1::2, 1::2] + 2) >> 2).astype(np.uint8)
    xi, xc = cv2_linear_table(w, new_w)
    r0, r1, yc = cv2_linear_rows(h, new_h)
    s = img.astype(np.int32)
    xi1 = np.minimum(xi + 1, w - 1)
    hor = s[:, xi, :] * xc[None, :, 0, None] + s[:, xi1, :] * xc[None, :, 1, None]          # [h, new_w, c] int32
    a = (yc[:, 0, None, None] * (hor[r0] >> 4)) >> 16
    b = (yc[:, 1, None, None] * (hor[r1] >> 4)) >> 16
    return ((a + b + 2) >> 2).astype(np.uint8)


# ---------------------------------------------------------------- Pillow bicubic, 8 bits per channel
def _bicubic(x: float) -> float:
    a = -0.5
    x = abs(x)
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


def pil_bicubic_table(src: int, dst: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """(first source index [dst], tap count [dst], 22-bit coefficients [dst, ksize]) of Resample.c precompute_coeffs +
    normalize_coeffs_8bpc."""
    scale = filterscale = src / dst
    if filterscale < 1.0:
        filterscale = 1.0
    support = 2.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    lo = np.zeros(dst, np.int32); cnt = np.zeros(dst, np.int32); kk = np.zeros((dst, ksize), np.int32)
    ss = 1.0 / filterscale
    for xx in range(dst):
        center = (xx + 0.5) * scale
        xmin = int(center - support + 0.5)
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > src:
            xmax = src
        xmax -= xmin
        ws = [_bicubic((x + xmin - center + 0.5) * ss) for x in range(xmax)]
        ww = 0.0
        for v in ws:
            ww += v
        for x in range(xmax):
            v = ws[x] / ww if ww != 0.0 else ws[x]
            kk[xx, x] = int(v * (1 << 22) - 0.5) if v < 0 else int(v * (1 << 22) + 0.5)
        lo[xx], cnt[xx] = xmin, xmax
    return lo, cnt, kk


def _pil_pass(img: np.ndarray, dst: int, axis: int) -> np.ndarray:
    src = img.shape[axis]
    lo, cnt, kk = pil_bicubic_table(src, dst)
    x = np.moveaxis(img, axis, 0).astype(np.int64)
    out = np.zeros((dst,) + x.shape[1:], np.int64)
    for d in range(dst):
        acc = np.full(x.shape[1:], 1 << 21, np.int64)
        for t in range(cnt[d]):
            acc += x[lo[d] + t] * int(kk[d, t])
        out[d] = np.clip(acc >> 22, 0, 255)
    return np.moveaxis(out, 0, axis).astype(np.uint8)


def pil_resize_bicubic(img: np.ndarray, new_w: int, new_h: int) -> np.ndarray:
    """img [h, w, c] uint8 -> [new_h, new_w, c], bit-identical to PIL.Image.resize((new_w, new_h)) (BICUBIC)."""
    h, w, _ = img.shape
    if h > w * 100 and new_h < h:
        # Pillow 12.2 Image.resize runs the vertical pass first for such pictures; the reference's pinned 10.4.0
        # (requirements.txt:36) predates that special case.  Version-dependent: refused here and by the kernel.
        raise ValueError("height > 100 x width: Pillow-version-dependent pass order")
    out = img
    if new_w != w:
        out = _pil_pass(out, new_w, 1)
    if new_h != h:
        out = _pil_pass(out, new_h, 0)
    return out.copy()


# ---------------------------------------------------------------- the two loaders
def ingest_frame(img_hwc: np.ndarray, flavour: str, out: int = OUT) -> np.ndarray:
    """One decoded frame [h, w, 3] uint8 -> the reference's [3, out, out] uint8 canvas.
    flavour 'cv2': BGR input, INTER_LINEAR, channel flip (test/inference.py:540-563);
    flavour 'pil': RGB input, bicubic (live_infer_for_video.py:107-119)."""
    h, w, _ = img_hwc.shape
    new_w, new_h, left, top = fit_geometry(w, h, out)
    if flavour == "cv2":
        r = cv2_resize_linear(img_hwc, new_w, new_h)[:, :, ::-1]
    elif flavour == "pil":
        r = pil_resize_bicubic(img_hwc, new_w, new_h)
    else:
        raise ValueError(flavour)
    canvas = np.zeros((out, out, 3), np.uint8)
    canvas[top:top + new_h, left:left + new_w] = r
    return np.ascontiguousarray(canvas.transpose(2, 0, 1))


def select_frames(input_fps: float, frame_count: float, output_fps: float, max_num_frames: Optional[int] = None,
                  frames_readable: Optional[int] = None) -> Tuple[List[int], float, float]:
    """Indices of the decoded frames load_video_for_testing keeps (test/inference.py:512-570), with the same float64
    accumulation of `cur_time += 1 / input_fps`.  Returns (indices, output_fps, video_duration)."""
    video_duration = frame_count / input_fps
    output_fps = output_fps if output_fps > 0 else max_num_frames / video_duration
    num_frames_total = math.ceil(video_duration * output_fps)
    frame_sec = [i / output_fps for i in range(num_frames_total)]
    kept: List[int] = []
    cur_time, frame_index = 0, 0
    n_read = int(frame_count) if frames_readable is None else frames_readable
    for true_index in range(n_read):
        if frame_index < len(frame_sec) and cur_time >= frame_sec[frame_index]:
            kept.append(true_index)
            frame_index += 1
        if max_num_frames and len(kept) >= max_num_frames:
            break
        cur_time += 1 / input_fps
    return kept, output_fps, video_duration
