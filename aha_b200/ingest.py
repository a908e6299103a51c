"""Frame ingest on the GPU, same names, argument meaning and return values as the reference:

    load_video_for_testing   test/inference.py:497-582       fps resampling + cv2 bilinear resize-keep-aspect + black
                                                             border to 384 x 384 + BGR->RGB + CHW
    load_one_frame           test/live_infer_for_video.py:98-119  PIL bicubic resize-keep-aspect + black border + CHW

Video DECODE stays with the caller (no NVDEC is assumed): the functions take the decoded frames ([T, h, w, 3] uint8,
BGR for the cv2 flavour as cv2.VideoCapture yields them, RGB for the PIL flavour) plus the container's fps and frame
count, and do everything after `cap.read()` on the device in one kernel launch per batch.  Outputs are bit-identical to
the reference's uint8 tensors (tests/golden/ingest.npz holds cv2's and PIL's own outputs).  No CPU fallback."""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib

FLAVOURS = {"cv2": 0, "pil": 1}
OUTPUT_RESOLUTION = 384          # test/inference.py:505, test/live_infer_for_video.py:101


def fit_geometry(width: int, height: int, out: int = OUTPUT_RESOLUTION) -> Tuple[int, int, int, int]:
    """(new_width, new_height, left, top): the longer side becomes `out` (test/inference.py:540-556)."""
    if width > height:
        new_w, new_h = out, int((height / width) * out)
    else:
        new_h, new_w = out, int((width / height) * out)
    return new_w, new_h, (out - new_w) // 2, (out - new_h) // 2


def ingest_frames(frames_hwc: torch.Tensor, flavour: str = "cv2", out: int = OUTPUT_RESOLUTION) -> torch.Tensor:
    """[n, h, w, 3] uint8 CUDA tensor -> [n, 3, out, out] uint8 CUDA tensor (the reference's canvas)."""
    if flavour not in FLAVOURS:
        raise ValueError(f"flavour must be one of {sorted(FLAVOURS)}")
    if frames_hwc.dtype != torch.uint8 or not frames_hwc.is_cuda:
        raise TypeError("frames must be a uint8 CUDA tensor")
    if frames_hwc.dim() != 4 or frames_hwc.shape[-1] != 3:
        raise ValueError("frames must be [n, height, width, 3]")
    src = frames_hwc.contiguous()
    n, h, w, _ = src.shape
    canvas = torch.empty(n, 3, out, out, dtype=torch.uint8, device=src.device)
    with torch.cuda.device(src.device):
        _lib.check(_lib.load().aha_ingest_frames(_lib.ptr(src), n, h, w, FLAVOURS[flavour], _lib.ptr(canvas), out,
                                                 _lib.stream_ptr()))
    return canvas


def select_frames(input_fps: float, frame_count: float, output_fps: float, max_num_frames: Optional[int] = None,
                  frames_readable: Optional[int] = None) -> Tuple[List[int], float, float]:
    """Which decoded frames load_video_for_testing keeps (test/inference.py:512-570): frame i is kept when the running
    `cur_time` (accumulated as `cur_time += 1 / input_fps`, float64) has reached `frame_index / output_fps`.
    Returns (kept indices, output_fps, video_duration)."""
    video_duration = frame_count / input_fps
    output_fps = output_fps if output_fps > 0 else max_num_frames / video_duration
    num_frames_total = math.ceil(video_duration * output_fps)
    n_read = int(frame_count) if frames_readable is None else int(frames_readable)
    kept: List[int] = []
    cur_time, target = 0, 0
    for i in range(n_read):
        if target < num_frames_total and cur_time >= target / output_fps:
            kept.append(i)
            target += 1
        if max_num_frames and len(kept) >= max_num_frames:
            break
        cur_time += 1 / input_fps
    return kept, output_fps, video_duration


def load_video_for_testing(decoded_bgr: torch.Tensor, input_fps: float, output_fps: float = 2,
                           return_true_frames: bool = False, max_num_frames: Optional[int] = None,
                           frame_count: Optional[float] = None, device="cuda", batch: int = 256):
    """The reference's loader after `cap.read()`: `decoded_bgr` [T, h, w, 3] uint8 (host or device) holds every frame
    the capture yields.  Returns (frames [K, 3, 384, 384] uint8 on `device`, output_fps, video_duration) and, with
    `return_true_frames`, the kept source indices - or the reference's all-None tuple when nothing is kept
    (test/inference.py:575-582).  Only the kept frames are copied to the device."""
    T = decoded_bgr.shape[0]
    kept, output_fps, duration = select_frames(input_fps, T if frame_count is None else frame_count, output_fps,
                                               max_num_frames, frames_readable=T)
    if not kept:
        return None, None, None, None
    outs = []
    for s in range(0, len(kept), batch):
        chunk = kept[s:s + batch]
        if decoded_bgr.is_cuda:
            part = decoded_bgr.index_select(0, torch.as_tensor(chunk, dtype=torch.long, device=decoded_bgr.device))
        else:
            # frame by frame straight into the device batch (asynchronous when the clip sits in pinned memory): no
            # gathered or re-pinned copy of the clip on the host
            part = torch.empty((len(chunk),) + tuple(decoded_bgr.shape[1:]), dtype=torch.uint8, device=device)
            for j, i in enumerate(chunk):
                part[j].copy_(decoded_bgr[i], non_blocking=True)
        outs.append(ingest_frames(part, "cv2"))
    frames = outs[0] if len(outs) == 1 else torch.cat(outs)
    if return_true_frames:
        return frames, output_fps, duration, kept
    return frames, output_fps, duration


def load_one_frame(frame_rgb: torch.Tensor, device="cuda", out: int = OUTPUT_RESOLUTION) -> torch.Tensor:
    """LiveInferForDemo.load_one_frame up to the image processor: one decoded RGB frame [h, w, 3] uint8 -> [1, 3, 384,
    384] uint8 on the device (feed it to `model.visual_embed`, which fuses the processor's rescale / normalise)."""
    if frame_rgb.dim() != 3:
        raise ValueError("frame must be [height, width, 3]")
    return ingest_frames(frame_rgb.to(device).unsqueeze(0), "pil", out)
