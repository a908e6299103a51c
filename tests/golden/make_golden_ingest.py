"""Generates tests/golden/ingest.npz by EXECUTING
  * the reference's own load_video_for_testing - its source is cut out of /root/reference/test/inference.py with `ast`
    (the module itself imports thop / h5py / matplotlib, absent here) and run unchanged against the real cv2 4.13, with
    only cv2.VideoCapture replaced by an in-memory capture that yields seeded frames;
  * PIL 12.2 with the call sequence of LiveInferForDemo.load_one_frame (/root/reference/test/live_infer_for_video.py:
    98-119; a method of a class that needs the model, so its five image lines are replayed here).
Stored: the libraries' outputs (full tensors for two cases, SHA-256 for the rest; the sources are regenerated from the
seeds with numpy's PCG64).  Run in the build container:   python tests/golden/make_golden_ingest.py"""
import ast
import hashlib
import math
import os
import types

import cv2
import numpy as np
import torch
from PIL import Image, ImageOps

REF = "/root/reference/test/inference.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ingest.npz")

GEOMETRIES = [(640, 360), (360, 640), (768, 768), (200, 150), (384, 384), (1001, 563), (1920, 1080), (97, 301),
              (768, 432), (500, 500), (385, 384), (1280, 720), (3, 2)]
FPS_CASES = [(29.97, 300, 2.0, None), (30.0, 300, 1.0, None), (25.0, 1000, 2.0, 40), (23.976, 500, 0.5, None),
             (30.0, 90, -1.0, 20), (10.0, 35, 2.0, None), (60.0, 600, 7.0, 33), (30.0, 30, 2.0, None)]


def source_frame(w, h, seed):
    return np.random.default_rng(seed).integers(0, 256, (h, w, 3), dtype=np.uint8)


class FakeCapture:
    """cv2.VideoCapture stand-in: frames from memory, properties as given."""
    registry = {}

    def __init__(self, name):
        self.frames, self.fps, self.count = FakeCapture.registry[name]
        self.i = 0

    def get(self, prop):
        if prop == cv2.CAP_PROP_FPS:
            return self.fps
        if prop == cv2.CAP_PROP_FRAME_COUNT:
            return self.count
        if prop == cv2.CAP_PROP_FRAME_WIDTH:
            return float(self.frames.shape[2])
        if prop == cv2.CAP_PROP_FRAME_HEIGHT:
            return float(self.frames.shape[1])
        raise KeyError(prop)

    def isOpened(self):
        return True

    def read(self):
        if self.i >= len(self.frames):
            return False, None
        f = self.frames[self.i]
        self.i += 1
        return True, f.copy()

    def release(self):
        pass


def reference_loader():
    tree = ast.parse(open(REF).read())
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "load_video_for_testing")
    shim = types.SimpleNamespace(**{k: getattr(cv2, k) for k in dir(cv2) if not k.startswith("__")})
    shim.VideoCapture = FakeCapture
    ns = {"cv2": shim, "math": math, "np": np, "torch": torch}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), REF, "exec"), ns)
    return ns["load_video_for_testing"]


def pil_one_frame(frame_rgb):
    frame = Image.fromarray(frame_rgb)
    output_resolution = 384
    width, height = frame.size
    if width > height:
        new_width = output_resolution
        new_height = int((height / width) * output_resolution)
    else:
        new_height = output_resolution
        new_width = int((width / height) * output_resolution)
    resized_frame = frame.resize((new_width, new_height))
    left = (output_resolution - new_width) // 2
    right = (output_resolution - new_width + 1) // 2
    top = (output_resolution - new_height) // 2
    bottom = (output_resolution - new_height + 1) // 2
    canvas = ImageOps.expand(resized_frame, border=(left, top, right, bottom), fill=(0, 0, 0))
    return np.transpose(np.array(canvas), (2, 0, 1))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    load = reference_loader()
    out = {"geometries": np.array(GEOMETRIES, np.int64), "fps_cases": np.array(
        [[a, b, c, -1 if d is None else d] for a, b, c, d in FPS_CASES], np.float64)}
    cv_hash, pil_hash = [], []
    for i, (w, h) in enumerate(GEOMETRIES):
        src = source_frame(w, h, 100 + i)
        FakeCapture.registry["one"] = (src[None], 1.0, 1.0)
        frames, _, _ = load("one", output_fps=1)
        a = frames.numpy()[0]
        b = pil_one_frame(src)
        cv_hash.append(sha(a)); pil_hash.append(sha(b))
        if i == 0:
            out["full_cv2"], out["full_pil"] = a, b
    out["cv2_sha256"], out["pil_sha256"] = np.array(cv_hash), np.array(pil_hash)
    # a whole clip: 45 frames of 320 x 180 at 29.97 fps resampled to 2 fps, through the reference's loop
    clip = np.random.default_rng(7).integers(0, 256, (45, 180, 320, 3), dtype=np.uint8)
    FakeCapture.registry["clip"] = (clip, 29.97, 45.0)
    frames, ofps, dur, true_idx = load("clip", output_fps=2, return_true_frames=True)
    out["clip_sha256"], out["clip_kept"] = np.array(sha(frames.numpy())), np.array(true_idx, np.int64)
    out["clip_meta"] = np.array([ofps, dur], np.float64)
    # the fps-resampling rule alone (tiny frames)
    for j, (ifps, count, ofps, cap) in enumerate(FPS_CASES):
        tiny = np.zeros((int(count), 4, 4, 3), np.uint8)
        FakeCapture.registry["fps"] = (tiny, ifps, float(count))
        _, o2, d2, idx = load("fps", output_fps=ofps, return_true_frames=True, max_num_frames=cap)
        out[f"fps_kept_{j}"] = np.array(idx, np.int64)
        out[f"fps_meta_{j}"] = np.array([o2, d2], np.float64)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
