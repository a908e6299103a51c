"""Times the frame-ingest kernels (N4) on one GPU: CUDA events on the launching stream, warm-up, inputs larger than L2
(64 frames of 1920 x 1080 x 3 = 398 MB).  Prints one JSON line per (flavour, geometry).
Algorithmic bytes per frame: PIL bicubic (down-scale): every source byte is under some tap -> h*w*3 read + 3*384*384
written; cv2 bilinear: 4 source pixels per output pixel -> min(h*w*3, new_w*new_h*4*3) read + 3*384*384 written."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import aha_b200 as A  # noqa: E402

PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))) \
    if os.path.exists("MEASURED_PEAKS.json") else {}


def main():
    dev = torch.device("cuda:0")
    for (w, h, n) in [(1920, 1080, 64), (1280, 720, 128), (640, 360, 512)]:
        src = torch.randint(0, 256, (n, h, w, 3), dtype=torch.uint8, device=dev)
        new_w, new_h, _, _ = A.fit_geometry(w, h)
        for flavour in ("cv2", "pil"):
            for _ in range(3):
                A.ingest_frames(src, flavour)
            torch.cuda.synchronize()
            reps = 10
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                A.ingest_frames(src, flavour)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            read = h * w * 3 if flavour == "pil" else min(h * w * 3, new_w * new_h * 4 * 3)
            alg = (read + 3 * 384 * 384) * n
            print(json.dumps({"kernel": f"ingest_{flavour}_kernel", "geometry": [w, h], "frames_per_launch": n,
                              "ms_per_launch": round(ms, 4), "frames_per_s": round(n / ms * 1e3, 1),
                              "algorithmic_GBps": round(alg / ms / 1e6, 1),
                              "source_GBps": round((h * w * 3 + 3 * 384 * 384) * n / ms / 1e6, 1)}))
        del src
    # CPU beside it: cv2 / PIL themselves on one host core, 1080p
    try:
        import cv2
        from PIL import Image
        f = np.random.default_rng(0).integers(0, 256, (1080, 1920, 3), dtype=np.uint8)
        t = time.perf_counter()
        for _ in range(50):
            cv2.cvtColor(cv2.copyMakeBorder(cv2.resize(f, (384, 216)), 84, 84, 0, 0, cv2.BORDER_CONSTANT, value=(0, 0, 0)),
                         cv2.COLOR_BGR2RGB)
        c = (time.perf_counter() - t) / 50
        im = Image.fromarray(f)
        t = time.perf_counter()
        for _ in range(50):
            im.resize((384, 216))
        p = (time.perf_counter() - t) / 50
        print(json.dumps({"cpu": "cv2.resize+border+cvtColor / PIL.resize, 1920x1080, one host process",
                          "cv2_frames_per_s": round(1 / c, 1), "pil_frames_per_s": round(1 / p, 1),
                          "cv2_threads": cv2.getNumThreads()}))
    except Exception as e:  # noqa: BLE001
        print(json.dumps({"cpu": f"unavailable: {e}"}))


if __name__ == "__main__":
    main()
