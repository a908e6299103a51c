"""CPU: the ingest restatement (oracle/ingest_oracle.py) and the host-side frame selection against the goldens that
cv2 / PIL / the reference's own load_video_for_testing produced (tests/golden/make_golden_ingest.py)."""
import hashlib
import os

import numpy as np
import pytest

from aha_b200 import _lib, ingest
from oracle import ingest_oracle as io

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ingest.npz"))


def source_frame(w, h, seed):
    return np.random.default_rng(seed).integers(0, 256, (h, w, 3), dtype=np.uint8)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("i", range(len(G["geometries"])))
def test_oracle_matches_cv2_and_pil(i):
    w, h = (int(v) for v in G["geometries"][i])
    src = source_frame(w, h, 100 + i)
    a, b = io.ingest_frame(src, "cv2"), io.ingest_frame(src, "pil")
    assert sha(a) == str(G["cv2_sha256"][i])
    assert sha(b) == str(G["pil_sha256"][i])
    if i == 0:
        assert np.array_equal(a, G["full_cv2"]) and np.array_equal(b, G["full_pil"])


def test_fit_geometry_matches_oracle():
    for w, h in [(640, 360), (360, 640), (1001, 563), (3, 2), (384, 384), (1, 1000)]:
        assert ingest.fit_geometry(w, h) == io.fit_geometry(w, h)


@pytest.mark.parametrize("select", [io.select_frames, ingest.select_frames])
def test_select_frames_matches_reference_loop(select):
    for j, (ifps, count, ofps, cap) in enumerate(G["fps_cases"]):
        cap = None if cap < 0 else int(cap)
        kept, o2, d2 = select(float(ifps), float(count), float(ofps), cap)
        assert kept == G[f"fps_kept_{j}"].tolist(), j
        assert [o2, d2] == G[f"fps_meta_{j}"].tolist()
    kept, o2, d2 = select(29.97, 45.0, 2, None)
    assert kept == G["clip_kept"].tolist() and [o2, d2] == G["clip_meta"].tolist()


def test_abi_rejects_bad_arguments_before_any_cuda_call():
    lib = _lib.load()
    assert lib.aha_ingest_frames(None, 1, 10, 10, 0, None, 384, None) == -1
    one = (np.zeros(4, np.uint8)).ctypes.data
    assert lib.aha_ingest_frames(one, 1, 10, 10, 7, one, 384, None) == -1 and b"flavour" in lib.aha_last_error()
    assert lib.aha_ingest_frames(one, 1, 0, 10, 0, one, 384, None) == -1 and b"geometry" in lib.aha_last_error()
    assert lib.aha_ingest_frames(one, 1, 10, 10, 0, one, 383, None) == -1 and b"multiple of 4" in lib.aha_last_error()
    assert lib.aha_ingest_frames(None, 0, 10, 10, 0, None, 384, None) == 0       # empty batch: nothing to do
