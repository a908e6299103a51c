"""Frame-ingest kernels (through the C-ABI) against cv2's / PIL's / the reference loader's own outputs
(tests/golden/ingest.npz) and against the oracle on seeded batches: bit-exact."""
import hashlib
import os

import numpy as np
import pytest
import torch

import aha_b200 as A
from oracle import ingest_oracle as io

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ingest.npz"))


def source_frame(w, h, seed):
    return np.random.default_rng(seed).integers(0, 256, (h, w, 3), dtype=np.uint8)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("i", range(len(G["geometries"])))
def test_kernels_match_cv2_and_pil(cuda, i):
    w, h = (int(v) for v in G["geometries"][i])
    src = torch.from_numpy(source_frame(w, h, 100 + i)).to(cuda)
    a = A.ingest_frames(src[None], "cv2")[0].cpu().numpy()
    b = A.load_one_frame(src)[0].cpu().numpy()
    assert sha(a) == str(G["cv2_sha256"][i]), f"cv2 flavour differs at {w}x{h}"
    assert sha(b) == str(G["pil_sha256"][i]), f"PIL flavour differs at {w}x{h}"
    if i == 0:
        assert np.array_equal(a, G["full_cv2"]) and np.array_equal(b, G["full_pil"])


def test_clip_through_load_video_for_testing(cuda):
    clip = torch.from_numpy(np.random.default_rng(7).integers(0, 256, (45, 180, 320, 3), dtype=np.uint8))
    frames, ofps, dur, kept = A.load_video_for_testing(clip, 29.97, output_fps=2, return_true_frames=True, device=cuda,
                                                       batch=2)
    assert kept == G["clip_kept"].tolist() and [ofps, dur] == G["clip_meta"].tolist()
    assert frames.is_cuda and tuple(frames.shape) == (len(kept), 3, 384, 384)
    assert sha(frames.cpu().numpy()) == str(G["clip_sha256"])
    assert A.load_video_for_testing(clip[:0], 29.97, output_fps=2, device=cuda) == (None, None, None, None)


@pytest.mark.parametrize("flavour", ["cv2", "pil"])
def test_batch_against_oracle_and_geometry_switches(cuda, flavour):
    # the table cache is per geometry: alternate geometries and flavours, batches of several frames
    for seed, (n, w, h) in enumerate([(3, 427, 240), (2, 240, 427), (3, 427, 240), (1, 1024, 1024), (2, 768, 432)]):
        src = np.random.default_rng(50 + seed).integers(0, 256, (n, h, w, 3), dtype=np.uint8)
        out = A.ingest_frames(torch.from_numpy(src).to(cuda), flavour).cpu().numpy()
        for k in range(n):
            assert np.array_equal(out[k], io.ingest_frame(src[k], flavour)), (seed, k)


@pytest.mark.parametrize("flavour", ["cv2", "pil"])
def test_full_size_batch_properties(cuda, flavour):
    """32 frames of 1920 x 1080: one frame against the oracle, the rest through size-independent properties."""
    g = torch.Generator(device="cpu").manual_seed(3)
    src = torch.randint(0, 256, (32, 1080, 1920, 3), dtype=torch.uint8, generator=g).to(cuda)
    out = A.ingest_frames(src, flavour)
    new_w, new_h, left, top = A.fit_geometry(1920, 1080)
    assert (new_w, new_h, left, top) == (384, 216, 0, 84)
    assert int(out[:, :, :top].max()) == 0 and int(out[:, :, top + new_h:].max()) == 0       # black border
    assert np.array_equal(out[5].cpu().numpy(), io.ingest_frame(src[5].cpu().numpy(), flavour))
    # channels are resampled independently: swapping source channels swaps output planes
    assert torch.equal(A.ingest_frames(src[:4].flip(-1), flavour), out[:4].flip(1))
    # frames are independent: a sub-batch gives the same bytes
    assert torch.equal(A.ingest_frames(src[7:9], flavour), out[7:9])
    # a constant image stays constant inside the frame (coefficients sum to one)
    flat = torch.full((1, 1080, 1920, 3), 201, dtype=torch.uint8, device=cuda)
    assert int(A.ingest_frames(flat, flavour)[0, :, top:top + new_h].min()) == 201


def test_ingest_feeds_visual_embed(cuda):
    """The canvas is what SigLipImageProcessor.preprocess receives: uint8 [n, 3, 384, 384]."""
    src = torch.from_numpy(source_frame(640, 360, 100)).to(cuda)
    canvas = A.ingest_frames(src[None], "cv2")
    assert canvas.dtype == torch.uint8 and tuple(canvas.shape) == (1, 3, 384, 384) and canvas.is_contiguous()


def test_demo_driver_takes_raw_pictures(cuda):
    """LiveInferForDemo.load_one_frame(frame_object=...) (test/live_infer_for_video.py:98-119): a raw RGB picture gives
    the same frame embeddings as the oracle's canvas handed in as `frame`."""
    import dataclasses
    base = A.config_preset("micro72")
    cfg = dataclasses.replace(base, vision=dataclasses.replace(base.vision, image_size=112), frame_resolution=112)
    w = A.random_init_state_dict(cfg, seed=1, device=cuda)
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, w, device=cuda, max_tokens=64, max_frames=8, logits_mode="none")
    start = torch.arange(5, dtype=torch.long).reshape(1, -1)
    a = A.LiveInferForDemo(model, start, window_length=64, num_sink_tokens=4, frame_fps=1.0)
    b = A.LiveInferForDemo(model, start, window_length=64, num_sink_tokens=4, frame_fps=1.0)
    for i, (wd, ht) in enumerate([(320, 180), (90, 200), (112, 112)]):
        pic = source_frame(wd, ht, 300 + i)
        a.load_one_frame(frame_object=pic)                                    # numpy HWC, as np.array(PIL image)
        b.load_one_frame(frame=torch.from_numpy(io.ingest_frame(pic, "pil", out=112)).to(cuda))
    for (ta, ea), (tb, eb) in zip(a.frame_embeds_queue, b.frame_embeds_queue):
        assert ta == tb and torch.equal(ea, eb)
    with pytest.raises(ValueError):
        a.load_one_frame()


def _library_canvas(pic, flavour):
    """cv2 / PIL themselves with the reference's call sequence (test/inference.py:540-563, live_infer_for_video.py:107-119)."""
    h, w = pic.shape[:2]
    nw, nh, left, top = io.fit_geometry(w, h)
    border = dict(top=top, bottom=(384 - nh + 1) // 2, left=left, right=(384 - nw + 1) // 2)
    if flavour == "cv2":
        cv2 = pytest.importorskip("cv2")
        c = cv2.copyMakeBorder(cv2.resize(pic, (nw, nh)), borderType=cv2.BORDER_CONSTANT, value=(0, 0, 0), **border)
        return np.transpose(cv2.cvtColor(c, cv2.COLOR_BGR2RGB), (2, 0, 1))
    Image = pytest.importorskip("PIL.Image")
    ImageOps = pytest.importorskip("PIL.ImageOps")
    r = Image.fromarray(pic).resize((nw, nh))
    c = ImageOps.expand(r, border=(border["left"], border["top"], border["right"], border["bottom"]), fill=(0, 0, 0))
    return np.transpose(np.array(c), (2, 0, 1))


@pytest.mark.parametrize("flavour", ["cv2", "pil"])
def test_random_geometries_against_the_libraries(cuda, flavour):
    """Seeded random and extreme geometries, compared with cv2 / PIL run on the box (skipped where they are absent)
    and with the oracle."""
    rng = np.random.default_rng(11)
    geos = [(1000, 3), (3, 290), (2, 2), (1, 1), (383, 384), (384, 383), (769, 768), (768, 769), (767, 767), (2500, 1407),
            (1536, 864), (192, 192), (191, 192)] + [(int(rng.integers(2, 1400)), int(rng.integers(2, 1400))) for _ in range(12)]
    for w, h in geos:
        pic = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        got = A.ingest_frames(torch.from_numpy(pic).to(cuda)[None], flavour)[0].cpu().numpy()
        assert np.array_equal(got, _library_canvas(pic, flavour)), (flavour, w, h)
        assert np.array_equal(got, io.ingest_frame(pic, flavour)), (flavour, w, h)


def test_refused_geometries(cuda):
    # cv2 raises on an empty target (int(2 / 1000 * 384) == 0); Pillow's pass order for h > 100 w depends on its version
    with pytest.raises(ValueError, match="empty image"):
        A.ingest_frames(torch.zeros(1, 2, 1000, 3, dtype=torch.uint8, device=cuda), "cv2")
    with pytest.raises(ValueError, match="version-dependent"):
        A.ingest_frames(torch.zeros(1, 1000, 3, 3, dtype=torch.uint8, device=cuda), "pil")
    with pytest.raises(ValueError):
        io.ingest_frame(np.zeros((1000, 3, 3), np.uint8), "pil")
    assert tuple(A.ingest_frames(torch.zeros(1, 1000, 3, 3, dtype=torch.uint8, device=cuda), "cv2").shape) == (1, 3, 384, 384)


def test_errors(cuda):
    with pytest.raises(ValueError):
        A.ingest_frames(torch.zeros(1, 8, 8, 3, dtype=torch.uint8, device=cuda), "lanczos")
    with pytest.raises(TypeError):
        A.ingest_frames(torch.zeros(1, 8, 8, 3, dtype=torch.uint8), "cv2")
    with pytest.raises(ValueError):
        A.ingest_frames(torch.zeros(1, 8, 8, 4, dtype=torch.uint8, device=cuda), "cv2")
    assert tuple(A.ingest_frames(torch.zeros(0, 8, 8, 3, dtype=torch.uint8, device=cuda), "pil").shape) == (0, 3, 384, 384)
