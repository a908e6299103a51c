"""The north-star parity bars AS WRITTEN (BASELINE.json): per-frame scores max-abs <= 2e-2, Spearman >= 0.999, identical
peak frames, cache indices / positions / eviction bit-exact - on score streams with real spread, at the specified
configurations:

  * micro72, 120 frames (fast);
  * BASELINE config 1: tiny = Qwen2-0.5B dims + the FULL SigLIP-so400m tower (1152 wide, 26 layers, 729 patches, head_dim 72),
    64 frames, sdpa mask rule, SinkCache W = 2048 and W = 512;
  * 7B dims at FULL DEPTH (28 decoder + 26 vision layers), steady state with eviction + re-rotation on every frame.

The reference implementation here is the oracle restatement in the reference's own dtype (bf16), run as plain PyTorch -
on the GPU for the two big cases, which is exactly how the reference runs (cuBLAS / eager attention); the same oracle in
fp32 supplies (a) the trained-like score heads (oracle.aligned_heads: random N(0, 0.02) heads make the scores of a
random-init model flat, and rank correlation then measures bf16 noise) and (b) the noise floor printed beside every bar.
"""
import dataclasses

import numpy as np
import pytest
import torch
from scipy.stats import spearmanr

import aha_b200 as A
from oracle import aha_oracle as o

pytestmark = pytest.mark.gpu

TOL, RHO = 2e-2, 0.999


def _series(data, key):
    x = np.array([d[key] for d in data], dtype=np.float64)
    return np.log(x) if key == "uncertainty_score" else x       # the head's own output is the log-variance


def _run_oracle(w, dims, frames, dtype, **kw):
    ww = w if dtype == torch.bfloat16 else {k: v.to(dtype) for k, v in w.items()}
    ref = o.LiveInferOracle(ww, dims, dtype=dtype, **kw)
    with torch.no_grad():
        ref.input_video_stream(frames)
        ref.input_query_stream()
        ref.inference()
    return ref


def _check_bars(tag, out, ref, ref32):
    report = {}
    for key in ("informative_score", "relevance_score", "uncertainty_score"):
        a, b, c = _series(out, key), _series(ref.debug_data_list, key), _series(ref32.debug_data_list, key)
        report[key] = dict(max_abs=float(np.abs(a - b).max()), spearman=float(spearmanr(a, b).statistic),
                           floor_max_abs=float(np.abs(b - c).max()), floor_spearman=float(spearmanr(b, c).statistic),
                           std=float(np.std(b)))
    print(f"{tag}: parity vs the bf16 reference (floor = bf16 reference vs fp32 reference):", report)
    for key, r in report.items():
        assert r["max_abs"] <= TOL, (tag, key, r)
        assert r["spearman"] >= RHO, (tag, key, r)
    for key in ("relevance_score", "informative_score"):        # the two series the reference picks peaks on
        mine = o.find_ticks([d[key] for d in out], fps=1.0)[0]
        theirs = o.find_ticks([d[key] for d in ref.debug_data_list], fps=1.0)[0]
        assert mine == theirs, (tag, key, mine, theirs)
    return report


def _cuda_stream(model, ref, frames, cuda, **kw):
    live = A.LiveInferForBenchmark(model, ref.start_ids, frame_fps=1.0, **kw)
    live.input_video_stream(frames.to(cuda))
    if ref.query_ids is not None:
        live.input_query_stream(ref.query_ids)
    lens = []
    while live.frame_embeds_queue:
        live.input_one_frame()
        lens.append(live.past_key_values.get_seq_length())
    live._flush_scores()
    return live, lens


def test_micro72_long_stream_bars_as_written(cuda):
    dims, cfg = o.dims_preset("micro72"), A.config_preset("micro72")
    w0 = o.init_weights(dims, seed=0, dtype=torch.bfloat16)
    frames = o.synthetic_frames_varied(120, image=cfg.vision.image_size, seed=9)
    kw = dict(window_length=128, num_sink_tokens=16, n_sys=12, n_query=6)
    w = o.aligned_heads(w0, _run_oracle(w0, dims, frames, torch.float32, **kw).hidden_last)
    ref, ref32 = _run_oracle(w, dims, frames, torch.bfloat16, **kw), _run_oracle(w, dims, frames, torch.float32, **kw)
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, w, device=cuda, max_tokens=64, max_frames=32, logits_mode="none")
    live, lens = _cuda_stream(model, ref, frames, cuda, window_length=128, num_sink_tokens=16)
    assert lens == [t["seq_len"] for t in ref.trace if t["kind"] == "frame"]
    assert torch.equal(live.past_key_values.provenance(0).cpu().long(), ref.cache.ids[0])
    _check_bars("micro72 x 120 frames", live.debug_data_list, ref, ref32)


@pytest.mark.parametrize("window", [2048, 512])
def test_config1_tiny_full_siglip(cuda, window):
    """BASELINE config 1 as specified: 64 frames at 1 fps, sdpa rule, sink cache 2048 / 32 (first eviction around frame 41)
    and 512 / 32 (evicts from frame 10 on); chunk sequence [query], [system prompt + frame], 63 x [frame]."""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dims, cfg = o.dims_preset("tiny"), A.config_preset("tiny")
    w0 = o.init_weights(dims, seed=0, dtype=torch.bfloat16, device="cuda")
    frames = o.synthetic_frames_varied(64, image=cfg.vision.image_size, seed=1)
    kw = dict(window_length=window, num_sink_tokens=32, n_sys=36, n_query=24)
    w = o.aligned_heads(w0, _run_oracle(w0, dims, frames, torch.float32, **kw).hidden_last)
    ref, ref32 = _run_oracle(w, dims, frames, torch.bfloat16, **kw), _run_oracle(w, dims, frames, torch.float32, **kw)
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, w, device=cuda, max_tokens=128, max_frames=32, logits_mode="none")
    live, lens = _cuda_stream(model, ref, frames, cuda, window_length=window, num_sink_tokens=32)
    assert lens == [t["seq_len"] for t in ref.trace if t["kind"] == "frame"]
    assert lens[-1] == window and torch.equal(live.past_key_values.provenance(0).cpu().long(), ref.cache.ids[0])
    assert live.past_key_values._seen_tokens == ref.cache.seen_tokens
    # vision tower at full size: the engine's frame embeddings against the reference's
    emb = model.visual_embed(frames[:4].to(cuda)).float()
    with torch.no_grad():
        emb_ref = o.visual_embed(w, dims, o.preprocess(frames[:4].to(cuda)).bfloat16()).float()
    assert (emb - emb_ref).abs().max().item() <= TOL * max(1.0, emb_ref.abs().max().item())
    _check_bars(f"config 1 (tiny + full SigLIP), W = {window}", live.debug_data_list, ref, ref32)


def test_7b_full_depth(cuda):
    """7B dims, 28 + 26 layers, every kernel at its production shape: 48 frames through a sink cache small enough that the
    stream is in steady state (eviction + re-rotation on every frame) from the fourth frame on.  (48 frames, not fewer: with
    24 points ONE swap of two neighbouring ranks already costs 8.7e-4 of rank correlation - the bf16 reference against its own
    fp32 run scored 0.99913 on such a stream - so the 0.999 bar would measure the length of the stream, not the kernels.)"""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dims = dataclasses.replace(o.dims_preset("7b"), vocab=2048)
    cfg = dataclasses.replace(A.config_preset("7b"), vocab_size=2048)
    w0 = o.init_weights(dims, seed=0, dtype=torch.bfloat16, device="cuda")
    frames = o.synthetic_frames_varied(48, image=cfg.vision.image_size, seed=2)
    kw = dict(window_length=160, num_sink_tokens=16, n_sys=12, n_query=6)
    r0 = _run_oracle(w0, dims, frames, torch.float32, **kw)
    w = o.aligned_heads(w0, r0.hidden_last)
    del r0
    ref = _run_oracle(w, dims, frames, torch.bfloat16, **kw)
    ref32 = _run_oracle(w, dims, frames, torch.float32, **kw)
    torch.cuda.empty_cache()
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, w, device=cuda, max_tokens=128, max_frames=8, logits_mode="none")
    live, lens = _cuda_stream(model, ref, frames, cuda, window_length=160, num_sink_tokens=16)
    assert lens == [t["seq_len"] for t in ref.trace if t["kind"] == "frame"]
    assert torch.equal(live.past_key_values.provenance(0).cpu().long(), ref.cache.ids[0])
    emb = model.visual_embed(frames[:2].to(cuda)).float()
    with torch.no_grad():
        emb_ref = o.visual_embed(w, dims, o.preprocess(frames[:2].to(cuda)).bfloat16()).float()
    assert (emb - emb_ref).abs().max().item() <= TOL * max(1.0, emb_ref.abs().max().item())
    _check_bars("7B full depth x 48 frames", live.debug_data_list, ref, ref32)
