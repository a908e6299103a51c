"""BASELINE configs 3 and 4 at the 7B width (depth cut to 2 decoder / 1 vision layer so the test stays in seconds):
size-independent properties of the full-size path.
  C3  sliding-window cache W = 8192, 8 streams batched into one prefill (M = 392): length pinned at W, provenance = the
      last W token ids, the batched step equals the single-stream step within bf16 noise (different GEMM orientation).
  C4  static cache: the first chunk (a 4096-token prefix prefilled in 256-token chunks is NOT allowed by the reference's
      TrulyStaticCache - only the FIRST update is kept), later frames attend the frozen prefix non-causally and leave it
      untouched; scores of a frame do not depend on the frames before it."""
import dataclasses

import pytest
import torch

import aha_b200 as A


@pytest.fixture(scope="module")
def wide(cuda):
    cfg = A.config_preset("7b")
    cfg = dataclasses.replace(cfg, num_hidden_layers=2, vision=dataclasses.replace(cfg.vision, num_hidden_layers=1))
    sd = A.random_init_state_dict(cfg, seed=0, device=cuda)
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, sd, device=cuda, max_tokens=512, max_frames=8, logits_mode="none")
    return cfg, model


@pytest.mark.gpu
def test_c3_sliding_window_8_streams(cuda, wide):
    cfg, model = wide
    q, W, ns = cfg.frame_num_tokens, 8192, 8
    torch.manual_seed(1)
    caches = [model.new_cache("sliding_window", W, 0, max_chunk=q + 8) for _ in range(ns)]
    single = model.new_cache("sliding_window", W, 0, max_chunk=q + 8)
    steps = W // q + 6                                        # fills the window and evicts for a few frames
    x = None
    for t in range(steps):
        x = (torch.randn(ns, q, cfg.hidden_size, device=cuda) * 0.05).bfloat16()
        out = model(inputs_embeds=x, past_key_values=caches)
        one = model(inputs_embeds=x[3:4].contiguous(), past_key_values=single)
        if t % 40 == 0 or t == steps - 1:
            # The two calls take different code paths: 392 tokens run the tall GEMM orientation behind a stand-alone
            # RMSNorm kernel, 49 tokens the skinny one with the norm folded into the GEMM epilogues.  Probabilities
            # within 2e-2; the uncertainty score is exp(log-variance) and the log-variance head carries 0.05 of bf16
            # rounding noise at this width (oracle bf16 against oracle fp32, test_7b_width_decoder_matches_oracle).
            a, b = out.frame_scores[3], one.frame_scores[0]
            assert (a[:2] - b[:2]).abs().max().item() <= 2e-2
            assert (a[2].log() - b[2].log()).abs().item() <= 0.1
    total = steps * q
    for c in caches + [single]:
        assert c.get_seq_length() == W
        ids = c.provenance(0).cpu().long()
        assert torch.equal(ids, torch.arange(total - W, total)), "window must hold exactly the last W tokens, in order"
    assert torch.isfinite(out.frame_scores).all()


@pytest.mark.gpu
def test_c4_static_prefix_is_frozen(cuda, wide):
    cfg, model = wide
    q = cfg.frame_num_tokens
    torch.manual_seed(2)
    cache = model.new_cache("static", 32768, 0, max_chunk=512)
    prefix = (torch.randn(1, 400, cfg.hidden_size, device=cuda) * 0.05).bfloat16()      # the query chunk (first update)
    model(inputs_embeds=prefix, past_key_values=cache)
    assert cache.get_seq_length() == 400
    k0 = [t.clone() for t in cache.key_cache[:2]]
    frames = (torch.randn(5, 1, q, cfg.hidden_size, device=cuda) * 0.05).bfloat16()
    first = [model(inputs_embeds=frames[i], past_key_values=cache).frame_scores.clone() for i in range(5)]
    assert cache.get_seq_length() == 400, "TrulyStaticCache ignores every update after the first"
    for a, b in zip(k0, cache.key_cache[:2]):
        assert torch.equal(a, b)
    again = [model(inputs_embeds=frames[i], past_key_values=cache).frame_scores.clone() for i in (4, 2, 0)]
    for s, i in zip(again, (4, 2, 0)):                         # order independence: every frame sees only the frozen prefix
        assert torch.equal(s, first[i])


@pytest.mark.gpu
def test_7b_width_decoder_matches_oracle(cuda):
    """Decoder parity at the 7B WIDTH (hidden 3584, 28 heads over 4 kv heads of 128, MLP 18944; depth cut to 2, a small
    vision tower that this test does not run): 49-token chunks against a sink cache small enough to evict and re-rotate
    on every chunk, per-chunk scores against the oracle in bf16.  The same oracle in fp32 measures how far bf16 rounding
    alone moves each score at this width; the CUDA path has to stay within max(2e-2, 2.5 x that noise)."""
    import numpy as np
    from oracle import aha_oracle as o
    vis_o = o.VisionDims(hidden=144, inter=304, layers=1, heads=2, patch=14, image=126)
    dims = dataclasses.replace(o.dims_preset("7b"), layers=2, vocab=1024, vision=vis_o)
    cfg = dataclasses.replace(A.config_preset("7b"), num_hidden_layers=2, vocab_size=1024, frame_resolution=126,
                              vision=A.VisionConfig(hidden_size=144, intermediate_size=304, num_hidden_layers=1,
                                                           num_attention_heads=2, patch_size=14, image_size=126))
    w = o.init_weights(dims, seed=3, dtype=torch.bfloat16)
    w32 = {k: v.float() for k, v in w.items()}
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, w, device=cuda, max_tokens=64, max_frames=2, logits_mode="none")
    q, W, S, steps = 49, 160, 16, 6                    # 160-token window: evicts from the 4th chunk on
    cache = model.new_cache("default_sink", W, S, max_chunk=q + 8)
    ref_c, ref32_c = o.make_cache("default_sink", W, S), o.make_cache("default_sink", W, S)
    torch.manual_seed(11)
    got, ref, ref32 = [], [], []
    for t in range(steps):
        x = (torch.randn(1, q, cfg.hidden_size) * 0.05).bfloat16()
        out = model(inputs_embeds=x.to(cuda), past_key_values=cache)
        got.append(out.frame_scores[0].float().cpu().numpy())
        ref.append(o.frame_scores(o.qwen2_prefill(w, dims, ref_c, x, want_logits=False)))
        ref32.append(o.frame_scores(o.qwen2_prefill(w32, dims, ref32_c, x.float(), want_logits=False)))
        assert cache.get_seq_length() == ref_c.get_seq_length()
    got, ref, ref32 = np.array(got, dtype=np.float64), np.array(ref), np.array(ref32)
    for col, name in enumerate(("informative", "relevance", "uncertainty")):
        a, b, c = got[:, col], ref[:, col], ref32[:, col]
        if name == "uncertainty":                      # exp(log-variance): the tolerance applies to the head's output
            a, b, c = np.log(a), np.log(b), np.log(c)
        noise = np.abs(b - c).max()
        err = np.abs(a - b).max()
        print(f"7b-width {name}: |cuda - oracle_bf16| = {err:.4f}, |oracle_bf16 - oracle_fp32| = {noise:.4f}, "
              f"|cuda - oracle_fp32| = {np.abs(a - c).max():.4f}")
        assert err <= max(2e-2, 2.0 ** -6 * np.abs(b).max(), 2.5 * noise), (name, err, noise)


@pytest.mark.gpu
def test_7b_width_two_streams_equal_single(cuda, wide):
    """Two streams batched into one call (98 tokens: 128-token GEMM tiles, two 16-token merge chunks per split more than
    the 49-token call) against the same stream alone: both run the folded-RMSNorm skinny path, with different tile and
    split geometry, through a sink cache that evicts and re-rotates."""
    cfg, model = wide
    q, W, S = cfg.frame_num_tokens, 160, 16
    torch.manual_seed(5)
    caches = [model.new_cache("default_sink", W, S, max_chunk=q + 8) for _ in range(2)]
    single = model.new_cache("default_sink", W, S, max_chunk=q + 8)
    for t in range(6):
        x = (torch.randn(2, q, cfg.hidden_size, device=cuda) * 0.05).bfloat16()
        out = model(inputs_embeds=x, past_key_values=caches)
        one = model(inputs_embeds=x[1:2].contiguous(), past_key_values=single)
        a, b = out.frame_scores[1], one.frame_scores[0]
        assert (a[:2] - b[:2]).abs().max().item() <= 2e-2, (t, a, b)
        assert (a[2].log() - b[2].log()).abs().item() <= 0.1, (t, a, b)
        assert caches[1].get_seq_length() == single.get_seq_length()
    assert torch.equal(caches[1].provenance(0), single.provenance(0))
