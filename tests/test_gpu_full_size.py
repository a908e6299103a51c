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
            assert (out.frame_scores[3] - one.frame_scores[0]).abs().max().item() <= 2e-2
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
