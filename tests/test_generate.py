"""N1 (greedy response decode): the device-side repetition penalty + argmax against torch (exact), and the decode loop of
the native engine against the oracle's restatement of fast_greedy_generate (teacher-forced, bf16 tolerance)."""
import numpy as np
import pytest
import torch

from oracle import aha_oracle as o


def test_oracle_repetition_penalty_is_applied_once_per_id():
    s = torch.tensor([[2.0, -2.0, 1.0, 0.5]])
    out = o.repetition_penalty_logits(s, torch.tensor([[0, 1, 0]]), 2.0)
    assert out.tolist() == [[1.0, -4.0, 1.0, 0.5]]


@pytest.mark.gpu
@pytest.mark.parametrize("vocab,n_gen,penalty", [(152064, 0, None), (152064, 37, 1.15), (1000, 5, 2.0), (33, 40, 1.3)])
def test_greedy_pick_matches_torch(cuda, vocab, n_gen, penalty):
    from aha_b200.generate import greedy_pick
    g = torch.Generator().manual_seed(vocab + n_gen)
    logits = torch.randn(1, 1, vocab, generator=g)
    logits[0, 0, vocab // 3] = logits.max() + 0.25          # a clear winner that the penalty may dethrone
    gen = torch.randint(0, vocab, (n_gen,), generator=g).tolist()
    if n_gen:
        gen[0] = vocab // 3
        gen.append(gen[1])                                   # duplicate id: penalised once
    ref = logits[:, -1, :].clone()
    if penalty is not None and gen:
        ref = o.repetition_penalty_logits(ref, torch.tensor(gen).unsqueeze(0), penalty)
    got = greedy_pick(logits.to(cuda), gen, penalty)
    assert got.shape == (1, 1) and int(got.item()) == int(ref.argmax(dim=-1))
    # ties: the first maximum wins
    tie = torch.zeros(1, 1, 64)
    tie[0, 0, [7, 9, 40]] = 3.0
    assert int(greedy_pick(tie.to(cuda)).item()) == 7


@pytest.mark.gpu
@pytest.mark.parametrize("penalty", [None, 1.15])
def test_fast_greedy_generate_vs_oracle(cuda, penalty):
    import aha_b200 as A
    from aha_b200.generate import fast_greedy_generate
    dims, cfg = o.dims_preset("micro72"), A.config_preset("micro72")
    w = o.init_weights(dims, seed=3, dtype=torch.bfloat16)
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, w, device=cuda, max_tokens=64, max_frames=4, logits_mode="none")
    cache = model.new_cache("default_sink", 48, 8, max_chunk=32)
    ref_cache = o.make_cache("default_sink", 48, 8)
    g = torch.Generator().manual_seed(5)
    prompt = torch.randint(0, dims.vocab, (1, 30), generator=g)
    emb = torch.nn.functional.embedding(prompt, w["model.embed_tokens.weight"])
    model(inputs_embeds=emb.to(cuda), past_key_values=cache)
    o.qwen2_prefill(w, dims, ref_cache, emb, "sdpa", want_logits=False)
    start = torch.randint(0, dims.vocab, (1, 4), generator=g)
    start_emb = torch.nn.functional.embedding(start, w["model.embed_tokens.weight"])
    n_new = 24                                               # 30 + 4 + 24 > 48: the decode crosses an eviction
    out_buf = torch.zeros(1, n_new, dtype=torch.long, device=cuda)
    ids, pkv, gen = fast_greedy_generate(model=model, inputs_embeds=start_emb.to(cuda), past_key_values=cache,
                                         eos_token_id=-1, inplace_output_ids=out_buf, repetition_penalty=penalty,
                                         generated_token_ids=[])
    ids = ids[0].tolist()
    assert len(ids) == n_new and pkv is cache
    assert gen == (ids if penalty is not None else [])
    ref_ids, _, step_logits = o.fast_greedy_generate(w, dims, ref_cache, start_emb, -1, n_new, penalty, [], "sdpa",
                                                     forced_ids=ids)
    assert ref_ids == ids
    for t, lg in zip(ids, step_logits):                      # the engine's choice is the oracle's argmax up to bf16 noise
        assert float(lg.max() - lg[t]) <= 2e-2 * max(1.0, float(lg.abs().max()))
    assert cache.get_seq_length() == ref_cache.get_seq_length() == 48
    assert torch.equal(cache.provenance(0).cpu().long(), ref_cache.ids[0])


@pytest.mark.gpu
def test_live_infer_generates_on_threshold(cuda):
    import aha_b200 as A
    dims, cfg = o.dims_preset("micro72"), A.config_preset("micro72")
    w = o.init_weights(dims, seed=0, dtype=torch.bfloat16)
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, w, device=cuda, max_tokens=64, max_frames=8, logits_mode="none")
    frames = o.synthetic_frames_varied(6, image=cfg.vision.image_size, seed=2).to(cuda)
    start = torch.arange(5, 15).reshape(1, -1)
    live = A.LiveInferForBenchmark(model, start, added_stream_prompt_ids=torch.tensor([[3, 4]]), window_length=64,
                                   num_sink_tokens=8, frame_fps=1.0, stream_end_score_sum_threshold=1.0,
                                   added_stream_generation_ids=torch.tensor([[7, 8, 9]]), eos_token_id=-1,
                                   repetition_penalty=1.15, max_new_tokens=5)
    live.input_video_stream(frames)
    out = live.inference()
    assert len(out) == 6 and live.model_response_list, "the score-sum threshold must have fired"
    for r in live.model_response_list:
        assert r["role"] == "assistant" and len(r["content"]) == 5
    assert live.last_role in ("stream", "assistant")
