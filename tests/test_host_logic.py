"""CPU: host-side mirror logic that needs no device."""
import pytest
import torch

import aha_b200 as A
from aha_b200.sharding import gather_stream_results, shard_streams


def test_config_presets():
    c = A.config_preset("7b")
    assert (c.hidden_size, c.num_hidden_layers, c.num_attention_heads, c.num_key_value_heads, c.head_dim) == (3584, 28, 28, 4, 128)
    assert c.vision.side == 27 and c.vision.num_patches == 729 and c.frame_num_tokens == 49
    t = A.config_preset("tiny")
    assert t.tie_word_embeddings and t.frame_num_tokens == 49 and t.head_dim == 64
    c.mm_spatial_pool_mode = "average"
    assert c.frame_num_tokens == 36
    with pytest.raises(ValueError):
        A.config_preset("nope")


def test_round_numbers():
    d = [dict(time=1.23456, informative_score=0.98765, n=3), 0.5557]
    assert A.round_numbers(d, 3) == [dict(time=1.235, informative_score=0.988, n=3), 0.556]


def test_cache_constructors_and_dispatch():
    s = A.SinkCache(window_length=2048, num_sink_tokens=32)
    assert s.window_length == 2048 and s.num_sink_tokens == 32 and bool(s.is_sliding) and s.is_sliding == True   # noqa: E712 (a truthy flag that transformers >= 4.54 can also search)
    assert s.get_seq_length() == 0 and s.get_max_length() == 2048 and s.get_max_cache_shape() == 2048 and s._seen_tokens == 0
    assert A.SlidingWindowCache(window_length=8192).window_length == 8192
    assert A.TrulyStaticCache(window_size=32768).window_size == 32768
    assert isinstance(A.make_cache("default_sink"), A.SinkCache)
    assert isinstance(A.make_cache("sliding_window", 100), A.SlidingWindowCache)
    assert isinstance(A.make_cache("static", 100), A.TrulyStaticCache)
    c = A.make_cache(sink_cache=True, window_length=2048, num_sink_tokens=32, n_sys=36)     # test/inference.py:140-147
    assert c.num_sink_tokens == 36 and c.window_length == 2048 + 32 - 36
    with pytest.raises(ValueError):
        A.make_cache("bogus")
    with pytest.raises(ValueError):
        A.SinkCache(window_length=8, num_sink_tokens=9)
    assert s._klen_after(0, 5000) == 5000 and s._klen_after(2000, 49) == 2048 and s._klen_after(100, 49) == 149
    assert A.TrulyStaticCache(10)._klen_after(0, 50) == 10 and A.TrulyStaticCache(10)._klen_after(7, 50) == 7


def test_cache_rejects_cpu_tensors():
    s = A.SinkCache(window_length=16, num_sink_tokens=2)
    with pytest.raises(TypeError):
        s.update(torch.zeros(1, 2, 4, 64, dtype=torch.bfloat16), torch.zeros(1, 2, 4, 64, dtype=torch.bfloat16), 0, {})


def test_model_needs_cuda():
    with pytest.raises(RuntimeError, match="no CPU path"):
        A.VideoHeadLiveLlavaQwenForCausalLM(A.config_preset("micro72"), {}, device="cpu")


def test_random_init_names_match_oracle():
    from oracle import aha_oracle as o
    cfg, dims = A.config_preset("micro72"), o.dims_preset("micro72")
    a = A.random_init_state_dict(cfg, device="cpu")
    b = o.init_weights(dims)
    assert list(a) == list(b)
    assert all(a[k].shape == b[k].shape for k in a)


def test_shard_streams():
    assert shard_streams(64, 8, 3) == list(range(3, 64, 8))
    assert sorted(sum((shard_streams(10, 4, r) for r in range(4)), [])) == list(range(10))
    assert shard_streams(2, 4, 3) == []
    with pytest.raises(ValueError):
        shard_streams(4, 2, 2)
    assert gather_stream_results({1: [0.5]}) == {1: [0.5]}
