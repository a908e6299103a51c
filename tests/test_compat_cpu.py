"""Host-side compatibility layer (aha_b200/compat.py): the reference's image processor arithmetic, the synthetic
tokenizer, checkpoint-directory loading, argument handling of the reference constructor form.  No GPU."""
import dataclasses
import json
import os

import numpy as np
import pytest
import torch

import aha_b200 as A
from aha_b200 import compat
from oracle import aha_oracle as o


def test_image_processor_matches_reference_arithmetic():
    """oracle.preprocess is pinned against the reference's SigLipImageProcessor (tests/golden/siglip.npz); the compat
    processor must give the same float32 bits, for tensors, arrays and lists of frames."""
    ip = compat.SigLipImageProcessor(size=(126, 126))
    fr = o.synthetic_frames(3, image=126)
    ref = o.preprocess(fr)
    assert torch.equal(ip.preprocess(fr, return_tensors="pt")["pixel_values"], ref)
    assert torch.equal(ip.preprocess([f.numpy() for f in fr], return_tensors="pt")["pixel_values"], ref)
    assert torch.equal(ip.preprocess(fr[0], return_tensors="pt")["pixel_values"], ref[:1])
    with pytest.raises(ValueError):
        ip.preprocess(torch.zeros(1, 3, 100, 126, dtype=torch.uint8))
    with pytest.raises(TypeError):
        ip.preprocess(torch.zeros(1, 3, 126, 126))


def test_synthetic_tokenizer_roles():
    tok = compat.SyntheticTokenizer(1000, n_sys=7, n_stream_prompt=3, n_generation=2, n_query=5)
    assert tok.apply_chat_template([{'role': 'system', 'content': 'x'}], return_tensors='pt').shape == (1, 7)
    assert tok.apply_chat_template([{}], add_stream_prompt=True, return_tensors='pt').shape == (1, 3)
    assert tok.apply_chat_template([{}], add_stream_generation_prompt=True, return_tensors='pt').shape == (1, 2)
    q = tok.apply_chat_template([{'role': 'user', 'content': 'what?'}], add_stream_query_prompt=True, add_stream_prompt=True,
                                return_tensors='pt')
    assert q.shape == (1, 5) and int(q.max()) < 1000
    assert isinstance(tok.decode(q[0]), str)


def test_config_from_checkpoint_json(tmp_path):
    cfg = dict(hidden_size=896, intermediate_size=4864, num_hidden_layers=24, num_attention_heads=14, num_key_value_heads=2,
               vocab_size=151936, rope_theta=1000000.0, rms_norm_eps=1e-6, tie_word_embeddings=True, eos_token_id=151645)
    p = tmp_path / "config.json"
    p.write_text(json.dumps(cfg))
    c = compat._config_from_hf_json(str(p))
    t = A.config_preset("tiny")
    assert dataclasses.asdict(c) == dataclasses.asdict(t)


def test_build_requires_local_checkpoint_and_inference_mode():
    with pytest.raises(NotImplementedError):
        A.build_model_and_tokenizer(is_training=True, llm_pretrained="x")
    with pytest.raises(FileNotFoundError):
        A.build_model_and_tokenizer(is_training=False, llm_pretrained="lmms-lab/llava-onevision-qwen2-7b-ov")
    with pytest.raises(TypeError):
        A.build_model_and_tokenizer(is_training=False, torch_dtype=torch.float16, state_dict={}, config=A.config_preset("micro72"))


def test_sink_cache_is_sliding_flag_for_both_transformers_generations():
    """The reference sets `is_sliding = True` (test/sink_cache.py:8); transformers >= 4.54 searches a per-layer list."""
    f = A.SinkCache.is_sliding
    assert f and f == True and (False not in f) and f.index(True) == 0      # noqa: E712
