"""Pin of the Qwen2 restatement (oracle.qwen2_prefill) to EXECUTED code:

  * live: the installed transformers `Qwen2Model(attn_implementation="eager")` + DynamicCache on the oracle's weights,
    chunk by chunk (growing regime) - bit-exact in fp32 and bf16;
  * golden (tests/golden/qwen2_hf.npz, written by tests/golden/make_golden_qwen2.py in the authoring container): the
    same, and the transformers-4.49-only semantics - HF's own decoder layers driven with the REFERENCE's cache classes
    (test/sink_cache.py, test/sliding_window_cache.py, test/static_cache.py), positions from get_seq_length(), the 4.49
    additive mask sliced to the Klen the cache returns: eviction + re-rotation, the bidirectional in-chunk regime of a
    full cache, the frozen static prefix.  The oracle must reproduce every hidden-state bit.
"""
import os

import numpy as np
import pytest
import torch

from oracle import aha_oracle as o
from tests.golden import make_golden_qwen2 as G

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "qwen2_hf.npz")
DTYPES = [(torch.float32, "f32"), (torch.bfloat16, "bf16")]


def _bits(t):
    return G.bits(t)


@pytest.mark.parametrize("dtype,tag", DTYPES)
def test_oracle_matches_live_hf_qwen2model(dtype, tag):
    dims = G.qwen2_dims()
    w = o.init_weights(dims, seed=G.W_SEED, dtype=dtype)
    hf = G.run_hf_growing(dims, w, dtype)
    cache = o.OracleKVCache("sink", 4096, 32)            # far from full: the growing regime
    for step, q in enumerate(G.GROWING_CHUNKS):
        x = G.chunk_inputs(dims, 99, step, q, dtype)
        out = o.qwen2_prefill(w, dims, cache, x, "sdpa", want_logits=False)
        assert torch.equal(out.hidden, hf[step]), (tag, step)
        assert out.positions[0].item() == sum(G.GROWING_CHUNKS[:step])


@pytest.mark.parametrize("dtype,tag", DTYPES)
def test_oracle_matches_hf_golden_growing(dtype, tag):
    gold = np.load(GOLD)
    dims = G.qwen2_dims()
    w = o.init_weights(dims, seed=G.W_SEED, dtype=dtype)
    cache = o.OracleKVCache("sliding", 4096)
    for step, q in enumerate(G.GROWING_CHUNKS):
        out = o.qwen2_prefill(w, dims, cache, G.chunk_inputs(dims, 99, step, q, dtype), "sdpa", want_logits=False)
        assert np.array_equal(_bits(out.hidden), gold[f"hf_growing_{tag}_{step}"]), (tag, step)


@pytest.mark.parametrize("dtype,tag", DTYPES)
@pytest.mark.parametrize("ci", range(len(G.REF_CASES)))
def test_oracle_matches_hf_layers_with_reference_caches(ci, dtype, tag):
    """Every chunk of every case, incl. the ones after the cache filled up (sink: eviction + re-rotation, all W keys visible
    to every row of the chunk; sliding: no re-rotation; static: later chunks attend the frozen prefix only)."""
    gold = np.load(GOLD)
    policy, kw, chunks = G.REF_CASES[ci]
    dims = G.qwen2_dims()
    w = o.init_weights(dims, seed=G.W_SEED, dtype=dtype)
    if policy == "sink":
        cache = o.OracleKVCache("sink", kw["window_length"], kw["num_sink_tokens"])
    elif policy == "sliding":
        cache = o.OracleKVCache("sliding", kw["window_length"])
    else:
        cache = o.OracleKVCache("static", kw["window_size"])
    lens = gold[f"ref{ci}_{tag}_lens"]
    saw_full = False
    for step, q in enumerate(chunks):
        L = cache.get_seq_length()
        out = o.qwen2_prefill(w, dims, cache, G.chunk_inputs(dims, ci, step, q, dtype), "sdpa", want_logits=False)
        assert [L, out.klen, cache.get_seq_length()] == lens[step].tolist(), (policy, step)
        assert np.array_equal(_bits(out.hidden), gold[f"ref{ci}_{tag}_{step}"]), (policy, tag, step)
        saw_full |= (L > 0 and out.klen < L + q)
    assert saw_full, "the case must reach the regime where the cache returns fewer keys than L + q"
