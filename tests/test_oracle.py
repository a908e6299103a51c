"""CPU: the oracle restatement against the golden fixtures produced by executing the reference itself
(tests/golden/make_golden.py) - cache policies bit-exact, image processor bit-exact, SigLIP tower, peak frames."""
import os

import numpy as np
import pytest
import torch

from oracle import aha_oracle as o
from tests.golden import make_golden as G

GOLD = os.path.dirname(os.path.abspath(G.__file__))


def _bits(t):
    return t.contiguous().view(torch.int16).numpy()


@pytest.mark.parametrize("ci", range(len(G.CACHE_CASES)))
def test_oracle_cache_matches_reference_golden(ci):
    gold = np.load(os.path.join(GOLD, "caches.npz"))
    policy, kw, chunks = G.CACHE_CASES[ci]
    c = o.OracleKVCache(policy, kw.get("window_length", kw.get("window_size")), kw.get("num_sink_tokens", 0))
    lens = []
    for step, q in enumerate(chunks):
        L = c.get_seq_length()
        k, v, cos, sin = G.cache_inputs(ci, step, q, L)
        K, V = c.update(k, v, 0, cos, sin)
        lens.append([L, K.shape[-2], c.get_seq_length()])
    assert np.array_equal(np.array(lens, dtype=np.int32), gold[f"case{ci}_lens"])
    assert np.array_equal(_bits(K), gold[f"case{ci}_K"])
    assert np.array_equal(_bits(V), gold[f"case{ci}_V"])


def test_oracle_preprocess_and_siglip_match_reference_golden():
    gold = np.load(os.path.join(GOLD, "siglip.npz"))
    assert np.array_equal(o.preprocess_lut(), gold["lut"][:256])
    frames = torch.from_numpy(gold["frames"])
    pv = o.preprocess(frames)
    assert np.array_equal(pv[0, :, :4, :4].numpy(), gold["pixel_values_sample"])
    dims = o.dims_preset("micro72")
    w = o.init_weights(dims, 3, dtype=torch.float32)
    hid = o.siglip_tower(w, dims, pv)
    assert np.abs(hid.numpy() - gold["hidden"]).max() < 1e-4      # fp32 GEMM order may differ between hosts


def test_oracle_find_ticks_matches_scipy_golden():
    gold = np.load(os.path.join(GOLD, "ticks.npz"))
    for k in range(8):
        x, fps = gold[f"s{k}_x"], float(gold[f"s{k}_fps"])
        ticks, sm, th = o.find_ticks(x, fps=fps)
        assert [int(round(t * fps)) for t in ticks] == gold[f"s{k}_peaks"].tolist()
        pk, sm2, th2 = o.find_ticks_restated(x, fps=fps)
        assert pk.tolist() == gold[f"s{k}_peaks"].tolist()
        assert np.abs(sm2 - gold[f"s{k}_smoothed"]).max() < 1e-12 and abs(th2 - float(gold[f"s{k}_thresh"])) < 1e-12
    with pytest.raises(ValueError):
        o.find_ticks_restated(np.zeros(10))


def test_oracle_savgol_table_is_scipy():
    from scipy.signal import savgol_coeffs
    assert np.array_equal(o.SAVGOL_15_3, savgol_coeffs(15, 3))


def test_oracle_stream_trace_sink_semantics():
    """SURVEY.md Appendix A.1 trace: W=40, S=4, chunks 7,9,9,... -> first shift evicts 3 tokens, then 9 per step."""
    c = o.OracleKVCache("sink", 40, 4)
    held = []
    for q in [7, 9, 9, 9, 9, 9, 9, 9]:
        L = c.get_seq_length()
        cos, sin = o.rope_cos_sin(torch.arange(L, L + q), 16, 1e6, torch.float32)
        c.update(torch.zeros(1, 1, q, 16), torch.zeros(1, 1, q, 16), 0, cos, sin)
        held.append(c.ids[0].tolist())
    assert held[4] == list(range(4)) + list(range(7, 43))
    assert held[7] == list(range(4)) + list(range(34, 70))
    assert c.seen_tokens == 70 and c.cos_table.shape[0] == 43 and list(c.rerot) == [9]


def test_oracle_mask_rules():
    """sdpa rule j <= L+i (bidirectional in-chunk once the cache is full) vs fa2 rule j <= i + Klen - q."""
    dims = o.dims_preset("micro")
    dims.layers = 1          # one layer: a row's output depends on the other rows only through what it may attend to
    w = o.init_weights(dims, 1, dtype=torch.float32)
    outs = {}
    for mode in ("sdpa", "fa2"):
        torch.manual_seed(0)
        c = o.OracleKVCache("sink", 12, 2)
        for _ in range(4):
            x = torch.randn(1, 5, dims.hidden) * 0.1
            out = o.qwen2_prefill(w, dims, c, x, mode, want_logits=False)
        outs[mode] = out.hidden
    assert not torch.allclose(outs["sdpa"][0, 0], outs["fa2"][0, 0])     # first row sees later in-chunk keys only under sdpa
    assert torch.allclose(outs["sdpa"][0, -1], outs["fa2"][0, -1], atol=1e-5)   # last row sees everything either way
