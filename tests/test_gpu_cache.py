"""Bit-exact parity of the native ring-buffer caches (through the HF Cache API wrappers -> C-ABI) with the oracle's
restatement of test/sink_cache.py, test/sliding_window_cache.py, test/static_cache.py: lengths, positions, provenance
(which global token sits in which logical slot) and every K/V byte after every update."""
import pytest
import torch

from aha_b200 import SinkCache, SlidingWindowCache, TrulyStaticCache
from oracle import aha_oracle as o

pytestmark = pytest.mark.gpu

CASES = [
    ("sink", dict(window_length=40, num_sink_tokens=4), [7, 9, 9, 9, 9, 9, 9, 9, 5, 9, 3]),
    ("sink", dict(window_length=64, num_sink_tokens=0), [10, 20, 20, 20, 20, 7]),
    ("sink", dict(window_length=30, num_sink_tokens=6), [40, 5, 5, 8]),          # first chunk longer than the window
    ("sink", dict(window_length=256, num_sink_tokens=32), [36 + 49] + [49] * 9),
    ("sliding", dict(window_length=20), [7, 9, 9, 30, 4]),
    ("sliding", dict(window_length=20), [33, 5, 5, 12]),                        # oversize first chunk, then compaction
    ("sliding", dict(window_length=128), [24, 85] + [49] * 6),
    ("static", dict(window_size=5), [7, 9, 4]),
    ("static", dict(window_size=50), [7, 9, 4]),
]


def _make(policy, kw):
    if policy == "sink":
        return SinkCache(**kw), o.OracleKVCache("sink", kw["window_length"], kw["num_sink_tokens"])
    if policy == "sliding":
        return SlidingWindowCache(**kw), o.OracleKVCache("sliding", kw["window_length"])
    return TrulyStaticCache(**kw), o.OracleKVCache("static", kw["window_size"])


@pytest.mark.parametrize("policy,kw,chunks", CASES)
@pytest.mark.parametrize("d", [64, 128])
def test_cache_bit_exact(cuda, policy, kw, chunks, d):
    torch.manual_seed(0)
    kvh, layers = 2, 3
    mine, ref = _make(policy, kw)
    for step, q in enumerate(chunks):
        L = ref.get_seq_length()
        assert mine.get_seq_length() == L, (step, q)
        pos = torch.arange(L, L + q)
        cos, sin = o.rope_cos_sin(pos, d, 1e6, torch.bfloat16)
        for layer in range(layers):
            k = torch.randn(1, kvh, q, d).bfloat16()
            v = torch.randn(1, kvh, q, d).bfloat16()
            Kr, Vr = ref.update(k, v, layer, cos, sin)
            # non-contiguous [1, kvh, q, d] views of a [q, kvh*d] projection, as Qwen2Attention hands them over
            kd = k[0].permute(1, 0, 2).contiguous().to(cuda).view(q, kvh, d).permute(1, 0, 2).unsqueeze(0)
            vd = v[0].permute(1, 0, 2).contiguous().to(cuda).view(q, kvh, d).permute(1, 0, 2).unsqueeze(0)
            Km, Vm = mine.update(kd, vd, layer, {"cos": cos.to(cuda), "sin": sin.to(cuda)})
            assert Km.shape == Kr.shape, (step, q, layer, Km.shape, Kr.shape)
            assert torch.equal(Km.cpu(), Kr), (policy, kw, step, q, layer, "K")
            assert torch.equal(Vm.cpu(), Vr), (policy, kw, step, q, layer, "V")
        assert mine.get_seq_length() == ref.get_seq_length()
        assert torch.equal(mine.provenance(0).cpu().long(), ref.ids[0]), (step, q)
        if policy == "sink":
            assert mine._seen_tokens == ref.seen_tokens
    assert mine.get_max_length() == ref.get_max_length()
    assert len(mine.key_cache) == layers
    assert torch.equal(mine.key_cache[1].cpu(), ref.keys[1])


def test_sink_without_rope_kwargs(cuda):
    """cache_kwargs without cos/sin: kept keys are shifted but not re-rotated (sink_cache.py:101,141)."""
    torch.manual_seed(1)
    mine, ref = _make("sink", dict(window_length=24, num_sink_tokens=2))
    for q in [10, 10, 10, 6]:
        k = torch.randn(1, 2, q, 64).bfloat16(); v = torch.randn(1, 2, q, 64).bfloat16()
        Kr, Vr = ref.update(k, v, 0, None, None)
        Km, Vm = mine.update(k.to(cuda), v.to(cuda), 0, {})
        assert torch.equal(Km.cpu(), Kr) and torch.equal(Vm.cpu(), Vr)


def test_cache_errors(cuda):
    with pytest.raises(ValueError):
        SinkCache(window_length=8, num_sink_tokens=8)
    c = SinkCache(window_length=16, num_sink_tokens=4)
    k = torch.zeros(1, 2, 8, 64, dtype=torch.bfloat16, device=cuda)
    c.update(k, k, 0, {})
    with pytest.raises(ValueError):          # q >= window - sinks once shifting
        c.update(torch.zeros(1, 2, 12, 64, dtype=torch.bfloat16, device=cuda),
                 torch.zeros(1, 2, 12, 64, dtype=torch.bfloat16, device=cuda), 0, {})
    with pytest.raises(TypeError):
        c.update(k.float(), k.float(), 0, {})
    assert c.get_seq_length(5) == 0


def test_full_size_roundtrip_properties(cuda):
    """BASELINE config sizes (7B geometry, W=2048, S=32, q=49): size-independent properties instead of an oracle run:
    length pinned at W, sinks never move, provenance is sinks + the last W-S token ids, values are pure moves."""
    torch.manual_seed(2)
    kvh, d, W, S, q = 4, 128, 2048, 32, 49
    c = SinkCache(window_length=W, num_sink_tokens=S)
    fed = 0
    tag = lambda n0, n: (torch.arange(n0, n0 + n, dtype=torch.float32) % 251).view(1, 1, n, 1).expand(1, kvh, n, d)
    first = None
    for step in range(48):
        n = 85 if step == 0 else q
        L = c.get_seq_length()
        cos, sin = o.rope_cos_sin(torch.arange(L, L + n), d, 1e6, torch.bfloat16)
        v = tag(fed, n).bfloat16().to(cuda).contiguous()
        k = torch.randn(1, kvh, n, d, device=cuda).bfloat16()
        K, V = c.update(k, v, 0, {"cos": cos.to(cuda), "sin": sin.to(cuda)})
        if first is None:
            first = K[:, :, :S].clone()
        fed += n
        assert K.shape[-2] == min(fed, W) if fed < W else K.shape[-2] == W
    ids = c.provenance(0).cpu().long()
    expect = torch.cat([torch.arange(S), torch.arange(fed - (W - S), fed)])
    assert torch.equal(ids, expect)
    assert torch.equal(K[:, :, :S], first)                                  # sinks untouched
    assert torch.equal(V[0, 0, :, 0].cpu().float(), (expect % 251).float())   # values: pure moves in eviction order
    assert c._seen_tokens == fed
