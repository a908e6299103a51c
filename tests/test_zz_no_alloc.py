"""aha_attention_noalloc (include/aha_b200.h): the attention entry point for callers that cannot allocate.  Its key-split
scratch is a fixed device array inside the library, so the call must (a) give aha_attention's result, call after call
(the array keeps the previous call's contents), (b) leave the device's free memory untouched, (c) reject problems beyond
the array's size with a ValueError.  Collected last on purpose (file name)."""
import pytest
import torch

from aha_b200 import _lib

pytestmark = pytest.mark.gpu

TOL = 2e-2      # the attention tests' bar against fp32 (tests/test_gpu_ops.py)


def _problem(cuda, heads, kvh, nq, klen, d, seed):
    torch.manual_seed(seed)
    G = heads // kvh
    q = torch.randn(heads, nq, d, device=cuda).bfloat16()
    k = torch.randn(kvh, klen, d, device=cuda).bfloat16()
    v = torch.randn(kvh, klen, d, device=cuda).bfloat16()
    qg = q.view(kvh, G, nq, d).permute(0, 2, 1, 3).contiguous()      # [kvh, nq, G, d]
    return q, k, v, qg


def _ref(q, k, v, off, scale):
    heads, nq, d = q.shape
    kvh, klen, _ = k.shape
    G = heads // kvh
    out = torch.empty(nq, heads * d, dtype=torch.float32, device=q.device)
    allowed = torch.arange(klen, device=q.device)[None, :] <= (off + torch.arange(nq, device=q.device)[:, None])
    for h in range(heads):
        s = (q[h].float() @ k[h // G].float().t()) * scale
        s = s.masked_fill(~allowed, float("-inf"))
        out[:, h * d:(h + 1) * d] = torch.softmax(s, -1) @ v[h // G].float()
    return out


def _call(name, qg, k, v, out, heads, kvh, nq, klen, d, off, scale):
    fn = getattr(_lib.load(), name)
    return fn(_lib.ptr(qg), _lib.ptr(k), _lib.ptr(v), _lib.ptr(out), heads, kvh, nq, klen, d, off, scale, _lib.stream_ptr())


@pytest.mark.parametrize("heads,kvh,nq,klen,d,off", [
    (28, 4, 49, 2048, 128, 2048),        # the single-stream chunk: 12 (query tile, kv head) groups x 8 key splits
    (14, 2, 49, 300, 64, 251),           # growing cache, causal limit inside the last tiles, ragged split
    (4, 4, 130, 100, 128, 100),          # one key tile: no split, no scratch
])
def test_no_alloc_attention_matches_the_allocating_entry_point(cuda, heads, kvh, nq, klen, d, off):
    scale = d ** -0.5
    outs = []
    for call in range(2):                 # two calls in a row on DIFFERENT inputs: the second finds the first one's scratch
        q, k, v, qg = _problem(cuda, heads, kvh, nq, klen, d, seed=40 + call)
        want = torch.full((nq, heads * d), float("nan"), dtype=torch.bfloat16, device=cuda)
        _lib.check(_call("aha_attention", qg, k, v, want, heads, kvh, nq, klen, d, off, scale))
        torch.cuda.synchronize()
        got = torch.full((nq, heads * d), float("nan"), dtype=torch.bfloat16, device=cuda)
        free0 = torch.cuda.mem_get_info()[0]
        reserved0 = torch.cuda.memory_reserved()
        _lib.check(_call("aha_attention_noalloc", qg, k, v, got, heads, kvh, nq, klen, d, off, scale))
        torch.cuda.synchronize()
        if call == 1:                     # (the first launch of a kernel may load its module; the second allocates nothing)
            assert torch.cuda.mem_get_info()[0] >= free0 and torch.cuda.memory_reserved() == reserved0
        assert not torch.isnan(got.float()).any()
        assert (got.float() - want.float()).abs().max().item() <= TOL
        assert (got.float() - _ref(q, k, v, off, scale)).abs().max().item() <= TOL
        outs.append(got)
    # the same inputs again: identical bits (the merge order is fixed, the scratch carries nothing over)
    q, k, v, qg = _problem(cuda, heads, kvh, nq, klen, d, seed=41)
    again = torch.empty_like(outs[1])
    _lib.check(_call("aha_attention_noalloc", qg, k, v, again, heads, kvh, nq, klen, d, off, scale))
    torch.cuda.synchronize()
    assert torch.equal(again, outs[1])


def test_no_alloc_attention_rejects_problems_beyond_its_scratch(cuda):
    # 37 groups x 4 key splits x (128 x 128 + 256) x 4 bytes = 9.4 MiB > the 8 MiB array
    heads = kvh = 37
    nq, klen, d = 100, 1024, 128
    q, k, v, qg = _problem(cuda, heads, kvh, nq, klen, d, seed=43)
    out = torch.zeros(nq, heads * d, dtype=torch.bfloat16, device=cuda)
    rc = _call("aha_attention_noalloc", qg, k, v, out, heads, kvh, nq, klen, d, klen, d ** -0.5)
    assert rc != 0
    with pytest.raises(ValueError, match="aha_attention_noalloc"):
        _lib.check(rc)
    torch.cuda.synchronize()
    assert not out.any()                                          # nothing was launched
    _lib.check(_call("aha_attention", qg, k, v, out, heads, kvh, nq, klen, d, klen, d ** -0.5))   # the allocating one takes it
    torch.cuda.synchronize()
    assert (out.float() - _ref(q, k, v, klen, d ** -0.5)).abs().max().item() <= TOL
