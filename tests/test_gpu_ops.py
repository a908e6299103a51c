"""Parity of the single operators of libaha_b200.so (called through the C-ABI) against torch / the oracle."""
import ctypes
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from aha_b200 import _lib
from oracle import aha_oracle as o

pytestmark = pytest.mark.gpu


def _st():
    return _lib.stream_ptr()


@pytest.mark.parametrize("rows,cols", [(49, 3584), (85, 896), (7, 256), (300, 1152)])
def test_rmsnorm(cuda, rows, cols):
    torch.manual_seed(0)
    x = (torch.randn(rows, cols) * 2).bfloat16()
    w = (1 + 0.1 * torch.randn(cols)).bfloat16()
    ref = o.rms_norm(x, w, 1e-6)
    out = torch.empty(rows, cols, dtype=torch.bfloat16, device=cuda)
    xd, wd = x.to(cuda), w.to(cuda)          # keep the device tensors alive across the raw-pointer call
    _lib.check(_lib.load().aha_rmsnorm(_lib.ptr(xd), _lib.ptr(wd), _lib.ptr(out), rows, cols, 1e-6, _st()))
    d = (out.cpu().float() - ref.float()).abs().max().item()
    # same op order as the reference; only the fp32 reduction order differs -> at most one bf16 ulp on a few elements
    assert d <= 2 ** -6 * ref.float().abs().max().item(), d
    assert (out.cpu() != ref).float().mean().item() < 0.01


@pytest.mark.parametrize("rows,cols", [(729, 1152), (81, 144), (10, 4304 - 4304 % 8), (5832, 1152), (1031, 72), (2048, 1280)])
def test_layernorm(cuda, rows, cols):
    torch.manual_seed(1)
    x = (torch.randn(rows, cols) * 3 + 0.5).bfloat16()
    w = (1 + 0.1 * torch.randn(cols)).bfloat16()
    b = (0.1 * torch.randn(cols)).bfloat16()
    ref = F.layer_norm(x.float(), (cols,), w.float(), b.float(), 1e-6)
    out = torch.empty(rows, cols, dtype=torch.bfloat16, device=cuda)
    xd, wd, bd = x.to(cuda), w.to(cuda), b.to(cuda)
    _lib.check(_lib.load().aha_layernorm(_lib.ptr(xd), _lib.ptr(wd), _lib.ptr(bd), _lib.ptr(out), rows, cols, 1e-6, _st()))
    assert (out.cpu().float() - ref).abs().max().item() <= 2e-2


def test_preprocess_bit_exact(cuda):
    u = torch.arange(256, dtype=torch.uint8).repeat(37)[: 256 * 37 - 5].contiguous()
    ref = o.preprocess(u).to(torch.bfloat16)
    out = torch.empty(u.numel(), dtype=torch.bfloat16, device=cuda)
    ud = u.to(cuda)
    _lib.check(_lib.load().aha_preprocess(_lib.ptr(ud), _lib.ptr(out), u.numel(), _st()))
    assert torch.equal(out.cpu(), ref)


@pytest.mark.parametrize("mode", ["bilinear", "average", "max"])
@pytest.mark.parametrize("side,ch", [(27, 3584), (9, 256)])
def test_pool(cuda, mode, side, ch):
    torch.manual_seed(2)
    dims = o.dims_preset("micro72")
    dims.vision.image, dims.vision.patch, dims.pool_mode = side * 14, 14, mode
    x = torch.randn(2, side * side, ch).bfloat16()
    ref = o.post_projector_pooling(dims, x)
    os_ = int(math.isqrt(ref.shape[1]))
    out = torch.empty(2, os_ * os_, ch, dtype=torch.bfloat16, device=cuda)
    xd = x.to(cuda)
    _lib.check(_lib.load().aha_pool(_lib.ptr(xd), _lib.ptr(out), 2, side, ch, 4, _lib.POOL_MODES[mode], _st()))
    d = (out.cpu().float() - ref.float()).abs().max().item()
    assert d <= (0 if mode == "max" else 2 ** -6 * 4), d


def test_pool_adaptive(cuda):
    torch.manual_seed(3)
    x = torch.randn(2, 27 * 27, 1152).bfloat16()
    ref = o.adaptive_pool_alt(x.float(), 27, (7, 7))
    out = torch.empty(2, 49, 1152, dtype=torch.bfloat16, device=cuda)
    xd = x.to(cuda)
    _lib.check(_lib.load().aha_pool(_lib.ptr(xd), _lib.ptr(out), 2, 27, 1152, 7, _lib.POOL_MODES["adaptive"], _st()))
    assert (out.cpu().float() - ref).abs().max().item() <= 2e-2


@pytest.mark.parametrize("pos0,q,d", [(0, 85, 128), (2048, 49, 128), (1999, 49, 64)])
def test_rope_table(cuda, pos0, q, d):
    cos, sin = o.rope_cos_sin(torch.arange(pos0, pos0 + q), d, 1e6, torch.bfloat16)
    c = torch.empty(q, d, dtype=torch.bfloat16, device=cuda)
    s = torch.empty_like(c)
    _lib.check(_lib.load().aha_rope_table(pos0, q, d, 1e6, _lib.ptr(c), _lib.ptr(s), _st()))
    # fp32 cos/sin of arguments up to ~2e3 rad: libm implementations differ by an ulp -> at most one bf16 ulp
    assert (c.cpu().float() - cos[0].float()).abs().max().item() <= 2 ** -7
    assert (s.cpu().float() - sin[0].float()).abs().max().item() <= 2 ** -7
    assert (c.cpu() != cos[0]).float().mean().item() < 0.02


def _attn_ref(q, k, v, off, scale):
    # q [heads, nq, d], k/v [kvh, klen, d] (fp32 math on bf16 inputs)
    heads, nq, d = q.shape
    kvh, klen, _ = k.shape
    G = heads // kvh
    kr = k.float().repeat_interleave(G, 0)
    vr = v.float().repeat_interleave(G, 0)
    s = torch.einsum("hqd,hkd->hqk", q.float(), kr) * scale
    allowed = torch.arange(klen, device=q.device)[None, :] <= (off + torch.arange(nq, device=q.device)[:, None])
    s = s.masked_fill(~allowed[None], float("-inf"))
    p = torch.softmax(s, -1)
    return torch.einsum("hqk,hkd->qhd", p, vr).reshape(nq, heads * d)


@pytest.mark.parametrize("heads,kvh,nq,klen,d,off", [
    (4, 2, 49, 49, 64, 0),          # first chunk: causal
    (28, 4, 49, 2048, 128, 2048),   # steady state: everything visible
    (28, 4, 85, 109, 128, 24),      # growing: query chunk already cached
    (14, 2, 49, 300, 64, 251),      # growing, ragged tiles
    (4, 4, 130, 700, 128, 570),     # two query tiles
    (28, 4, 49, 8192, 128, 8192),   # sliding window 8k
])
def test_attention_vs_torch(cuda, heads, kvh, nq, klen, d, off):
    torch.manual_seed(4)
    G = heads // kvh
    q = torch.randn(heads, nq, d, device=cuda).bfloat16()
    k = torch.randn(kvh, klen, d, device=cuda).bfloat16()
    v = torch.randn(kvh, klen, d, device=cuda).bfloat16()
    scale = d ** -0.5
    ref = _attn_ref(q, k, v, off, scale)
    qg = q.view(kvh, G, nq, d).permute(0, 2, 1, 3).contiguous()      # [kvh, nq, G, d]
    out = torch.full((nq, heads * d), float("nan"), dtype=torch.bfloat16, device=cuda)
    _lib.check(_lib.load().aha_attention(_lib.ptr(qg), _lib.ptr(k), _lib.ptr(v), _lib.ptr(out), heads, kvh, nq, klen, d,
                                         off, scale, _st()))
    torch.cuda.synchronize()
    err = (out.float() - ref).abs().max().item()
    assert not torch.isnan(out.float()).any()
    assert err <= 2e-2, err


def test_attention_lazy_rescale(cuda):
    """Scores that keep growing along the key axis force the running-max correction of O in TMEM."""
    torch.manual_seed(5)
    heads, kvh, nq, klen, d = 2, 2, 64, 1024, 128
    q = torch.ones(heads, nq, d, device=cuda).bfloat16()
    ramp = torch.linspace(-2.0, 2.0, klen, device=cuda)
    k = (ramp[None, :, None] * torch.ones(kvh, klen, d, device=cuda)).bfloat16()
    v = torch.randn(kvh, klen, d, device=cuda).bfloat16()
    ref = _attn_ref(q, k, v, klen, 1.0)
    out = torch.empty(nq, heads * d, dtype=torch.bfloat16, device=cuda)
    qg = q.view(kvh, 1, nq, d).permute(0, 2, 1, 3).contiguous()
    _lib.check(_lib.load().aha_attention(_lib.ptr(qg), _lib.ptr(k), _lib.ptr(v), _lib.ptr(out), heads, kvh, nq, klen, d,
                                         klen, 1.0, _st()))
    assert (out.float() - ref).abs().max().item() <= 3e-2


@pytest.mark.parametrize("T,F_,K,act", [(49, 3584, 3584, 0), (392, 37888, 3584, 3), (49, 37888, 3584, 3), (729, 4304, 1152, 1),
                                       (2187, 1152, 4304, 0), (1, 152064, 3584, 0)])
def test_gemm_vs_torch(cuda, T, F_, K, act):
    torch.manual_seed(6)
    x = (torch.randn(T, K, device=cuda) * 0.5).bfloat16()
    w = (torch.randn(F_, K, device=cuda) / K ** 0.5).bfloat16()
    Fo = F_ // 2 if act == 3 else F_
    out = torch.full((T, Fo), float("nan"), dtype=torch.bfloat16, device=cuda)
    _lib.check(_lib.load().aha_gemm_bf16(_lib.ptr(x), T, K, _lib.ptr(w), F_, K, K, _lib.ptr(out), Fo, 0, 0, act, None, None,
                                         0, 0, -1, 0, 0, _st()))
    ref = (x.float() @ w.float().t()).bfloat16().float()
    if act == 1:
        ref = F.gelu(ref, approximate="tanh")
    if act == 3:
        ref = (F.silu(ref[:, 0::2]).bfloat16().float() * ref[:, 1::2])
    err = (out.float() - ref.bfloat16().float()).abs().max().item()
    assert err <= 0.02 * max(1.0, ref.abs().max().item()), err


@pytest.mark.parametrize("T,F_,K,act,bn", [(5832, 3840, 1152, 0, 0), (5832, 4304, 1152, 1, 0), (5832, 1152, 4304, 0, 0),
                                          (5000, 1152, 1152, 2, 256), (3001, 520, 200, 0, 192), (5832, 4304, 1152, 1, 1128),
                                          (3001, 520, 200, 2, 1128),
                                          # CTA pairs (tcgen05.mma.cta_group::2): odd number of 128-row blocks, ragged columns
                                          (5832, 3456, 1152, 0, 2192), (5832, 4304, 1152, 1, 2256), (5832, 1152, 4304, 0, 2256),
                                          (3001, 520, 200, 2, 2192), (2945, 1000, 328, 0, 2256), (23328, 1152, 1152, 0, 2192)])
def test_gemm_persistent_tall(cuda, T, F_, K, act, bn):
    """The persistent 128 x {192, 256} kernel (two TMEM accumulators): bias, activation and an in-place residual, ragged
    token / feature / K tails.  bn = 1128: the 128-column variant; 2192 / 2256: the same tiles on CTA pairs."""
    torch.manual_seed(8)
    x = (torch.randn(T, K, device=cuda) * 0.5).bfloat16()
    w = (torch.randn(F_, K, device=cuda) / K ** 0.5).bfloat16()
    bias = (torch.randn(F_, device=cuda) * 0.1).bfloat16()
    res = (torch.randn(T, F_, device=cuda) * 0.5).bfloat16()
    out = res.clone()                                    # residual aliases the output, as in the SigLIP tower
    _lib.check(_lib.load().aha_gemm_bf16(_lib.ptr(x), T, K, _lib.ptr(w), F_, K, K, _lib.ptr(out), F_, 0, 0, act,
                                         _lib.ptr(bias), _lib.ptr(out), F_, 0, 0, bn, 0, _st()))
    y = (x.float() @ w.float().t() + bias.float()).bfloat16().float()
    if act == 1:
        y = F.gelu(y, approximate="tanh").bfloat16().float()
    if act == 2:
        y = F.gelu(y).bfloat16().float()
    ref = (res.float() + y).bfloat16().float()
    err = (out.float() - ref).abs().max().item()
    assert err <= 0.03 * max(1.0, ref.abs().max().item()), err


@pytest.mark.parametrize("splits", [2, 3, 5, 8])
def test_gemm_split_k_workspace_is_deterministic(cuda, splits):
    """Split-K through the L2 workspace: any split count gives the same bits run after run, and matches torch."""
    torch.manual_seed(9)
    T, F_, K = 49, 3584, 18944
    x = (torch.randn(T, K, device=cuda) * 0.5).bfloat16()
    w = (torch.randn(F_, K, device=cuda) / K ** 0.5).bfloat16()
    outs = []
    for _ in range(3):
        out = torch.full((T, F_), float("nan"), dtype=torch.bfloat16, device=cuda)
        _lib.check(_lib.load().aha_gemm_bf16(_lib.ptr(x), T, K, _lib.ptr(w), F_, K, K, _lib.ptr(out), F_, 0, 0, 0, None, None,
                                             0, 0, -1, 0, splits, _st()))
        outs.append(out)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    ref = (x.float() @ w.float().t()).bfloat16().float()
    assert (outs[0].float() - ref).abs().max().item() <= 0.02 * max(1.0, ref.abs().max().item())
