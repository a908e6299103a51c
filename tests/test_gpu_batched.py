"""The batched multi-stream regime (BASELINE config 3): the persistent skinny GEMM (256 < tokens <= 832), one attention
launch for all streams (per-segment tensor maps, last-arriver key-split merge), one re-rotation launch for all streams,
the stream-group driver, and the robustness items that came with them.  Everything goes through the C-ABI."""
import dataclasses

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import aha_b200 as A
from aha_b200 import _lib

pytestmark = pytest.mark.gpu


def _gemm(x, w, out, act=0, bias=None, res=None, splits=0, swap=-1, bn=0):
    T, K = x.shape
    F_ = w.shape[0]
    _lib.check(_lib.load().aha_gemm_bf16(_lib.ptr(x), T, K, _lib.ptr(w), F_, K, K, _lib.ptr(out), out.shape[1], 0, 0, act,
                                         _lib.ptr(bias), _lib.ptr(res), out.shape[1] if res is not None else 0, 0, swap, bn,
                                         splits, _lib.stream_ptr()))


@pytest.mark.parametrize("T,F_,K,act,splits", [
    (392, 4608, 3584, 0, 0),      # qkv of 8 streams (auto split)
    (392, 3584, 18944, 0, 0),     # down: long K, auto split-K with the last-arriver merge
    (392, 3584, 3584, 0, 1),      # o_proj unsplit
    (392, 3584, 3584, 0, 5),
    (392, 37888, 3584, 3, 0),     # gate|up with the SiLU*mul epilogue
    (300, 1000, 2000, 0, 3),      # ragged everywhere: 92 tokens in the second column tile, 104 valid rows in the last
                                  # feature tile, K not a multiple of 64
    (257, 2048, 1024, 3, 2),
    (832, 3584, 3584, 0, 0),      # four column tiles
    (600, 1536, 8960, 3, 4),
    (392, 896, 3584, 0, 2),       # 7 weight tiles: the last CTA pair has only one valid tile
    (430, 3584, 2048, 0, 1),      # second token group with a single 208-token half
    (600, 1536, 8960, 3, 0),      # three token tiles, stream-K ranges across tile boundaries (49 ranks per token tile)
    (392, 3584, 18944, 3, 0),
    (300, 256, 4096, 0, -1),      # two feature tiles only: fewer ranks than SMs (stream-K)
    (392, 3584, 18944, 0, -1),    # stream-K: 296 k-blocks per tile shared by ranges of 112
    (600, 1536, 8960, 3, -1),
])
@pytest.mark.parametrize("bn", [416, 208, 112, 2208, 2112], ids=["cta_pair_416", "single_cta_208", "single_cta_112", "cta_pair_208", "cta_pair_112"])
def test_gemm_persistent_skinny(cuda, T, F_, K, act, splits, bn):
    """bn = 2208 / 2112: the persistent kernel on CTA pairs (tcgen05.mma.cta_group::2: two feature tiles x one token tile per
    cluster, two accumulators, each CTA loads half of the token rows); bn = 416: the first CTA-pair kernel (256 weight rows x
    416 tokens per pair, one accumulator); bn = 208: the
    single-CTA persistent kernel with 128 x 208 tiles (two accumulators), bn = 112: the same with 128 x 112 tiles (small
    matrices: more work items instead of split-K).  splits = -1: stream-K decomposition."""
    torch.manual_seed(21)
    x = (torch.randn(T, K, device=cuda) * 0.5).bfloat16()
    w = (torch.randn(F_, K, device=cuda) / K ** 0.5).bfloat16()
    Fo = F_ // 2 if act == 3 else F_
    bias = None if act == 3 else (torch.randn(F_, device=cuda) * 0.1).bfloat16()
    res = None if act == 3 else (torch.randn(T, Fo, device=cuda) * 0.5).bfloat16()
    outs = []
    for _ in range(2):
        out = torch.full((T, Fo), float("nan"), dtype=torch.bfloat16, device=cuda) if res is None else res.clone()
        _gemm(x, w, out, act, bias, out if res is not None else None, splits, bn=bn)       # residual aliases the output
        outs.append(out)
    torch.cuda.synchronize()
    assert torch.equal(outs[0], outs[1]), "the split merge must be deterministic"
    y = x.float() @ w.float().t()
    if bias is not None:
        y = y + bias.float()
    y = y.bfloat16().float()
    if act == 3:
        ref = (F.silu(y[:, 0::2]).bfloat16().float() * y[:, 1::2]).bfloat16().float()
    else:
        ref = (res.float() + y).bfloat16().float()
    err = (outs[0].float() - ref).abs().max().item()
    assert not torch.isnan(outs[0].float()).any()
    assert err <= 0.03 * max(1.0, ref.abs().max().item()), err


def test_gemm_persistent_skinny_matches_one_shot_rows(cuda):
    """The same rows through the one-shot skinny kernel (<= 256 tokens) and the persistent one: identical bits when
    neither splits K (one fp32 accumulator per element, same epilogue arithmetic)."""
    torch.manual_seed(22)
    T, F_, K = 392, 1024, 1536
    x = (torch.randn(T, K, device=cuda) * 0.5).bfloat16()
    w = (torch.randn(F_, K, device=cuda) / K ** 0.5).bfloat16()
    big = torch.empty(T, F_, dtype=torch.bfloat16, device=cuda)
    _gemm(x, w, big, splits=1)
    small = torch.empty(200, F_, dtype=torch.bfloat16, device=cuda)
    _gemm(x[:200].contiguous(), w, small, splits=1)
    assert torch.equal(big[:200], small)


def _attn_ref(q, k, v, off, scale):
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


@pytest.mark.parametrize("nq,klen,off", [(392, 8192, 8192), (392, 4096 + 77, 4096 + 77 - 392), (200, 32768, 32768)])
def test_attention_long_keys_last_arriver_merge(cuda, nq, klen, off):
    """Long key axes with more (query tile, head, split) CTAs than SMs: the splits do not wait for each other, the last
    one to arrive merges.  Run twice: the self-resetting counters must be clean for the second launch."""
    torch.manual_seed(23)
    heads, kvh, d = 28, 4, 128
    G = heads // kvh
    q = torch.randn(heads, nq, d, device=cuda).bfloat16()
    k = torch.randn(kvh, klen, d, device=cuda).bfloat16()
    v = torch.randn(kvh, klen, d, device=cuda).bfloat16()
    scale = d ** -0.5
    ref = _attn_ref(q, k, v, off, scale)
    qg = q.view(kvh, G, nq, d).permute(0, 2, 1, 3).contiguous()
    outs = []
    for _ in range(2):
        out = torch.full((nq, heads * d), float("nan"), dtype=torch.bfloat16, device=cuda)
        _lib.check(_lib.load().aha_attention(_lib.ptr(qg), _lib.ptr(k), _lib.ptr(v), _lib.ptr(out), heads, kvh, nq, klen, d,
                                             off, scale, _lib.stream_ptr()))
        outs.append(out)
    torch.cuda.synchronize()
    assert torch.equal(outs[0], outs[1])
    assert not torch.isnan(outs[0].float()).any()
    assert (outs[0].float() - ref).abs().max().item() <= 2e-2


@pytest.mark.parametrize("heads,kvh,nq,klen,d,off", [
    (64, 64, 729, 729, 80, 100000),     # whole-group mode: 384 (head, query tile) groups, ragged last key tile, all visible
    (256, 256, 729, 729, 80, 100000),   # ten groups per CTA: every pipeline runs across the group boundaries
    (60, 60, 700, 600, 128, 0),         # causal, head_dim 128
    (160, 160, 260, 77, 64, 100000),    # one key tile per group: Q.K^T of the next group is issued before this group's P.V
    (28, 4, 392, 8192, 128, 8192),      # stream-K mode: 88 groups x 64 key tiles over 148 CTAs, groups shared by up to 3 CTAs
    (28, 4, 392, 2500, 128, 2500 - 392),  # stream-K with the causal limit inside the last tiles
    (14, 2, 520, 3000, 64, 3000),
])
def test_attention_persistent_kernel(cuda, monkeypatch, heads, kvh, nq, klen, d, off):
    """The persistent kernel (attention_persistent.cu) against fp32 torch and against the one-CTA-per-tile kernel
    (AHA_ATTN_PERSIST=0) on the same inputs; twice, for the self-resetting counters of the stream-K merge."""
    torch.manual_seed(29)
    G = heads // kvh
    q = torch.randn(heads, nq, d, device=cuda).bfloat16()
    k = torch.randn(kvh, klen, d, device=cuda).bfloat16()
    v = torch.randn(kvh, klen, d, device=cuda).bfloat16()
    scale = d ** -0.5
    qg = q.view(kvh, G, nq, d).permute(0, 2, 1, 3).contiguous()

    def run():
        out = torch.full((nq, heads * d), float("nan"), dtype=torch.bfloat16, device=cuda)
        _lib.check(_lib.load().aha_attention(_lib.ptr(qg), _lib.ptr(k), _lib.ptr(v), _lib.ptr(out), heads, kvh, nq, klen, d,
                                             off, scale, _lib.stream_ptr()))
        torch.cuda.synchronize()
        return out

    monkeypatch.setenv("AHA_ATTN_PERSIST", "1")
    n0 = _lib.load().aha_attention_persistent_launch_count()
    a, b = run(), run()
    assert _lib.load().aha_attention_persistent_launch_count() == n0 + 2
    monkeypatch.setenv("AHA_ATTN_PERSIST", "0")
    c = run()
    assert _lib.load().aha_attention_persistent_launch_count() == n0 + 2
    assert torch.equal(a, b)
    assert not torch.isnan(a.float()).any()
    assert (a.float() - c.float()).abs().max().item() <= 1e-2
    step = max(1, heads // 8)                                   # fp32 reference on a subset of the heads (memory)
    for h in range(0, heads, step):
        s = (q[h].float() @ k[h // G].float().t()) * scale
        allowed = torch.arange(klen, device=cuda)[None, :] <= (off + torch.arange(nq, device=cuda)[:, None])
        s = s.masked_fill(~allowed, float("-inf"))
        ref = torch.softmax(s, -1) @ v[h // G].float()
        assert (a[:, h * d:(h + 1) * d].float() - ref).abs().max().item() <= 2e-2, h


@pytest.fixture(scope="module")
def wide(cuda):
    cfg = A.config_preset("7b")
    cfg = dataclasses.replace(cfg, num_hidden_layers=2, vision=dataclasses.replace(cfg.vision, num_hidden_layers=1))
    sd = A.random_init_state_dict(cfg, seed=0, device=cuda)
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, sd, device=cuda, max_tokens=32 * 49, max_frames=8, logits_mode="none")
    return cfg, model


@pytest.mark.parametrize("ns", [6, 8, 16, 24])
def test_batched_sink_streams_equal_single(cuda, wide, ns):
    """ns streams in one call (one attention launch with per-stream tensor maps, one re-rotation launch, the persistent
    skinny GEMM with the folded RMSNorm from 6 streams on, the tall GEMM + stand-alone norm above 832 tokens) against
    stream 2 alone, through a sink cache that evicts and re-rotates: cache bytes of the compared stream identical in
    layer 0 (same qkv inputs), scores within the bf16 bars."""
    cfg, model = wide
    q, W, S = cfg.frame_num_tokens, 160, 16
    torch.manual_seed(31)
    caches = [model.new_cache("default_sink", W, S, max_chunk=q + 8) for _ in range(ns)]
    single = model.new_cache("default_sink", W, S, max_chunk=q + 8)
    for t in range(6):
        x = (torch.randn(ns, q, cfg.hidden_size, device=cuda) * 0.05).bfloat16()
        out = model(inputs_embeds=x, past_key_values=caches)
        one = model(inputs_embeds=x[2:3].contiguous(), past_key_values=single)
        a, b = out.frame_scores[2], one.frame_scores[0]
        assert torch.isfinite(out.frame_scores).all()
        assert (a[:2] - b[:2]).abs().max().item() <= 2e-2, (t, a, b)
        assert (a[2].log() - b[2].log()).abs().item() <= 0.1, (t, a, b)
    for c in caches:
        assert c.get_seq_length() == single.get_seq_length() == W
        assert torch.equal(c.provenance(0), single.provenance(0))
    # values are never modified by the cache; layer-0 values depend only on the chunk itself
    v_b, v_s = caches[2].value_cache[0].float(), single.value_cache[0].float()
    assert (v_b - v_s).abs().max().item() <= 2 ** -6 * max(1.0, v_s.abs().max().item())


def test_batched_streams_out_of_phase(cuda, wide):
    """Streams of one call at different cache fill levels (growing / full, different ring bases): every segment of the
    batched attention launch carries its own mask geometry."""
    cfg, model = wide
    q, W = cfg.frame_num_tokens, 400
    torch.manual_seed(32)
    ns = 6
    caches = [model.new_cache("sliding_window", W, 0, max_chunk=q + 8) for _ in range(ns)]
    singles = [model.new_cache("sliding_window", W, 0, max_chunk=q + 8) for _ in range(ns)]
    # pre-fill stream s with s * 2 chunks, alone (both copies)
    for s in range(ns):
        for _ in range(2 * s):
            x = (torch.randn(1, q, cfg.hidden_size, device=cuda) * 0.05).bfloat16()
            model(inputs_embeds=x, past_key_values=caches[s])
            model(inputs_embeds=x, past_key_values=singles[s])
    for t in range(4):
        x = (torch.randn(ns, q, cfg.hidden_size, device=cuda) * 0.05).bfloat16()
        out = model(inputs_embeds=x, past_key_values=caches)
        for s in range(ns):
            one = model(inputs_embeds=x[s:s + 1].contiguous(), past_key_values=singles[s])
            a, b = out.frame_scores[s], one.frame_scores[0]
            assert (a[:2] - b[:2]).abs().max().item() <= 2e-2, (t, s, a, b)
            assert caches[s].get_seq_length() == singles[s].get_seq_length()
    for s in range(ns):
        assert torch.equal(caches[s].provenance(0), singles[s].provenance(0))


@pytest.mark.parametrize("kind,sinks", [("sliding_window", 0), ("default_sink", 32)])
def test_batched_long_windows_out_of_phase(cuda, wide, kind, sinks):
    """Eight streams against windows of 2304 keys at different fill levels (294 ... 2304 keys: growing caches with the causal
    limit inside the last tiles, a full one whose ring has wrapped, a sink cache that evicts and re-rotates): the batched
    call runs the PERSISTENT attention kernel in its stream-K mode (one team of CTAs per query tile, segments of different
    lengths, (stream, head) groups shared by up to four CTAs), the single-stream calls it is compared with the one-CTA-per-
    tile kernel with co-resident key splits."""
    cfg, model = wide
    q, W, ns = cfg.frame_num_tokens, 2304, 8
    torch.manual_seed(33)
    caches = [model.new_cache(kind, W, sinks, max_chunk=q + 8) for _ in range(ns)]
    singles = [model.new_cache(kind, W, sinks, max_chunk=q + 8) for _ in range(ns)]
    for s in range(ns):
        for _ in range(6 * (s + 1)):                       # stream 7: 48 chunks = 2352 tokens > W
            x = (torch.randn(1, q, cfg.hidden_size, device=cuda) * 0.05).bfloat16()
            model(inputs_embeds=x, past_key_values=caches[s])
            model(inputs_embeds=x, past_key_values=singles[s])
    for s in range(ns):
        assert torch.equal(caches[s].provenance(0), singles[s].provenance(0))
    count = _lib.load().aha_attention_persistent_launch_count
    for t in range(3):
        x = (torch.randn(ns, q, cfg.hidden_size, device=cuda) * 0.05).bfloat16()
        n0 = count()
        out = model(inputs_embeds=x, past_key_values=caches)
        assert count() == n0 + cfg.num_hidden_layers, "the batched call must run the persistent kernel"
        for s in range(ns):
            n1 = count()
            one = model(inputs_embeds=x[s:s + 1].contiguous(), past_key_values=singles[s])
            assert count() == n1, "a single stream at this window stays on the one-CTA-per-tile kernel"
            a, b = out.frame_scores[s], one.frame_scores[0]
            assert not torch.isnan(a).any()
            assert (a[:2] - b[:2]).abs().max().item() <= 2e-2, (t, s, a, b)
            assert caches[s].get_seq_length() == singles[s].get_seq_length()
    for s in range(ns):
        assert torch.equal(caches[s].provenance(0), singles[s].provenance(0))


def test_rejected_call_leaves_caches_untouched(cuda, wide):
    """Every argument check of aha_prefill_chunk runs before the first cache mutation."""
    cfg, model = wide
    q = cfg.frame_num_tokens
    a = model.new_cache("default_sink", 160, 16, max_chunk=q + 8)
    b = model.new_cache("default_sink", 320, 16, max_chunk=q + 8)      # different capacity
    x = (torch.randn(2, q, cfg.hidden_size, device=cuda) * 0.05).bfloat16()
    model(inputs_embeds=x[:1].contiguous(), past_key_values=a)
    len_a, seen_a = a.get_seq_length(), a._seen_tokens
    with pytest.raises(ValueError):
        model(inputs_embeds=x, past_key_values=[a, b])
    assert a.get_seq_length() == len_a and a._seen_tokens == seen_a and b.get_seq_length() == 0
    with pytest.raises(ValueError):
        model(inputs_embeds=x, past_key_values=[a, a])
    assert a.get_seq_length() == len_a and a._seen_tokens == seen_a


def test_cache_slabs_do_not_inherit_nan(cuda):
    """Fresh cache slabs are zeroed: the P.V product multiplies P = 0 with the unwritten V rows of a partly filled
    128-key tile, so recycled device memory holding NaN / Inf patterns would poison the output."""
    cfg = A.config_preset("micro72")
    sd = A.random_init_state_dict(cfg, seed=1, device=cuda)
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, sd, device=cuda, max_tokens=64, max_frames=4, logits_mode="none")
    q = cfg.frame_num_tokens
    for _ in range(3):
        # park NaN bit patterns in freed memory of the sizes the cache is about to allocate
        junk = [torch.full((cfg.num_key_value_heads * 256 * cfg.head_dim,), float("nan"), dtype=torch.bfloat16, device=cuda)
                for _ in range(8)]
        torch.cuda.synchronize()
        del junk
        torch.cuda.empty_cache()
        cache = model.new_cache("default_sink", 256, 8, max_chunk=q + 8)
        x = (torch.randn(1, q + 3, cfg.hidden_size, device=cuda) * 0.05).bfloat16()
        out = model(inputs_embeds=x, past_key_values=cache)
        out2 = model(inputs_embeds=x[:, :q].contiguous(), past_key_values=cache)
        assert torch.isfinite(out.frame_scores).all() and torch.isfinite(out2.frame_scores).all()
        del cache


def test_sink_cache_accepts_a_longer_later_chunk(cuda):
    """HF-compatible update(): the cos / sin bookkeeping table grows on demand, a later chunk may be longer than the first
    (the reference's SinkCache accepts any q: test/sink_cache.py:109-121)."""
    from oracle import aha_oracle as o
    torch.manual_seed(33)
    W, S, kvh, d = 96, 4, 2, 64
    c = A.SinkCache(W, S)
    ref = o.OracleKVCache("sink", W, S)
    pos = 0
    for qlen in (5, 40, 60, 9, 9):
        cos, sin = o.rope_cos_sin(torch.arange(pos, pos + qlen), d, 1e6, torch.bfloat16)
        k = torch.randn(1, kvh, qlen, d).bfloat16()
        v = torch.randn(1, kvh, qlen, d).bfloat16()
        Kr, Vr = ref.update(k, v, 0, cos, sin)
        K, V = c.update(k.to(cuda), v.to(cuda), 0, {"cos": cos.to(cuda), "sin": sin.to(cuda)})
        assert torch.equal(K.cpu(), Kr) and torch.equal(V.cpu(), Vr)
        pos = ref.get_seq_length()


def test_stream_group_equals_separate_drivers(cuda):
    """LiveInferGroup (one batched prefill per step for all streams) against one LiveInferForBenchmark per stream: same
    cache provenance, scores within the bf16 bars, identical record layout - including streams of different length and a
    stream with a query."""
    cfg = A.config_preset("micro72")
    sd = A.random_init_state_dict(cfg, seed=2, device=cuda)
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, sd, device=cuda, max_tokens=256, max_frames=8, logits_mode="none")
    torch.manual_seed(34)
    ns, T = 5, [20, 20, 17, 20, 9]
    start_ids = torch.randint(0, cfg.vocab_size, (1, 10), device=cuda)
    query = torch.randint(0, cfg.vocab_size, (1, 6), device=cuda)
    videos = [A.synthetic_frames(T[s], image=cfg.vision.image_size, seed=40 + s) for s in range(ns)]

    def mk():
        return A.LiveInferForBenchmark(model, start_ids, window_length=64, num_sink_tokens=8, frame_fps=1.0, max_frames=64)

    sep = []
    for s in range(ns):
        lv = mk()
        lv.input_video_stream(videos[s].to(cuda))
        if s == 1:
            lv.input_query_stream(query, video_time=3.0)
        sep.append((lv, lv.inference()))
    lives = [mk() for _ in range(ns)]
    group = A.LiveInferGroup(lives, max_frames=64)
    group.input_video_streams(videos)
    lives[1].input_query_stream(query, video_time=3.0)
    res = group.inference()
    for s in range(ns):
        a, b = res[s], sep[s][1]
        assert len(a) == len(b) == T[s]
        assert [r["time"] for r in a] == [r["time"] for r in b]
        for key in ("informative_score", "relevance_score"):
            d = np.abs(np.array([r[key] for r in a]) - np.array([r[key] for r in b])).max()
            assert d <= 2e-2, (s, key, d)
        assert torch.equal(lives[s].past_key_values.provenance(0), sep[s][0].past_key_values.provenance(0))


def test_lora_merged_engine_equals_unmerged_oracle(cuda):
    """N2 (models/modeling_live.py:153-173): the reference runs its r = 16 / alpha = 32 adapters UNMERGED through peft on
    q/k/v/o/gate/up/down of every decoder layer; the engine streams one folded matrix per projection
    (merge_lora_state_dict).  Merged engine against the oracle running the unmerged adapters, through a stream that
    evicts and re-rotates."""
    from oracle import aha_oracle as o
    dims, cfg = o.dims_preset("micro72"), A.config_preset("micro72")
    w = o.init_weights(dims, seed=4, dtype=torch.bfloat16)
    g = torch.Generator().manual_seed(8)
    adapter, lora = {}, {}
    r, alpha = 16, 32.0
    for li in range(dims.layers):
        for mod in ("self_attn.q_proj", "self_attn.k_proj", "self_attn.v_proj", "self_attn.o_proj", "mlp.gate_proj",
                    "mlp.up_proj", "mlp.down_proj"):
            name = f"model.layers.{li}.{mod}"
            out_f, in_f = w[name + ".weight"].shape
            A_ = (torch.randn(r, in_f, generator=g) * 0.05).bfloat16()
            B_ = (torch.randn(out_f, r, generator=g) * 0.05).bfloat16()
            adapter[f"base_model.model.{name}.lora_A.default.weight"] = A_
            adapter[f"base_model.model.{name}.lora_B.default.weight"] = B_
            lora[name] = (A_, B_, alpha / r)
    merged = A.merge_lora_state_dict(w, adapter, lora_alpha=alpha, r=r)
    assert not torch.equal(merged["model.layers.0.mlp.down_proj.weight"], w["model.layers.0.mlp.down_proj.weight"])
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, merged, device=cuda, max_tokens=64, max_frames=2, logits_mode="none")
    q, W, S = cfg.frame_num_tokens, 40, 4
    cache = model.new_cache("default_sink", W, S, max_chunk=32)
    ref_c = o.make_cache("default_sink", W, S)
    torch.manual_seed(12)
    for t in range(8):
        x = (torch.randn(1, q + (3 if t == 0 else 0), cfg.hidden_size) * 0.5).bfloat16()
        out = model(inputs_embeds=x.to(cuda), past_key_values=cache)
        ref = o.frame_scores(o.qwen2_prefill(w, dims, ref_c, x, want_logits=False, lora=lora))
        got = out.frame_scores[0].float().cpu().numpy()
        assert cache.get_seq_length() == ref_c.get_seq_length()
        assert abs(got[0] - ref[0]) <= 2e-2 and abs(got[1] - ref[1]) <= 2e-2, (t, got, ref)
        assert abs(np.log(got[2]) - np.log(ref[2])) <= 3e-2, (t, got, ref)
    assert torch.equal(cache.provenance(0).cpu().long(), ref_c.ids[0])
