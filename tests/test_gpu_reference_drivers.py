"""Boundary: the reference's driver code runs against the native objects AS IT IS WRITTEN.

  * `_ReferenceStyleDriver` below is the statement sequence of test/inference.py:39-97 (constructor), :169-187
    (input_video_stream), :196-229 (_encode_frame), :283-334 (inference) with only the three imports replaced - the
    image processor comes from model.get_vision_tower(), frames are preprocessed on the host side, embeddings are parked on
    the CPU, the three scores are read with .item() - and gives the scores of aha_b200's own fast driver;
  * the reference constructor form LiveInferForBenchmark(args, peft_model_id, sink_cache, alt_cache) and LiveInferForDemo;
  * HF interop: the installed transformers Qwen2Model runs with an aha_b200 cache as `past_key_values` (Cache.update /
    get_seq_length / get_mask_sizes) and matches the native engine on the same weights.
"""
import collections
import dataclasses
import math

import numpy as np
import pytest
import torch

import aha_b200 as A
from oracle import aha_oracle as o

pytestmark = pytest.mark.gpu


@dataclasses.dataclass
class _Args:                       # the fields of LiveTestArguments the driver reads (models/arguments_live.py:40-75)
    bf16: bool = True
    fp16: bool = False
    llm_pretrained: str = "lmms-lab/llava-onevision-qwen2-7b-ov"
    lora_pretrained: str = None
    frame_fps: float = 1.0
    system_prompt: str = "A multimodal AI assistant is helping users with some activities."
    stream_end_prob_threshold: float = None
    threshold_z: float = None
    stream_end_score_sum_threshold: float = 1e30
    running_list_length: int = 20
    score_heads: str = "relevance_score,informative_score"
    remove_assistant_turns: bool = False
    repetition_penalty: float = None
    grounding_mode: bool = True
    attn_implementation: str = "sdpa"


class _ReferenceStyleDriver:
    """test/inference.py statement by statement (see the module docstring)."""

    def __init__(self, model, tokenizer, args, sink_cache=False, alt_cache="default_sink", window_length=64, num_sink_tokens=8):
        self.model, self.tokenizer = model, tokenizer
        self.model.eval()
        self.model.currently_training = False
        self.torch_dtype = torch.bfloat16
        self.image_processor = self.model.get_vision_tower().image_processor
        self.hidden_size = self.model.config.hidden_size
        self.frame_fps = args.frame_fps
        self.frame_num_tokens = self.model.config.frame_num_tokens
        self.score_heads = args.score_heads.split(',')
        self._start_ids = self.tokenizer.apply_chat_template([{'role': 'system', 'content': args.system_prompt}], return_tensors='pt').to('cuda')
        self.sink_cache, self.alt_cache = sink_cache, alt_cache
        self.window_length, self.num_sink_tokens = window_length, num_sink_tokens
        self.reset()

    def _init_cache(self):
        if self.alt_cache == "default_sink":
            self.past_key_values = A.SinkCache(window_length=self.window_length, num_sink_tokens=self.num_sink_tokens)
        elif self.alt_cache == "sliding_window":
            self.past_key_values = A.SlidingWindowCache(window_length=self.window_length)
        elif self.alt_cache == "static":
            self.past_key_values = A.TrulyStaticCache(window_size=self.window_length)

    def reset(self):
        self.query_queue = collections.deque()
        self.frame_embeds_queue = collections.deque()
        self.video_time = 0
        self.last_role = 'system'
        self.last_ids = torch.tensor([[]], device='cuda', dtype=torch.long)
        self._init_cache()
        self.debug_data_list = list()
        self.init_vision_time = False

    def input_video_stream(self, video_frames):
        video_frames = self.image_processor.preprocess(video_frames, return_tensors='pt')['pixel_values'].to('cuda').to(self.torch_dtype)
        batch_size = 32
        for batch_i in range(0, math.ceil(len(video_frames) / batch_size)):
            video_frames_batch = video_frames[batch_i * batch_size: batch_i * batch_size + batch_size]
            frame_embeds = self.model.visual_embed(video_frames_batch).split(self.frame_num_tokens)
            self.frame_embeds_queue.extend([((r + batch_i * batch_size) / self.frame_fps, f.to('cpu')) for r, f in enumerate(frame_embeds)])

    def _encode_frame(self):
        video_time, frame_embeds = self.frame_embeds_queue.popleft()
        if not self.init_vision_time:
            self.last_ids = self._start_ids
            self.init_vision_time = True
        else:
            self.last_ids = torch.tensor([[]], device='cuda', dtype=torch.long)
        inputs_embeds = torch.cat([
            self.model.get_input_embeddings()(self.last_ids).view(1, -1, self.hidden_size),
            frame_embeds.view(1, -1, self.hidden_size).to(self.last_ids.device),
        ], dim=1)
        outputs = self.model(inputs_embeds=inputs_embeds, use_cache=True, past_key_values=self.past_key_values,
                             cache_implementation="static", max_new_tokens=2048, output_attentions=False, return_dict=True)
        self.past_key_values = outputs.past_key_values
        informative_score = outputs.informative_logits[0, -1].softmax(dim=-1)[1].item()
        relevance_score = outputs.relevance_logits[0, -1].item()
        uncertainty_score = torch.exp(outputs.uncertainty[0, -1]).item()
        self.last_role = 'stream'
        return {"informative_score": informative_score, "relevance_score": relevance_score}, uncertainty_score

    def inference(self):
        while self.frame_embeds_queue:
            video_scores, uncertainty_score = self._encode_frame()
            self.debug_data_list.append(dict(time=self.video_time, **video_scores, uncertainty_score=uncertainty_score))
            self.video_time += 1 / self.frame_fps


@pytest.fixture(scope="module")
def built(cuda):
    cfg = A.config_preset("micro72")
    w = o.init_weights(o.dims_preset("micro72"), seed=0, dtype=torch.bfloat16)
    tok = A.SyntheticTokenizer(cfg.vocab_size, n_sys=10, n_query=6)
    model, tok = A.build_model_and_tokenizer(is_training=False, set_vision_inside=True, torch_dtype=torch.bfloat16,
                                             state_dict=w, config=cfg, tokenizer=tok, device=cuda, max_tokens=64, max_frames=8)
    return cfg, w, model, tok


@pytest.mark.parametrize("alt_cache", ["default_sink", "sliding_window", "static"])
def test_reference_driver_statements_run_unchanged(cuda, built, alt_cache):
    cfg, w, model, tok = built
    frames = o.synthetic_frames_varied(20, image=cfg.vision.image_size, seed=3)       # uint8 CPU tensor, as a loader returns it
    ref = _ReferenceStyleDriver(model, tok, _Args(), alt_cache=alt_cache)
    ref.input_video_stream(frames)
    ref.inference()
    fast = A.LiveInferForBenchmark(model, ref._start_ids, alt_cache=alt_cache, window_length=64, num_sink_tokens=8, frame_fps=1.0)
    fast.input_video_stream(frames.to(cuda))
    out = fast.inference()
    assert len(out) == len(ref.debug_data_list) == 20
    for a, b in zip(out, ref.debug_data_list):
        assert a["time"] == b["time"]
        for k in ("informative_score", "relevance_score", "uncertainty_score"):
            assert a[k] == pytest.approx(b[k], rel=1e-6, abs=1e-7), (k, a[k], b[k])     # same kernels, same inputs
    assert ref.past_key_values.get_seq_length() == fast.past_key_values.get_seq_length()
    if alt_cache == "default_sink":
        assert ref.past_key_values._seen_tokens == fast.past_key_values._seen_tokens


def test_reference_constructor_signature(cuda, built):
    """LiveInferForBenchmark(args, peft_model_id, sink_cache, alt_cache) + the conversation form of input_query_stream +
    LiveInferForDemo.load_one_frame / input_one_frame (test/live_infer_for_video.py:98-176)."""
    cfg, w, model, tok = built
    frames = o.synthetic_frames_varied(12, image=cfg.vision.image_size, seed=4)
    infer = A.LiveInferForBenchmark(_Args(), None, False, "default_sink", state_dict=w, config=cfg, tokenizer=tok, device=cuda,
                                    max_tokens=64, max_frames=8)
    assert infer.image_processor is infer.model.get_vision_tower().image_processor and infer.tokenizer is tok
    infer.reset()
    infer.set_fps(1.0)
    infer.input_video_stream(infer.image_processor.preprocess(frames, return_tensors='pt')['pixel_values'])
    infer.input_query_stream([{"role": "system", "content": "s"}, {'role': 'user', 'content': 'what happens?', 'time': 0}])
    out = infer.inference(verbose=False, total=12)
    assert len(out) == 12 and infer.past_key_values.get_seq_length() == 6 + 10 + 12 * cfg.frame_num_tokens
    with pytest.raises(ValueError):
        A.LiveInferForBenchmark(dataclasses.replace(_Args(), stream_end_prob_threshold=0.5), state_dict=w, config=cfg,
                                tokenizer=tok, device=cuda)
    # (the device resize needs a canvas side that is a multiple of 4: a 112-pixel variant of the small config)
    cfg4 = dataclasses.replace(cfg, vision=dataclasses.replace(cfg.vision, image_size=112), frame_resolution=112)
    w4 = A.random_init_state_dict(cfg4, seed=1, device=cuda)
    demo = A.LiveInferForDemo(_Args(), None, state_dict=w4, config=cfg4, tokenizer=tok, device=cuda, max_tokens=64, max_frames=8)
    from PIL import Image
    rng = np.random.default_rng(0)
    recs = []
    for i in range(3):
        demo.load_one_frame(frame_object=Image.fromarray(rng.integers(0, 255, (90, 160, 3), dtype=np.uint8)))
        recs.append(demo.input_one_frame())
    assert [r["frame_idx"] for r in recs] == [1, 2, 3] and all(r["response"] is None for r in recs)
    assert all(0.0 <= r["relevance_score"] <= 1.0 and r["uncertainty_score"] > 0 for r in recs)


def test_hf_qwen2model_runs_on_native_caches(cuda, built):
    """transformers' own Qwen2Model with an aha_b200 cache as past_key_values: HF computes q / k / v and the attention, the
    cache object stores and returns K / V through Cache.update() and sizes the mask through get_mask_sizes().  Growing
    regime (the installed transformers no longer hands cos / sin to update(), so a full SinkCache could not re-rotate -
    SURVEY.md 8(b)); compared with the native engine's final hidden states on the same chunks."""
    from tests.golden import make_golden_qwen2 as G
    cfg, w, model, tok = built
    dims = o.dims_preset("micro72")
    hf = G.hf_model(dims, {k: v.float() for k, v in w.items()}, torch.bfloat16).to(cuda)
    torch.manual_seed(6)
    for cache_cls, kw in ((A.SinkCache, dict(window_length=256, num_sink_tokens=8)), (A.SlidingWindowCache, dict(window_length=256))):
        hf_cache, eng_cache = cache_cls(**kw), model.new_cache("default_sink" if cache_cls is A.SinkCache else "sliding_window", 256, 8, max_chunk=64)
        for q in (17, 9, 9, 1, 9):
            x = (torch.randn(1, q, cfg.hidden_size, device=cuda) * 0.5).bfloat16()
            with torch.no_grad():
                h_hf = hf(inputs_embeds=x, past_key_values=hf_cache, use_cache=True).last_hidden_state
            h_eng = model(inputs_embeds=x, past_key_values=eng_cache, output_hidden_states=True).hidden_states
            assert hf_cache.get_seq_length() == eng_cache.get_seq_length()
            err = (h_hf.float() - h_eng.float()).abs().max().item()
            assert err <= 2e-2 * max(1.0, h_hf.float().abs().max().item()), (cache_cls.__name__, q, err)
        # the cache the HF layers filled holds what the engine's own RoPE + append kernel wrote (layer 0: same inputs)
        k_hf, k_eng = hf_cache.key_cache[0].float(), eng_cache.key_cache[0].float()
        assert (k_hf - k_eng).abs().max().item() <= 2 ** -6 * max(1.0, k_eng.abs().max().item())
