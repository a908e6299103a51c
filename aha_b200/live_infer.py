"""Host-side mirror of the reference's live drivers, restricted to ingest-and-score:

    LiveInferForBenchmark      test/inference.py:38-348   (reset / set_fps / input_video_stream / input_query_stream /
                                                            _encode_frame / _encode_query / inference)
    LiveInferForDemo           test/live_infer_for_video.py:80-229 (load_one_frame / input_one_frame / find_ticks)

Differences that are the point of the rewrite: frame embeddings stay on the device (the reference parks them in host
RAM, test/inference.py:185,213); the three per-frame scores are written by the heads kernel into a device score buffer
and read back once per stream instead of three `.item()` syncs per frame (:222-227) - unless a response threshold is
active, in which case one read per frame is needed, exactly where the reference needs its decision.  With
`overlap_vision=True`, `input_video_stream` only ENQUEUES the SigLIP batches, on a second CUDA stream, so the
tensor-bound vision tower of later batches runs underneath the HBM-bound decoder steps of earlier ones (`inference()`
waits per batch on an event); off by default, see DESIGN.md section 3.
Response generation ("next" row N1) runs when the driver is given `added_stream_generation_ids` (the reference gets them
from its tokenizer's chat template): a fired threshold then calls `_generate_response` (greedy decode against the same
cache, aha_b200/generate.py) and the token ids are recorded in `model_response_list`; without them the frame is only
recorded in `response_needed` and the stream goes on.
"""
from __future__ import annotations

import os
from collections import deque
from typing import List, Optional

import torch

from . import _lib
from .model import VideoHeadLiveLlavaQwenForCausalLM
from .postproc import find_ticks as _find_ticks


def round_numbers(data, n):
    """test/inference.py:365-375."""
    if isinstance(data, list):
        return [round_numbers(d, n) for d in data]
    if isinstance(data, dict):
        return {k: round_numbers(v, n) for k, v in data.items()}
    if isinstance(data, float):
        return round(data, n)
    return data


class LiveInferForBenchmark:
    def __init__(self, model_or_args, *args, **kwargs):
        """Two constructor forms:
          LiveInferForBenchmark(model, start_ids, ...)                                  an engine handle + prompt ids
          LiveInferForBenchmark(args, peft_model_id=None, sink_cache=False, alt_cache="default_sink")
                                                                                        the reference's own signature
        (test/inference.py:39): `args` is a LiveTestArguments-like object (models/arguments_live.py:40-75); the model and
        the tokenizer come from aha_b200.build_model_and_tokenizer(**asdict(args)), extra keyword arguments
        (state_dict=, config=, tokenizer=, device=, max_tokens= ...) are passed through to it."""
        if isinstance(model_or_args, VideoHeadLiveLlavaQwenForCausalLM):
            self._init_native(model_or_args, *args, **kwargs)
        else:
            self._init_from_args(model_or_args, *args, **kwargs)

    def _init_from_args(self, args, peft_model_id=None, sink_cache=False, alt_cache="default_sink", **build_kw):
        import dataclasses
        from .compat import build_model_and_tokenizer
        kw = dataclasses.asdict(args) if dataclasses.is_dataclass(args) else dict(vars(args))
        assert not (kw.get("bf16") and kw.get("fp16")), "only one of --bf16 true and --fp16 true can be set"
        kw = {k: v for k, v in kw.items() if k in ("llm_pretrained", "lora_pretrained", "lora_alpha", "lora_r",
                                                     "attn_implementation", "video_pooling_stride", "frame_resolution")}
        kw.update(build_kw)
        if peft_model_id and os.path.isdir(str(peft_model_id)) and not kw.get("lora_pretrained"):
            kw["lora_pretrained"] = peft_model_id
        self.peft_model_id = peft_model_id
        model, self.tokenizer = build_model_and_tokenizer(is_training=False, set_vision_inside=True,
                                                          torch_dtype=torch.bfloat16, **kw)
        self.image_processor = model.get_vision_tower().image_processor
        g = lambda name, default=None: getattr(args, name, default)
        self.system_prompt = g("system_prompt", "")
        thr = (g("threshold_z"), g("stream_end_prob_threshold"), g("stream_end_score_sum_threshold"))
        if sum(t is not None for t in thr) != 1:          # test/inference.py:84-85
            raise ValueError("only one of --stream_end_prob_threshold, --threshold_z and --stream_end_score_sum_threshold "
                             f"can be set. However, they are: {thr[1]}, {thr[0]}, {thr[2]}")
        tok, dev = self.tokenizer, model.device
        start_ids = tok.apply_chat_template([{'role': 'system', 'content': self.system_prompt}], return_tensors='pt').to(dev)
        stream_ids = tok.apply_chat_template([{}], add_stream_prompt=True, return_tensors='pt').to(dev)
        gen_ids = tok.apply_chat_template([{}], add_stream_generation_prompt=True, return_tensors='pt').to(dev)
        heads = g("score_heads", "relevance_score,informative_score")
        self._init_native(model, start_ids, added_stream_prompt_ids=stream_ids, sink_cache=sink_cache, alt_cache=alt_cache,
                          frame_fps=g("frame_fps", 2) if g("frame_fps", 2) > 0 else 2.0,
                          score_heads=heads.split(',') if isinstance(heads, str) else heads,
                          stream_end_prob_threshold=g("stream_end_prob_threshold"),
                          stream_end_score_sum_threshold=g("stream_end_score_sum_threshold"),
                          running_list_length=g("running_list_length", 20),
                          added_stream_generation_ids=None if g("grounding_mode", False) else gen_ids,
                          repetition_penalty=g("repetition_penalty"), remove_assistant_turns=g("remove_assistant_turns", False))

    def _init_native(self, model: VideoHeadLiveLlavaQwenForCausalLM, start_ids: torch.Tensor,
                 added_stream_prompt_ids: Optional[torch.Tensor] = None, sink_cache: bool = False,
                 alt_cache: str = "default_sink", window_length: int = 2048, num_sink_tokens: int = 32,
                 frame_fps: float = 2.0, score_heads=("relevance_score", "informative_score"),
                 stream_end_prob_threshold: Optional[float] = None, stream_end_score_sum_threshold: Optional[float] = None,
                 running_list_length: int = 20, max_frames: int = 4096, overlap_vision: bool = False,
                 added_stream_generation_ids: Optional[torch.Tensor] = None, eos_token_id: Optional[int] = None,
                 repetition_penalty: Optional[float] = None, remove_assistant_turns: bool = False,
                 max_new_tokens: int = 200):
        self.model = model
        self.sink_cache, self.alt_cache = sink_cache, alt_cache
        if not hasattr(self, "tokenizer"):
            self.tokenizer = None
            self.image_processor = model.get_vision_tower().image_processor
        self.overlap_vision = overlap_vision and os.environ.get("AHA_OVERLAP_VISION", "1") != "0"
        self.model.eval()
        cfg = model.config
        self.hidden_size = cfg.hidden_size
        self.frame_fps = frame_fps
        self.frame_interval = 1 / frame_fps
        self.frame_resolution = cfg.frame_resolution
        self.frame_num_tokens = cfg.frame_num_tokens
        self.score_heads = list(score_heads)
        self.stream_end_prob_threshold = stream_end_prob_threshold
        self.stream_end_score_sum_threshold = stream_end_score_sum_threshold
        self.running_list_length = running_list_length
        self._start_ids = start_ids.reshape(1, -1).to(model.device)
        self._added_stream_prompt_ids = None if added_stream_prompt_ids is None else \
            added_stream_prompt_ids.reshape(1, -1).to(model.device)
        self._added_stream_generation_ids = None if added_stream_generation_ids is None else \
            added_stream_generation_ids.reshape(1, -1).to(model.device)
        self.eos_token_id = cfg.eos_token_id if eos_token_id is None else eos_token_id
        self.repetition_penalty = repetition_penalty
        self.remove_assistant_turns = remove_assistant_turns
        self.inplace_output_ids = torch.zeros(1, max_new_tokens, device=model.device, dtype=torch.long)   # test/inference.py:73
        self._cache_args = dict(alt_cache=alt_cache, window_length=window_length, num_sink_tokens=num_sink_tokens,
                                sink_cache=sink_cache, n_sys=self._start_ids.shape[1])
        self._max_frames = max_frames
        self._score_buf = torch.zeros(max_frames, 3, dtype=torch.float32, device=model.device)
        self.reset()

    # ---- state (test/inference.py:112-155) ----
    def _init_cache(self):
        max_chunk = max(self._start_ids.shape[1] + self.frame_num_tokens + 8, 256)
        self.past_key_values = self.model.new_cache(max_chunk=max_chunk, **self._cache_args)

    def reset(self):
        self.query_queue = deque()
        self.frame_embeds_queue = deque()
        self._ready_events = {}
        self.video_time = 0
        self.last_frame_idx = -1
        self.video_tensor = None
        self.last_ids = torch.tensor([[]], device=self.model.device, dtype=torch.long)
        self._init_cache()
        self.debug_data_list = []
        self.generated_token_ids = []
        self.num_frames_no_reply = 0
        self.stream_end_prob_list = []
        self.stream_end_score_sum = 0
        self.consec_frames = 0
        self.init_vision_time = False
        self.response_needed = []
        self.model_response_list = []
        self.last_role = None
        self._n_scored = 0
        self._times = []

    def set_fps(self, fps=None, frame_interval=None):
        assert fps is not None or frame_interval is not None
        self.frame_fps = fps or 1 / frame_interval
        self.frame_interval = 1 / self.frame_fps

    # ---- vision (test/inference.py:169-187, test/live_infer_for_video.py:98-128) ----
    def input_video_stream(self, video_frames: torch.Tensor, batch_size: int = 32):
        """uint8 [T, 3, S, S]; embeds stay on the device.  With `overlap_vision` the batches are enqueued on the model's
        vision stream and this call returns at once; every batch carries an event that `_encode_frame` waits for."""
        if video_frames.dtype == torch.float32:           # image_processor.preprocess(...)['pixel_values'] (test/inference.py:176)
            video_frames = video_frames.to(self.model.device).to(torch.bfloat16)
        if not (self.overlap_vision and video_frames.is_cuda):
            for b in range(0, len(video_frames), batch_size):
                emb = self.model.visual_embed(video_frames[b:b + batch_size]).split(self.frame_num_tokens)
                for r, f in enumerate(emb):
                    self.frame_embeds_queue.append(((r + b) / self.frame_fps, f))
            return
        cur, side = torch.cuda.current_stream(), self.model.vision_stream()
        side.wait_stream(cur)                      # the frames (and their H2D copy) were produced on the caller's stream
        video_frames.record_stream(side)
        _lib.load().aha_set_sm_sharing(1)      # vision kernels that leave room for the decoder's CTAs on every SM
        with torch.cuda.stream(side):
            for b in range(0, len(video_frames), batch_size):
                batch = self.model.visual_embed(video_frames[b:b + batch_size])
                batch.record_stream(cur)           # consumed by the decoder on the caller's stream ...
                batch.record_stream(self.model.decoder_stream())   # ... or on the high-priority decoder stream
                ev = torch.cuda.Event()
                ev.record(side)
                for r, f in enumerate(batch.split(self.frame_num_tokens)):
                    self.frame_embeds_queue.append(((r + b) / self.frame_fps, f))      # the reference's (time, embeds) entries
                    if r == 0:
                        self._ready_events[id(f)] = (f, ev)                            # first frame of a batch carries its event
        _lib.load().aha_set_sm_sharing(0)

    def load_one_frame(self, frame: Optional[torch.Tensor] = None, frame_object=None, frame_path=None):
        """`frame`: a canvas [3, 384, 384] uint8 that is already resized and padded.  `frame_object` (the reference's
        keyword, test/live_infer_for_video.py:98-119): a decoded RGB picture of any size - PIL image, HWC numpy array or
        HWC uint8 tensor - resized (PIL bicubic, bit-identical), padded and transposed on the device (aha_b200/ingest.py)."""
        if frame_path is not None and frame_object is None and frame is None:
            from PIL import Image
            frame_object = Image.open(frame_path).convert("RGB")        # test/live_infer_for_video.py:102-103
        if (frame is None) == (frame_object is None):
            raise ValueError("give exactly one of frame / frame_object / frame_path")
        if frame_object is not None:
            from .ingest import load_one_frame as _ingest_one
            if not isinstance(frame_object, torch.Tensor):
                import numpy as np
                frame_object = torch.from_numpy(np.ascontiguousarray(np.asarray(frame_object)))
            frame = _ingest_one(frame_object, device=self.model.device, out=self.frame_resolution)[0]
        emb = self.model.visual_embed(frame.reshape(1, *frame.shape[-3:]))
        self.frame_embeds_queue.append((self.video_time if not self.frame_embeds_queue else
                                        self.frame_embeds_queue[-1][0] + self.frame_interval, emb))

    def input_query_stream(self, query_ids, video_time: float = 0.0):
        """Token ids of one query, or the reference's `conversation` list of turns ({'role': 'user', 'content', 'time'},
        test/inference.py:188-191) - those are tokenised by the chat template when they are encoded (:250)."""
        if isinstance(query_ids, (list, tuple)) and query_ids and isinstance(query_ids[0], dict):
            for turn in query_ids:
                if turn.get('role') == 'user':
                    self.query_queue.append((turn.get('time', 0), turn['content']))
            return
        self.query_queue.append((video_time, query_ids.reshape(1, -1).to(self.model.device)))

    # ---- one chunk (test/inference.py:196-262) ----
    def _next_frame_chunk(self):
        """First half of _encode_frame (test/inference.py:196-214): pops the next frame and assembles its chunk
        [1, q, H] = cat(embed_tokens(last_ids), frame_embeds).  Returns (video_time, chunk, had_ids)."""
        video_time, frame_embeds = self.frame_embeds_queue.popleft()
        ready = self._ready_events.pop(id(frame_embeds), None)
        if ready is not None:
            torch.cuda.current_stream().wait_event(ready[1])
        if not self.init_vision_time:
            self.last_ids = self._start_ids
            self.init_vision_time = True
        elif self.last_role == 'assistant' and not self.remove_assistant_turns and self._added_stream_prompt_ids is not None:
            self.last_ids = torch.cat([self.last_ids, self._added_stream_prompt_ids], dim=1)     # test/inference.py:205-206
        else:       # last_role is stream, now we just input another frame
            self.last_ids = torch.tensor([[]], device=self.model.device, dtype=torch.long)
        if self.last_ids.numel() == 0:
            return video_time, frame_embeds.view(1, -1, self.hidden_size), False
        return video_time, self.model.assemble_chunk(self.last_ids, frame_embeds), True

    def _frame_done(self, video_time, past_key_values):
        """Second half of _encode_frame (test/inference.py:218-229): adopts the cache, records the frame."""
        self.past_key_values = past_key_values
        self.last_role = 'stream'
        self._times.append(video_time)
        self._n_scored += 1

    def _encode_frame(self):
        video_time, x, _ = self._next_frame_chunk()
        if self._n_scored >= self._max_frames:
            raise RuntimeError("score buffer full: raise max_frames")
        slot = self._score_buf[self._n_scored:self._n_scored + 1]
        out = self.model(inputs_embeds=x, use_cache=True, past_key_values=self.past_key_values, return_dict=True,
                         scores_out=slot)
        self._frame_done(video_time, out.past_key_values)
        return out

    def _encode_query(self):
        _, ids = self.query_queue.popleft()
        if isinstance(ids, str):
            if self.tokenizer is None:
                raise RuntimeError("a text query needs a tokenizer (construct the driver from args, or pass token ids)")
            ids = self.tokenizer.apply_chat_template([{'role': 'user', 'content': ids}], add_stream_prompt=True,
                                                     add_stream_query_prompt=self.last_role == 'stream',
                                                     return_tensors='pt').to(self.model.device)
        x = self.model.assemble_chunk(ids, None)
        out = self.model(inputs_embeds=x, use_cache=True, past_key_values=self.past_key_values, return_dict=True)
        self.past_key_values = out.past_key_values
        self.last_query_output = out
        self.last_role = 'user'
        return out

    def _generate_response(self):
        """test/inference.py:264-281 without the tokenizer: returns the generated token ids."""
        from .generate import fast_greedy_generate
        self.last_ids = self._added_stream_generation_ids
        inputs_embeds = self.model.get_input_embeddings()(self.last_ids)
        output_ids, past_key_values, self.generated_token_ids = fast_greedy_generate(
            model=self.model, inputs_embeds=inputs_embeds, past_key_values=self.past_key_values,
            eos_token_id=self.eos_token_id, inplace_output_ids=self.inplace_output_ids,
            repetition_penalty=self.repetition_penalty, generated_token_ids=self.generated_token_ids)
        if not self.remove_assistant_turns:
            self.past_key_values = past_key_values
            self.last_ids = output_ids[:, -1:].clone()
        else:
            self.last_ids = torch.tensor([[]], device=self.model.device, dtype=torch.long)
        self.num_frames_no_reply = 0
        self.last_role = 'assistant'
        return output_ids[0].tolist()

    def _after_frame_scores(self, inf: float, rel: float):
        """Response trigger of the loop (test/inference.py:306-319) for one scored frame."""
        scores = dict(informative_score=inf, relevance_score=rel)
        stream_end_score = sum(v for k, v in scores.items() if k in self.score_heads)
        self.stream_end_prob_list.append(stream_end_score)
        self.stream_end_score_sum += stream_end_score
        if isinstance(self.running_list_length, int) and self.running_list_length > 0:
            self.stream_end_prob_list = self.stream_end_prob_list[-self.running_list_length:]
        need = False
        if self.stream_end_score_sum_threshold is not None and self.stream_end_score_sum > self.stream_end_score_sum_threshold:
            need = True
            self.stream_end_score_sum = 0
        if self.stream_end_prob_threshold is not None and stream_end_score > self.stream_end_prob_threshold:
            need = True
        if need:
            self.response_needed.append(self.video_time)
            if self._added_stream_generation_ids is not None:
                response = self._generate_response()
                self.model_response_list.append({'time': self.video_time, 'content': response, 'role': 'assistant'})

    @property
    def _wants_decisions(self) -> bool:
        return self.stream_end_prob_threshold is not None or self.stream_end_score_sum_threshold is not None

    def input_one_frame(self):
        """One iteration of the loop (test/live_infer_for_video.py:135-176)."""
        if self.query_queue and self.video_time >= self.query_queue[0][0]:
            self._encode_query()
        self._encode_frame()
        if self._wants_decisions:
            inf, rel, unc = self._score_buf[self._n_scored - 1].tolist()       # one sync, only when a decision is due
            self._after_frame_scores(inf, rel)
        self.video_time += 1 / self.frame_fps

    def inference(self, verbose=False, total=None):
        """test/inference.py:283-334; fills `debug_data_list` with one dict per frame."""
        if self.overlap_vision and self._ready_events:
            # vision batches are in flight on the vision stream: run the decoder steps on the high-priority stream so the
            # block scheduler interleaves them with the (much larger) vision grids instead of queueing behind them
            cur, hp = torch.cuda.current_stream(), self.model.decoder_stream()
            hp.wait_stream(cur)
            _lib.load().aha_set_sm_sharing(1)
            try:
                with torch.cuda.stream(hp):
                    while self.frame_embeds_queue:
                        self.input_one_frame()
                    self._flush_scores()
            finally:
                _lib.load().aha_set_sm_sharing(0)
            cur.wait_stream(hp)
            return self.debug_data_list
        while self.frame_embeds_queue:
            self.input_one_frame()
        self._flush_scores()
        return self.debug_data_list

    def _flush_scores(self):
        n0 = len(self.debug_data_list)
        if self._n_scored > n0:
            host = self._score_buf[n0:self._n_scored].cpu().tolist()          # the single device->host read
            for t, (inf, rel, unc) in zip(self._times[n0:self._n_scored], host):
                self.debug_data_list.append(dict(time=t, informative_score=inf, relevance_score=rel, uncertainty_score=unc))

    # ---- post-processing (test/live_infer_for_video.py:195-228) ----
    def find_ticks(self, scores, fps, min_separation=10, prominence=0.02, thresh=False, verbose=False):
        return _find_ticks(scores, fps, min_separation, prominence, thresh, verbose, device=self.model.device)


class LiveInferForDemo(LiveInferForBenchmark):
    """test/live_infer_for_video.py:80-229: the one-frame-at-a-time driver of the interactive demo."""

    def __init__(self, model_or_args, peft_model_id=None, query=None, **kw):
        if isinstance(model_or_args, VideoHeadLiveLlavaQwenForCausalLM):
            super().__init__(model_or_args, peft_model_id, **kw)       # (model, start_ids, ...)
        else:
            super().__init__(model_or_args, peft_model_id, **kw)
        self.system_prompt = ("A multimodal AI assistant is helping users with some activities. Below is their conversation, "
                              "interleaved with the list of video frames received by the assistant.")
        self.frame_idx = 0

    def encode_given_query(self, query):
        """test/live_infer_for_video.py:89-95: a user turn typed into the demo page, encoded at once."""
        self.query_queue.appendleft((self.video_time, query))
        out = self._encode_query()
        self.last_ids = out.logits[:, -1:].argmax(dim=-1)
        return out

    def input_one_frame(self):
        """test/live_infer_for_video.py:135-176: one frame in, its record out (one device->host read per frame - the demo
        decides after every frame whether to answer)."""
        self._encode_frame()
        inf, rel, unc = self._score_buf[self._n_scored - 1].tolist()
        self.frame_idx += 1
        ret = dict(frame_idx=self.frame_idx, time=round(self.video_time, 1), uncertainty_score=unc,
                   informative_score=inf, relevance_score=rel)
        n_resp = len(self.model_response_list)
        if self._wants_decisions:
            self._after_frame_scores(inf, rel)
        ret['response'] = self.model_response_list[-1]['content'] if len(self.model_response_list) > n_resp else None
        self.video_time += 1 / self.frame_fps
        return ret

    def input_video(self, video_path, query, fps=None):
        """test/live_infer_for_video.py:179-193: decode (OpenCV, on the host), then resize / border / colour order on the
        device (aha_b200.ingest, bit-identical to the reference's cv2 calls), the whole clip through the loop."""
        import cv2
        import numpy as np
        from .ingest import load_video_for_testing
        cap = cv2.VideoCapture(video_path)
        input_fps = cap.get(cv2.CAP_PROP_FPS)
        frames = []
        while cap.isOpened():
            ok, fr = cap.read()
            if not ok:
                break
            frames.append(fr)
        cap.release()
        if not frames:
            raise ValueError(f"no frames decoded from {video_path}")
        raw = torch.from_numpy(np.stack(frames))
        out_fps = fps if fps and fps > 0 else 400 / (len(frames) / input_fps)
        video_frames, out_fps, _ = load_video_for_testing(raw, input_fps, output_fps=out_fps, max_num_frames=400,
                                                          device=self.model.device)
        self.set_fps(fps=out_fps)
        self.input_video_stream(video_frames)
        self.input_query_stream([{"role": "system", "content": self.system_prompt},
                                 {'role': 'user', 'content': query, 'time': 0}])
        self.inference(verbose=True, total=len(video_frames))
        return round_numbers(self.debug_data_list, 3), self.model_response_list
