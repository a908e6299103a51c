"""aha_b200: B200-native (sm_100a) implementation of Aha!'s per-frame streaming ingest-and-score step.
Importing the package does not need a GPU; every compute call goes through libaha_b200.so and raises without it."""
from . import _lib
from .config import AhaConfig, VisionConfig, config_preset
from .caches import SinkCache, SlidingWindowCache, TrulyStaticCache, make_cache
from .model import VideoHeadLiveLlavaQwenForCausalLM, VideoHeadCausalLMOutputWithPast, build_model_from_state_dict
from .postproc import find_ticks, find_ticks_batched
from .highlight import (fuse_scores, fuse_scores_batched, knapsack_selection, knapsack_select_batched, highlight_indices,
                        highlight_mask_batched)
from .ingest import ingest_frames, select_frames, load_video_for_testing, load_one_frame, fit_geometry
from .live_infer import LiveInferForBenchmark, LiveInferForDemo, round_numbers
from .stream_group import LiveInferGroup
from .compat import build_model_and_tokenizer, SigLipImageProcessor, SyntheticTokenizer
from .weights import random_init_state_dict, synthetic_frames, merge_lora_state_dict

__all__ = ["AhaConfig", "VisionConfig", "config_preset", "SinkCache", "SlidingWindowCache", "TrulyStaticCache",
           "make_cache", "VideoHeadLiveLlavaQwenForCausalLM", "VideoHeadCausalLMOutputWithPast",
           "build_model_from_state_dict", "find_ticks", "find_ticks_batched", "LiveInferForBenchmark", "LiveInferForDemo",
           "LiveInferGroup", "build_model_and_tokenizer", "SigLipImageProcessor", "SyntheticTokenizer", "round_numbers", "random_init_state_dict", "synthetic_frames", "fuse_scores", "fuse_scores_batched",
           "knapsack_selection", "knapsack_select_batched", "highlight_indices", "highlight_mask_batched",
           "merge_lora_state_dict", "ingest_frames", "select_frames", "load_video_for_testing", "load_one_frame",
           "fit_geometry"]
