"""CUDA-event time of one vision batch (32 frames through the SigLIP tower + projector + pooling), with per-class GEMM /
attention launch times from the library's profiler.  A/B knobs come from the environment (AHA_ATTN_PERSIST=0/1, ...)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import aha_b200 as A

dev = torch.device("cuda", 0)
cfg = A.config_preset("7b")
sd = A.random_init_state_dict(cfg, seed=0, device=dev)
model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, sd, device=dev, max_tokens=256, max_frames=32, logits_mode="none")
del sd
frames = A.synthetic_frames(32, image=cfg.vision.image_size, seed=1).to(dev)
for _ in range(2):
    model.visual_embed(frames)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(4):
    model.visual_embed(frames)
e1.record(); torch.cuda.synchronize()
print("AHA_ATTN_PERSIST=%s: vision batch of 32 frames: %.3f ms  (%.3f ms per frame)" % (os.environ.get("AHA_ATTN_PERSIST", "default"),
      e0.elapsed_time(e1) / 4, e0.elapsed_time(e1) / 128))
