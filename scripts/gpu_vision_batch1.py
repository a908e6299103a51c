"""SigLIP tower + projector + pooling for ONE frame (the streaming path's vision step): CUDA-event time, and a profiler
range for `ncu --profile-from-start off --metrics gpu__time_duration.sum`.  Bring-up aid."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import aha_b200 as A

dev = torch.device("cuda", 0)
cfg = A.config_preset("7b")
sd = A.random_init_state_dict(cfg, seed=0, device=dev)
model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, sd, device=dev, max_tokens=256, max_frames=32, logits_mode="none")
del sd
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1
frames = A.synthetic_frames(n, image=cfg.vision.image_size, seed=1).to(dev)
for _ in range(3):
    model.visual_embed(frames)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    model.visual_embed(frames)
e1.record(); torch.cuda.synchronize()
print("visual_embed of %d frame(s): %.3f ms" % (n, e0.elapsed_time(e1) / 20))
if os.environ.get("AHA_NCU"):
    torch.cuda.profiler.start()
    model.visual_embed(frames)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
