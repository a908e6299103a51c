"""One SigLIP batch (32 frames at 7B dims, vision only) inside a profiler range, for `ncu --profile-from-start off`;
without AHA_NCU it times the batch with CUDA events.   python scripts/gpu_vision_ncu.py [frames]"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import aha_b200 as A
dev = torch.device("cuda", 0)
import dataclasses
cfg = dataclasses.replace(A.config_preset("7b"), num_hidden_layers=1, vocab_size=1024)
sd = A.random_init_state_dict(cfg, seed=0, device=dev)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, sd, device=dev, max_tokens=64, max_frames=n, logits_mode="none")
frames = A.synthetic_frames(n, image=384, seed=1).to(dev)
for _ in range(3):
    model.visual_embed(frames)
torch.cuda.synchronize()
if os.environ.get("AHA_NCU"):
    torch.cuda.profiler.start()
    model.visual_embed(frames)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
else:
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        model.visual_embed(frames)
    e1.record(); torch.cuda.synchronize()
    print(f"visual_embed({n} frames): {e0.elapsed_time(e1) / 10:.3f} ms = {e0.elapsed_time(e1) / 10 / n:.4f} ms/frame")
