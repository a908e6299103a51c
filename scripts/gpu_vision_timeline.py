"""Phase stamps of the SigLIP attention kernel (AHA_TS=1) and CUDA-event time of one vision batch.  Bring-up aid."""
import ctypes, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("AHA_TS", "1")
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
print("vision batch of 32 frames: %.3f ms  (%.3f ms per frame)" % (e0.elapsed_time(e1) / 4, e0.elapsed_time(e1) / 128))
a = (ctypes.c_ulonglong * 16)()
A._lib.check(A._lib.load().aha_debug_attention_timeline(a))
t = list(a)
names = ["entry", "setup", "pdl_wait", "S0 ready", "P0 written", "S(last) ready", "P(last) written", "O done", "it2 S ready",
         "it2 P written", "it1 P written", "epilogue stores issued", "exit", "epilogue begins", "O block 0 loaded", "O block 2 loaded"]
print({n: round((x - t[0]) / 1.965e3, 2) for n, x in zip(names, t) if x})
