"""Timing breakdown of one steady-state 7B decoder step (no vision): run once per AHA_SKIP / AHA_PDL setting.
Bring-up aid, never a bench value.   python scripts/gpu_decoder_breakdown.py [frames]"""
import os, sys, json, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import aha_b200 as A

dev = torch.device("cuda", 0)
cfg = A.config_preset(os.environ.get("AHA_PRESET", "7b"))
sd = A.random_init_state_dict(cfg, seed=0, device=dev)
model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, sd, device=dev, max_tokens=256, max_frames=1, logits_mode="none")
del sd
q, W = cfg.frame_num_tokens, int(os.environ.get("AHA_WINDOW", "2048"))
cache = model.new_cache(os.environ.get("AHA_CACHE", "default_sink"), W, 32, max_chunk=q + 8)
x = (torch.randn(1, q, cfg.hidden_size, device=dev) * 0.02).to(torch.bfloat16)
scores = torch.zeros(1, 3, device=dev)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
for _ in range(W // q + 8):
    model(inputs_embeds=x, past_key_values=[cache], scores_out=scores)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n):
    model(inputs_embeds=x, past_key_values=[cache], scores_out=scores)
e1.record(); torch.cuda.synchronize()
print(json.dumps(dict(skip=os.environ.get("AHA_SKIP", "0"), pdl=os.environ.get("AHA_PDL", "1"),
                      ms_per_frame=e0.elapsed_time(e1) / n)))
