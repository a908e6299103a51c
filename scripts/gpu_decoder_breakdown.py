"""Timing breakdown of one steady-state 7B decoder step (no vision): run once per AHA_SKIP / AHA_PDL setting.
Bring-up aid, never a bench value.   python scripts/gpu_decoder_breakdown.py [frames]"""
import os, sys, json, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import aha_b200 as A

dev = torch.device("cuda", 0)
cfg = A.config_preset(os.environ.get("AHA_PRESET", "7b"))
sd = A.random_init_state_dict(cfg, seed=0, device=dev)
ns = int(os.environ.get("AHA_STREAMS", "1"))      # streams batched into one prefill call
q = cfg.frame_num_tokens
model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, sd, device=dev, max_tokens=max(256, ns * q), max_frames=1, logits_mode="none")
del sd
W = int(os.environ.get("AHA_WINDOW", "2048"))
cache = [model.new_cache(os.environ.get("AHA_CACHE", "default_sink"), W, 32, max_chunk=q + 8) for _ in range(ns)]
x = (torch.randn(ns, q, cfg.hidden_size, device=dev) * 0.02).to(torch.bfloat16)
scores = torch.zeros(ns, 3, device=dev)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
for _ in range(W // q + 8):
    model(inputs_embeds=x, past_key_values=cache, scores_out=scores)
torch.cuda.synchronize()
if os.environ.get("AHA_NCU"):      # ncu --profile-from-start off: one steady-state frame inside the profiler range
    torch.cuda.profiler.start()
    model(inputs_embeds=x, past_key_values=cache, scores_out=scores)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    sys.exit(0)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n):
    model(inputs_embeds=x, past_key_values=cache, scores_out=scores)
e1.record(); torch.cuda.synchronize()
print(json.dumps(dict(skip=os.environ.get("AHA_SKIP", "0"), pdl=os.environ.get("AHA_PDL", "1"), streams=ns,
                      pair=os.environ.get("AHA_PS_PAIR", "1"),
                      ms_per_frame=e0.elapsed_time(e1) / n)))

if os.environ.get("AHA_TS") == "1":
    import ctypes
    lib = A._lib.load()
    GHZ = 1.965
    a = (ctypes.c_ulonglong * 16)()
    A._lib.check(lib.aha_debug_attention_timeline(a))
    t = list(a)
    names = ["entry", "setup", "pdl_wait", "S0 ready", "P0 written", "S1 ready", "P1 written", "O done", "O parked",
             "cluster.sync", "merged", "cluster.sync2", "exit"]
    print("attention CTA(0,0,0) us since entry:", {n: round((x - t[0]) / GHZ / 1e3, 2) for n, x in zip(names, t) if x})
    n, gx, gz = ctypes.c_int(), (ctypes.c_int * 64)(), (ctypes.c_int * 64)()
    out = (ctypes.c_ulonglong * 1024)()
    A._lib.check(lib.aha_debug_gemm_timeline(ctypes.byref(n), gx, gz, out))
    gn = ["entry", "pre-wait", "pdl_wait", "accum done", "partial parked", "all arrived", "epilogue done", "exit",
          "c0 ld", "c0 end", "c1 ld", "c1 end", "c2 ld", "c2 end", "c3 ld", "c3 end"]
    for i in range(n.value):
        t = list(out[i * 16:(i + 1) * 16])
        print(f"gemm grid.x={gx[i]} splits={gz[i]} us since entry:", {k: round((x - t[0]) / GHZ / 1e3, 2) for k, x in zip(gn, t) if x})
    # tail skew of each GEMM grid: when the accumulators of its CTAs complete and when the CTAs exit, relative to the first
    import numpy as np
    for i in range(n.value):
        buf = (ctypes.c_ulonglong * 1024)()
        A._lib.check(lib.aha_debug_gemm_tails(i, buf))
        a = np.array(list(buf), dtype=np.int64).reshape(512, 2)
        a = a[a[:, 0] > 0]
        if not len(a):
            continue
        t0 = a[:, 0].min()
        acc, ex = (a[:, 0] - t0) / 1e3, (a[:, 1] - t0) / 1e3
        q = lambda v: [round(float(x), 2) for x in np.percentile(v, [0, 25, 50, 75, 100])]
        print(f"gemm grid.x={gx[i]} splits={gz[i]} ({len(a)} CTAs) us after the first accumulator: accum done min/q1/med/q3/max {q(acc)}  exit {q(ex)}")

if os.environ.get("AHA_TS") == "2":
    import ctypes
    lib = A._lib.load()
    a = (ctypes.c_ulonglong * 32)()
    A._lib.check(lib.aha_debug_attention_persistent_timeline(a))
    t = list(a)
    names = ["ready", "pdl"] + [f"S0 of item {i}" for i in range(6)] + [f"epilogue {i} done" for i in range(6)] + ["o_full(2) seen", "stats(2) seen"]
    print("persistent attention CTA 0, cycles:", {k: x - t[0] for k, x in zip(names, t[:16]) if x})
    nt = max(1, t[18])
    print(f"   per tile ({nt} tiles): softmax thread 128 waits {t[16] / nt:.0f} cycles for scores, works {t[17] / nt:.0f}; MMA thread: early waits "
          f"{t[19] / nt:.0f}, waits for P {t[20] / nt:.0f}, issues P.V + Q.K^T {t[21] / nt:.0f}")
