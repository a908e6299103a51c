"""bench.py - BASELINE.json's metric (aggregate frames/s over N B200s; p50 per-frame latency) on its config 2:
Aha 7B (Qwen2-7B + SigLIP-so400m, random init) bf16, one stream per GPU, sink cache W=2048 / S=32, synthetic 384x384
frames.  A "step" is one frame of every resident stream through the whole ingest-and-score path (SigLIP tower ->
mm_projector -> pooling -> KV append / eviction / re-rotation -> Qwen2 prefill -> score heads).

    python bench.py --gpus N --steps K --warmup W            our CUDA path (rank 0 prints one JSON line)
    python bench.py --impl reference ...                      the reference's CPU path (oracle port) on the host cores

Keys beyond the base contract: roofline (dominant kernel = the weight-streaming decoder GEMMs, HBM-bound), cpu_baseline,
e2e (host frames in pinned memory -> public API -> scores back on the host), latency (streaming, one sync per frame).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

NCU_TRAFFIC_BYTES_PER_LAUNCH = (33.48e6 + 26.48e6 + 275.48e6 + 141.32e6) / 4
VISION_BATCH = 32   # the reference embeds frames in batches of 32 (test/inference.py:181)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=256)
    ap.add_argument("--warmup", type=int, default=32)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--preset", default="7b", choices=["7b", "tiny", "micro72"])
    ap.add_argument("--streams-per-gpu", type=int, default=1)
    ap.add_argument("--window", type=int, default=2048)
    ap.add_argument("--sinks", type=int, default=32)
    ap.add_argument("--cache", default="default_sink", choices=["default_sink", "sliding_window", "static"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--overlap", action="store_true",
                    help="enqueue the vision batches on a second stream under the decoder steps (measured at 7B: 4.62 ms per "
                         "frame against 4.57 ms without - the persistent vision GEMMs need whole SMs - so it is off by default)")
    ap.add_argument("--cpu-threads", type=int, default=0)
    return ap.parse_args()


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(hbm=float(p["hbm_gbs"]), tf_burst=float(p["bf16_tflops"]), tf_sustained=float(p["bf16_tflops_sustained"]),
                    source="measured")
    except Exception:
        return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 100 ms DURING the timed region (B200_PROFILING.md).  The
    process is started - and its first sample awaited - BEFORE the warm-up steps: attaching nvidia-smi to the GPU stalls
    kernel launches for tens of milliseconds, which used to land inside the timed region (a 10 % error at K = 64).
    `mark()` is called where the timed region starts; only samples taken after it are reported."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines, self.first = index, None, [], 0

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)

            def pump():
                for ln in self.proc.stdout:
                    self.lines.append(ln)
            threading.Thread(target=pump, daemon=True).start()
            t0 = time.time()
            while not self.lines and time.time() - t0 < 10.0 and self.proc.poll() is None:
                time.sleep(0.01)
        except Exception:
            self.proc = None

    def mark(self):
        self.first = len(self.lines)

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        lines = self.lines[self.first:] or self.lines[-1:]      # a region shorter than the sampling period: nearest sample
        for ln in lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return dict(sm_mhz=(sm[len(sm) // 2] if sm else None), sm_max_mhz=mx, reasons=sorted(reasons), samples=len(sm))


# ------------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port (oracle/aha_oracle.py) timed on the host cores.  Every sampled frame runs the FULL-depth path
# at the preset's full width (26 SigLIP layers, projector, pooling, 28 decoder layers against a cache of `window` keys
# with one eviction + re-rotation, score heads).  To keep host memory bounded the layers cycle through a few distinct
# weight sets (4 vision, 2 decoder: one fp32 7B decoder layer is 0.93 GB, far beyond any L3, so every layer still
# streams its weights from DRAM exactly as 28 distinct layers would).
# ------------------------------------------------------------------------------------------------------------------
class CpuFrameSampler:
    def __init__(self, preset, window, sinks, threads):
        import copy
        import torch
        from oracle import aha_oracle as o
        torch.set_num_threads(threads)
        self.o, self.torch = o, torch
        full = o.dims_preset(preset)
        self.dims = full
        dt = torch.float32      # fp32 on the host: bf16 GEMMs have no fast CPU path in this torch build; favours the reference
        small = copy.deepcopy(full)
        nv, nd = min(4, full.vision.layers), min(2, full.layers)
        small.layers, small.vision.layers, small.vocab = nd, nv, 1024
        w = o.init_weights(small, seed=0, dtype=dt)
        for i in range(full.vision.layers):            # alias layer i -> weight set i mod n
            src, dst = o.VT + f"encoder.layers.{i % nv}.", o.VT + f"encoder.layers.{i}."
            for k in [k for k in w if k.startswith(src)]:
                w[dst + k[len(src):]] = w[k]
        for i in range(full.layers):
            src, dst = f"model.layers.{i % nd}.", f"model.layers.{i}."
            for k in [k for k in w if k.startswith(src)]:
                w[dst + k[len(src):]] = w[k]
        self.w = w
        self.run_dims = copy.deepcopy(full)
        self.run_dims.vocab = 1024
        cache = o.OracleKVCache("sink", window, sinks)
        g = torch.Generator().manual_seed(3)
        for _ in range(full.layers):                   # steady state: every layer's cache already holds `window` keys
            cache.keys.append(torch.randn(1, full.kv_heads, window, full.head_dim, generator=g, dtype=dt))
            cache.vals.append(torch.randn(1, full.kv_heads, window, full.head_dim, generator=g, dtype=dt))
            cache.ids.append(torch.arange(window))
        q = full.frame_num_tokens
        cos, sin = o.rope_cos_sin(torch.arange(window + q), full.head_dim, full.rope_theta, dt)
        cache.cos_table, cache.sin_table = cos[0], sin[0]
        self.cache, self.q, self.dt = cache, q, dt
        self.pv = o.preprocess(o.synthetic_frames(1, image=full.vision.image)).to(dt)

    def frame(self):
        """one frame through the whole path; returns (vision_s, decoder_s)"""
        o, torch = self.o, self.torch
        with torch.no_grad():
            t0 = time.perf_counter()
            emb = o.visual_embed(self.w, self.run_dims, self.pv)
            t1 = time.perf_counter()
            out = o.qwen2_prefill(self.w, self.run_dims, self.cache, emb.view(1, self.q, self.run_dims.hidden), "sdpa",
                                  want_logits=False)
            o.frame_scores(out)
            t2 = time.perf_counter()
        return t1 - t0, t2 - t1


def cpu_sample(args, threads, max_frames, budget_s, warmup_frames=1):
    smp = CpuFrameSampler(args.preset, args.window, args.sinks, threads)
    for _ in range(warmup_frames):
        smp.frame()
    vs, ds = [], []
    t0 = time.perf_counter()
    while len(vs) < max_frames and (not vs or time.perf_counter() - t0 < budget_s):
        v, d = smp.frame()
        vs.append(v); ds.append(d)
    total = sum(vs) + sum(ds)
    return dict(frames=len(vs), sec_per_frame=total / len(vs), vision_s=sum(vs) / len(vs), decoder_s=sum(ds) / len(ds),
                cpu_seconds=total)


def cpu_baseline(args):
    threads = args.cpu_threads or (os.cpu_count() or 1)
    r = cpu_sample(args, threads, max_frames=256, budget_s=12.0)
    return dict(value=1.0 / r["sec_per_frame"], unit="frames/s", cores=threads, kind="port",
                sample=(f"{r['frames']} steady-state frames (Klen={args.window}, eviction + re-rotation) at {args.preset} dims, "
                        f"full depth, fp32, through the oracle port; {r['cpu_seconds']:.1f} s of CPU work"),
                parts=dict(vision_s_per_frame=round(r["vision_s"], 4), decoder_s_per_frame=round(r["decoder_s"], 4)))


# ------------------------------------------------------------------------------------------------------------------
# reference arm: the reference's own CPU implementation of the path (oracle port; the reference is pure Python on
# torch and does not import here), all host threads, rank 0 only.  A step = one frame; the run is
# bounded to about a minute of CPU work, `steps` reports how many frames were actually timed.
# ------------------------------------------------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = args.cpu_threads or (os.cpu_count() or 1)
    warm = max(1, min(args.warmup, 2))
    r = cpu_sample(args, threads, max_frames=max(1, args.steps), budget_s=45.0, warmup_frames=warm)
    fps = 1.0 / r["sec_per_frame"]
    line = dict(impl="reference", metric="aggregate frames/sec (streams sharded over N B200); p50 per-frame latency",
                value=fps, unit="frames/s", n_gpus=args.gpus, steps=r["frames"], warmup=warm,
                ms_per_step=r["sec_per_frame"] * 1e3, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32",
                data="synthetic", config=workload_config(args, 1),
                cpu_baseline=dict(value=fps, unit="frames/s", cores=threads, kind="port",
                                  sample=(f"{r['frames']} steady-state frames at full depth and width in fp32 through the "
                                          f"oracle port; {r['cpu_seconds']:.1f} s of CPU work"),
                                  parts=dict(vision_s_per_frame=round(r["vision_s"], 4),
                                             decoder_s_per_frame=round(r["decoder_s"], 4))),
                e2e=dict(value=fps, unit="frames/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                latency=dict(p50_ms=r["sec_per_frame"] * 1e3, mode="mean per-frame time of the CPU port"))
    print(json.dumps(line), flush=True)


def workload_config(args, n_gpus):
    return dict(workload=f"Aha {args.preset} (Qwen2 + SigLIP-so400m, random init) bf16, {args.streams_per_gpu} stream(s)/GPU, "
                         f"{args.cache} cache W={args.window} S={args.sinks}, synthetic 384x384 frames, steady state "
                         f"(cache full, one eviction + re-rotation per frame)",
                streams=args.streams_per_gpu * n_gpus, window=args.window, sinks=args.sinks,
                frame_tokens=49, vision_batch=VISION_BATCH, parallelism=f"stream-sharded x{n_gpus} (no collective)",
                vision_overlap=("SigLIP batch b+1 on a second CUDA stream under the decoder steps of batch b" if args.overlap else "off"),
                l2="per-step working set (>= 14 GB of weights at 7B) exceeds the 126 MB L2; no flush needed")


# ------------------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import aha_b200 as A
    from aha_b200 import _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    lib = _lib.load()
    cfg = A.config_preset(args.preset)
    ns, q = args.streams_per_gpu, cfg.frame_num_tokens
    n_sys = 36
    image = cfg.vision.image_size

    sd = A.random_init_state_dict(cfg, seed=0, device=dev)
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, sd, device=dev, max_tokens=max(256, ns * (n_sys + q)),
                                                 max_frames=VISION_BATCH, logits_mode="none")
    del sd
    torch.cuda.empty_cache()
    start_ids = torch.randint(0, cfg.vocab_size, (1, n_sys), device=dev)
    K, W = args.steps, max(args.warmup, 3)
    prime = (args.window // q) + 4 if args.cache != "static" else 2      # frames needed to fill the cache
    total_frames = prime + W + K
    # synthetic frames: a pool of VISION_BATCH distinct frames per stream, resident in HBM (device leg) and pinned (e2e leg)
    pool_host = A.synthetic_frames(VISION_BATCH, image=image, seed=100 + rank).pin_memory()
    pool_dev = pool_host.to(dev)

    caches = [model.new_cache(args.cache, args.window, args.sinks, max_chunk=n_sys + q + 8) for _ in range(ns)]
    scores = torch.zeros(total_frames + 4096, ns, 3, dtype=torch.float32, device=dev)
    overlap = args.overlap
    side = model.vision_stream() if overlap else None
    embeds = [[None] * ns, [None] * ns]            # double buffer: batch b lives in slot b % 2
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    # vision batches: `start`/`n` = the batch the decoder is consuming (slot `slot`), `next_n` = frames of the batch already
    # enqueued ahead (overlap only), `stop_at` = frame at which the timed region ends
    state = dict(frame=0, overlap=overlap, start=0, n=0, slot=1, next_n=0, stop_at=None)

    def vision_batch(slot, n):
        """SigLIP -> projector -> pooling for the next n (<= VISION_BATCH) frames of every stream.  With overlap it is
        enqueued on the vision stream one batch AHEAD of the decoder, so it runs underneath the decoder steps of the
        current batch."""
        cur = torch.cuda.current_stream()
        if not state["overlap"]:
            for s in range(ns):
                embeds[slot][s] = model.visual_embed(pool_dev[:n]).view(n, q, cfg.hidden_size)
            return
        side.wait_stream(cur)                      # the slot's previous embeddings have been consumed on `cur`
        with torch.cuda.stream(side):
            for s in range(ns):
                e = model.visual_embed(pool_dev[:n])
                e.record_stream(cur)
                embeds[slot][s] = e.view(n, q, cfg.hidden_size)
            ready[slot].record(side)

    def batch_size(first):
        """Frames in the vision batch that starts at frame `first`: VISION_BATCH, cut so that a batch ends exactly where
        the timed region ends and nothing is computed ahead for frames beyond it - the K timed frames then carry exactly
        K frames of SigLIP work whatever K is."""
        stop = state["stop_at"]
        if stop is None:
            return VISION_BATCH
        return max(0, min(VISION_BATCH, stop - first))

    def step():
        """one frame of every resident stream, inputs already in HBM"""
        f = state["frame"]
        if f == state["start"] + state["n"]:        # the current batch is used up (or this is the first frame)
            slot = 1 - state["slot"]
            if state["overlap"] and state["next_n"]:
                n = state["next_n"]
            else:
                n = batch_size(f) or VISION_BATCH
                vision_batch(slot, n)
            if state["overlap"]:
                torch.cuda.current_stream().wait_event(ready[slot])
            state.update(start=f, n=n, slot=slot, next_n=0)
            if state["overlap"]:
                nn = batch_size(f + n)
                if nn:
                    vision_batch(1 - slot, nn)     # the NEXT batch: one vision batch per VISION_BATCH decoder steps
                state["next_n"] = nn
        slot, i = state["slot"], f - state["start"]
        if f == 0:
            x = torch.cat([model.assemble_chunk(start_ids, embeds[slot][s][i]) for s in range(ns)], 0)
        else:
            x = torch.stack([embeds[slot][s][i] for s in range(ns)], 0) if ns > 1 else embeds[slot][0][i:i + 1]
        model(inputs_embeds=x, past_key_values=caches, scores_out=scores[f])
        state["frame"] = f + 1

    def begin_region(frames):
        """Start a fresh vision batch at the next frame and end the last one `frames` later (None: open-ended)."""
        state.update(n=state["frame"] - state["start"], next_n=0,
                     stop_at=None if frames is None else state["frame"] + frames)

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    if overlap:
        # decoder steps on the model's high-priority stream (what LiveInferForBenchmark.inference does as well)
        hp = model.decoder_stream()
        hp.wait_stream(torch.cuda.current_stream())
        torch.cuda.set_stream(hp)
        lib.aha_set_sm_sharing(1)
    for _ in range(prime):       # untimed priming: fills the cache so every timed frame runs the steady-state eviction path
        step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()          # before the warm-up: nvidia-smi's start-up perturbs the GPU, its steady sampling does not
    for _ in range(W):
        step()
    assert all(c.get_seq_length() == (args.window if args.cache != "static" else c.get_seq_length()) for c in caches)
    barrier()
    sampler.mark()
    launches0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    begin_region(K)              # the timed frames carry exactly their own K frames of vision work
    e0.record()
    for _ in range(K):
        step()
    e1.record()
    barrier()
    begin_region(None)
    ms = e0.elapsed_time(e1)
    launches = _lib.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    if dist is not None:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    fps = world * ns * K / (ms / 1e3)
    return dict(rank=rank, world=world, dev=dev, model=model, caches=caches, cfg=cfg, ms=ms, fps=fps, launches=launches,
                clocks=clocks, pool_host=pool_host, start_ids=start_ids, step=step, state=state, begin_region=begin_region,
                scores=scores, dist=dist,
                lib=lib, K=K, W=W, ns=ns)


def roofline_leg(r, args, peaks):
    """CUDA-event timing of every weight-streaming GEMM launch over a dedicated timed pass of the same steps."""
    import ctypes
    import torch
    lib, cfg = r["lib"], r["cfg"]
    n = max(1, min(64, r["K"]))
    # kernel quality, not stream sharing: this pass runs the vision batches on the decoder's stream (no concurrent kernels)
    torch.cuda.synchronize(r["dev"])
    lib.aha_set_sm_sharing(0)
    was, r["state"]["overlap"] = r["state"]["overlap"], False
    r["begin_region"](n)
    lib.aha_profile_gemm(1)
    for _ in range(n):
        r["step"]()
    torch.cuda.synchronize(r["dev"])
    out = {}
    for cls, name in ((0, "skinny"), (1, "tall"), (2, "rerot"), (3, "vattn"), (4, "cattn")):
        ms, cnt, by, fl = ctypes.c_double(), ctypes.c_longlong(), ctypes.c_double(), ctypes.c_double()
        lib.aha_profile_gemm_read(cls, ctypes.byref(ms), ctypes.byref(cnt), ctypes.byref(by), ctypes.byref(fl))
        out[name] = dict(ms=ms.value, launches=cnt.value, bytes=by.value, flops=fl.value)
    lib.aha_profile_gemm(0)
    # decoder-only step time WITHOUT the per-launch events (they break the overlap of a launch's prologue with the kernel in
    # front of it): 24 consecutive frames of one vision batch, i.e. pure prefill steps
    r["begin_region"](None)
    r["step"]()                                   # computes the batch's embeddings
    d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(r["dev"])
    d0.record()
    for _ in range(24):
        r["step"]()
    d1.record()
    torch.cuda.synchronize(r["dev"])
    decoder_ms = d0.elapsed_time(d1) / 24
    r["begin_region"](None)
    r["state"]["overlap"] = was
    sk, tall, rr = out["skinny"], out["tall"], out["rerot"]
    achieved = sk["bytes"] / (sk["ms"] / 1e3) / 1e9 if sk["ms"] > 0 else 0.0
    # DRAM traffic per launch from the committed `ncu --set full` capture of the same four launches
    # (profiles/r01_ncu_full_decoder_v5.txt, capture with AHA_PF_MB=0 so that a launch is charged its own bytes only:
    # qkv 33.5, o_proj 26.5, gate|up 272.0 + 3.5, down 138.1 + 3.2 MB read + write)
    traffic = NCU_TRAFFIC_BYTES_PER_LAUNCH if (args.preset == "7b" and args.streams_per_gpu == 1) else None
    roof = dict(bound="hbm", kernel="gemm_tcgen05_kernel (decoder weight-streaming launches: qkv, o_proj, gate|up, down)",
                achieved=achieved, peak=peaks["hbm"], unit="GB/s", frac=achieved / peaks["hbm"], traffic=traffic,
                traffic_source="ncu dram__bytes_read.sum + dram__bytes_write.sum, mean over the four launches of a layer",
                peak_source=f"{peaks['source']} copy bandwidth", launches=sk["launches"],
                avg_launch_us=(sk["ms"] * 1e3 / sk["launches"]) if sk["launches"] else None,
                algorithmic_bytes_per_launch=(sk["bytes"] / sk["launches"]) if sk["launches"] else None,
                ms_per_frame=sk["ms"] / n)
    if sk["launches"] and decoder_ms > 0:
        # every byte the prefill step has to move (weights + activations of the four GEMMs of every layer) over the WHOLE
        # decoder step - attention, RoPE / append, eviction, heads and kernel boundaries included
        gbs = (sk["bytes"] / n) / (decoder_ms / 1e3) / 1e9
        roof["decoder_step"] = dict(bound="hbm", what="GEMM bytes of one frame / duration of the whole prefill step (all kernels)",
                                    achieved=gbs, peak=peaks["hbm"], unit="GB/s", frac=gbs / peaks["hbm"], ms_per_frame=decoder_ms)
    if rr["ms"] > 0 and rr["launches"]:
        gbs = rr["bytes"] / (rr["ms"] / 1e3) / 1e9
        roof["kv_eviction"] = dict(bound="hbm", kernel="rerotate_kernel (sink-cache eviction: every kept key of all layers, one launch)",
                                   achieved=gbs, peak=peaks["hbm"], unit="GB/s", frac=gbs / peaks["hbm"],
                                   launches=rr["launches"], avg_launch_us=rr["ms"] * 1e3 / rr["launches"],
                                   algorithmic_bytes_per_launch=rr["bytes"] / rr["launches"])
    if tall["ms"] > 0:
        tf = tall["flops"] / (tall["ms"] / 1e3) / 1e12
        roof["vision_gemm"] = dict(bound="tensor", achieved=tf, peak=peaks["tf_sustained"], unit="TFLOP/s",
                                   frac=tf / peaks["tf_sustained"], launches=tall["launches"], ms_per_frame=tall["ms"] / n)
    va, ca = out["vattn"], out["cattn"]
    if va["ms"] > 0 and va["launches"]:
        tf = va["flops"] / (va["ms"] / 1e3) / 1e12
        roof["vision_attention"] = dict(bound="tensor", kernel="attention_kernel (SigLIP: 729 x 729 x 72, 16 heads x 32 frames per launch)",
                                        achieved=tf, peak=peaks["tf_sustained"], unit="TFLOP/s", frac=tf / peaks["tf_sustained"],
                                        launches=va["launches"], avg_launch_us=va["ms"] * 1e3 / va["launches"],
                                        ms_per_frame=va["ms"] / n)
    if ca["ms"] > 0 and ca["launches"]:
        roof["chunk_attention"] = dict(kernel="attention_kernel (49-token frame chunk x 7 query heads per kv head against the cache, key-split)",
                                       launches=ca["launches"], avg_launch_us=ca["ms"] * 1e3 / ca["launches"],
                                       achieved=ca["bytes"] / (ca["ms"] / 1e3) / 1e9, unit="GB/s", ms_per_frame=ca["ms"] / n)
    return roof


def e2e_leg(r, args):
    """Same metric through the public API with HOST frames: pinned uint8 frames -> H2D -> visual_embed -> per-frame
    prefill -> scores read back to the host, all inside the timed region."""
    import torch
    import aha_b200 as A
    model, dev, cfg, ns = r["model"], r["dev"], r["cfg"], r["ns"]
    K = r["K"] - r["K"] % VISION_BATCH if r["K"] >= VISION_BATCH else r["K"]
    lives = [A.LiveInferForBenchmark(model, r["start_ids"], alt_cache=args.cache, window_length=args.window,
                                     num_sink_tokens=args.sinks, frame_fps=1.0, max_frames=K + 4096,
                                     overlap_vision=args.overlap) for _ in range(ns)]
    pool = r["pool_host"]
    prime = (args.window // cfg.frame_num_tokens) + 4
    # the whole clip is handed over at once, as test/inference.py does with a video: K frames in pinned host memory
    reps = (max(K, prime + VISION_BATCH) + VISION_BATCH - 1) // VISION_BATCH
    clip = pool.repeat(reps, 1, 1, 1).pin_memory()

    def run(nframes):
        for lv in lives:
            lv.input_video_stream(clip[:nframes].to(dev, non_blocking=True), batch_size=VISION_BATCH)   # H2D from pinned memory
        for lv in lives:
            lv.inference()                                                                           # ends with the D2H read

    run(prime + VISION_BATCH - prime % VISION_BATCH)
    torch.cuda.synchronize(dev)
    if r["dist"] is not None:
        r["dist"].barrier()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run(K)
    e1.record()
    torch.cuda.synchronize(dev)
    wall = time.perf_counter() - t0
    ms = max(e0.elapsed_time(e1), wall * 1e3)
    if r["dist"] is not None:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        r["dist"].all_reduce(t, op=r["dist"].ReduceOp.MAX)
        ms = float(t.item())
    frame_bytes = 3 * cfg.vision.image_size * cfg.vision.image_size
    out = dict(value=r["world"] * ns * K / (ms / 1e3), unit="frames/s", h2d_bytes_per_step=ns * frame_bytes,
               d2h_bytes_per_step=ns * 12, api="LiveInferForBenchmark.input_video_stream + inference", frames=K)
    if r["rank"] == 0 and ns == 1 and cfg.vision.image_size == 384:
        # the same stream fed from DECODED 1920 x 1080 BGR frames in pinned host memory (N4 in front of the path):
        # H2D of the raw frames, resize / border / BGR->RGB / CHW on the device, then the loop above.  One stream, rank 0.
        kr = 2 * VISION_BATCH
        raw = torch.randint(0, 256, (kr, 1080, 1920, 3), dtype=torch.uint8).pin_memory()

        def run_raw():
            frames, _, _ = A.load_video_for_testing(raw, 1.0, output_fps=1, device=dev, batch=VISION_BATCH)
            lives[0].input_video_stream(frames, batch_size=VISION_BATCH)
            lives[0].inference()

        run_raw()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        run_raw()
        torch.cuda.synchronize(dev)
        out["raw_video"] = dict(value=kr / (time.perf_counter() - t0), unit="frames/s", frames=kr,
                                h2d_bytes_per_step=1080 * 1920 * 3, source="decoded 1920x1080 BGR frames, pinned host memory",
                                api="load_video_for_testing + input_video_stream + inference", timing="host wall clock")
        del raw
    return out


def latency_leg(r, args):
    """Streaming path (test/live_infer_for_video.py:98-176): one frame at a time, vision batch 1, one host sync per frame."""
    import torch
    import aha_b200 as A
    model, dev, cfg = r["model"], r["dev"], r["cfg"]
    live = A.LiveInferForBenchmark(model, r["start_ids"], alt_cache=args.cache, window_length=args.window,
                                   num_sink_tokens=args.sinks, frame_fps=1.0, max_frames=8192)
    pool = r["pool_host"]
    prime = (args.window // cfg.frame_num_tokens) + 4
    lat = []
    for i in range(prime + 100):
        t0 = time.perf_counter()
        live.load_one_frame(pool[i % VISION_BATCH].to(dev, non_blocking=True))
        live.input_one_frame()
        live._flush_scores()
        if i >= prime:
            lat.append((time.perf_counter() - t0) * 1e3)
    lat.sort()
    return dict(p50_ms=lat[len(lat) // 2], p99_ms=lat[min(len(lat) - 1, int(len(lat) * 0.99))], frames=len(lat),
                mode="streaming: H2D frame -> vision (batch 1) -> prefill -> D2H scores, synchronised every frame")


def ingest_leg(r, peaks):
    """N4 (frame ingest, upstream of the timed path): both resize flavours on 32 decoded 1920 x 1080 frames per launch,
    CUDA events on the launching stream after warm-up; the 199 MB source exceeds the 126 MB L2."""
    import torch
    import aha_b200 as A
    dev, n, h, w = r["dev"], 32, 1080, 1920
    src = torch.randint(0, 256, (n, h, w, 3), dtype=torch.uint8, device=dev)
    new_w, new_h, _, _ = A.fit_geometry(w, h)
    out = {}
    for flavour, read in (("cv2", min(h * w * 3, new_w * new_h * 12)), ("pil", h * w * 3)):
        for _ in range(3):
            A.ingest_frames(src, flavour)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(dev)
        e0.record()
        for _ in range(10):
            A.ingest_frames(src, flavour)
        e1.record()
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1) / 10
        by = (read + 3 * 384 * 384) * n
        gbs = by / ms / 1e6
        out[flavour] = dict(kernel=f"ingest_{flavour}_kernel (1920x1080 -> 384x384 canvas, {n} frames per launch)",
                            bound="hbm" if flavour == "cv2" else "instruction issue (21x21-tap 22-bit fixed-point filter)",
                            achieved=gbs, peak=peaks["hbm"], unit="GB/s", frac=gbs / peaks["hbm"], avg_launch_us=ms * 1e3,
                            frames_per_s=n / ms * 1e3, algorithmic_bytes_per_launch=by)
    del src
    return out


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: aha_b200 has no CPU path (use --impl reference for the CPU arm)")
    peaks = measured_peaks()
    r = run_ours(args)
    if os.environ.get("AHA_NCU"):
        r["state"]["overlap"] = False
        r["begin_region"](None)
        # profiling aid (never a bench value): one vision batch + 32 steady-state frames inside a cudaProfiler range,
        # for `ncu --profile-from-start off`
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        for _ in range(int(os.environ.get("AHA_NCU_STEPS", "32"))):
            r["step"]()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        return
    roof = roofline_leg(r, args, peaks)
    e2e = e2e_leg(r, args)
    line = None
    if r["rank"] == 0:
        roof["frame_ingest"] = ingest_leg(r, peaks)
        lat = None if args.no_latency else latency_leg(r, args)
        cpu = None if (args.no_cpu_baseline or r["world"] > 1) else cpu_baseline(args)
        line = dict(metric="aggregate frames/sec (streams sharded over N B200); p50 per-frame latency", value=r["fps"],
                    unit="frames/s", n_gpus=r["world"], steps=r["K"], warmup=r["W"], ms_per_step=r["ms"] / r["K"],
                    higher_is_better=True, scaling="weak", vs_baseline=None, dtype="bf16", data="synthetic",
                    config=workload_config(args, r["world"]), clocks=r["clocks"], e2e=e2e, gpu_launches=int(r["launches"]),
                    roofline=roof, cpu_baseline=cpu, latency=lat)
        print(json.dumps(line), flush=True)
    if r["dist"] is not None:
        r["dist"].barrier()
        r["dist"].destroy_process_group()


if __name__ == "__main__":
    main()
