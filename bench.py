"""bench.py - BASELINE.json's metric (aggregate frames/s over N B200s; p50 per-frame latency).

Headline workload = BASELINE config 3: Aha 7B (Qwen2-7B + SigLIP-so400m, random init) bf16, SlidingWindowCache W = 8192,
8 concurrent streams per GPU (64 streams over 8 B200), synthetic 384x384 frames, steady state (window full).  A "step"
is one frame of EVERY resident stream through the whole ingest-and-score path (SigLIP tower -> mm_projector -> pooling ->
KV append / eviction -> batched Qwen2 prefill (M = 8 x 49) -> score heads).  `--config 2` runs BASELINE config 2 (one
stream per GPU, sink cache 2048 / 32) as the headline instead; with the default it is measured as the `single_stream`
sub-leg, which also carries the streaming p50 latency.

    python bench.py --gpus N --steps K --warmup W            our CUDA path (rank 0 prints one JSON line)
    python bench.py --impl reference ...                      the reference's CPU path (oracle port) on the host cores
    python bench.py --impl reference-gpu ...                  the oracle as PyTorch eager + SDPA on the B200 ("reference on this box")

Keys beyond the base contract: roofline (dominant kernel + per-kernel sub-legs), cpu_baseline, gpu_eager_baseline, e2e
(host frames in pinned memory -> public API -> scores back on the host), latency, single_stream (config 2), c4_stress
(static cache, 32 768-key prefix), c5_postproc (4096 x 10 000 score streams), repeats (the K-step region is timed R
times; value is the median).
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

# DRAM bytes per launch of the four single-stream decoder GEMMs, from the committed ncu --set full capture
# (profiles/r01_ncu_full_decoder_v5.txt)
NCU_TRAFFIC_BYTES_PER_LAUNCH_C2 = (33.48e6 + 26.48e6 + 275.48e6 + 141.32e6) / 4
# the same for the four batched launches of a config-3 layer (392 tokens; profiles/r02_ncu_full_gemm_c3.txt: qkv, o_proj,
# gate|up, down - read + write)
NCU_TRAFFIC_BYTES_PER_LAUNCH_C3 = (36.25e6 + 31.41e6 + 281.26e6 + 157.08e6) / 4
VISION_BATCH = 32   # the reference embeds frames in batches of 32 (test/inference.py:181)
N_SYS = 36
METRIC = "aggregate frames/sec (streams sharded over N B200); p50 per-frame latency"
CONFIGS = {
    2: dict(streams_per_gpu=1, cache="default_sink", window=2048, sinks=32),
    3: dict(streams_per_gpu=8, cache="sliding_window", window=8192, sinks=0),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=32)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--repeats", type=int, default=5, help="the K-step timed region is repeated R times; value = median")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference-gpu"])
    ap.add_argument("--preset", default="7b", choices=["7b", "tiny", "micro72"])
    ap.add_argument("--config", type=int, default=3, choices=[2, 3], help="BASELINE config of the headline value")
    ap.add_argument("--streams-per-gpu", type=int, default=None)
    ap.add_argument("--window", type=int, default=None)
    ap.add_argument("--sinks", type=int, default=None)
    ap.add_argument("--cache", default=None, choices=["default_sink", "sliding_window", "static"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the sub-legs (single_stream, c4, c5, gpu eager, ingest)")
    ap.add_argument("--cpu-threads", type=int, default=0)
    args = ap.parse_args()
    base = CONFIGS[args.config]
    for k, v in base.items():
        if getattr(args, k) is None:
            setattr(args, k, v)
    if args.cache != "default_sink":
        args.sinks = 0
    return args


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
    kernel launches for tens of milliseconds.  `mark()` is called where the timed region starts."""
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


def workload_text(args):
    cache = {"default_sink": f"SinkCache W={args.window} S={args.sinks}", "sliding_window": f"SlidingWindowCache W={args.window}",
             "static": f"TrulyStaticCache W={args.window}"}[args.cache]
    return (f"BASELINE config {args.config}: Aha {args.preset} (Qwen2 + SigLIP-so400m, random init) bf16, {args.streams_per_gpu} "
            f"stream(s)/GPU batched into one prefill per step, {cache}, synthetic 384x384 frames, steady state (cache full: "
            f"one eviction per frame and stream)")


def workload_config(args, n_gpus):
    return dict(workload=workload_text(args), baseline_config=args.config, streams=args.streams_per_gpu * n_gpus,
                streams_per_gpu=args.streams_per_gpu, cache=args.cache, window=args.window, sinks=args.sinks,
                frame_tokens=49, tokens_per_step_per_gpu=49 * args.streams_per_gpu, vision_batch=VISION_BATCH,
                parallelism=f"stream-sharded x{n_gpus} (no collective)",
                l2="per-step working set (>= 14 GB of weights + the KV windows at 7B) exceeds the 126 MB L2; no flush needed")


# ------------------------------------------------------------------------------------------------------------------
# Oracle-port samplers (CPU arm and the GPU-eager arm).  Every sampled frame runs the FULL-depth path at the preset's
# full width (26 SigLIP layers, projector, pooling, 28 decoder layers against a cache of `window` keys with one
# eviction (+ re-rotation for the sink policy), score heads).
# ------------------------------------------------------------------------------------------------------------------
def _oracle_policy(cache):
    return {"default_sink": "sink", "sliding_window": "sliding", "static": "static"}[cache]


class CpuFrameSampler:
    """fp32 on the host cores.  To keep host memory bounded the layers cycle through a few distinct weight sets (4
    vision, 2 decoder: one fp32 7B decoder layer is 0.93 GB, far beyond any L3, so every layer still streams its weights
    from DRAM exactly as 28 distinct layers would)."""

    def __init__(self, preset, cache, window, sinks, threads):
        import copy
        import torch
        from oracle import aha_oracle as o
        torch.set_num_threads(threads)
        self.o, self.torch = o, torch
        full = o.dims_preset(preset)
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
        self.q = full.frame_num_tokens
        self.cache = _steady_cache(o, torch, full, _oracle_policy(cache), window, sinks, dt, "cpu")
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


def _steady_cache(o, torch, dims, policy, window, sinks, dt, device):
    """An OracleKVCache that already holds `window` keys per layer (steady state of a long stream)."""
    cache = o.OracleKVCache(policy, window, sinks if policy == "sink" else 0)
    g = torch.Generator(device=device).manual_seed(3)
    for _ in range(dims.layers):
        cache.keys.append(torch.randn(1, dims.kv_heads, window, dims.head_dim, generator=g, dtype=dt, device=device))
        cache.vals.append(torch.randn(1, dims.kv_heads, window, dims.head_dim, generator=g, dtype=dt, device=device))
        cache.ids.append(torch.arange(window))
    if policy == "sink":
        cos, sin = o.rope_cos_sin(torch.arange(window + dims.frame_num_tokens, device=device), dims.head_dim, dims.rope_theta, dt)
        cache.cos_table, cache.sin_table = cos[0], sin[0]
    return cache


def cpu_sample(args, threads, max_frames, budget_s, warmup_frames=1):
    smp = CpuFrameSampler(args.preset, args.cache, args.window, args.sinks, threads)
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
                sample=(f"{r['frames']} steady-state frames of one stream (Klen={args.window}, {args.cache}) at {args.preset} dims, "
                        f"full depth, fp32, through the oracle port (the reference runs its streams one at a time, B = 1); "
                        f"{r['cpu_seconds']:.1f} s of CPU work"),
                parts=dict(vision_s_per_frame=round(r["vision_s"], 4), decoder_s_per_frame=round(r["decoder_s"], 4)))


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path (oracle port; the reference is pure Python
    on torch and does not import here), all host threads, rank 0 only.  A step = one frame of one stream
    (the reference has no batching across streams); bounded to about a minute of CPU work."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = args.cpu_threads or (os.cpu_count() or 1)
    warm = max(1, min(args.warmup, 2))
    r = cpu_sample(args, threads, max_frames=max(1, args.steps), budget_s=45.0, warmup_frames=warm)
    fps = 1.0 / r["sec_per_frame"]
    line = dict(impl="reference", metric=METRIC, value=fps, unit="frames/s", n_gpus=args.gpus, steps=r["frames"], warmup=warm,
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


def gpu_eager_sample(preset, cache, window, sinks, streams, frames_per_stream, dev):
    """The "reference on this box" arm: the oracle port moved to the B200 as the reference runs it -
    PyTorch eager in bf16, SDPA, `torch.cat` caches (test/sink_cache.py:124-164), vision in batches of 32
    (test/inference.py:181), one prefill per frame and stream, three `.item()` reads per frame (:222-227).  Full depth,
    28 + 26 distinct layers.  Streams run one after the other, as separate LiveInferForBenchmark objects would."""
    import torch
    import aha_b200 as A
    from oracle import aha_oracle as o
    dims, cfg = o.dims_preset(preset), A.config_preset(preset)
    dt = torch.bfloat16
    w = A.random_init_state_dict(cfg, seed=0, device=dev)
    q, H = dims.frame_num_tokens, dims.hidden
    caches = [_steady_cache(o, torch, dims, _oracle_policy(cache), window, sinks, dt, dev) for _ in range(streams)]
    frames = A.synthetic_frames(VISION_BATCH, image=dims.vision.image, seed=100).to(dev)
    n = frames_per_stream

    def run(nf):
        with torch.no_grad():
            for c in caches:
                for b in range(0, nf, VISION_BATCH):
                    nb = min(VISION_BATCH, nf - b)
                    pv = o.preprocess(frames[:nb]).to(dt)                       # image_processor.preprocess + .to(bf16)
                    emb = o.visual_embed(w, dims, pv).view(nb, q, H)
                    for f in range(nb):
                        out = o.qwen2_prefill(w, dims, c, emb[f:f + 1], "sdpa", want_logits=False, attn_impl="sdpa")
                        o.frame_scores(out)                                       # three .item() syncs
    run(2)
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    run(n)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3)
    del w, caches
    torch.cuda.empty_cache()
    return dict(value=streams * n / (ms / 1e3), unit="frames/s", frames=streams * n, ms_per_frame=ms / (streams * n),
                kind="oracle port as PyTorch eager (bf16, SDPA, torch.cat caches, vision batch 32, 3 .item() per frame) on this GPU",
                config=f"{streams} stream(s) one after the other, {cache} W={window}, steady state, full depth")


def run_reference_gpu(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    n = max(8, min(64, args.steps))
    r = gpu_eager_sample(args.preset, args.cache, args.window, args.sinks, args.streams_per_gpu, n, dev)
    line = dict(impl="reference-gpu", metric=METRIC, value=r["value"], unit="frames/s", n_gpus=1, steps=n, warmup=2,
                ms_per_step=r["ms_per_frame"] * args.streams_per_gpu, higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="bf16", data="synthetic", config=workload_config(args, 1), gpu_eager_baseline=r,
                e2e=dict(value=r["value"], unit="frames/s", h2d_bytes_per_step=0, d2h_bytes_per_step=12 * args.streams_per_gpu))
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------------------
class Workload:
    """`ns` resident streams on one GPU stepped with inputs already in HBM: vision in batches of VISION_BATCH frames per
    stream, one batched prefill per step."""

    def __init__(self, model, cfg, ns, cache, window, sinks, pool_dev, start_ids, total_frames):
        import torch
        self.torch, self.model, self.cfg, self.ns = torch, model, cfg, ns
        self.q = cfg.frame_num_tokens
        self.pool, self.start_ids = pool_dev, start_ids
        self.caches = [model.new_cache(cache, window, sinks, max_chunk=N_SYS + self.q + 8) for _ in range(ns)]
        self.scores = torch.zeros(total_frames, ns, 3, dtype=torch.float32, device=model.device)
        self.embeds = [None, None]      # double buffer: [n, ns, q, H] (frame-major, stream-minor) of batch b in slot b % 2
        self.state = dict(frame=0, start=0, n=0, slot=1, stop_at=None)
        self.prime_steps = (window // self.q) + 4 if cache != "static" else 2
        self.window, self.cache = window, cache

    def vision_batch(self, slot, n):
        """SigLIP -> projector -> pooling for the next n frames of every stream (frames of the streams interleaved, so that
        frame f of all streams is one contiguous [ns, q, H] block)."""
        torch, ns = self.torch, self.ns
        outs = []
        for f0 in range(0, n, max(1, VISION_BATCH // ns)):
            nf = min(max(1, VISION_BATCH // ns), n - f0)
            fr = self.pool[f0:f0 + nf].repeat_interleave(ns, dim=0) if ns > 1 else self.pool[f0:f0 + nf]
            outs.append(self.model.visual_embed(fr).view(nf, ns, self.q, self.cfg.hidden_size))
        self.embeds[slot] = torch.cat(outs, 0) if len(outs) > 1 else outs[0]

    def batch_size(self, first):
        stop = self.state["stop_at"]
        if stop is None:
            return VISION_BATCH
        return max(0, min(VISION_BATCH, stop - first))

    def step(self):
        st = self.state
        f = st["frame"]
        if f == st["start"] + st["n"]:            # the current batch is used up (or this is the first frame)
            slot = 1 - st["slot"]
            n = self.batch_size(f) or VISION_BATCH
            self.vision_batch(slot, n)
            st.update(start=f, n=n, slot=slot)
        e = self.embeds[st["slot"]][f - st["start"]]          # [ns, q, H]
        if f == 0:
            x = self.torch.cat([self.model.assemble_chunk(self.start_ids, e[s]) for s in range(self.ns)], 0)
        else:
            x = e
        self.model(inputs_embeds=x, past_key_values=self.caches, scores_out=self.scores[f % self.scores.shape[0]])
        st["frame"] = f + 1

    def begin_region(self, frames):
        """Start a fresh vision batch at the next frame and end the last one `frames` later (None: open-ended): a timed
        region of K steps then carries exactly K frames of SigLIP work per stream."""
        st = self.state
        st.update(n=st["frame"] - st["start"], stop_at=None if frames is None else st["frame"] + frames)

    def prime(self):
        for _ in range(self.prime_steps):
            self.step()


def timed_regions(wl, K, R, dev, dist):
    """R repeats of a K-step region, each bracketed by barrier + synchronise, CUDA events, max over ranks."""
    import torch
    out = []
    for _ in range(R):
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        wl.begin_region(K)
        e0.record()
        for _ in range(K):
            wl.step()
        e1.record()
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)
        wl.begin_region(None)
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        out.append(ms)
    return out


def profile_pass(wl, lib, n):
    """CUDA-event timing of every GEMM / attention / re-rotation launch over a dedicated pass of n steps."""
    import ctypes
    torch = wl.torch
    torch.cuda.synchronize()
    wl.begin_region(n)
    lib.aha_profile_gemm(1)
    for _ in range(n):
        wl.step()
    torch.cuda.synchronize()
    out = {}
    for cls, name in ((0, "skinny"), (1, "tall"), (2, "rerot"), (3, "vattn"), (4, "cattn"), (5, "batched")):
        ms, cnt, by, fl = ctypes.c_double(), ctypes.c_longlong(), ctypes.c_double(), ctypes.c_double()
        lib.aha_profile_gemm_read(cls, ctypes.byref(ms), ctypes.byref(cnt), ctypes.byref(by), ctypes.byref(fl))
        out[name] = dict(ms=ms.value, launches=cnt.value, bytes=by.value, flops=fl.value)
    lib.aha_profile_gemm(0)
    wl.begin_region(None)
    return out


def decoder_step_ms(wl, n=24):
    """decoder-only step time WITHOUT per-launch events: n consecutive steps inside one vision batch"""
    torch = wl.torch
    wl.begin_region(None)
    wl.step()                                   # computes the batch's embeddings
    n = min(n, VISION_BATCH - 2)
    d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    d0.record()
    for _ in range(n):
        wl.step()
    d1.record()
    torch.cuda.synchronize()
    wl.begin_region(None)
    return d0.elapsed_time(d1) / n


def roofline_leg(wl, lib, K, peaks, single_stream):
    n = max(1, min(VISION_BATCH, K))
    out = profile_pass(wl, lib, n)
    dec_ms = decoder_step_ms(wl)
    ns = wl.ns
    sk, tall, rr, va, ca, bt = (out[k] for k in ("skinny", "tall", "rerot", "vattn", "cattn", "batched"))
    roof = {}
    if single_stream or bt["launches"] == 0:
        achieved = sk["bytes"] / (sk["ms"] / 1e3) / 1e9 if sk["ms"] > 0 else 0.0
        roof = dict(bound="hbm", kernel="gemm_tcgen05_kernel (decoder weight-streaming launches: qkv, o_proj, gate|up, down)",
                    achieved=achieved, peak=peaks["hbm"], unit="GB/s", frac=achieved / peaks["hbm"],
                    traffic=NCU_TRAFFIC_BYTES_PER_LAUNCH_C2 if wl.cfg.hidden_size == 3584 else None,
                    traffic_source="ncu dram__bytes_read.sum + dram__bytes_write.sum, mean over the four launches of a layer "
                                   "(profiles/r01_ncu_full_decoder_v5.txt)",
                    peak_source=f"{peaks['source']} copy bandwidth", launches=sk["launches"],
                    avg_launch_us=(sk["ms"] * 1e3 / sk["launches"]) if sk["launches"] else None,
                    algorithmic_bytes_per_launch=(sk["bytes"] / sk["launches"]) if sk["launches"] else None,
                    ms_per_step=sk["ms"] / n)
        if sk["launches"] and dec_ms > 0:
            gbs = (sk["bytes"] / n) / (dec_ms / 1e3) / 1e9
            roof["decoder_step"] = dict(bound="hbm", what="GEMM bytes of one frame / duration of the whole prefill step (all kernels)",
                                        achieved=gbs, peak=peaks["hbm"], unit="GB/s", frac=gbs / peaks["hbm"], ms_per_step=dec_ms)
    else:
        tf = bt["flops"] / (bt["ms"] / 1e3) / 1e12 if bt["ms"] > 0 else 0.0
        roof = dict(bound="tensor", kernel="gemm_swap_persistent_kernel (batched decoder GEMMs: qkv, o_proj, gate|up, down at "
                                           f"M = {49 * ns} tokens)",
                    achieved=tf, peak=peaks["tf_sustained"], unit="TFLOP/s", frac=tf / peaks["tf_sustained"],
                    traffic=NCU_TRAFFIC_BYTES_PER_LAUNCH_C3 if (wl.cfg.hidden_size == 3584 and ns == 8) else None,
                    traffic_source="ncu dram__bytes_read.sum + dram__bytes_write.sum, mean over the four launches of a layer "
                                   "(profiles/r02_ncu_full_gemm_c3.txt); algorithmic bytes are the weights + activations once",
                    peak_source=f"{peaks['source']} sustained cuBLAS bf16", launches=bt["launches"],
                    avg_launch_us=(bt["ms"] * 1e3 / bt["launches"]) if bt["launches"] else None,
                    algorithmic_flops_per_launch=(bt["flops"] / bt["launches"]) if bt["launches"] else None,
                    algorithmic_bytes_per_launch=(bt["bytes"] / bt["launches"]) if bt["launches"] else None,
                    hbm_gbs=(bt["bytes"] / (bt["ms"] / 1e3) / 1e9) if bt["ms"] > 0 else None,
                    ms_per_step=bt["ms"] / n)
        if dec_ms > 0:
            fl = (bt["flops"] + ca["flops"]) / n
            roof["decoder_step"] = dict(bound="tensor", what="decoder GEMM + chunk-attention FLOPs of one step / duration of the "
                                        "whole batched prefill step (all kernels, no per-launch events)",
                                        achieved=fl / (dec_ms / 1e3) / 1e12, peak=peaks["tf_sustained"], unit="TFLOP/s",
                                        frac=fl / (dec_ms / 1e3) / 1e12 / peaks["tf_sustained"], ms_per_step=dec_ms)
    if rr["ms"] > 0 and rr["launches"]:
        gbs = rr["bytes"] / (rr["ms"] / 1e3) / 1e9
        roof["kv_eviction"] = dict(bound="hbm", kernel="rerotate kernel (sink-cache eviction: every kept key of all layers, one launch)",
                                   achieved=gbs, peak=peaks["hbm"], unit="GB/s", frac=gbs / peaks["hbm"],
                                   launches=rr["launches"], avg_launch_us=rr["ms"] * 1e3 / rr["launches"],
                                   algorithmic_bytes_per_launch=rr["bytes"] / rr["launches"],
                                   note="algorithmic read + write bytes over the launch time; part of the written lines is still "
                                        "dirty in L2 when the kernel ends and drains under the next kernel")
    if tall["ms"] > 0:
        tf = tall["flops"] / (tall["ms"] / 1e3) / 1e12
        roof["vision_gemm"] = dict(bound="tensor", kernel="gemm_tall_persistent_kernel (SigLIP + projector GEMMs)", achieved=tf,
                                   peak=peaks["tf_sustained"], unit="TFLOP/s", frac=tf / peaks["tf_sustained"],
                                   launches=tall["launches"], ms_per_frame=tall["ms"] / (n * ns))
    if va["ms"] > 0 and va["launches"]:
        tf = va["flops"] / (va["ms"] / 1e3) / 1e12
        roof["vision_attention"] = dict(bound="tensor", kernel="attention_persistent_kernel (SigLIP: 729 x 729 x 72, 16 heads x 32 frames per launch; attention_kernel for batches under two waves)",
                                        achieved=tf, peak=peaks["tf_sustained"], unit="TFLOP/s", frac=tf / peaks["tf_sustained"],
                                        launches=va["launches"], avg_launch_us=va["ms"] * 1e3 / va["launches"],
                                        ms_per_frame=va["ms"] / (n * ns))
    if ca["ms"] > 0 and ca["launches"]:
        tf = ca["flops"] / (ca["ms"] / 1e3) / 1e12
        roof["chunk_attention"] = dict(bound="tensor" if not single_stream else "latency",
                                       kernel=f"{'attention_persistent_kernel' if ns > 1 else 'attention_kernel'} (49-token frame chunks x 7 query heads per kv head of {ns} stream(s) "
                                              f"against {wl.window} cached keys, one launch per layer)",
                                       launches=ca["launches"], avg_launch_us=ca["ms"] * 1e3 / ca["launches"],
                                       achieved=tf, peak=peaks["tf_sustained"], unit="TFLOP/s", frac=tf / peaks["tf_sustained"],
                                       kv_gbs=ca["bytes"] / (ca["ms"] / 1e3) / 1e9, ms_per_step=ca["ms"] / n)
    return roof


def e2e_leg(model, cfg, args, ns, K, pool_host, start_ids, dev, dist, world):
    """Same metric through the public API with HOST frames: per stream a clip of pinned uint8 frames -> H2D ->
    visual_embed -> batched per-frame prefill (LiveInferGroup over one LiveInferForBenchmark per stream) -> scores read
    back to the host, all inside the timed region."""
    import torch
    import aha_b200 as A
    K = K - K % VISION_BATCH if K >= VISION_BATCH else K
    lives = [A.LiveInferForBenchmark(model, start_ids, alt_cache=args.cache, window_length=args.window,
                                     num_sink_tokens=args.sinks, frame_fps=1.0, max_frames=8) for _ in range(ns)]
    prime = (args.window // cfg.frame_num_tokens) + 4 if args.cache != "static" else 2
    prime += (-prime) % VISION_BATCH
    group = A.LiveInferGroup(lives, max_frames=max(K, prime) + 8)
    reps = (max(K, prime) + VISION_BATCH - 1) // VISION_BATCH
    clip = pool_host.repeat(reps, 1, 1, 1).pin_memory()      # handed over at once, as test/inference.py does with a video

    def run(nframes):
        group.reset_scores()
        for lv in lives:
            lv.debug_data_list, lv._times, lv._n_scored = [], [], 0
        group.input_video_streams([clip[:nframes]] * ns, batch_size=VISION_BATCH)      # H2D from pinned memory inside
        return group.inference()                                                      # ends with the D2H read

    run(prime)
    torch.cuda.synchronize(dev)
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    res = run(K)
    e1.record()
    torch.cuda.synchronize(dev)
    wall = time.perf_counter() - t0
    assert all(len(r) == K for r in res)
    ms = max(e0.elapsed_time(e1), wall * 1e3)
    if dist is not None:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    frame_bytes = 3 * cfg.vision.image_size * cfg.vision.image_size
    del group, lives
    return dict(value=world * ns * K / (ms / 1e3), unit="frames/s", h2d_bytes_per_step=ns * frame_bytes,
                d2h_bytes_per_step=ns * 12, api="LiveInferGroup(LiveInferForBenchmark x streams).input_video_streams + inference",
                frames_per_stream=K, ms_per_step=ms / K)


def latency_leg(model, cfg, args, pool_host, start_ids, dev):
    """Streaming path (test/live_infer_for_video.py:98-176): one stream, one frame at a time, vision batch 1, one host
    sync per frame; sink cache 2048 / 32 (BASELINE config 2)."""
    import aha_b200 as A
    live = A.LiveInferForBenchmark(model, start_ids, alt_cache="default_sink", window_length=2048, num_sink_tokens=32,
                                   frame_fps=1.0, max_frames=8192)
    prime = (2048 // cfg.frame_num_tokens) + 4
    lat = []
    for i in range(prime + 100):
        t0 = time.perf_counter()
        live.load_one_frame(pool_host[i % VISION_BATCH].to(dev, non_blocking=True))
        live.input_one_frame()
        live._flush_scores()
        if i >= prime:
            lat.append((time.perf_counter() - t0) * 1e3)
    lat.sort()
    return dict(p50_ms=lat[len(lat) // 2], p99_ms=lat[min(len(lat) - 1, int(len(lat) * 0.99))], frames=len(lat),
                mode="streaming, one stream (config 2: sink cache 2048/32): H2D frame -> vision (batch 1) -> prefill -> D2H "
                     "scores, synchronised every frame")


def c4_stress_leg(model, cfg, lib, peaks, dev):
    """BASELINE config 4, stress flow: TrulyStaticCache holding a 32 768-key prefix (filled through
    Cache.update(), the first update is the frozen prefix: test/static_cache.py:26-33), then per-frame q = 49 against all
    32 768 keys, non-causal: 644 GFLOP of attention per frame."""
    import ctypes
    import torch
    P, q = 32768, cfg.frame_num_tokens
    cache = model.new_cache("static", P, 0, max_chunk=256)
    g = torch.Generator(device=dev).manual_seed(5)
    for l in range(cfg.num_hidden_layers):
        k = torch.randn(1, cfg.num_key_value_heads, P, cfg.head_dim, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
        v = torch.randn(1, cfg.num_key_value_heads, P, cfg.head_dim, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
        cache.update(k, v, l)
        del k, v
    assert cache.get_seq_length() == P
    x = (torch.randn(1, q, cfg.hidden_size, generator=g, device=dev, dtype=torch.float32) * 0.05).to(torch.bfloat16)
    sc = torch.zeros(1, 3, dtype=torch.float32, device=dev)
    for _ in range(3):
        model(inputs_embeds=x, past_key_values=cache, scores_out=sc)
    n = 16
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    e0.record()
    for _ in range(n):
        model(inputs_embeds=x, past_key_values=cache, scores_out=sc)
    e1.record()
    torch.cuda.synchronize(dev)
    step_ms = e0.elapsed_time(e1) / n
    lib.aha_profile_gemm(1)
    for _ in range(4):
        model(inputs_embeds=x, past_key_values=cache, scores_out=sc)
    torch.cuda.synchronize(dev)
    ms, cnt, by, fl = ctypes.c_double(), ctypes.c_longlong(), ctypes.c_double(), ctypes.c_double()
    lib.aha_profile_gemm_read(4, ctypes.byref(ms), ctypes.byref(cnt), ctypes.byref(by), ctypes.byref(fl))
    lib.aha_profile_gemm(0)
    assert cache.get_seq_length() == P          # later chunks never extend the frozen prefix
    tf = fl.value / (ms.value / 1e3) / 1e12 if ms.value > 0 else 0.0
    del cache
    torch.cuda.empty_cache()
    return dict(workload="BASELINE config 4 stress flow: TrulyStaticCache with a 32 768-key prefix, q = 49 per frame, one stream",
                frames_per_s=1e3 / step_ms, ms_per_frame=step_ms,
                roofline=dict(bound="tensor", kernel="attention_kernel (343 query rows per kv head x 32 768 keys, key-split)",
                              achieved=tf, peak=peaks["tf_sustained"], unit="TFLOP/s", frac=tf / peaks["tf_sustained"],
                              launches=cnt.value, avg_launch_us=ms.value * 1e3 / max(1, cnt.value),
                              algorithmic_flops_per_frame=fl.value / 4, kv_gbs=by.value / (ms.value / 1e3) / 1e9 if ms.value else None,
                              attention_ms_per_frame=ms.value / 4))


def c5_scores(n_streams, T, seed=7):
    """0.5 + 0.25 sin(2 pi t / T_k) + 0.1 AR(1)(rho = 0.9) + 0.02 N(0, 1), clipped to [0, 1]."""
    import numpy as np
    from scipy.signal import lfilter
    rng = np.random.default_rng(seed)
    t = np.arange(T)[None, :]
    period = rng.uniform(200.0, 2000.0, size=(n_streams, 1))
    ar = lfilter([1.0], [1.0, -0.9], rng.standard_normal((n_streams, T)) * np.sqrt(1 - 0.81), axis=1)
    x = 0.5 + 0.25 * np.sin(2 * np.pi * t / period) + 0.1 * ar + 0.02 * rng.standard_normal((n_streams, T))
    return np.clip(x, 0.0, 1.0)


def _scipy_ticks(rows):
    from scipy.signal import find_peaks, savgol_filter
    n = 0
    for r in rows:
        sm = savgol_filter(r, 15, 3)
        p, _ = find_peaks(sm, height=sm.mean() + 0.5 * sm.std(), prominence=0.02, distance=10)
        n += len(p)
    return n


def c5_leg(peaks, dev):
    """BASELINE config 5: find_ticks over 4096 score streams of 10 000 frames (float64), one kernel launch; scipy on the
    host cores beside it (the reference's own find_ticks body, test/live_infer_for_video.py:209-221)."""
    import multiprocessing as mp
    import numpy as np
    import torch
    import aha_b200 as A
    N, T = 4096, 10000
    x = c5_scores(N, T)
    xd = torch.from_numpy(x).to(dev)
    for _ in range(2):
        peaks_d, n_d = A.find_ticks_batched(xd, fps=1.0)
    reps = 5
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    e0.record()
    for _ in range(reps):
        peaks_d, n_d = A.find_ticks_batched(xd, fps=1.0)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / reps
    t0 = time.perf_counter()
    host = xd.cpu()
    p2, n2 = A.find_ticks_batched(host.to(dev), fps=1.0)
    n_host = n2.cpu()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    total_peaks = int(n_d.sum().item())
    # scipy: single thread on a bounded sample, then a Pool over all cores on all streams
    sample = 256
    t0 = time.perf_counter()
    ref_peaks = _scipy_ticks(x[:sample])
    t1 = time.perf_counter() - t0
    assert ref_peaks == int(n_d[:sample].sum().item()), "peak counts differ from scipy on the sampled streams"
    cores = os.cpu_count() or 1
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(cores) as pool:
        chunks = np.array_split(x, cores * 4)
        tot = sum(pool.map(_scipy_ticks, chunks))
    tp = time.perf_counter() - t0
    assert tot == total_peaks, "peak counts differ from scipy over all streams"
    gbs = N * T * 8 / (ms / 1e3) / 1e9
    return dict(workload="BASELINE config 5: Savitzky-Golay(15,3) + find_peaks(height=mean+0.5 std, prominence 0.02, distance 10) "
                         "over 4096 streams x 10 000 float64 scores",
                value=N / (ms / 1e3), unit="streams/s", ms_per_launch=ms, peaks_found=total_peaks,
                roofline=dict(bound="hbm", kernel="find_ticks_kernel (one block per stream)", achieved=gbs, peak=peaks["hbm"],
                              unit="GB/s", frac=gbs / peaks["hbm"], algorithmic_bytes_per_launch=N * T * 8),
                e2e=dict(value=N / (e2e_ms / 1e3), unit="streams/s", h2d_bytes=N * T * 8, d2h_bytes=N * 4,
                         note="host scores -> device -> peak counts back on the host, wall clock"),
                cpu_baseline=dict(single_thread=dict(value=sample / t1, unit="streams/s", cores=1, kind="reference",
                                                     sample=f"{sample} streams through scipy savgol_filter + find_peaks"),
                                  pool=dict(value=N / tp, unit="streams/s", cores=cores, kind="reference",
                                            sample=f"all {N} streams, multiprocessing.Pool({cores})")))


def ingest_leg(peaks, dev):
    """N4 (frame ingest, upstream of the timed path): both resize flavours on 32 decoded 1920 x 1080 frames per launch."""
    import torch
    import aha_b200 as A
    n, h, w = 32, 1080, 1920
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
    peaks = measured_peaks()
    cfg = A.config_preset(args.preset)
    ns, q = args.streams_per_gpu, cfg.frame_num_tokens
    image = cfg.vision.image_size

    sd = A.random_init_state_dict(cfg, seed=0, device=dev)
    model = A.VideoHeadLiveLlavaQwenForCausalLM(cfg, sd, device=dev, max_tokens=max(256, ns * (N_SYS + q)),
                                                 max_frames=VISION_BATCH, logits_mode="none")
    del sd
    torch.cuda.empty_cache()
    start_ids = torch.randint(0, cfg.vocab_size, (1, N_SYS), device=dev)
    K, W, R = args.steps, max(args.warmup, 3), max(1, args.repeats)
    pool_host = A.synthetic_frames(VISION_BATCH, image=image, seed=100 + rank).pin_memory()
    pool_dev = pool_host.to(dev)

    wl = Workload(model, cfg, ns, args.cache, args.window, args.sinks, pool_dev, start_ids, total_frames=4096)
    wl.prime()                   # untimed: fills the caches so every timed frame runs the steady-state eviction path
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()          # before the warm-up: nvidia-smi's start-up perturbs the GPU, its steady sampling does not
    for _ in range(W):
        wl.step()
    if args.cache != "static":
        assert all(c.get_seq_length() == args.window for c in wl.caches)
    torch.cuda.synchronize(dev)
    sampler.mark()
    launches0 = _lib.launch_count()
    if os.environ.get("AHA_NCU"):
        # profiling aid (never a bench value): steady-state steps inside a cudaProfiler range for `ncu --profile-from-start off`
        wl.begin_region(None)
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        for _ in range(int(os.environ.get("AHA_NCU_STEPS", "8"))):
            wl.step()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        if rank == 0:
            sampler.stop()
        return
    times = timed_regions(wl, K, R, dev, dist)
    launches = (_lib.launch_count() - launches0) // R
    clocks = sampler.stop() if rank == 0 else None
    ms = sorted(times)[len(times) // 2]
    fps = world * ns * K / (ms / 1e3)

    roof = roofline_leg(wl, lib, K, peaks, single_stream=(ns == 1))
    del wl
    torch.cuda.empty_cache()
    e2e = e2e_leg(model, cfg, args, ns, max(K, VISION_BATCH), pool_host, start_ids, dev, dist, world)
    line = None
    if rank == 0:
        extras = {}
        lat = None
        if not args.no_extras:
            lat = latency_leg(model, cfg, args, pool_host, start_ids, dev)
            if ns != 1 and args.preset == "7b":
                # BASELINE config 2 (one stream, sink cache 2048 / 32) on the same model handle
                w2 = Workload(model, cfg, 1, "default_sink", 2048, 32, pool_dev, start_ids, total_frames=4096)
                w2.prime()
                for _ in range(W):
                    w2.step()
                t2 = timed_regions(w2, max(K, VISION_BATCH), 3, dev, None)
                m2 = sorted(t2)[1]
                r2 = roofline_leg(w2, lib, VISION_BATCH, peaks, single_stream=True)
                extras["single_stream"] = dict(workload="BASELINE config 2: one stream per GPU, SinkCache W=2048 S=32, steady state "
                                                        "(one eviction + re-rotation per frame)",
                                               value=max(K, VISION_BATCH) / (m2 / 1e3), unit="frames/s",
                                               ms_per_frame=m2 / max(K, VISION_BATCH), region_ms=t2, roofline=r2,
                                               latency=lat)
                if "kv_eviction" in r2:
                    roof["kv_eviction"] = r2["kv_eviction"]
                del w2
                torch.cuda.empty_cache()
            if args.preset == "7b":
                extras["c4_stress"] = c4_stress_leg(model, cfg, lib, peaks, dev)
            extras["c5_postproc"] = c5_leg(peaks, dev)
            roof["frame_ingest"] = ingest_leg(peaks, dev)
        del model
        torch.cuda.empty_cache()
        if not args.no_extras and args.preset == "7b":
            ge = {"config3" if ns != 1 else "headline": gpu_eager_sample(args.preset, args.cache, args.window, args.sinks, ns,
                                                                        VISION_BATCH if ns > 1 else 64, dev)}
            if ns != 1:
                ge["config2"] = gpu_eager_sample(args.preset, "default_sink", 2048, 32, 1, 64, dev)
            extras["gpu_eager_baseline"] = ge
        cpu = None if (args.no_cpu_baseline or world > 1) else cpu_baseline(args)
        line = dict(metric=METRIC, value=fps, unit="frames/s", n_gpus=world, steps=K, warmup=W, ms_per_step=ms / K,
                    higher_is_better=True, scaling="weak", vs_baseline=None, dtype="bf16", data="synthetic",
                    config=workload_config(args, world), clocks=clocks, e2e=e2e, gpu_launches=int(launches),
                    repeats=dict(n=R, region_ms=times, statistic="median of the R timed K-step regions (max over ranks each)"),
                    roofline=roof, cpu_baseline=cpu, latency=lat, **extras)
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: aha_b200 has no CPU path (use --impl reference for the CPU arm)")
    if args.impl == "reference-gpu":
        run_reference_gpu(args)
        return
    run_ours(args)


if __name__ == "__main__":
    main()
