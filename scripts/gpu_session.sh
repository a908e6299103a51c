#!/bin/bash
# One gpurun call: the GPU parity suite, the bench line, and an ncu launch list of steady-state config-3 steps.
# usage: gpurun --timeout 1500 -- 'bash scripts/gpu_session.sh <tag> [pytest-args]'
TAG=${1:-s}; shift
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $OUT/${TAG}_gpu.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q --maxfail 8 --timeout 600 "$@" > $OUT/${TAG}_pytest.log 2>&1
echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
tail -5 $OUT/${TAG}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench rc=$?"
tail -c 600 $OUT/${TAG}_bench.err
python - <<PY
import json
try:
    d = json.loads(open("$OUT/${TAG}_bench.json").read().strip().splitlines()[-1])
    r = d["roofline"]
    print("value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "ms/step", round(d["ms_per_step"], 3), "regions", [round(x, 2) for x in d["repeats"]["region_ms"]])
    print("main roofline", r.get("kernel", "")[:40], round(r["achieved"], 1), r["unit"], round(r["frac"], 3), "avg us", r.get("avg_launch_us"))
    for k in ("decoder_step", "chunk_attention", "vision_gemm", "vision_attention", "kv_eviction"):
        if k in r:
            print(k, {a: (round(b, 3) if isinstance(b, float) else b) for a, b in r[k].items() if a in ("achieved", "frac", "avg_launch_us", "ms_per_step", "ms_per_frame", "unit")})
    for k in ("single_stream", "c4_stress", "c5_postproc", "gpu_eager_baseline", "latency", "cpu_baseline"):
        if k in d and d[k]:
            print(k, json.dumps(d[k])[:700])
except Exception as e:
    print("no bench line:", e)
PY
