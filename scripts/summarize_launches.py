"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel (and grid for the GEMM):
python scripts/summarize_launches.py gpurun_out/launches.csv [frames]"""
import collections
import csv
import re
import sys

path = sys.argv[1]
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 32
lines = [l for l in open(path) if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
seq = []
for row in csv.DictReader(lines):
    try:
        v = float(row["Metric Value"].replace(",", ""))
    except ValueError:
        continue
    v = v / 1000 if row["Metric Unit"] == "ns" else v * 1000 if row["Metric Unit"] == "ms" else v
    name = row["Kernel Name"]
    m = re.search(r"::(\w+)(<[^>]*>)?\(", name) or re.search(r"(\w+)(<[^>]*>)?\(", name)
    key = (m.group(1) + (m.group(2) or "")) if m else name[:60]
    if "gemm" in key or "attention" in key:
        key += " grid=" + row["Grid Size"].replace(" ", "")
    agg[key][0] += 1
    agg[key][1] += v
    seq.append((key, v))
tot = sum(v[1] for v in agg.values())
print(f"launches {sum(v[0] for v in agg.values())}  total {tot / 1000:.2f} ms  = {tot / 1000 / frames:.3f} ms/frame over {frames} frames "
      "(ncu times are cold-cache and serialised: compare shares)")
print(f"{'us total':>10} {'share':>6} {'n':>6} {'avg us':>8}  kernel")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v[1]:10.0f} {100 * v[1] / tot:5.1f}% {v[0]:6d} {v[1] / v[0]:8.1f}  {k}")

if len(sys.argv) > 4:      # launches [a, b) in stream order (one decoder layer is about seven launches)
    a, b = int(sys.argv[3]), int(sys.argv[4])
    print(f"launches {a}..{b - 1} in order:")
    for i, (k, v) in enumerate(seq[a:b]):
        print(f"  {a + i:4d} {v:8.1f} us  {k}")
