"""Top stalled SASS instructions of one kernel in an .ncu-rep:  python scripts/ncu_top_stalls.py rep regex [launch-skip] [n]"""
import csv, subprocess, sys, io
rep, rx = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
n = int(sys.argv[4]) if len(sys.argv) > 4 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{rx}", "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
si, src, ie = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = []
for k, r in enumerate(rows[2:]):
    try:
        data.append((int(r[si]), k, r))
    except (ValueError, IndexError):
        pass
tot = sum(s for s, _, _ in data)
print(rows[0][1][:100], "| samples", tot, "| SASS instructions", len(data))
agg = {}
for s, _, r in data:
    for i in stall:
        if r[i] not in ("", "0"):
            agg[hdr[i][6:]] = agg.get(hdr[i][6:], 0) + int(r[i])
print("stall totals:", sorted(agg.items(), key=lambda kv: -kv[1])[:8])
for s, k, r in sorted(data, key=lambda x: -x[0])[:n]:
    st = {hdr[i][6:]: int(r[i]) for i in stall if r[i] not in ("", "0")}
    print(f"{s:6d} {100 * s / max(tot, 1):5.1f}% #{k:5d} exec={r[ie]:>8} {r[src][:64]:64s} {sorted(st.items(), key=lambda kv: -kv[1])[:3]}")
