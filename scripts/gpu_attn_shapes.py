"""The two heavy attention shapes as bare aha_attention launches (for ncu / timing):  python scripts/gpu_attn_shapes.py [vision|chunk] [reps]
vision: 32 frames x 16 heads x 729 x 729, head_dim 80 (72 padded); chunk: 28 / 4 heads, 392 query tokens x 8192 keys."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import aha_b200 as A
from aha_b200 import _lib

dev = torch.device("cuda", 0)
lib = _lib.load()
which = sys.argv[1] if len(sys.argv) > 1 else "vision"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
heads, kvh, nq, klen, d, off = (512, 512, 729, 729, 80, 1 << 20) if which == "vision" else (28, 4, 392, 8192, 128, 8192)
G = heads // kvh
torch.manual_seed(0)
q = torch.randn(kvh, nq, G, d, device=dev).bfloat16()
k = torch.randn(kvh, klen, d, device=dev).bfloat16()
v = torch.randn(kvh, klen, d, device=dev).bfloat16()
out = torch.empty(nq, heads * d, dtype=torch.bfloat16, device=dev)


def go():
    _lib.check(lib.aha_attention(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(out), heads, kvh, nq, klen, d, off, d ** -0.5,
                                 _lib.stream_ptr()))


for _ in range(3):
    go()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    go()
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / reps
fl = 4.0 * heads * nq * klen * d
print(f"{which}: {us:.1f} us per launch, {fl / us / 1e6:.1f} TFLOP/s (algorithmic, padded d = {d})")
