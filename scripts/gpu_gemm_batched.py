"""Times the four decoder GEMMs of a batched step (T tokens) with weights rotating through copies larger than the L2:
    python scripts/gpu_gemm_batched.py [T] [bn ...]      bn: 416 = CTA-pair kernel, 208 = single-CTA persistent kernel"""
import sys
import torch
sys.path.insert(0, ".")
from aha_b200 import _lib

T = int(sys.argv[1]) if len(sys.argv) > 1 else 392
bns = [int(a) for a in sys.argv[2:]] or [416, 208]
dev = torch.device("cuda:0")
lib = _lib.load()
shapes = [("qkv", 4608, 3584, 0), ("o", 3584, 3584, 0), ("gate|up", 37888, 3584, 3), ("down", 3584, 18944, 0)]
for bn in bns:
    tot = 0.0
    for name, F_, K, act in shapes:
        copies = max(2, int(300e6 // (F_ * K * 2)) + 1)
        ws = [(torch.randn(F_, K, device=dev) / K ** 0.5).bfloat16() for _ in range(copies)]
        x = (torch.randn(T, K, device=dev) * 0.5).bfloat16()
        Fo = F_ // 2 if act == 3 else F_
        out = torch.empty(T, Fo, dtype=torch.bfloat16, device=dev)
        for splits in ([0] if name == "gate|up" else [0, 1, 2, 3, 4, 5, 6, 8]):
            def run(i):
                _lib.check(lib.aha_gemm_bf16(_lib.ptr(x), T, K, _lib.ptr(ws[i % copies]), F_, K, K, _lib.ptr(out), Fo, 0, 0, act,
                                             None, None, 0, 0, -1, bn, splits, _lib.stream_ptr()))
            for i in range(3):
                run(i)
            n = 20
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for i in range(n):
                run(i)
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / n * 1e3
            tf = 2.0 * T * F_ * K / us / 1e6
            print(f"bn={bn:3d} {name:8s} T={T} splits={splits}: {us:7.1f} us  {tf:7.1f} TFLOP/s", flush=True)
            if splits == 0:
                tot += us
        del ws
    print(f"bn={bn}: sum of the four (auto splits) = {tot:.1f} us per layer, {2.0 * T * 233.05e6 / tot / 1e6:.1f} TFLOP/s", flush=True)
