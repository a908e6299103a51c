"""Timing sweep of the skinny (weight-streaming) GEMM: python scripts/gpu_gemm_sweep.py  (B200 box)."""
import ctypes, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from aha_b200 import _lib
lib = _lib.load()
P = ctypes.c_void_p


def gemm(x, w, out, act=0, bias=None, residual=None, swap=-1, bn=0, splits=0):
    T, K = x.shape
    F = w.shape[0]
    _lib.check(lib.aha_gemm_bf16(P(x.data_ptr()), T, x.stride(0), P(w.data_ptr()), F, w.stride(0), K, P(out.data_ptr()),
                                 out.stride(0), 0, 0, act, P(bias.data_ptr()) if bias is not None else None,
                                 P(residual.data_ptr()) if residual is not None else None,
                                 residual.stride(0) if residual is not None else 0, 0, swap, bn, splits,
                                 P(torch.cuda.current_stream().cuda_stream)))


def time_case(T, F, K, act=0, res=False, swap=-1, bn=0, splits=0, n=32):
    dev = "cuda"
    x = (torch.randn(T, K, device=dev) * 0.5).bfloat16()
    ncopy = max(2, min(24, int(400e6 // (F * K * 2)) + 1))
    ws = [(torch.randn(F, K, device=dev) / K ** 0.5).bfloat16() for _ in range(ncopy)]
    Fo = F // 2 if act == 3 else F
    out = torch.zeros(T, Fo, device=dev, dtype=torch.bfloat16)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for i in range(ncopy):
            gemm(x, ws[i], out, act=act, residual=out if res else None, swap=swap, bn=bn, splits=splits)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for i in range(n):
                gemm(x, ws[i % ncopy], out, act=act, residual=out if res else None, swap=swap, bn=bn, splits=splits)
        g.replay()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); g.replay(); e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / n
    return dict(T=T, F=F, K=K, act=act, res=res, swap=swap, bn=bn, splits=splits, us=round(us, 2),
                w_gbs=round(F * K * 2 / us / 1e3, 1), tflops=round(2.0 * T * F * K / us / 1e6, 1))


cases = []
for sp in (1, 2, 4, 8):
    cases.append(dict(T=49, F=3584, K=3584, splits=sp))
cases += [dict(T=49, F=3584, K=3584, res=True), dict(T=49, F=128, K=64, splits=1), dict(T=49, F=3584, K=64, splits=1),
          dict(T=49, F=3584, K=448, splits=1), dict(T=49, F=4608, K=3584), dict(T=49, F=37888, K=3584, act=3),
          dict(T=49, F=3584, K=18944, res=True), dict(T=49, F=3584, K=18944, splits=1), dict(T=49, F=3584, K=18944, splits=4),
          dict(T=49, F=3584, K=3584, swap=0, splits=1), dict(T=49, F=3584, K=3584, swap=0, bn=32),
          dict(T=392, F=37888, K=3584, act=3), dict(T=392, F=3584, K=18944, res=True), dict(T=392, F=4608, K=3584),
          dict(T=23328, F=1152, K=1152), dict(T=23328, F=3840, K=1152), dict(T=23328, F=4304, K=1152, act=1),
          dict(T=23328, F=1152, K=4304, res=True), dict(T=23328, F=3584, K=3584), dict(T=729, F=1152, K=1152)]
if len(sys.argv) > 1 and sys.argv[1] == "splits":     # split-K factor of the three split decoder GEMMs (0 = the host's choice)
    cases = []
    for shape in (dict(T=49, F=4608, K=3584), dict(T=49, F=3584, K=3584, res=True), dict(T=49, F=3584, K=18944, res=True)):
        for sp in (0, 2, 3, 4, 5, 6, 8, 10):
            cases.append(dict(shape, splits=sp))
if len(sys.argv) > 1 and sys.argv[1] == "vision1":    # the four SigLIP GEMM shapes at ONE frame (729 tokens): tile width x split-K
    cases = []
    for shape in (dict(T=729, F=1152, K=1152, res=True), dict(T=729, F=1152, K=4304, res=True), dict(T=729, F=3840, K=1152),
                  dict(T=729, F=4304, K=1152, act=1)):
        for bn in (0, 64, 128, 192, 256):
            for sp in ((0,) if bn in (0, 192, 256) else (1, 2, 4)):
                cases.append(dict(shape, bn=bn, splits=sp, swap=0))
res = []
for c in cases:
    try:
        r = time_case(**c)
    except Exception as e:
        r = dict(c, error=str(e))
    print(json.dumps(r), flush=True)
    res.append(r)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "gemm_sweep.json"), "w"), indent=1)
