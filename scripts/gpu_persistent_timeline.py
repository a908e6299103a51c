"""Epilogue-warp phase stamps of the persistent tall GEMM (AHA_TS=1): per tile 'accumulator full' and 'epilogue done'."""
import ctypes, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["AHA_TS"] = "1"
from aha_b200 import _lib
lib = _lib.load()
dev = torch.device("cuda", 0)
T = 23328
for (F_, K, act, res) in [(1152, 1152, 0, True), (1152, 4304, 0, True), (3840, 1152, 0, False), (4304, 1152, 1, False)]:
    x = (torch.randn(T, K, device=dev) * 0.5).bfloat16()
    w = (torch.randn(F_, K, device=dev) / K ** 0.5).bfloat16()
    bias = torch.zeros(F_, device=dev).bfloat16()
    out = torch.zeros(T, F_, device=dev).bfloat16()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    def run():
        _lib.check(lib.aha_gemm_bf16(_lib.ptr(x), T, K, _lib.ptr(w), F_, K, K, _lib.ptr(out), F_, 0, 0, act, _lib.ptr(bias),
                                     _lib.ptr(out) if res else None, F_, 0, 0, 0, 0, st))
    for _ in range(3): run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(); e1.record(); torch.cuda.synchronize()
    n, gx, gz = ctypes.c_int(), (ctypes.c_int * 64)(), (ctypes.c_int * 64)()
    buf = (ctypes.c_ulonglong * 1024)()
    _lib.check(lib.aha_debug_gemm_timeline(ctypes.byref(n), gx, gz, buf))
    i = [k for k in range(n.value) if gx[k] == 148][0]
    t = list(buf[i * 16:(i + 1) * 16])
    us = [round((v - t[0]) / 1.965e3, 1) if v else None for v in t]
    print(f"F={F_} K={K} act={act} res={res}: {e0.elapsed_time(e1)*1e3:.0f} us, {2*T*F_*K/e0.elapsed_time(e1)/1e9:.0f} TFLOP/s; "
          f"(full, done) per tile us: {[(us[1+2*k], us[2+2*k]) for k in range(7)]}")
