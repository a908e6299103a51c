"""Phase stamps (AHA_TS=2) of CTA 0 of the persistent attention kernel on the two heavy shapes.  Bring-up aid."""
import ctypes, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["AHA_TS"] = "2"
import aha_b200 as A
from aha_b200 import _lib

dev = torch.device("cuda", 0)
lib = _lib.load()
for which, (heads, kvh, nq, klen, d, off) in (("vision", (512, 512, 729, 729, 80, 1 << 20)), ("chunk", (28, 4, 392, 8192, 128, 8192))):
    G = heads // kvh
    q = torch.randn(kvh, nq, G, d, device=dev).bfloat16()
    k = torch.randn(kvh, klen, d, device=dev).bfloat16()
    v = torch.randn(kvh, klen, d, device=dev).bfloat16()
    out = torch.empty(nq, heads * d, dtype=torch.bfloat16, device=dev)
    for _ in range(3):
        _lib.check(lib.aha_attention(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(out), heads, kvh, nq, klen, d, off, d ** -0.5,
                                     _lib.stream_ptr()))
    torch.cuda.synchronize()
    a = (ctypes.c_ulonglong * 32)()
    _lib.check(lib.aha_debug_attention_persistent_timeline(a))
    t = list(a)
    names = ["ready", "pdl"] + [f"S0 of item {i}" for i in range(6)] + [f"epilogue {i} done" for i in range(6)] + ["o_full(2) seen", "stats(2) seen"]
    print(which, {n: x - t[0] for n, x in zip(names, t[:16]) if x})
    n = max(1, t[18])
    print(f"   per tile ({n} tiles): softmax thread 128 waits {t[16] / n:.0f} cycles for scores, works {t[17] / n:.0f}; MMA thread: early waits {t[19] / n:.0f}, "
          f"waits for P {t[20] / n:.0f}, issues P.V + Q.K^T {t[21] / n:.0f}")
