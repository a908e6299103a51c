import sys, os; sys.path.insert(0, os.getcwd()); __file__ = os.path.join(os.getcwd(), "scripts", "x.py")
exec(open('scripts/gpu_gemm_sweep.py').read().split("cases = []")[0])
for F, K, res in ((4608, 3584, False), (3584, 3584, True), (3584, 18944, True)):
    for sp in (2, 3, 4, 5, 6, 7, 8):
        r = time_case(T=49, F=F, K=K, res=res, splits=sp)
        print(F, K, "splits", sp, r["us"], r["w_gbs"], flush=True)
for bn in (64, 128):
    r = time_case(T=49, F=37888, K=3584, act=3, bn=bn)
    print("gateup bn", bn, r["us"], r["w_gbs"], flush=True)
