import sys, os; sys.path.insert(0, os.getcwd()); __file__ = os.path.join(os.getcwd(), "scripts", "x.py")
exec(open('scripts/gpu_gemm_sweep.py').read().split("cases = []")[0])
import os
for c in [dict(T=49,F=128,K=64,splits=1), dict(T=49,F=3584,K=3584,splits=1), dict(T=49,F=3584,K=3584,splits=4), dict(T=128,F=128,K=64,swap=0,splits=1)]:
    print(os.environ.get("AHA_GEMM_DBG"), time_case(**c), flush=True)
