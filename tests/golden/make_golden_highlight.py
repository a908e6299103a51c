"""Generates tests/golden/highlight.npz by EXECUTING the reference's own knapsack_selection (its source is cut out of
/root/reference/test/highlight_generator.py, which cannot be imported as a module here: it needs cv2 and test.inference)
and the per-frame fusion loop of /root/reference/test/grid_search.py:58-65 restated verbatim.  Run in the build container:
    python tests/golden/make_golden_highlight.py"""
import ast
import os

import numpy as np

REF = "/root/reference/test/highlight_generator.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "highlight.npz")


def reference_knapsack():
    src = open(REF).read()
    tree = ast.parse(src)
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "knapsack_selection")
    ns = {}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), REF, "exec"), ns)
    return ns["knapsack_selection"]


def streams(seed, n, T, quantise):
    g = np.random.default_rng(seed)
    inf = g.random((n, T)).astype(np.float32).astype(np.float64)     # float32 scores read back as Python floats
    rel = g.random((n, T)).astype(np.float32).astype(np.float64)
    unc = (g.random((n, T)) * 0.2).astype(np.float32).astype(np.float64)
    if quantise:                                                     # many exact ties: exercises the backtracking rule
        inf, rel, unc = np.round(inf * 4) / 4, np.round(rel * 4) / 4, np.round(unc * 20) / 20
    return inf, rel, unc


def main():
    knap = reference_knapsack()
    out = {}
    cases = [("a", 1, 6, 60, False, 20, 0.0, -1.0, -5.0), ("b", 2, 5, 120, False, 20, 1.0, 0.65, -0.3),
             ("ties", 3, 6, 90, True, 7, 1.0, 1.0, 1.0), ("neg", 4, 3, 40, False, 31, -1.0, 0.5, -5.0),
             ("zero_budget", 5, 2, 30, False, 0, 1.0, 1.0, 0.0)]
    for name, seed, n, T, q, budget, alpha, beta, eps in cases:
        inf, rel, unc = streams(seed, n, T, q)
        sel = np.zeros((n, T), np.uint8)
        fused = np.zeros((n, T))
        thr = 0.04
        for s in range(n):
            frames = [dict(idx=i, informative_score=float(inf[s, i]), relevance_score=float(rel[s, i]),
                           uncertainty_score=float(unc[s, i])) for i in range(T)]
            for i in knap(frames, budget, "relevance_score", alpha, beta, eps):
                sel[s, i] = 1
            for i, e in enumerate(frames):        # grid_search.py:58-65
                cur = alpha * e["informative_score"] + beta * e["relevance_score"]
                if e["uncertainty_score"] >= thr:
                    diff = e["uncertainty_score"] - thr
                    penalty = diff * eps
                    cur -= penalty
                fused[s, i] = cur
        out.update({f"{name}_inf": inf, f"{name}_rel": rel, f"{name}_unc": unc, f"{name}_sel": sel, f"{name}_fused": fused,
                    f"{name}_par": np.array([budget, alpha, beta, eps, thr])})
    np.savez_compressed(OUT, names=np.array([c[0] for c in cases]), **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
