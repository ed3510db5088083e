"""GPU probe of the two-level preconditioner against block-Jacobi alone: PCG iterations and wall time of one damped
solve, and a full LM run, per config.   python tools/two_level_probe.py [C3 C4 C2 C5]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from graphslam_b200 import graph, synth  # noqa: E402

SIZES = [int(a[2:]) for a in sys.argv[1:] if a.startswith("T=")] or [0]
for cid in ([a for a in sys.argv[1:] if not a.startswith("T=")] or ["C3", "C4", "C2", "C5"]):
    g = synth.config(cid)
    rec = {"config": g.name, "n": g.n, "e": g.e}
    for name, flags, size in [(f"two_level_T{t}", 0, t) for t in SIZES] + [("block_jacobi", 32, 0)]:
        with graph.Graph(solver_flags=flags, coarse_aggregate_size=size) as G:
            G.add_graph(g)
            G.prepare()
            rec[name + "_prepare_ms"] = G.structure_timing()
            G.solve_linear(1e-5)
            t0 = time.perf_counter()
            d, it, rr = G.solve_linear(1e-5)
            dt = time.perf_counter() - t0
            rec[name] = {"pcg_iterations": it, "solve_ms": round(dt * 1e3, 2), "us_per_iteration": round(dt / it * 1e6, 2), "relres": rr}
            if name.startswith("two_level"):
                rec[name]["kernels_us"] = {k: round(v[0] * 1e3, 2) for k, v in G.profile_pcg(1e-5, 40).items()}
            t0 = time.perf_counter()
            st = G.optimize()
            rec[name]["lm"] = {"seconds": round(time.perf_counter() - t0, 3), "iterations": st["iterations"],
                               "pcg_iterations_total": st["pcg_iterations_total"], "chi2_final": st["chi2_final"],
                               "ms_solve": round(st["ms_solve"], 1), "ms_linearize": round(st["ms_linearize"], 1)}
    print(json.dumps(rec), flush=True)
