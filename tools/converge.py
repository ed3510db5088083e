"""LM to convergence on the BASELINE.json configs (C1 .. C5 at full size; --n cuts them), one JSON line per
config: outer iterations, PCG iterations, device time, chi2 before / after, per-kernel event times.  GTSAM's default
stopping rule (relative / absolute error decrease 1e-5), FP64, one B200, from the dead-reckoned start — the number a
user of solve() sees; optionally the same solve on the CPU port with its multifrontal Cholesky (oracle/cpu_ref.c +
cpu_chol.c, all host threads) for the ratio.

    python tools/converge.py [C1 C2 C4 ...] [--cpu] [--n POSES] > profiles/rN_converge.jsonl
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from graphslam_b200 import graph, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("configs", nargs="*", default=["C1", "C2", "C3", "C4", "C5"])
    ap.add_argument("--n", type=int, default=None)
    ap.add_argument("--cpu", action="store_true", help="also run the CPU port (sparse Cholesky LM) to convergence")
    args = ap.parse_args()
    for cid in args.configs:
        g = synth.config(cid, n=args.n)
        with graph.Graph() as G:
            G.add_graph(g)
            G.prepare()
            t0 = time.perf_counter()
            st = G.optimize()
            wall = time.perf_counter() - t0
            prof = G.profile_pcg(1e-5, 20)
            outer = G.profile_outer(3)
        rec = {"config": g.name, "n_poses": int(g.n), "n_factors": int(g.e), "lm_iterations": int(st["iterations"]),
               "linear_solves": int(st["inner_iterations"]), "pcg_iterations": int(st["pcg_iterations_total"]),
               "converged": bool(st["converged"]), "chi2_initial": st["chi2_initial"], "chi2_final": st["chi2_final"],
               "ms_device": round(st["ms_device"], 3), "ms_wall": round(wall * 1e3, 3),
               "gn_iters_per_sec": st["iterations"] / max(st["ms_device"] * 1e-3, 1e-12),
               "us_per_pcg_iteration": round(st["ms_solve"] * 1e3 / max(st["pcg_iterations_total"], 1), 2),
               "gpu_launches": int(st["gpu_launches"]),
               "kernel_us": {k: round(v[0] * 1e3, 2) for k, v in {**prof, **outer}.items()},
               "kernel_frac_of_hbm_peak": {k: round(v[1] / (v[0] * 1e-3) / 1e9 / 6557.8, 3) for k, v in {**prof, **outer}.items()},
               "ms_structure": round(st["ms_structure"], 1)}
        if args.cpu:
            from oracle import cpu_ref
            cpu_ref.set_threads()
            t0 = time.perf_counter()
            _, cst, _, _ = cpu_ref.optimize(g, solver=cpu_ref.SOLVER_CHOLESKY)
            rec["cpu_port"] = {"threads": int(cst["threads"]), "solver": "multifrontal Cholesky", "lm_iterations": int(cst["iterations"]),
                               "ms_total": round(cst["ms_total"], 1), "ms_ordering": round(cst["ms_analyze"], 1),
                               "ms_structure": round(cst["ms_structure"], 1),
                               "chi2_final": cst["chi2_final"], "ms_wall": round((time.perf_counter() - t0) * 1e3, 1)}
        print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
