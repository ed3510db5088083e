"""How far apart are the converged poses of the GPU LM (PCG to a relative residual) and the CPU LM with the exact
Cholesky solve, as a function of pcg_rel_tol?  python tools/tol_probe.py [C3] [n]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from graphslam_b200 import graph, synth  # noqa: E402
from oracle import cpu_ref  # noqa: E402

cid = sys.argv[1] if len(sys.argv) > 1 else "C3"
n = int(sys.argv[2]) if len(sys.argv) > 2 and int(sys.argv[2]) > 0 else None
g = synth.config(cid, n=n)
cpu_ref.set_threads()
cx, cst, cerrs, ctries = cpu_ref.optimize(g, solver=1)
print(json.dumps({"cpu_cholesky": {"iterations": cst["iterations"], "chi2": cst["chi2_final"], "tries": ctries.tolist()}}), flush=True)


def wrap(t):
    return np.arctan2(np.sin(t), np.cos(t))


for flags, tol in ((0, 1e-12), (0, 1e-13), (0, 1e-14), (0, 1e-15), (32, 1e-12)):
    with graph.Graph(solver_flags=flags, pcg_rel_tol=tol) as G:
        G.add_graph(g)
        st = G.optimize()
        _, x = G.poses()
        errors, lambdas, tries = G.trace()
    dxy = np.abs(x[:, :2] - cx[:, :2]).max()
    dth = np.abs(wrap(x[:, 2] - cx[:, 2])).max()
    # the same after removing the rigid motion about the prior pose that best explains the difference in theta
    rot = np.median(wrap(x[:, 2] - cx[:, 2]))
    c, s = np.cos(-rot), np.sin(-rot)
    xr = np.c_[c * x[:, 0] - s * x[:, 1], s * x[:, 0] + c * x[:, 1]]
    sh = np.median(xr - cx[:, :2], axis=0)
    print(json.dumps({"flags": flags, "tol": tol, "iterations": st["iterations"], "tries": tries.tolist(), "pcg": st["pcg_iterations_total"],
                      "status": st["status"], "chi2_rel": abs(st["chi2_final"] - cst["chi2_final"]) / cst["chi2_final"],
                      "max_dxy": dxy, "max_dtheta": dth, "global_rotation": rot,
                      "max_dxy_after_rigid_alignment": float(np.abs(xr - sh - cx[:, :2]).max()),
                      "max_dtheta_after": float(np.abs(wrap(x[:, 2] - cx[:, 2] - rot)).max()),
                      "seconds": st["ms_total"] * 1e-3}), flush=True)
