"""Small target for compute-sanitizer (memcheck / racecheck): every kernel family once, on graphs a sanitizer run
finishes in a minute — the persistent small-graph kernel (C1 cut to 1500 poses), the two-level path (C3 cut to 9000
poses: aggregation, coarse matrix, restriction, dense coarse product, prolongation), the plain three-kernel path,
marginals and closest_keyframe.   python tools/sanitizer_target.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graphslam_b200 import graph, synth  # noqa: E402

for name, g, flags, iters in (("persistent", synth.config("C1", n=1500), 0, 2),
                              ("two-level", synth.config("C3", n=9000), 0, 2),
                              ("three-kernel block-Jacobi", synth.config("C3", n=9000), 32 | 4, 1)):
    with graph.Graph(solver_flags=flags, max_iterations=iters, pcg_max_iters=20000) as G:
        G.add_graph(g)
        st = G.optimize()
        key, dist = G.closest_keyframe(int(g.keys[-1]), 10)
        cov = G.marginal_covariances(np.array([g.keys[3]], dtype=np.uint64)) if name != "two-level" else None
        print(name, "N", g.n, "LM iterations", st["iterations"], "PCG", st["pcg_iterations_total"], "chi2", st["chi2_final"],
              "closest", key, flush=True)
print("SANITIZER_TARGET_OK")
