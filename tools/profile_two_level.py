"""Launch-bounded target for ncu: the two-level PCG iteration on one config with the coarse matrix set up beforehand
(its cuSOLVER launches stay out of the capture window when -s skips them).
  python tools/profile_two_level.py [C3] [iters]"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graphslam_b200 import _lib, graph, synth  # noqa: E402

cid = sys.argv[1] if len(sys.argv) > 1 else "C3"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 6
g = synth.config(cid)
with graph.Graph(pcg_max_iters=iters, pcg_rel_tol=1e-300, solver_flags=1) as G:   # no CUDA graph: every kernel is a launch
    G.add_graph(g)
    G.prepare()
    print("outer", G.profile_outer(2))
    d = np.zeros((g.n, 3))
    it, rr = C.c_int(), C.c_double()
    for _ in range(2):   # the second solve reuses the coarse matrix: only PCG kernels
        G._lib.gs_solve_linear(G._h, 1e-5, d.ctypes.data_as(_lib.c_double_p), C.byref(it), C.byref(rr))
    print("iterations", it.value)
