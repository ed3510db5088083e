"""Regenerates tests/golden/*.npz from the oracle (run from the repo root: python tests/golden/make_golden.py).
The reference holds no golden vectors for this path; these freeze OUR restatement."""
import math
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from graphslam_b200 import synth            # noqa: E402
from oracle import gs_oracle as O           # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

# compose_reference table (graph.hpp:30-43), computed by hand-checkable formula in plain python floats
rng = np.random.default_rng(2026)
a = rng.uniform(-4, 4, size=(64, 3))
b = rng.uniform(-4, 4, size=(64, 3))
a[:4] = [[0, 0, 0], [10, -4, 0], [0, 0, -3.0], [0, 0, 3.0]]
b[:4] = [[1, 2, 0.5], [1, 2, 0.5], [0, 0, -0.5], [0, 0, 0.5]]
out = np.zeros_like(a)
for k in range(a.shape[0]):
    c, s = math.cos(a[k, 2]), math.sin(a[k, 2])
    out[k, 0] = (c * b[k, 0]) + (-s * b[k, 1])
    out[k, 1] = (s * b[k, 0]) + (c * b[k, 1])
    out[k, 2] = math.fmod(a[k, 2] + b[k, 2] + math.pi, 2 * math.pi) - math.pi
np.savez(os.path.join(HERE, "compose_reference.npz"), a=a, b=b, out=out)

g = synth.config("C1")
prob = O.Problem.from_graph(g)
x, tr = O.lm_optimize(prob, g.init)
np.savez_compressed(os.path.join(HERE, "c1_lm.npz"), n=g.n, e=g.e, iterations=tr.iterations, tries=np.array(tr.tries),
                    errors=np.array(tr.errors), lambdas=np.array(tr.lambdas), poses=x.astype(np.float64))

# small graphs with full covariances / exact jacobians / duplicates for the GPU parity tests
for name, graph, mode in (("c3_small", synth.config("C3", n=2500, full_cov=True), 0),
                          ("c4_small_exact", synth.config("C4", n=1200), 1)):
    prob = O.Problem.from_graph(graph)
    x, tr = O.lm_optimize(prob, graph.init, O.LMParams(jacobian_mode=mode))
    np.savez_compressed(os.path.join(HERE, name + ".npz"), n=graph.n, e=graph.e, iterations=tr.iterations,
                        tries=np.array(tr.tries), errors=np.array(tr.errors), poses=x, mode=mode)
print("golden written")
