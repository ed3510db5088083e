"""The C/OpenMP port (timed CPU baseline, block-Jacobi PCG like the CUDA path) against the exact-solve oracle."""
import numpy as np
import pytest

from graphslam_b200 import synth
from oracle import cpu_ref
from oracle import gs_oracle as O


@pytest.mark.parametrize("cid,n,kw,okw", [("C1", None, {}, {}), ("C3", 3000, dict(full_cov=True), {}),
                                          ("C4", 1000, {}, dict(jacobian_mode=1)), ("C2", 8000, {}, dict(diagonal_damping=1))])
def test_cpu_port_matches_oracle(cid, n, kw, okw):
    g = synth.config(cid, n=n, **kw)
    prob = O.Problem.from_graph(g)
    ox, otr = O.lm_optimize(prob, g.init, O.LMParams(jacobian_mode=okw.get("jacobian_mode", 0),
                                                      diagonal_damping=bool(okw.get("diagonal_damping", 0))))
    x, st, errs, tries = cpu_ref.optimize(g, **okw)
    assert st["status"] == 0 and st["iterations"] == otr.iterations and tries.tolist() == otr.tries
    rel = np.abs(errs - np.array(otr.errors)) / np.array(otr.errors)
    assert rel.max() <= 1e-6 and rel[-1] <= 1e-9, rel
    assert np.abs(x[:, :2] - ox[:, :2]).max() <= 1e-6 and np.abs(O.wrap_pi(x[:, 2] - ox[:, 2])).max() <= 1e-6


def test_cpu_port_bounded_sample():
    g = synth.config("C3", n=20000)
    x, st, errs, tries = cpu_ref.optimize(g, pcg_iter_cap_total=50, max_iterations=1)
    assert st["pcg_iterations_total"] == 50 and st["outer_loops"] == 1 and st["ms_solve"] > 0
