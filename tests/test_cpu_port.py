"""The C/OpenMP port (timed CPU baseline) against the exact-solve numpy oracle, with both of its linear solvers:
block-Jacobi PCG (the CUDA path's algorithm) and the sparse multifrontal Cholesky of oracle/cpu_chol.c — an exact
solve written independently of gs_oracle.py's SuperLU call, i.e. a second witness for the LM restatement."""
import numpy as np
import pytest

from graphslam_b200 import synth
from oracle import cpu_ref
from oracle import gs_oracle as O


@pytest.mark.parametrize("solver", [cpu_ref.SOLVER_PCG, cpu_ref.SOLVER_CHOLESKY], ids=["pcg", "cholesky"])
@pytest.mark.parametrize("cid,n,kw,okw", [("C1", None, {}, {}), ("C3", 3000, dict(full_cov=True), {}),
                                          ("C4", 1000, {}, dict(jacobian_mode=1)), ("C2", 8000, {}, dict(diagonal_damping=1))])
def test_cpu_port_matches_oracle(cid, n, kw, okw, solver):
    """Same iterations, same accept/reject sequence, final chi2 within 1e-9 and poses within 1e-6 of the oracle."""
    g = synth.config(cid, n=n, **kw)
    prob = O.Problem.from_graph(g)
    ox, otr = O.lm_optimize(prob, g.init, O.LMParams(jacobian_mode=okw.get("jacobian_mode", 0),
                                                      diagonal_damping=bool(okw.get("diagonal_damping", 0))))
    x, st, errs, tries = cpu_ref.optimize(g, solver=solver, **okw)
    assert st["status"] == 0 and st["iterations"] == otr.iterations and tries.tolist() == otr.tries
    rel = np.abs(errs - np.array(otr.errors)) / np.array(otr.errors)
    assert rel.max() <= 1e-6 and rel[-1] <= 1e-9, rel
    assert np.abs(x[:, :2] - ox[:, :2]).max() <= 1e-6 and np.abs(O.wrap_pi(x[:, 2] - ox[:, 2])).max() <= 1e-6


def test_cpu_port_bounded_sample():
    g = synth.config("C3", n=20000)
    x, st, errs, tries = cpu_ref.optimize(g, pcg_iter_cap_total=50, max_iterations=1)
    assert st["pcg_iterations_total"] == 50 and st["outer_loops"] == 1 and st["ms_solve"] > 0


def _graph_from_edges(n, ei, ej, seed):
    """A consistent-ish pose graph over arbitrary edges: random ground truth, measurements = gt between + noise."""
    rng = np.random.default_rng(seed)
    gt = np.stack([rng.uniform(-20, 20, n), rng.uniform(-20, 20, n), rng.uniform(-3, 3, n)], axis=1)
    gt[0] = 0.0
    ei, ej = np.asarray(ei, np.int64), np.asarray(ej, np.int64)
    z = synth.compose(synth.between(gt[ei], gt[ej]), rng.normal(0, [0.05, 0.05, 0.02], (len(ei), 3)))
    cov = np.tile(np.diag([0.05 ** 2, 0.05 ** 2, 0.02 ** 2]).reshape(9), (len(ei), 1))
    init = gt + rng.normal(0, [0.3, 0.3, 0.05], (n, 3))
    init[0] = 0.0
    return synth.PoseGraph(f"edges{seed}", np.arange(1, n + 1, dtype=np.uint64), init, gt, (ei + 1).astype(np.uint64),
                           (ej + 1).astype(np.uint64), z, cov)


def test_cholesky_solve_is_exact_on_irregular_graphs():
    """One damped LM step with the multifrontal Cholesky against the oracle's SuperLU step, on graphs whose elimination
    trees differ in kind: a hub (one vertex joined to 600 others), several connected components hanging off a short
    spine, duplicate factors, a clique too dense to dissect, and the dense C4 pattern."""
    rng = np.random.default_rng(7)
    n = 900
    chain_i, chain_j = np.arange(0, n - 1), np.arange(1, n)
    hub = _graph_from_edges(n, np.r_[chain_i, np.zeros(600, int), [5, 5, 5]],
                            np.r_[chain_j, rng.choice(np.arange(2, n), 600, replace=False), [6, 6, 6]], 11)
    # a 60-clique (no level structure deeper than two) tied to a chain
    ci, cj = np.triu_indices(60, 1)
    clique = _graph_from_edges(400, np.r_[ci, np.arange(59, 399)], np.r_[cj, np.arange(60, 400)], 12)
    # a spine of 20 vertices with 20 separate 40-vertex chains hanging off it: removing a spine vertex leaves components
    si, sj = [], []
    for k in range(19):
        si.append(k); sj.append(k + 1)
    for k in range(20):
        base = 20 + 40 * k
        si.append(k); sj.append(base)
        for q in range(39):
            si.append(base + q); sj.append(base + q + 1)
    comb = _graph_from_edges(20 + 40 * 20, si, sj, 13)
    p = O.LMParams(max_iterations=1, lambda_initial=1e-3, rel_error_tol=0.0, abs_error_tol=0.0)
    for g in (hub, clique, comb, synth.config("C4", n=1500)):
        prob = O.Problem.from_graph(g)
        ox, otr = O.lm_optimize(prob, g.init, p)
        x, st, errs, tries = cpu_ref.optimize(g, solver=cpu_ref.SOLVER_CHOLESKY, max_iterations=1, lambda_initial=1e-3,
                                              rel_error_tol=0.0, abs_error_tol=0.0)
        assert tries.tolist() == otr.tries, g.name
        # two exact solvers agree to rounding times the conditioning of the damped system
        assert np.abs(x[:, :2] - ox[:, :2]).max() <= 1e-7 and np.abs(O.wrap_pi(x[:, 2] - ox[:, 2])).max() <= 1e-7, g.name
        assert errs[-1] == pytest.approx(otr.errors[-1], rel=1e-9)
