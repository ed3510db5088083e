"""CPU ORACLE for the pose-graph optimisation path (numpy + scipy, FP64).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package; the product (libgraphslam_cuda + graphslam_b200) never does and has no CPU fallback.

PARITY UNPINNED.  The reference (`/root/reference/src/graph/src/graph.cpp:115-132`, `solve()`) does
its arithmetic inside GTSAM, an external dependency that is neither vendored nor version-pinned
(`src/graph/CMakeLists.txt:9,18`; absent from `package.xml:8-12`) and is not installable here
(no network; no Eigen/Boost).  The reference ships no tests, fixtures or golden vectors
(SURVEY.md §4, §8(c)).  This file therefore RESTATES the published GTSAM 4.0.x algorithm that the
reference's call sites select, and is pinned only by our own known-answer tests (tests/golden/):

  call site (reference)                          restated here
  graph.cpp:57   PriorFactor<Pose2>             prior_residual()      e = -Local(x^-1 o prior), H = I
  graph.cpp:87-92,106-111 BetweenFactor<Pose2>  between_residual()    e = Local(z^-1 o (xi^-1 o xj)), H1, H2 = I
  graph.cpp:45,83,103 noiseModel::Gaussian::Covariance(Q)   information()   Omega = Q^-1 (SPD check)
  graph.cpp:119  LevenbergMarquardtOptimizer(graph, initial).optimize()  lm_optimize()  default params
  graph.cpp:122-125 poses_opti.at<Pose2>(id).x()/y()/theta()           theta reported in (-pi, pi]
  graph.hpp:30-43 compose                        compose_reference()   (both quirks kept)
  graph.hpp:45-58 covariance_to_eigen            row-major 9 -> 3x3

The damped normal equations are solved EXACTLY (sparse LU, scipy SuperLU), as GTSAM's multifrontal
Cholesky does; the CUDA path solves the same system by block-Jacobi PCG to 1e-12.

GTSAM semantics restated (Pose2 without SLOW_BUT_CORRECT_EXPMAP / _BETWEENFACTOR):
  Local(p) = (x, y, theta);  p.retract(v) = p o Pose2(v)  (body-frame update)
  between(p1,p2) = p1^-1 o p2 with H1 = [[-c,-s,dt1],[s,-c,dt2],[0,0,-1]], H2 = I
  LM: lambdaInitial 1e-5, lambdaFactor 10, lambdaUpperBound 1e5, lambdaLowerBound 0,
      minModelFidelity 1e-3, diagonalDamping false, useFixedLambdaFactor true,
      maxIterations 100, relativeErrorTol 1e-5, absoluteErrorTol 1e-5, errorTol 0.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

JAC_GTSAM_DEFAULT = 0
JAC_EXACT = 1


@dataclass
class LMParams:
    """GTSAM LevenbergMarquardtParams defaults (what graph.cpp:119 gets)."""
    max_iterations: int = 100
    rel_error_tol: float = 1e-5
    abs_error_tol: float = 1e-5
    error_tol: float = 0.0
    lambda_initial: float = 1e-5
    lambda_factor: float = 10.0
    lambda_upper: float = 1e5
    lambda_lower: float = 0.0
    min_model_fidelity: float = 1e-3
    diagonal_damping: bool = False
    jacobian_mode: int = JAC_GTSAM_DEFAULT


@dataclass
class LMTrace:
    errors: list = field(default_factory=list)        # graph.error (= chi2/2) after each outer iteration; [0] = initial
    lambdas: list = field(default_factory=list)       # lambda used by the accepted try of each outer iteration
    tries: list = field(default_factory=list)         # number of tryLambda calls per outer iteration
    iterations: int = 0
    lambda_final: float = 0.0
    inner_iterations: int = 0

    @property
    def chi2(self):
        return [2.0 * e for e in self.errors]


# --------------------------------------------------------------------------- SE(2) helpers

def wrap_pi(t):
    """Rot2::theta(): atan2(s, c) in (-pi, pi]."""
    return np.arctan2(np.sin(t), np.cos(t))


def compose_reference(a, b):
    """Reference `compose` (graph.hpp:30-43) bit-for-bit: input_1's translation is NOT added
    (graph.hpp:37-38) and theta is wrapped with a one-sided fmod (graph.hpp:40)."""
    cos_th = math.cos(a[2])
    sin_th = math.sin(a[2])
    dx, dy, dth = b[0], b[1], b[2]
    x = (cos_th * dx) + (-sin_th * dy)
    y = (sin_th * dx) + (cos_th * dy)
    th = a[2] + dth
    th = math.fmod(th + math.pi, 2 * math.pi) - math.pi
    return np.array([x, y, th])


def compose(a, b):
    """Correct SE(2) a (+) b (Pose2::compose), theta wrapped by atan2."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    c, s = np.cos(a[..., 2]), np.sin(a[..., 2])
    return np.stack([a[..., 0] + c * b[..., 0] - s * b[..., 1],
                     a[..., 1] + s * b[..., 0] + c * b[..., 1],
                     wrap_pi(a[..., 2] + b[..., 2])], axis=-1)


def covariance_to_matrix(cov9):
    """graph.hpp:45-58: float64[9] row-major -> 3x3."""
    return np.asarray(cov9, dtype=np.float64).reshape(-1, 3, 3)


def information(cov9):
    """noiseModel::Gaussian::Covariance(Q): Omega = Q^-1; raises (GTSAM throws) if Q is not SPD."""
    q = covariance_to_matrix(cov9)
    q = 0.5 * (q + np.transpose(q, (0, 2, 1)))
    try:
        np.linalg.cholesky(q)
    except np.linalg.LinAlgError as exc:
        raise ValueError("covariance is not symmetric positive definite") from exc
    om = np.linalg.inv(q)
    return 0.5 * (om + np.transpose(om, (0, 2, 1)))


def retract(poses, delta):
    """Values::retract: Pose2 body-frame update p o (dx, dy, dth)."""
    c, s = np.cos(poses[:, 2]), np.sin(poses[:, 2])
    out = np.empty_like(poses)
    out[:, 0] = poses[:, 0] + c * delta[:, 0] - s * delta[:, 1]
    out[:, 1] = poses[:, 1] + s * delta[:, 0] + c * delta[:, 1]
    out[:, 2] = poses[:, 2] + delta[:, 2]
    return out


def between_residual(pi, pj, z, jacobian_mode=JAC_GTSAM_DEFAULT):
    """BetweenFactor<Pose2>::evaluateError.  Returns e (E,3), A_i (E,3,3), A_j (E,3,3)."""
    ci, si = np.cos(pi[:, 2]), np.sin(pi[:, 2])
    cj, sj = np.cos(pj[:, 2]), np.sin(pj[:, 2])
    dx, dy = pj[:, 0] - pi[:, 0], pj[:, 1] - pi[:, 1]
    hx = ci * dx + si * dy
    hy = -si * dx + ci * dy
    c = ci * cj + si * sj
    s = -si * cj + ci * sj
    scale = c * c + s * s                        # gtsam::Rot2::normalize(): only when off 1 by more than 1e-10
    fix = np.abs(scale - 1.0) > 1e-10
    if fix.any():
        inv = np.where(fix, 1.0 / np.sqrt(np.where(fix, scale, 1.0)), 1.0)
        c, s = c * inv, s * inv
    cm, sm = np.cos(z[:, 2]), np.sin(z[:, 2])
    rx, ry = hx - z[:, 0], hy - z[:, 1]
    e = np.empty((pi.shape[0], 3))
    e[:, 0] = cm * rx + sm * ry
    e[:, 1] = -sm * rx + cm * ry
    e[:, 2] = np.arctan2(s * cm - c * sm, c * cm + s * sm)
    n = pi.shape[0]
    a_i = np.zeros((n, 3, 3))
    a_i[:, 0, 0] = -c
    a_i[:, 0, 1] = -s
    a_i[:, 0, 2] = -sj * dx + cj * dy
    a_i[:, 1, 0] = s
    a_i[:, 1, 1] = -c
    a_i[:, 1, 2] = -cj * dx - sj * dy
    a_i[:, 2, 2] = -1.0
    a_j = np.zeros((n, 3, 3))
    a_j[:, 0, 0] = a_j[:, 1, 1] = a_j[:, 2, 2] = 1.0
    if jacobian_mode == JAC_EXACT:
        d = np.zeros((n, 3, 3))
        ce, se = np.cos(e[:, 2]), np.sin(e[:, 2])
        d[:, 0, 0] = ce
        d[:, 0, 1] = -se
        d[:, 1, 0] = se
        d[:, 1, 1] = ce
        d[:, 2, 2] = 1.0
        a_i = d @ a_i
        a_j = d @ a_j
    return e, a_i, a_j


def prior_residual(p, prior, jacobian_mode=JAC_GTSAM_DEFAULT):
    """PriorFactor<Pose2>::evaluateError (GTSAM 4.0): e = -Local(x^-1 o prior), H = I."""
    c, s = np.cos(p[:, 2]), np.sin(p[:, 2])
    dx, dy = prior[:, 0] - p[:, 0], prior[:, 1] - p[:, 1]
    e = np.empty((p.shape[0], 3))
    e[:, 0] = -(c * dx + s * dy)
    e[:, 1] = -(-s * dx + c * dy)
    e[:, 2] = -wrap_pi(prior[:, 2] - p[:, 2])
    a = np.zeros((p.shape[0], 3, 3))
    a[:, 0, 0] = a[:, 1, 1] = a[:, 2, 2] = 1.0
    if jacobian_mode == JAC_EXACT:
        a[:, 0, 2] = e[:, 1]
        a[:, 1, 2] = -e[:, 0]
    return e, a


# --------------------------------------------------------------------------- the problem

class Problem:
    """A pose graph over dense indices 0..N-1 (dense index = rank of the key; = key-1 for the
    reference's ids, graph.cpp:31,48,68,75)."""

    def __init__(self, n, ei, ej, z, cov9, prior_idx, prior_pose, prior_cov9):
        self.n = int(n)
        self.ei = np.asarray(ei, dtype=np.int64)
        self.ej = np.asarray(ej, dtype=np.int64)
        self.z = np.asarray(z, dtype=np.float64).reshape(-1, 3)
        self.omega = information(cov9) if self.ei.size else np.zeros((0, 3, 3))
        self.prior_idx = np.atleast_1d(np.asarray(prior_idx, dtype=np.int64))
        self.prior_pose = np.asarray(prior_pose, dtype=np.float64).reshape(-1, 3)
        self.prior_omega = information(prior_cov9) if self.prior_idx.size else np.zeros((0, 3, 3))

    @classmethod
    def from_graph(cls, g):
        keys = np.asarray(g.keys)
        order = np.argsort(keys, kind="stable")
        skeys = keys[order]
        def idx(k):
            return np.searchsorted(skeys, np.asarray(k))
        return cls(keys.shape[0], idx(g.k1), idx(g.k2), g.z, g.cov, idx([g.prior_key]),
                   np.asarray(g.prior_pose).reshape(1, 3), np.asarray(g.prior_cov).reshape(1, 9))

    # ---- graph.error(values) = 0.5 * chi2
    def error(self, poses, mode=JAC_GTSAM_DEFAULT):
        return 0.5 * self.chi2(poses)

    def chi2(self, poses):
        tot = 0.0
        if self.ei.size:
            e, _, _ = between_residual(poses[self.ei], poses[self.ej], self.z)
            tot += float(np.einsum("ea,eab,eb->", e, self.omega, e))
        if self.prior_idx.size:
            e, _ = prior_residual(poses[self.prior_idx], self.prior_pose)
            tot += float(np.einsum("ea,eab,eb->", e, self.prior_omega, e))
        return tot

    def edge_terms(self, poses, mode=JAC_GTSAM_DEFAULT):
        """Per-edge e, A_i, A_j, and the five normal-equation pieces."""
        e, ai, aj = between_residual(poses[self.ei], poses[self.ej], self.z, mode)
        om = self.omega
        oe = np.einsum("eab,eb->ea", om, e)
        return dict(e=e, a_i=ai, a_j=aj,
                    h_ii=np.einsum("eba,ebc,ecd->ead", ai, om, ai),
                    h_ij=np.einsum("eba,ebc,ecd->ead", ai, om, aj),
                    h_jj=np.einsum("eba,ebc,ecd->ead", aj, om, aj),
                    b_i=np.einsum("eba,eb->ea", ai, oe), b_j=np.einsum("eba,eb->ea", aj, oe),
                    chi2=np.einsum("ea,ea->e", e, oe))

    def linearize(self, poses, mode=JAC_GTSAM_DEFAULT):
        """H = sum J^T Omega J (3N x 3N CSR, duplicates merged), b = sum J^T Omega e, chi2."""
        n = self.n
        rows, cols, vals = [], [], []
        b = np.zeros((n, 3))
        chi2 = 0.0
        k3 = np.arange(3)

        def push(bi, bj, blk):
            r = (3 * bi)[:, None, None] + k3[None, :, None]
            c = (3 * bj)[:, None, None] + k3[None, None, :]
            rows.append(np.broadcast_to(r, blk.shape).ravel())
            cols.append(np.broadcast_to(c, blk.shape).ravel())
            vals.append(blk.ravel())

        if self.ei.size:
            t = self.edge_terms(poses, mode)
            push(self.ei, self.ei, t["h_ii"])
            push(self.ei, self.ej, t["h_ij"])
            push(self.ej, self.ei, np.transpose(t["h_ij"], (0, 2, 1)))
            push(self.ej, self.ej, t["h_jj"])
            np.add.at(b, self.ei, t["b_i"])
            np.add.at(b, self.ej, t["b_j"])
            chi2 += float(t["chi2"].sum())
        if self.prior_idx.size:
            e, a = prior_residual(poses[self.prior_idx], self.prior_pose, mode)
            om = self.prior_omega
            oe = np.einsum("eab,eb->ea", om, e)
            push(self.prior_idx, self.prior_idx, np.einsum("eba,ebc,ecd->ead", a, om, a))
            np.add.at(b, self.prior_idx, np.einsum("eba,eb->ea", a, oe))
            chi2 += float(np.einsum("ea,ea->", e, oe))
        h = sp.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))),
                          shape=(3 * n, 3 * n)).tocsr()
        h.sum_duplicates()
        return h, b, chi2

    def pattern(self):
        """Sorted unique 3x3 block pattern of H incl. every diagonal block: (rowptr, colidx) int32.
        This is the integer output that must match the CUDA structure build bit-exactly."""
        n = self.n
        r = np.concatenate([np.arange(n), self.ei, self.ej])
        c = np.concatenate([np.arange(n), self.ej, self.ei])
        key = np.unique(r.astype(np.int64) * n + c.astype(np.int64))
        rows = (key // n).astype(np.int64)
        colidx = (key % n).astype(np.int32)
        rowptr = np.zeros(n + 1, dtype=np.int32)
        np.cumsum(np.bincount(rows, minlength=n), out=rowptr[1:])
        return rowptr, colidx

    def bsr(self, h):
        """CSR H -> (rowptr, colidx, blocks[nnzb,3,3]) on the full pattern() (missing blocks = 0)."""
        rowptr, colidx = self.pattern()
        hb = sp.bsr_matrix(h, blocksize=(3, 3))
        hb.sort_indices()
        dense = {}
        out = np.zeros((colidx.shape[0], 3, 3))
        ip = hb.indptr
        for r in range(self.n):
            src = {int(c): hb.data[k] for k, c in zip(range(ip[r], ip[r + 1]), hb.indices[ip[r]:ip[r + 1]])}
            for k in range(rowptr[r], rowptr[r + 1]):
                blk = src.get(int(colidx[k]))
                if blk is not None:
                    out[k] = blk
        return rowptr, colidx, out


def solve_damped(h, b, lam, diag=None):
    """(H + lam*I) delta = -b, exact sparse LU (stands in for GTSAM's multifrontal Cholesky).
    diagonal_damping: lam * diag(H) clipped to [1e-6, 1e32] (GTSAM min/max_diagonal)."""
    n3 = h.shape[0]
    if diag is None:
        damp = sp.identity(n3, format="csc") * lam
    else:
        damp = sp.diags(lam * diag, format="csc")
    lu = spla.splu((h + damp).tocsc(), permc_spec="MMD_AT_PLUS_A", diag_pivot_thresh=0.0,
                   options=dict(SymmetricMode=True))
    return lu.solve(-b.ravel()).reshape(-1, 3)


def marginal_covariances(prob: Problem, poses, idx, mode=JAC_GTSAM_DEFAULT):
    """gtsam::Marginals(graph, values).marginalCovariance(key) for each dense index in idx: the (key, key) 3x3 block
    of the inverse of the Hessian J^T Omega J of the graph linearised at `poses` (tangent space of each pose).
    The reference has the call commented out (src/graph/src/graph.cpp:120, 126-127).  Exact sparse LU."""
    h, _, _ = prob.linearize(np.asarray(poses, dtype=np.float64), mode)
    lu = spla.splu(h.tocsc(), permc_spec="MMD_AT_PLUS_A", diag_pivot_thresh=0.0, options=dict(SymmetricMode=True))
    idx = np.atleast_1d(idx)
    out = np.zeros((idx.shape[0], 3, 3))
    for k, i in enumerate(idx):
        rhs = np.zeros((h.shape[0], 3))
        rhs[3 * i:3 * i + 3] = np.eye(3)
        sol = lu.solve(rhs)
        blk = sol[3 * i:3 * i + 3]
        out[k] = 0.5 * (blk + blk.T)
    return out


def lm_optimize(prob: Problem, init, params: LMParams | None = None, solver=None):
    """LevenbergMarquardtOptimizer(graph, initial).optimize() restated (GTSAM 4.0.x
    NonlinearOptimizer::defaultOptimize + LevenbergMarquardtOptimizer::iterate/tryLambda).
    Returns (poses with theta in (-pi,pi], LMTrace)."""
    p = params or LMParams()
    solver = solver or solve_damped
    x = np.array(init, dtype=np.float64, copy=True)
    mode = p.jacobian_mode
    error = prob.error(x)
    lam = p.lambda_initial
    tr = LMTrace()
    tr.errors.append(error)
    iterations = 0
    eps = np.finfo(np.float64).eps
    # NonlinearOptimizer::defaultOptimize: return before iterating when the initial error is already <= errorTol
    nothing_to_do = error <= p.error_tol
    while not nothing_to_do:
        if iterations >= p.max_iterations:
            break
        current_error = error
        # ---- iterate(): linearize once, then tryLambda until a step is taken or we give up
        h, b, _ = prob.linearize(x, mode)
        diag = None
        if p.diagonal_damping:
            diag = np.clip(h.diagonal(), 1e-6, 1e32)
        tries = 0
        while True:
            tries += 1
            tr.inner_iterations += 1
            delta = solver(h, b, lam, diag)
            d = delta.ravel()
            # linear.error(0) - linear.error(delta), linear.error(d) = 0.5*|A d - b_gtsam|^2 (undamped)
            old_lin = error
            lin_change = -float(d @ b.ravel()) - 0.5 * float(d @ (h @ d))
            step_ok = False
            stop_search = False
            new_x, new_error = None, math.inf
            if lin_change >= 0.0:
                new_x = retract(x, delta)
                new_error = prob.error(new_x)
                cost_change = error - new_error
                if lin_change > eps * old_lin:
                    fidelity = cost_change / lin_change
                    step_ok = fidelity > p.min_model_fidelity
                if abs(cost_change) < p.rel_error_tol * error:
                    stop_search = True
            if step_ok:
                tr.lambdas.append(lam)
                x, error = new_x, new_error
                lam = max(p.lambda_lower, lam / p.lambda_factor)
                iterations += 1
                break
            elif not stop_search:
                lam *= p.lambda_factor
                if lam >= p.lambda_upper:
                    break
            else:
                break
        tr.tries.append(tries)
        tr.errors.append(error)
        # ---- checkConvergence(relTol, absTol, errTol, currentError, newError)
        if iterations >= p.max_iterations:
            break
        if error <= p.error_tol:
            break
        abs_dec = current_error - error
        rel_dec = abs_dec / current_error if current_error != 0 else 0.0
        if (p.rel_error_tol and rel_dec <= p.rel_error_tol) or abs_dec <= p.abs_error_tol:
            break
        if not math.isfinite(current_error):
            break
    tr.iterations = iterations
    tr.lambda_final = lam
    out = x.copy()
    out[:, 2] = wrap_pi(out[:, 2])
    return out, tr
