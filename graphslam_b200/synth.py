"""Seeded synthetic SE(2) pose graphs: the five BASELINE.json configs.

Every consumer (oracle, CPU port, CUDA path, bench) reads graphs produced here, so they all see
identical bytes.  PRNG = splitmix64 over (seed, stream, counter) -> Box-Muller FP64, fully
vectorised and reproducible.

Graph construction mirrors the reference node's call order (reference
`src/graph/src/graph.cpp:180-200`): for keyframe k the odometry edge (k-1 -> k) is added first
(`new_factor`, graph.cpp:63-95; id_1 = last, id_2 = new), then any loop edges (k -> older j,
`loop_factor`, graph.cpp:97-113; id_1 = last keyframe, id_2 = closest older keyframe, so i > j).
Keys are 1..N (graph.cpp:31,68,207), one prior on key 1 at the origin with sigma 0.1
(graph.cpp:13-14,34-45).  Covariances are full row-major 3x3 (`Pose2DWithCovariance.msg:2`).

The initial guess is dead reckoning over the noisy odometry edges with the CORRECT SE(2) compose
(SURVEY.md §0.11); `reference_compose=True` replays the reference's `compose`
(graph.hpp:30-43, translation dropped, one-sided wrap) for small faithful-mode tests.
`anchor_every=R` restarts the dead reckoning from the ground truth every R poses: the reference
re-solves after keyframes and warm starts (`initial = poses_opti`, graph.cpp:130), so a batch of R
new keyframes is dead-reckoned from an already optimised pose; 0 = pure dead reckoning.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser, vectorised over uint64."""
    with np.errstate(over="ignore"):
        z = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def _u01(seed: int, stream: int, n: int, offset: int = 0) -> np.ndarray:
    """n uniforms in (0,1): counter-based, so any slice is reproducible on its own."""
    ctr = np.arange(offset, offset + n, dtype=np.uint64)
    with np.errstate(over="ignore"):
        key = splitmix64(np.uint64(seed) ^ (np.uint64(stream) << np.uint64(48)))
        bits = splitmix64(key ^ (ctr * np.uint64(0xD1342543DE82EF95)))
    return ((bits >> np.uint64(11)).astype(np.float64) + 0.5) * (1.0 / 9007199254740992.0)


def _normal(seed: int, stream: int, n: int) -> np.ndarray:
    """n standard normals by Box-Muller over two uniform streams."""
    u1 = _u01(seed, 2 * stream, n)
    u2 = _u01(seed, 2 * stream + 1, n)
    return np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * math.pi * u2)


def wrap_pi(t):
    return np.arctan2(np.sin(t), np.cos(t))


def compose(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """Correct SE(2) a (+) b, rows of (x, y, theta)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    c, s = np.cos(a[..., 2]), np.sin(a[..., 2])
    out = np.empty(np.broadcast(a, b).shape, dtype=np.float64)
    out[..., 0] = a[..., 0] + c * b[..., 0] - s * b[..., 1]
    out[..., 1] = a[..., 1] + s * b[..., 0] + c * b[..., 1]
    out[..., 2] = wrap_pi(a[..., 2] + b[..., 2])
    return out


def between(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """a^-1 o b."""
    c, s = np.cos(a[..., 2]), np.sin(a[..., 2])
    dx, dy = b[..., 0] - a[..., 0], b[..., 1] - a[..., 1]
    out = np.empty(np.broadcast(a, b).shape, dtype=np.float64)
    out[..., 0] = c * dx + s * dy
    out[..., 1] = -s * dx + c * dy
    out[..., 2] = wrap_pi(b[..., 2] - a[..., 2])
    return out


def compose_reference_scalar(a, b):
    """Reference `compose`, graph.hpp:30-43, restated with its quirks (host scalar)."""
    cos_th, sin_th = math.cos(a[2]), math.sin(a[2])
    x = (cos_th * b[0]) + (-sin_th * b[1])
    y = (sin_th * b[0]) + (cos_th * b[1])
    th = a[2] + b[2]
    th = math.fmod(th + math.pi, 2 * math.pi) - math.pi
    return (x, y, th)


@dataclass
class PoseGraph:
    """Plain host container.  Keys are 1-based as in the reference; ids index into `keys`."""
    name: str
    keys: np.ndarray            # (N,) uint64, 1..N
    init: np.ndarray            # (N,3) initial guess x,y,theta
    gt: np.ndarray              # (N,3) ground truth
    k1: np.ndarray              # (E,) uint64 key of id_1
    k2: np.ndarray              # (E,) uint64 key of id_2
    z: np.ndarray               # (E,3) measured delta
    cov: np.ndarray             # (E,9) row-major covariance
    prior_key: int = 1
    prior_pose: np.ndarray = field(default_factory=lambda: np.zeros(3))
    prior_cov: np.ndarray = field(default_factory=lambda: np.diag([0.01, 0.01, 0.01]).reshape(9))
    meta: dict = field(default_factory=dict)

    @property
    def n(self) -> int:
        return int(self.keys.shape[0])

    @property
    def e(self) -> int:
        return int(self.k1.shape[0])

    def dense_edges(self):
        """(i, j) int32 dense indices (= key-1 for the reference's ids)."""
        return (self.k1.astype(np.int64) - 1).astype(np.int32), (self.k2.astype(np.int64) - 1).astype(np.int32)

    def prefix(self, n: int) -> "PoseGraph":
        """Sub-graph over the first n poses (the graph as it was when keyframe n arrived)."""
        m = (self.k1 <= n) & (self.k2 <= n)
        return PoseGraph(self.name + f"[:{n}]", self.keys[:n].copy(), self.init[:n].copy(), self.gt[:n].copy(),
                         self.k1[m].copy(), self.k2[m].copy(), self.z[m].copy(), self.cov[m].copy(),
                         self.prior_key, self.prior_pose.copy(), self.prior_cov.copy(), dict(self.meta))


# ------------------------------------------------------------------ trajectories

def _lawnmower_gt(n: int, width: int) -> np.ndarray:
    k = np.arange(n)
    r, q = k // width, k % width
    c = np.where(r % 2 == 0, q, width - 1 - q)
    gt = np.empty((n, 3))
    gt[:, 0] = c
    gt[:, 1] = r
    heading = np.where(r % 2 == 0, 0.0, math.pi)
    # the last cell of a row already faces the next row (a step in +y follows)
    heading = np.where((q == width - 1), math.pi / 2, heading)
    gt[:, 2] = wrap_pi(heading)
    return gt


def _lawn_index(r, c, width):
    """pose index of grid cell (row r, column c)."""
    return r * width + np.where(r % 2 == 0, c, width - 1 - c)


def _manhattan_gt(n: int, seed: int, box: int = 40) -> np.ndarray:
    """Unit steps on the integer lattice, turn +-90deg w.p. 0.25 after a step; reflect at a box wall
    so that the walk revisits itself (loop closures)."""
    u = _u01(seed, 100, n)
    sgn = _u01(seed, 101, n)
    gt = np.zeros((n, 3))
    x = y = 0
    h = 0
    dirs = ((1, 0), (0, 1), (-1, 0), (0, -1))
    for k in range(1, n):
        dx, dy = dirs[h]
        if abs(x + dx) > box or abs(y + dy) > box:      # wall: turn (left or right) until free
            h = (h + (1 if sgn[k] < 0.5 else 3)) % 4
            dx, dy = dirs[h]
            if abs(x + dx) > box or abs(y + dy) > box:
                h = (h + 2) % 4
                dx, dy = dirs[h]
        x, y = x + dx, y + dy
        if u[k] < 0.25:
            h_next = (h + (1 if sgn[k] < 0.5 else 3)) % 4
        else:
            h_next = h
        gt[k] = (x, y, wrap_pi(h_next * math.pi / 2))
        h = h_next
    return gt


# ------------------------------------------------------------------ assembly of a graph

def _finish(name, seed, gt, loops_from, loops_to, sigma_xy, sigma_th, full_cov, anchor_every,
            reference_compose, meta) -> PoseGraph:
    n = gt.shape[0]
    # odometry edges (k-1 -> k), loop edges (k -> j), interleaved in keyframe arrival order
    odo_i = np.arange(0, n - 1, dtype=np.int64)
    odo_j = odo_i + 1
    li = np.asarray(loops_from, dtype=np.int64)
    lj = np.asarray(loops_to, dtype=np.int64)
    ei = np.concatenate([odo_i, li])
    ej = np.concatenate([odo_j, lj])
    arrival = np.concatenate([odo_j, li])                 # keyframe at which the factor is added
    kind = np.concatenate([np.zeros(n - 1, np.int64), np.ones(li.shape[0], np.int64)])
    seq = np.concatenate([np.zeros(n - 1, np.int64), np.arange(li.shape[0], dtype=np.int64)])
    order = np.lexsort((seq, kind, arrival))
    ei, ej = ei[order], ej[order]
    ne = ei.shape[0]

    # measurement = gt_between (+) noise, noise ~ N(0, diag(sxy^2, sxy^2, sth^2))
    h = between(gt[ei], gt[ej])
    noise = np.stack([sigma_xy * _normal(seed, 1, ne), sigma_xy * _normal(seed, 2, ne),
                      sigma_th * _normal(seed, 3, ne)], axis=1)
    if full_cov:
        # rotate the xy noise ellipse by a random angle per edge -> dense 3x3 Q
        phi = 2 * math.pi * _u01(seed, 20, ne)
        scale = 1.0 + 2.0 * _u01(seed, 21, ne)             # anisotropy: sigma_x = scale*sigma_xy
        c, s = np.cos(phi), np.sin(phi)
        nx, ny = scale * noise[:, 0], noise[:, 1]
        noise[:, 0], noise[:, 1] = c * nx - s * ny, s * nx + c * ny
        sx2, sy2 = (scale * sigma_xy) ** 2, np.full(ne, sigma_xy ** 2)
        cov = np.zeros((ne, 9))
        cov[:, 0] = c * c * sx2 + s * s * sy2
        cov[:, 1] = cov[:, 3] = c * s * (sx2 - sy2)
        cov[:, 4] = s * s * sx2 + c * c * sy2
        rho = 0.3 * (2 * _u01(seed, 22, ne) - 1)            # couple theta with x a little
        cov[:, 8] = sigma_th ** 2
        cov[:, 2] = cov[:, 6] = rho * np.sqrt(cov[:, 0]) * sigma_th
    else:
        cov = np.zeros((ne, 9))
        cov[:, 0] = cov[:, 4] = sigma_xy ** 2
        cov[:, 8] = sigma_th ** 2
    z = compose(h, noise)

    # initial guess: dead reckoning over the noisy odometry edges (graph.cpp:71)
    odo_mask = (ej == ei + 1)
    # first odometry factor per target keyframe, in arrival order
    zo = np.zeros((n, 3))
    idx = np.flatnonzero(odo_mask)
    _, first = np.unique(ej[idx], return_index=True)
    zo[ej[idx][first]] = z[idx][first]
    init = np.zeros((n, 3))
    if reference_compose:
        cur = (0.0, 0.0, 0.0)
        for k in range(1, n):
            cur = compose_reference_scalar(cur, zo[k])
            init[k] = cur
    else:
        seg = np.zeros(n, dtype=np.int64) if not anchor_every else (np.arange(n) // anchor_every)
        starts = np.flatnonzero(np.r_[True, seg[1:] != seg[:-1]])
        # theta_k = theta_start + sum dtheta ; t_k = t_start + sum R(theta_{k-1}) dxy  (vectorised scan)
        dth = zo[:, 2].copy()
        dth[starts] = 0.0
        cs = np.cumsum(dth)
        theta = gt[starts[seg], 2] + (cs - cs[starts[seg]])
        prev_theta = np.r_[0.0, theta[:-1]]
        c, s = np.cos(prev_theta), np.sin(prev_theta)
        sx = c * zo[:, 0] - s * zo[:, 1]
        sy = s * zo[:, 0] + c * zo[:, 1]
        sx[starts] = 0.0
        sy[starts] = 0.0
        cx, cy = np.cumsum(sx), np.cumsum(sy)
        init[:, 0] = gt[starts[seg], 0] + (cx - cx[starts[seg]])
        init[:, 1] = gt[starts[seg], 1] + (cy - cy[starts[seg]])
        init[:, 2] = wrap_pi(theta)
        if anchor_every:
            # an anchor is "a previously optimised pose": ground truth plus a small residual error
            na = starts.shape[0]
            pert = np.stack([0.5 * sigma_xy * _normal(seed, 30, na), 0.5 * sigma_xy * _normal(seed, 31, na),
                             0.5 * sigma_th * _normal(seed, 32, na)], axis=1)
            pert[0] = 0.0
            off = pert[seg]
            init[:, 0] += off[:, 0]
            init[:, 1] += off[:, 1]
            init[:, 2] = wrap_pi(init[:, 2] + off[:, 2])
    meta = dict(meta)
    meta.update(seed=seed, sigma_xy=sigma_xy, sigma_th=sigma_th, full_cov=bool(full_cov),
                anchor_every=int(anchor_every), reference_compose=bool(reference_compose),
                n=n, e=ne, n_loops=int(li.shape[0]))
    return PoseGraph(name, np.arange(1, n + 1, dtype=np.uint64), init, gt,
                     (ei + 1).astype(np.uint64), (ej + 1).astype(np.uint64), z, cov, meta=meta)


def manhattan(n: int = 3500, seed: int = 3500, target_edges: int | None = 5500, sigma_xy=0.05, sigma_th=0.02,
              full_cov=False, anchor_every=0, reference_compose=False, skip: int = 10, radius: float = 2.0,
              name="C1-manhattan") -> PoseGraph:
    """C1: Manhattan world; loop edge k -> nearest earlier pose j <= k-skip within `radius`
    (mirrors `closest_keyframe`, graph.cpp:15,146-178), thinned to about target_edges total."""
    gt = _manhattan_gt(n, seed)
    cells: dict[tuple[int, int], list[int]] = {}
    cand_from, cand_to = [], []
    rr = int(math.ceil(radius))
    xs, ys = gt[:, 0].astype(np.int64), gt[:, 1].astype(np.int64)
    for k in range(n):
        kk = k - skip
        if kk >= 0:
            cells.setdefault((int(xs[kk]), int(ys[kk])), []).append(kk)
        best, bd = -1, radius * radius + 1e-9
        for dx in range(-rr, rr + 1):
            for dy in range(-rr, rr + 1):
                d2 = dx * dx + dy * dy
                if d2 > radius * radius:
                    continue
                lst = cells.get((int(xs[k]) + dx, int(ys[k]) + dy))
                if lst and (d2 < bd or (d2 == bd and lst[0] < best)):   # strict '<' then lowest index
                    best, bd = lst[0], d2
        if best >= 0:
            cand_from.append(k)
            cand_to.append(best)
    cand_from = np.asarray(cand_from, dtype=np.int64)
    cand_to = np.asarray(cand_to, dtype=np.int64)
    if target_edges is not None and cand_from.shape[0] > 0:
        p = min(1.0, max(0.0, (target_edges - (n - 1)) / cand_from.shape[0]))
        keep = _u01(seed, 110, cand_from.shape[0]) < p
        cand_from, cand_to = cand_from[keep], cand_to[keep]
    return _finish(name, seed, gt, cand_from, cand_to, sigma_xy, sigma_th, full_cov, anchor_every,
                   reference_compose, dict(kind="manhattan"))


def lawnmower(n: int, width: int, seed: int, offsets, probs, sigma_xy=0.05, sigma_th=0.02, full_cov=False,
              anchor_every=100, name="lawnmower") -> PoseGraph:
    """Lawn-mower sweep; loop edges from pose k to the earlier pose at grid offset (dr, dc) w.p. probs."""
    gt = _lawnmower_gt(n, width)
    k = np.arange(n, dtype=np.int64)
    r, q = k // width, k % width
    c = np.where(r % 2 == 0, q, width - 1 - q)
    lf, lt = [], []
    for s, ((dr, dc), p) in enumerate(zip(offsets, probs)):
        rr, cc = r + dr, c + dc
        ok = (rr >= 0) & (cc >= 0) & (cc < width)
        j = _lawn_index(np.maximum(rr, 0), np.clip(cc, 0, width - 1), width)
        ok &= (j < k - 1) & (j >= 0)                        # strictly earlier, and not the odometry pair
        if p < 1.0:
            ok &= _u01(seed, 200 + s, n) < p
        lf.append(k[ok])
        lt.append(j[ok])
    lf = np.concatenate(lf) if lf else np.zeros(0, np.int64)
    lt = np.concatenate(lt) if lt else np.zeros(0, np.int64)
    o = np.lexsort((lt, lf))
    return _finish(name, seed, gt, lf[o], lt[o], sigma_xy, sigma_th, full_cov, anchor_every, False,
                   dict(kind="lawnmower", width=width))


def config(cid: str, n: int | None = None, **kw) -> PoseGraph:
    """The BASELINE.json configs.  `n` overrides the pose count (scaled-down parity cases)."""
    cid = cid.upper()
    if cid == "C1":
        return manhattan(n or 3500, seed=3500, **kw)
    if cid == "C2":
        return lawnmower(n or 100_000, 317, 100000, [(-1, 0)], [0.20], name="C2-grid100k", **kw)
    if cid == "C3":
        nn = n or 1_000_000
        width = 1000 if nn >= 1_000_000 else max(8, int(round(math.sqrt(nn))))
        return lawnmower(nn, width, 1000000, [(-1, 0), (-1, -1)], [1.0, 1.0], name="C3-grid1M", **kw)
    if cid == "C4":
        offs = [(dr, dc) for dr in (0, -1, -2, -3) for dc in range(-3, 4)]
        return lawnmower(n or 200_000, 448, 200000, offs, [5.0 / 6.0] * len(offs), name="C4-dense200k", **kw)
    if cid == "C5":
        return lawnmower(n or 50_000, 224, 50000, [(-1, 0)], [0.20], name="C5-incremental", **kw)
    raise ValueError(f"unknown config {cid!r}")
