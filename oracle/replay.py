"""TEST INFRASTRUCTURE: replay of a recorded registration stream (graphslam_b200/regstream.py format) with the numpy
oracle instead of the library — the reference node's bookkeeping (graph.cpp:27-132,180-200) restated over
gs_oracle.lm_optimize: prior on the first keyframe (sigma 0.1, graph.cpp:13-14,38-42), new keyframe = compose(last
optimised pose, delta) (graph.cpp:72; the correct SE(2) compose, as the library's default), one LM per SOLVE
warm-started from the previous optimum (graph.cpp:119,130)."""
from __future__ import annotations

import numpy as np

from . import gs_oracle as O


def _compose(a, b):
    c, s = np.cos(a[2]), np.sin(a[2])
    th = a[2] + b[2]
    return np.array([a[0] + c * b[0] - s * b[1], a[1] + s * b[0] + c * b[1], np.arctan2(np.sin(th), np.cos(th))])


def replay(msgs, sigma_xy_prior=0.1, sigma_th_prior=0.1):
    """Returns (ids, poses) of the Keyframes message after the last event, and the LM traces of every solve."""
    poses, ei, ej, z, cov = [], [], [], [], []
    traces = []
    for m in msgs:
        if isinstance(m, str):
            n = len(poses)
            prob = O.Problem(n, np.array(ei, dtype=np.int64), np.array(ej, dtype=np.int64),
                             np.array(z, dtype=np.float64).reshape(-1, 3), np.array(cov, dtype=np.float64).reshape(-1, 9),
                             np.array([0]), np.zeros((1, 3)),
                             np.diag([sigma_xy_prior ** 2, sigma_xy_prior ** 2, sigma_th_prior ** 2]).reshape(1, 9))
            x, tr = O.lm_optimize(prob, np.array(poses))
            poses = [p for p in x]
            traces.append(tr)
            continue
        if m.get("first_frame_flag"):
            poses.append(np.zeros(3))
        elif m.get("keyframe_flag"):
            f = m["factor_new"]
            poses.append(_compose(poses[-1], f["delta"]))
            ei.append(f["id_1"] - 1)
            ej.append(len(poses) - 1)
            z.append(f["delta"])
            cov.append(f["covariance"])
            if m.get("loop_closure_flag"):
                f = m["factor_loop"]
                ei.append(f["id_1"] - 1)
                ej.append(f["id_2"] - 1)
                z.append(f["delta"])
                cov.append(f["covariance"])
    return np.arange(1, len(poses) + 1, dtype=np.uint64), np.array(poses), traces
