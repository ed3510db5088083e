"""Recorded `common/Registration` streams (reference src/common/msg/Registration.msg) as text files, and their replay
through the graph node's verbs (registration_callback, graph.cpp:180-200) — ROS stripped.

One line per message the graph node receives, in arrival order; only the fields the node reads are kept
(first_frame_flag, keyframe_flag, loop_closure_flag, factor_new, factor_loop; `keyframe_last` is filled at replay
time from the node's own last_keyframe service, as the scanner does before it publishes, scanner.cpp / graph.cpp:134-144):

    REG <first> <keyframe> <loop> [NEW id_1 dx dy dtheta c0..c8] [LOOP id_1 id_2 dx dy dtheta c0..c8]
    SOLVE                     # where the node's solve() runs (the call the reference has commented out, graph.cpp:195)

Numbers are written with %.17g, so a replay is bit-for-bit the recorded run.  The same format is read by the C++
shim (include/graphslam_cuda.hpp: gs::read_registration_stream) and replayed by examples/graph_node_noros.cpp.
"""
from __future__ import annotations

import numpy as np


def _fmt(v):
    return " ".join("%.17g" % float(x) for x in np.asarray(v, dtype=np.float64).ravel())


def from_graph(g, solve_every: int = 0):
    """The message sequence a scanner would have published for synth.PoseGraph `g`: keyframe k arrives with its
    odometry factor (k-1 -> k) and at most one loop factor (k -> older, as scanner.cpp:151-152 orders them); further
    loop factors of the same keyframe are carried by keyframe-less messages?  No: the reference adds one loop factor
    per keyframe, so graphs with more are refused."""
    i, j = g.dense_edges()
    by_arrival = {}
    for k in range(g.e):
        by_arrival.setdefault(int(max(i[k], j[k])), []).append(k)
    msgs = [dict(first_frame_flag=True, keyframe_flag=False, loop_closure_flag=False)]
    for kf in range(1, g.n):
        reg = dict(first_frame_flag=False, keyframe_flag=True, loop_closure_flag=False)
        for k in by_arrival.get(kf, []):
            f = dict(id_1=int(g.k1[k]), id_2=int(g.k2[k]), delta=np.array(g.z[k], dtype=np.float64),
                     covariance=np.array(g.cov[k], dtype=np.float64))
            if i[k] < j[k]:
                if "factor_new" in reg:
                    raise ValueError("two odometry factors into one keyframe")
                reg["factor_new"] = f
            else:
                if reg["loop_closure_flag"]:
                    raise ValueError("the reference carries one loop factor per Registration message")
                reg["loop_closure_flag"] = True
                reg["factor_loop"] = f
        msgs.append(reg)
        if solve_every and kf % solve_every == 0:
            msgs.append("SOLVE")
    msgs.append("SOLVE")
    return msgs


def write(path, msgs):
    with open(path, "w") as f:
        f.write("# registration stream v1 (graphslam_b200/regstream.py): REG first keyframe loop [NEW id_1 dx dy dth c0..c8] "
                "[LOOP id_1 id_2 dx dy dth c0..c8] | SOLVE\n")
        for m in msgs:
            if m == "SOLVE":
                f.write("SOLVE\n")
                continue
            parts = ["REG", str(int(bool(m.get("first_frame_flag")))), str(int(bool(m.get("keyframe_flag")))),
                     str(int(bool(m.get("loop_closure_flag"))))]
            if "factor_new" in m:
                fn = m["factor_new"]
                parts += ["NEW", str(int(fn["id_1"])), _fmt(fn["delta"]), _fmt(fn["covariance"])]
            if m.get("loop_closure_flag"):
                fl = m["factor_loop"]
                parts += ["LOOP", str(int(fl["id_1"])), str(int(fl["id_2"])), _fmt(fl["delta"]), _fmt(fl["covariance"])]
            f.write(" ".join(parts) + "\n")


def read(path):
    msgs = []
    with open(path) as f:
        for ln, line in enumerate(f, 1):
            t = line.split("#", 1)[0].split()
            if not t:
                continue
            if t[0] == "SOLVE":
                msgs.append("SOLVE")
                continue
            if t[0] != "REG" or len(t) < 4:
                raise ValueError(f"{path}:{ln}: not a registration line")
            m = dict(first_frame_flag=t[1] == "1", keyframe_flag=t[2] == "1", loop_closure_flag=t[3] == "1")
            k = 4
            while k < len(t):
                if t[k] == "NEW":
                    v = [float(x) for x in t[k + 2:k + 14]]
                    m["factor_new"] = dict(id_1=int(t[k + 1]), delta=np.array(v[:3]), covariance=np.array(v[3:]))
                    k += 14
                elif t[k] == "LOOP":
                    v = [float(x) for x in t[k + 3:k + 15]]
                    m["factor_loop"] = dict(id_1=int(t[k + 1]), id_2=int(t[k + 2]), delta=np.array(v[:3]), covariance=np.array(v[3:]))
                    k += 15
                else:
                    raise ValueError(f"{path}:{ln}: unknown field {t[k]}")
            msgs.append(m)
    return msgs


def replay(node, msgs):
    """Play the stream through a graph.GraphNode; returns the stats of every solve().  The Keyframes the node would
    publish (graph.cpp:18-25) are node.keyframes afterwards."""
    stats = []
    for m in msgs:
        if m == "SOLVE":
            stats.append(node.solve())
            continue
        if m.get("keyframe_flag"):
            m = dict(m, keyframe_last=node.last_keyframe())     # what the scanner gets from the last_keyframe service
        node.registration_callback(m)
    return stats
