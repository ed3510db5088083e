"""The ROS-free replay driver: a RECORDED stream of common/Registration messages
(tests/golden/registration_stream_manhattan400.reg) goes through the graph node's verbs — registration_callback ->
prior_factor / new_factor / loop_factor, solve (reference graph.cpp:27-132,180-200) — and the Keyframes message the
node would publish (graph.cpp:18-25) is compared between the C++ GraphNode (examples/graph_node_noros.cpp, built here
with g++ against the library), the Python GraphNode and the oracle's replay (oracle/replay.py)."""
import os
import subprocess

import numpy as np
import pytest

from graphslam_b200 import build, graph, regstream
from oracle import gs_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STREAM = os.path.join(ROOT, "tests", "golden", "registration_stream_manhattan400.reg")
GOLD = os.path.join(ROOT, "tests", "golden", "registration_stream_manhattan400_keyframes.npz")


def test_stream_file_round_trips_bit_for_bit(tmp_path):
    msgs = regstream.read(STREAM)
    assert sum(1 for m in msgs if m == "SOLVE") == 4 and sum(1 for m in msgs if m != "SOLVE") == 400
    assert msgs[0]["first_frame_flag"] and not msgs[0]["keyframe_flag"]
    out = tmp_path / "again.reg"
    regstream.write(out, msgs)
    again = regstream.read(out)
    assert len(again) == len(msgs)
    for a, b in zip(msgs, again):
        if a == "SOLVE":
            assert b == "SOLVE"
            continue
        assert {k: a[k] for k in ("first_frame_flag", "keyframe_flag", "loop_closure_flag")} == \
               {k: b[k] for k in ("first_frame_flag", "keyframe_flag", "loop_closure_flag")}
        for f in ("factor_new", "factor_loop"):
            assert (f in a) == (f in b)
            if f in a:
                assert a[f]["id_1"] == b[f]["id_1"] and a[f].get("id_2") == b[f].get("id_2")
                assert np.array_equal(a[f]["delta"], b[f]["delta"]) and np.array_equal(a[f]["covariance"], b[f]["covariance"])


def test_oracle_replay_matches_the_committed_keyframes():
    from oracle import replay as oreplay
    gold = np.load(GOLD)
    ids, poses, traces = oreplay.replay(regstream.read(STREAM))
    assert np.array_equal(ids, gold["ids"]) and [t.iterations for t in traces] == gold["iterations"].tolist()
    assert np.abs(poses[:, :2] - gold["poses"][:, :2]).max() <= 1e-9
    assert np.abs(O.wrap_pi(poses[:, 2] - gold["poses"][:, 2])).max() <= 1e-9


def _build_replayer(tmp_path):
    build.build()
    exe = str(tmp_path / "graph_node_noros")
    libdir = os.path.dirname(build.LIB)
    subprocess.run(["g++", "-std=c++17", "-O2", "-Wall", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "examples", "graph_node_noros.cpp"), "-L", libdir, "-lgraphslam_cuda",
                    f"-Wl,-rpath,{libdir}", "-o", exe], check=True, capture_output=True, text=True)
    return exe


@pytest.mark.gpu
def test_cpp_graph_node_replays_the_recorded_stream(tmp_path):
    gold = np.load(GOLD)
    # C++ node
    exe = _build_replayer(tmp_path)
    out = tmp_path / "keyframes.txt"
    res = subprocess.run([exe, STREAM, str(out)], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    rows = [ln.split() for ln in open(out) if ln.startswith("KEYFRAME")]
    cids = np.array([int(r[1]) for r in rows], dtype=np.uint64)
    cposes = np.array([[float(v) for v in r[2:5]] for r in rows])
    # Python node, same library, same call sequence
    node = graph.GraphNode()
    stats = regstream.replay(node, regstream.read(STREAM))
    pids, pposes = node.publish_keyframes()
    node.graph.close()
    assert np.array_equal(cids, gold["ids"]) and np.array_equal(pids, gold["ids"])
    assert np.array_equal(cposes, pposes)                       # same calls into the same library: same bits
    assert [s["iterations"] for s in stats] == gold["iterations"].tolist()
    # against the oracle's replay: the north_star's pose tolerance
    assert np.abs(cposes[:, :2] - gold["poses"][:, :2]).max() <= 1e-6
    assert np.abs(O.wrap_pi(cposes[:, 2] - gold["poses"][:, 2])).max() <= 1e-6
    assert "closest keyframe to 400:" in res.stderr
