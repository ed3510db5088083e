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

# the same function run as the REFERENCE'S OWN CODE: graph.hpp compiled from /root/reference by `make -C oracle ref`
# (oracle/_ref/libgraph_hpp_ref.so).  2048 pairs incl. the theta wrap on both sides and large translations; this
# table is what pins compose_reference on machines where the reference build is absent.
from oracle import ref_graph_hpp as R      # noqa: E402
if R.available():
    assert all(np.array_equal(R.compose(x, y), w) for x, y, w in zip(a, b, out)), "python restatement != reference build"
    rng2 = np.random.default_rng(30_43)
    a2 = rng2.uniform(-50, 50, size=(2048, 3))
    b2 = rng2.uniform(-50, 50, size=(2048, 3))
    a2[:, 2] = rng2.uniform(-2 * math.pi, 2 * math.pi, size=2048)
    b2[:, 2] = rng2.uniform(-2 * math.pi, 2 * math.pi, size=2048)
    a2[:6, 2] = [math.pi, -math.pi, 3.0, -3.0, 0.0, math.pi / 2]
    b2[:6, 2] = [0.0, 0.0, 0.2, -0.2, -math.pi, math.pi / 2]
    out2 = np.array([R.compose(x, y) for x, y in zip(a2, b2)])
    cov = rng2.uniform(-1, 1, size=(64, 9))
    q = np.array([R.covariance_roundtrip(c)[0].reshape(9) for c in cov])
    back = np.array([R.covariance_roundtrip(c)[1] for c in cov])
    np.savez_compressed(os.path.join(HERE, "graph_hpp_reference_build.npz"), a=a2, b=b2, out=out2, cov=cov, q=q, back=back,
                        source="reference src/graph/include/graph.hpp compiled by oracle/Makefile target ref "
                               "(g++ -O2 -ffp-contract=off, stand-in headers oracle/ref_stubs)")
else:
    print("oracle/_ref/libgraph_hpp_ref.so absent: graph_hpp_reference_build.npz not regenerated")

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

# a recorded registration stream (common/Registration messages, graphslam_b200/regstream.py format) and the Keyframes
# message the ORACLE's replay of it ends with (oracle/replay.py): the fixture of the ROS-free replay driver
# (examples/graph_node_noros.cpp, graph.GraphNode) — solve() after every 100 keyframes and at the end
from graphslam_b200 import regstream      # noqa: E402
from oracle import replay as oreplay       # noqa: E402
gm = synth.manhattan(400, seed=5, target_edges=560)
msgs = regstream.from_graph(gm, solve_every=100)
regstream.write(os.path.join(HERE, "registration_stream_manhattan400.reg"), msgs)
ids, kposes, traces = oreplay.replay(regstream.read(os.path.join(HERE, "registration_stream_manhattan400.reg")))
np.savez_compressed(os.path.join(HERE, "registration_stream_manhattan400_keyframes.npz"), ids=ids, poses=kposes,
                    iterations=np.array([t.iterations for t in traces]), final_errors=np.array([t.errors[-1] for t in traces]))
print("registration stream written:", len(msgs), "events,", len(ids), "keyframes, LM iterations per solve",
      [t.iterations for t in traces])
