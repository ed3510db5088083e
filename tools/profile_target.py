"""Small, launch-bounded target for ncu: build one config, run a few PCG iterations and one outer pass.
  python tools/profile_target.py [C3] [n] [pcg_iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graphslam_b200 import graph, synth  # noqa: E402

cid = sys.argv[1] if len(sys.argv) > 1 else "C3"
n = int(sys.argv[2]) if len(sys.argv) > 2 and int(sys.argv[2]) > 0 else None
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 8
g = synth.config(cid, n=n)
with graph.Graph() as G:
    G.add_graph(g)
    G.prepare()
    print("outer", G.profile_outer(2))
    print("pcg", G.profile_pcg(1e-5, iters))
