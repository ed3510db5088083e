"""gs_closest_keyframe_noalloc (include/graphslam_cuda.h): the closest_keyframe query with its scratch in a fixed
file-scope device array, for callers that cannot call cudaMalloc.  No counterpart in the reference.
Collected last on purpose (file name).

Checked through the C ABI against the existing entry point (gs_closest_keyframe, reference graph.cpp:146-178) on two
calls in a row — the shared array keeps the last call's contents, so the second call is the one that would show a
stale counter or partial — with different candidate counts, exact ties included.  The results must be identical
(same kernel, same reduction tree): equality is the tolerance the closest_keyframe tests use."""
import numpy as np
import pytest

from graphslam_b200 import _lib, graph

pytestmark = pytest.mark.gpu


def _graph(n, seed):
    rng = np.random.default_rng(seed)
    xy = rng.uniform(-300, 300, (n, 2))
    xy[[11, n // 2]] = xy[-1] + [3.0, 4.0]            # two candidates at exactly the same distance: lowest index wins
    poses = np.c_[xy, rng.uniform(-3, 3, n)]
    G = graph.Graph()
    G.add_batch(np.arange(1, n + 1, dtype=np.uint64), poses, np.zeros(0, np.uint64), np.zeros(0, np.uint64),
                np.zeros((0, 3)), np.zeros((0, 9)))
    G.add_prior(1, poses[0], np.eye(3).reshape(9))
    return G, poses


def test_noalloc_entry_point_matches_the_allocating_one_twice_in_a_row():
    n = 50_000                                         # 25 CTAs of 2048 candidates
    G, poses = _graph(n, 5)
    with G:
        with pytest.raises(graph.GraphSlamError) as e:   # not prepared yet: it must refuse, not allocate
            G.closest_keyframe_noalloc(n, 10)
        assert e.value.status == _lib.GS_EINVAL and "allocates nothing" in str(e.value)
        G.prepare()
        queries = [(n, 10), (n, 30_000), (123, 0), (n, n - 1)]     # many CTAs, fewer CTAs, self, count == 1
        for rnd in range(2):                           # two calls in a row per query, and the whole list twice
            for key, skip in queries:
                want = G.closest_keyframe(key, skip)
                got1 = G.closest_keyframe_noalloc(key, skip)
                got2 = G.closest_keyframe_noalloc(key, skip)
                assert got1 == want and got2 == want, (rnd, key, skip, want, got1, got2)
        # against numpy too (graph.cpp:149-165: strict '<', so the lowest index among equal distances)
        d = np.sqrt(((poses[:n - 10, :2] - poses[-1, :2]) ** 2).sum(1))
        key, dist = G.closest_keyframe_noalloc(n, 10)
        assert key == int(np.argmin(d)) + 1 and dist == pytest.approx(float(d.min()), rel=1e-14)   # (the device contracts to FMA)
        # too few keyframes: the service's "false"
        with pytest.raises(graph.GraphSlamError) as e:
            G.closest_keyframe_noalloc(n, n)
        assert e.value.status == _lib.GS_ENOKEY


def test_noalloc_entry_point_after_an_optimisation_and_across_graphs():
    """The array is shared by every graph on the GPU: calls on two graphs, one after the other, do not disturb each
    other; and it works on optimised values (the usual caller: a loop-closure search after solve())."""
    from graphslam_b200 import synth
    g = synth.config("C3", n=9000)
    with graph.Graph() as A:
        A.add_graph(g)
        A.optimize()
        B, _ = _graph(4000, 6)
        with B:
            B.prepare()
            for _ in range(2):
                assert A.closest_keyframe_noalloc(int(g.keys[-1]), 10) == A.closest_keyframe(int(g.keys[-1]), 10)
                assert B.closest_keyframe_noalloc(4000, 10) == B.closest_keyframe(4000, 10)
