"""CPU-side checks of the product's own host code: the C-ABI library loads and exports every symbol the
header declares, the host structure build is bit-exact against the oracle, the host-evaluated SE(2) inline
functions (the same ones the kernels compile) agree with the oracle, and nothing falls back to the CPU."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from graphslam_b200 import _lib, graph, synth
from oracle import gs_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _cuda_present():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_header_symbols_all_exported(lib, dbg):
    hdr = open(os.path.join(ROOT, "include", "graphslam_cuda.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(gs_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    for name in declared:
        assert hasattr(lib, name), f"{name} not exported"
    assert lib.gs_version() == 1
    # the test-only header too
    dhdr = open(os.path.join(ROOT, "include", "graphslam_cuda_debug.h")).read()
    dhdr = re.sub(r"/\*.*?\*/", "", dhdr, flags=re.S)
    declared_dbg = set(re.findall(r"\b(gs_[a-z0-9_]+)\s*\(", dhdr))
    assert declared_dbg == set(_lib.DEBUG_SIGNATURES), declared_dbg ^ set(_lib.DEBUG_SIGNATURES)
    for name in declared_dbg:
        assert hasattr(dbg, name), f"{name} not exported by the debug library"
    # the product library carries no test scaffolding (emulator, host hooks)
    for name in declared_dbg:
        assert not hasattr(lib, name), f"{name} leaked into libgraphslam_cuda.so"


def test_default_params_are_gtsam_defaults(lib):
    p = graph.default_params()
    assert (p.max_iterations, p.rel_error_tol, p.abs_error_tol, p.error_tol) == (100, 1e-5, 1e-5, 0.0)
    assert (p.lambda_initial, p.lambda_factor, p.lambda_upper, p.lambda_lower) == (1e-5, 10.0, 1e5, 0.0)
    assert p.min_model_fidelity == 1e-3 and p.diagonal_damping == 0 and p.jacobian_mode == 0
    want = O.LMParams()
    for f in ("max_iterations", "rel_error_tol", "abs_error_tol", "lambda_initial", "lambda_factor", "lambda_upper",
              "min_model_fidelity"):
        assert getattr(p, f) == getattr(want, f)


def _host_pattern(lib, n, ei, ej):
    ei = np.ascontiguousarray(ei, np.int32)
    ej = np.ascontiguousarray(ej, np.int32)
    nn = C.c_size_t()
    p = _lib.c_i32_p
    assert lib.gs_host_pattern(n, len(ei), ei.ctypes.data_as(p), ej.ctypes.data_as(p), None, None, C.byref(nn), None) == 0
    rp, ci, perm = np.zeros(n + 1, np.int32), np.zeros(nn.value, np.int32), np.zeros(len(ei), np.int32)
    assert lib.gs_host_pattern(n, len(ei), ei.ctypes.data_as(p), ej.ctypes.data_as(p), rp.ctypes.data_as(p),
                               ci.ctypes.data_as(p), C.byref(nn), perm.ctypes.data_as(p)) == 0
    return rp, ci, perm


@pytest.mark.parametrize("cid,n", [("C1", None), ("C2", 20000), ("C3", 50000), ("C4", 4000)])
def test_structure_bit_exact_vs_oracle(dbg, cid, n):
    g = synth.config(cid, n=n)
    prob = O.Problem.from_graph(g)
    rp, ci, perm = _host_pattern(dbg, g.n, prob.ei, prob.ej)
    orp, oci = prob.pattern()
    assert rp.dtype == orp.dtype == np.int32
    assert np.array_equal(rp, orp) and np.array_equal(ci, oci)
    assert sorted(perm.tolist()) == list(range(g.e))


def test_structure_duplicates_isolated_and_hub(dbg):
    rng = np.random.default_rng(5)
    n = 300
    ei = np.r_[rng.integers(0, n - 10, 900), np.zeros(600, np.int64)]        # vertex 0 is a hub (600 edges)
    ej = np.r_[rng.integers(0, n - 10, 900), rng.integers(1, n - 10, 600)]   # last 10 vertices isolated
    m = ei != ej
    ei, ej = ei[m], ej[m]
    prob = O.Problem(n, ei, ej, np.zeros((len(ei), 3)), np.tile(np.eye(3).reshape(9), (len(ei), 1)), [0],
                     np.zeros((1, 3)), np.eye(3).reshape(1, 9))
    rp, ci, _ = _host_pattern(dbg, n, ei, ej)
    orp, oci = prob.pattern()
    assert np.array_equal(rp, orp) and np.array_equal(ci, oci)
    # canonical order does not depend on the insertion order of the factors
    sh = rng.permutation(len(ei))
    rp2, ci2, _ = _host_pattern(dbg, n, ei[sh], ej[sh])
    assert np.array_equal(rp, rp2) and np.array_equal(ci, ci2)


@pytest.mark.parametrize("mode", [0, 1])
def test_host_between_matches_oracle(dbg, mode):
    g = synth.config("C1", full_cov=True)
    prob = O.Problem.from_graph(g)
    t = prob.edge_terms(g.init, mode)
    rng = np.random.default_rng(1)
    for k in rng.integers(0, g.e, 300):
        outs = [np.zeros(3), np.zeros(9), np.zeros(9), np.zeros(9), np.zeros(9), np.zeros(9), np.zeros(3), np.zeros(3)]
        args = [np.ascontiguousarray(a) for a in (g.init[prob.ei[k]], g.init[prob.ej[k]], g.z[k], g.cov[k])]
        st = dbg.gs_host_between(*[graph._dp(a) for a in args], mode, *[graph._dp(o) for o in outs])
        assert st == 0
        ref = [t["e"][k], t["a_i"][k].ravel(), t["a_j"][k].ravel(), t["h_ii"][k].ravel(), t["h_ij"][k].ravel(),
               t["h_jj"][k].ravel(), t["b_i"][k], t["b_j"][k]]
        for got, want in zip(outs, ref):
            assert np.allclose(got, want, rtol=1e-12, atol=1e-12 * (1 + np.abs(want).max()))


def test_compose_helpers_bitwise(lib):
    tab = np.load(os.path.join(ROOT, "tests", "golden", "compose_reference.npz"))
    for a, b, want in zip(tab["a"], tab["b"], tab["out"]):
        assert np.array_equal(graph.compose_reference(a, b), want)       # graph.hpp:30-43 bit-for-bit
        assert np.allclose(graph.compose(a, b), O.compose(a, b), atol=1e-15)
    # the dropped translation and the one-sided wrap
    assert np.array_equal(graph.compose_reference([10, -4, 0], [1, 2, 0.5])[:2], [1.0, 2.0])
    assert graph.compose_reference([0, 0, -3.0], [0, 0, -0.5])[2] == pytest.approx(-3.5)


@pytest.mark.skipif(_cuda_present(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback(lib):
    """Without a device the product must fail loudly, not compute on the CPU."""
    with pytest.raises(graph.GraphSlamError) as ei:
        graph.Graph()
    assert ei.value.status == _lib.GS_ECUDA and "no CPU fallback" in str(ei.value)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "graphslam_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"(from|import)\s+oracle|oracle[/.]\w|libgs_oracle", src), f"{f} reaches into oracle/"


def test_header_is_plain_c99(tmp_path):
    """The boundary is a C ABI: include/graphslam_cuda.h (and the debug header) must compile as strict C99."""
    src = tmp_path / "abi.c"
    src.write_text('#include "graphslam_cuda.h"\n#include "graphslam_cuda_debug.h"\n'
                   "int main(void) { gs_params p; gs_default_params(&p); return (int)GS_ENOMEM - 9 + p.max_iterations * 0; }\n")
    inc = os.path.join(ROOT, "include")
    res = subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-fsyntax-only", "-I", inc, str(src)],
                         capture_output=True, text=True)
    assert res.returncode == 0, res.stderr


def test_cpp_shim_and_ros_free_example_build_against_the_library(lib, tmp_path):
    """include/graphslam_cuda.hpp (the drop-in for graph.cpp's verbs) and examples/graph_node_noros.cpp compile and
    link against the built .so — what a maintainer of the reference node would do."""
    from graphslam_b200 import build
    exe = tmp_path / "graph_node_noros"
    res = subprocess.run(["g++", "-std=c++17", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                          os.path.join(ROOT, "examples", "graph_node_noros.cpp"), "-o", str(exe), build.LIB,
                          "-Wl,-rpath," + os.path.dirname(build.LIB)], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    # without a CUDA device the example must fail loudly (no CPU fallback), not crash
    stream = os.path.join(ROOT, "tests", "golden", "registration_stream_manhattan400.reg")
    run = subprocess.run([str(exe), stream], capture_output=True, text=True)
    if not _cuda_present():
        assert run.returncode != 0 and run.returncode > 0, (run.returncode, run.stderr)
        assert "CUDA" in (run.stderr + run.stdout)


@pytest.mark.parametrize("cid,n", [("C1", None), ("C5", 6000), ("C3", 40000), ("C4", 12000)])
def test_coarse_space_aggregation_invariants(dbg, cid, n):
    """The aggregates of the two-level preconditioner (structure.hpp: build_aggregation; from 8,192 rows the eight fixed
    segments run on eight host threads): every row in exactly one aggregate of at most `target` rows, aggregates numbered
    segment by segment and never across a segment boundary (so they belong to one rank for any number of ranks), a valid
    distance-2 colouring of the aggregate graph (what lets one product H w serve all aggregates of a colour), and the same
    answer on every call."""
    g = synth.config(cid, n=n)
    prob = O.Problem.from_graph(g)
    ei = np.ascontiguousarray(prob.ei, np.int32)
    ej = np.ascontiguousarray(prob.ej, np.int32)
    p = _lib.c_i32_p

    def run():
        counts = np.zeros(3, np.int32)
        agg_of, seg = np.zeros(g.n, np.int32), np.zeros(9, np.int32)
        assert dbg.gs_host_aggregation(g.n, len(ei), ei.ctypes.data_as(p), ej.ctypes.data_as(p), 0, counts.ctypes.data_as(p),
                                       agg_of.ctypes.data_as(p), None, seg.ctypes.data_as(p)) == 0
        color = np.zeros(counts[0], np.int32)
        assert dbg.gs_host_aggregation(g.n, len(ei), ei.ctypes.data_as(p), ej.ctypes.data_as(p), 0, counts.ctypes.data_as(p),
                                       None, color.ctypes.data_as(p), None) == 0
        return counts, agg_of, color, seg

    counts, agg_of, color, seg = run()
    n_agg, n_colors, target = (int(v) for v in counts)
    assert target == max(48, int(np.ceil(np.sqrt(g.n))))
    sizes = np.bincount(agg_of, minlength=n_agg)
    assert agg_of.min() == 0 and agg_of.max() == n_agg - 1 and sizes.min() >= 1 and sizes.max() <= target
    # segments: fixed multiples of 256 rows, aggregates numbered segment by segment
    tiles = (g.n + 255) // 256
    seg_row = np.minimum(g.n, (tiles * np.arange(9) // 8) * 256)
    assert seg[0] == 0 and seg[8] == n_agg and np.all(np.diff(seg) >= 0)
    for s in range(8):
        rows = agg_of[seg_row[s]:seg_row[s + 1]]
        if len(rows):
            assert rows.min() == seg[s] and rows.max() == seg[s + 1] - 1
    # distance-2 colouring: two aggregates of one colour share no neighbour (closed neighbourhoods)
    a, b = agg_of[prob.ei], agg_of[prob.ej]
    nb = [set([k]) for k in range(n_agg)]
    for x, y in zip(a.tolist(), b.tolist()):
        nb[x].add(y)
        nb[y].add(x)
    assert color.min() == 0 and color.max() == n_colors - 1
    for k in range(n_agg):
        seen = {}
        for m in nb[k]:
            assert seen.setdefault(int(color[m]), m) == m, "two aggregates of one colour next to aggregate %d" % k
    again = run()
    assert np.array_equal(again[1], agg_of) and np.array_equal(again[2], color)
