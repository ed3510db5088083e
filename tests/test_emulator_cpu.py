"""CPU checks of the device path's host logic through the test-only emulator (gs_emu_run,
include/graphslam_cuda_debug.h): the structure build, the symmetric block storage, the vertex-range partition and
its halo lists, and the per-row functions the CUDA kernels call (compiled for the host), against the oracle —
and the design property that results do not depend on the number of ranks (bitwise)."""
import ctypes as C

import numpy as np
import pytest

from graphslam_b200 import _lib, synth
from oracle import gs_oracle as O


@pytest.fixture(scope="module")
def lib(dbg):
    return dbg   # every entry point used here lives in the test-only debug library


def _upper6(om):
    return np.ascontiguousarray(np.stack([om[:, 0, 0], om[:, 0, 1], om[:, 0, 2], om[:, 1, 1], om[:, 1, 2], om[:, 2, 2]], 1))


def emu_run(lib, prob, xyt, world=1, mode=0, lam=1e-5, max_iters=5000, tol=1e-12, want_blocks=True, form=0):
    n, e = prob.n, len(prob.ei)
    ei = np.ascontiguousarray(prob.ei, np.int32)
    ej = np.ascontiguousarray(prob.ej, np.int32)
    z = np.ascontiguousarray(prob.z, np.float64)
    om = _upper6(prob.omega) if e else np.zeros((0, 6))
    pidx = np.ascontiguousarray(prob.prior_idx, np.int32)
    pxyt = np.ascontiguousarray(prob.prior_pose, np.float64)
    pom = _upper6(prob.prior_omega) if len(pidx) else np.zeros((0, 6))
    xyt = np.ascontiguousarray(xyt, np.float64)
    dp, ip = _lib.c_double_p, _lib.c_i32_p
    pb = _lib.GsEmuProblem(n, e, len(pidx), ei.ctypes.data_as(ip), ej.ctypes.data_as(ip), z.ctypes.data_as(dp),
                           om.ctypes.data_as(dp), pidx.ctypes.data_as(ip), pxyt.ctypes.data_as(dp), pom.ctypes.data_as(dp),
                           xyt.ctypes.data_as(dp))
    rowptr, colidx = prob.pattern()
    blocks = np.zeros((len(colidx), 3, 3)) if want_blocks else None
    b, delta, xnew = np.zeros((n, 3)), np.zeros((n, 3)), np.zeros((n, 3))
    res = _lib.GsEmuResult()
    res.blocks = blocks.ctypes.data_as(dp) if want_blocks else None
    res.b, res.delta, res.xyt_new = b.ctypes.data_as(dp), delta.ctypes.data_as(dp), xnew.ctypes.data_as(dp)
    rc = lib.gs_emu_run_form(C.byref(pb), world, mode, lam, max_iters, tol, form, C.byref(res))
    assert rc == 0, rc
    return dict(blocks=blocks, b=b, delta=delta, xyt_new=xnew, chi2=res.chi2, chi2_check=res.chi2_check,
                chi2_new=res.chi2_new, dHd=res.dHd, db=res.db, rel_residual=res.rel_residual, pcg_iters=res.pcg_iters,
                pcg_done=res.pcg_done, n_ghost=res.n_ghost_total, n_pairs=res.n_pairs_total, halo=res.halo_send_total,
                fcopies=res.n_factor_copies, bcopies=res.n_block_copies)


def _check_against_oracle(lib, g, mode=0, world=1, lam=1e-3):
    prob = O.Problem.from_graph(g)
    out = emu_run(lib, prob, g.init, world=world, mode=mode, lam=lam)
    h, ob, ochi2 = prob.linearize(g.init, mode)
    _, _, oblocks = prob.bsr(h)
    assert np.abs(out["blocks"] - oblocks).max() <= 1e-11 * np.abs(oblocks).max()
    assert np.abs(out["b"] - ob).max() <= 1e-11 * np.abs(ob).max()
    assert out["chi2"] == pytest.approx(ochi2, rel=1e-12)
    assert out["chi2_check"] == pytest.approx(prob.chi2(g.init), rel=1e-12)
    want = O.solve_damped(h, ob, lam)
    assert out["pcg_done"] == 1 and out["rel_residual"] <= 1e-12
    assert np.abs(out["delta"] - want).max() <= 1e-7 * max(1.0, np.abs(want).max())
    xn = O.retract(np.asarray(g.init, float), out["delta"])
    d = np.abs(out["xyt_new"][:, :2] - xn[:, :2]).max()
    assert max(d, np.abs(O.wrap_pi(out["xyt_new"][:, 2] - xn[:, 2])).max()) <= 1e-12
    assert out["chi2_new"] == pytest.approx(prob.chi2(out["xyt_new"]), rel=1e-11)
    dd = out["delta"].ravel()
    assert out["dHd"] == pytest.approx(float(dd @ (h @ dd)), rel=1e-9)
    assert out["db"] == pytest.approx(float(dd @ ob.ravel()), rel=1e-9)
    return out


@pytest.mark.parametrize("cid,n,kw,mode", [("C1", None, {}, 0), ("C3", 3000, dict(full_cov=True), 0), ("C4", 1500, {}, 1),
                                           ("C2", 4000, {}, 0)])
def test_emulated_path_matches_oracle(lib, cid, n, kw, mode):
    _check_against_oracle(lib, synth.config(cid, n=n, **kw), mode=mode)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_emulated_multi_rank_matches_oracle_and_is_bitwise_rank_invariant(lib, world):
    g = synth.config("C3", n=5000, full_cov=True)
    one = _check_against_oracle(lib, g, world=1)
    many = _check_against_oracle(lib, g, world=world)
    assert many["halo"] > 0 and many["n_ghost"] == many["halo"]        # send lists and ghost lists pair up
    assert many["n_pairs"] > one["n_pairs"]                              # cut pairs are stored by both owners
    for k in ("blocks", "b", "delta", "xyt_new"):
        assert np.array_equal(one[k], many[k]), k                        # not one bit depends on the partition
    for k in ("chi2", "chi2_new", "dHd", "db", "pcg_iters"):
        assert one[k] == many[k], k


def test_share_or_copy_decision(lib):
    """Regular graphs share every block/factor between its two rows (symmetric storage, one DRAM read);
    irregular ones keep copies where the slice's gather would not coalesce.  Either way the numbers are the same."""
    reg = _check_against_oracle(lib, synth.config("C3", n=4000))
    assert reg["bcopies"] == 0 and reg["fcopies"] == 0
    irr = _check_against_oracle(lib, synth.config("C4", n=3000))
    assert irr["bcopies"] > 0 and irr["fcopies"] > 0


def test_emulated_hub_duplicates_and_tiny_graphs(lib):
    rng = np.random.default_rng(9)
    n = 700
    base = synth.manhattan(n, seed=77, target_edges=1000)
    i, j = base.dense_edges()
    hub_j = rng.integers(1, n - 5, 500)
    dup = np.full(60, 7)
    ei = np.r_[i, np.zeros(500, np.int64), dup, np.full(40, 9), np.full(30, 650)]
    ej = np.r_[j, hub_j, np.full(60, 3), np.full(40, 7), rng.integers(0, 200, 30)]    # long-range pairs cross ranks
    ne = len(ei)
    z = synth.between(base.gt[ei], base.gt[ej]) + 0.01 * rng.normal(size=(ne, 3))
    cov = np.tile(np.diag([0.01, 0.02, 0.005]).reshape(9), (ne, 1))
    g = synth.PoseGraph("hub", base.keys, base.init, base.gt, (ei + 1).astype(np.uint64), (ej + 1).astype(np.uint64), z, cov)
    one = _check_against_oracle(lib, g, world=1)
    for world in (2, 4):
        many = _check_against_oracle(lib, g, world=world)
        assert np.array_equal(one["delta"], many["delta"]) and one["chi2_new"] == many["chi2_new"]
    # fewer tiles than segments: some ranks own no rows at all
    small = synth.manhattan(300, seed=3, target_edges=420)
    a = _check_against_oracle(lib, small, world=1)
    b = _check_against_oracle(lib, small, world=8)
    assert np.array_equal(a["delta"], b["delta"]) and a["chi2"] == b["chi2"]


@pytest.mark.parametrize("world", [1, 2, 8])
def test_direction_formed_on_the_fly_is_bitwise_the_same(lib, world):
    """The two-kernel peer-memory iteration and the persistent small-graph kernel never store p before the product:
    p(c) = z(c) + beta p_prev(c) is formed per touched column and z is what ranks exchange.  Same rounding, same bits."""
    g = synth.config("C3", n=3000, full_cov=True)
    prob = O.Problem.from_graph(g)
    a = emu_run(lib, prob, g.init, world=world, want_blocks=False, form=0)
    b = emu_run(lib, prob, g.init, world=world, want_blocks=False, form=1)
    assert a["pcg_iters"] == b["pcg_iters"] and a["pcg_done"] == b["pcg_done"] == 1
    assert np.array_equal(a["delta"], b["delta"]) and np.array_equal(a["xyt_new"], b["xyt_new"])
    assert a["chi2_new"] == b["chi2_new"] and a["dHd"] == b["dHd"]


@pytest.mark.parametrize("cid,n,mode", [("C3", 3000, 0), ("C2", 5000, 0), ("C3", 1500, 1)])
def test_assemble_without_run_accumulator_is_bitwise_the_same(lib, cid, n, mode):
    """k_assemble_nodup's row code (no run accumulator: a block is stored as soon as it is formed) on graphs where
    every vertex pair carries one factor: H, b, chi2 and the PCG solution equal the general form bit for bit; and the
    form refuses a graph that has a duplicated factor."""
    g = synth.config(cid, n=n)
    prob = O.Problem.from_graph(g)
    a = emu_run(lib, prob, g.init, mode=mode, form=0)
    b = emu_run(lib, prob, g.init, mode=mode, form=2)
    for k in a:
        if isinstance(a[k], np.ndarray):
            assert np.array_equal(a[k], b[k]), k
        else:
            assert a[k] == b[k], k
    # a graph with a duplicated factor is refused (GS_EINVAL = 1): it needs the run accumulator
    dup = synth.config("C3", n=600)
    for name in ("k1", "k2"):
        setattr(dup, name, np.concatenate([getattr(dup, name), getattr(dup, name)[-1:]]))
    dup.z = np.concatenate([dup.z, dup.z[-1:]])
    dup.cov = np.concatenate([dup.cov, dup.cov[-1:]])
    with pytest.raises(AssertionError):
        emu_run(lib, O.Problem.from_graph(dup), dup.init, form=2)


@pytest.mark.parametrize("cid,n", [("C3", 20000), ("C4", 18000)])
def test_structure_passes_on_host_threads(lib, cid, n):
    """From 16,384 owned rows the per-row / per-slice passes of build_local_structure run on several host threads
    (structure.hpp: for_parts): the assembled system must still be the oracle's, shared cells and copies included."""
    g = synth.config(cid, n=n)
    prob = O.Problem.from_graph(g)
    out = emu_run(lib, prob, g.init, max_iters=2)
    h, ob, ochi2 = prob.linearize(g.init, 0)
    _, _, oblocks = prob.bsr(h)
    assert np.abs(out["blocks"] - oblocks).max() <= 1e-11 * np.abs(oblocks).max()
    assert np.abs(out["b"] - ob).max() <= 1e-11 * np.abs(ob).max()
    assert out["chi2"] == pytest.approx(ochi2, rel=1e-12)
    assert out["chi2_check"] == pytest.approx(prob.chi2(g.init), rel=1e-12)
    if cid == "C4":
        assert out["bcopies"] > 0 and out["fcopies"] > 0
