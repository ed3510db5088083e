"""Python view of the C ABI (tests and bench drive the library through this; the C++ drop-in for
graph.cpp is include/graphslam_cuda.hpp).  Method names follow the reference's verbs:

  reference (src/graph/src/graph.cpp)              here
  initial.insert(id, Pose2)            :58,86       Graph.add_vertex
  graph.add(PriorFactor<Pose2>)        :57          Graph.add_prior
  graph.add(BetweenFactor<Pose2>)      :87-92,106   Graph.add_between
  LevenbergMarquardtOptimizer(...).optimize() :119  Graph.optimize
  poses_opti.at<Pose2>(id)             :123-125     Graph.pose / Graph.poses
  prior_factor / new_factor / loop_factor / solve   GraphNode.* (same call order as registration_callback :180-200)
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import GsParams, GsStats, c_double_p, c_i32_p, c_u64_p


class GraphSlamError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"{_lib.STATUS_NAMES[status] if 0 <= status < len(_lib.STATUS_NAMES) else status}: {msg}")
        self.status = status


def _dp(a):
    return a.ctypes.data_as(c_double_p)


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        a = a.reshape(shape)
    return a


def default_params(**overrides) -> GsParams:
    p = GsParams()
    _lib.load().gs_default_params(C.byref(p))
    for k, v in overrides.items():
        if not hasattr(p, k):
            raise AttributeError(f"gs_params has no field {k!r}")
        setattr(p, k, v)
    return p


class Graph:
    """One gs_graph.  Requires a CUDA device: there is no CPU fallback."""

    def __init__(self, params: GsParams | None = None, nccl_id: bytes | None = None, **overrides):
        """`nccl_id` (from `nccl_unique_id()`, the same bytes on every rank) + rank / world_size / device overrides
        create one rank of a multi-GPU graph (gs_create_distributed); every rank then makes the same calls."""
        self._lib = _lib.load()
        self._h = C.c_void_p()
        self.params = params if params is not None else default_params(**overrides)
        if params is not None:
            for k, v in overrides.items():
                setattr(self.params, k, v)
        if nccl_id is not None or self.params.world_size != 1:
            buf = C.create_string_buffer(bytes(nccl_id or b""), len(nccl_id or b""))
            st = self._lib.gs_create_distributed(C.byref(self._h), C.byref(self.params), buf, len(nccl_id or b""))
        else:
            st = self._lib.gs_create(C.byref(self._h), C.byref(self.params))
        if st != _lib.GS_OK:
            msg = self._lib.gs_last_error(self._h).decode() if self._h else "gs_create failed"
            if self._h:
                self._lib.gs_destroy(self._h)
                self._h = C.c_void_p()
            raise GraphSlamError(st, msg)

    # -- lifecycle
    def close(self):
        if getattr(self, "_h", None):
            self._lib.gs_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _ck(self, st, ok=(0,)):
        if st not in ok:
            raise GraphSlamError(st, self._lib.gs_last_error(self._h).decode())
        return st

    def set_params(self, **overrides):
        for k, v in overrides.items():
            if not hasattr(self.params, k):
                raise AttributeError(f"gs_params has no field {k!r}")
            setattr(self.params, k, v)
        self._ck(self._lib.gs_set_params(self._h, C.byref(self.params)))

    # -- building
    def add_vertex(self, key: int, pose):
        p = _f64(pose, 3)
        self._ck(self._lib.gs_add_vertex(self._h, int(key), _dp(p)))

    def add_prior(self, key: int, pose, cov9):
        p, q = _f64(pose, 3), _f64(cov9, 9)
        self._ck(self._lib.gs_add_prior(self._h, int(key), _dp(p), _dp(q)))

    def add_between(self, k1: int, k2: int, delta, cov9):
        d, q = _f64(delta, 3), _f64(cov9, 9)
        self._ck(self._lib.gs_add_between(self._h, int(k1), int(k2), _dp(d), _dp(q)))

    def add_batch(self, keys, poses, k1, k2, deltas, cov9s):
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        poses = _f64(poses, (-1, 3))
        k1 = np.ascontiguousarray(k1, dtype=np.uint64)
        k2 = np.ascontiguousarray(k2, dtype=np.uint64)
        deltas = _f64(deltas, (-1, 3))
        cov9s = _f64(cov9s, (-1, 9))
        assert poses.shape[0] == keys.shape[0] and deltas.shape[0] == k1.shape[0] == k2.shape[0] == cov9s.shape[0]
        self._ck(self._lib.gs_add_batch(self._h, keys.shape[0], keys.ctypes.data_as(c_u64_p), _dp(poses), k1.shape[0],
                                        k1.ctypes.data_as(c_u64_p), k2.ctypes.data_as(c_u64_p), _dp(deltas), _dp(cov9s)))

    def add_graph(self, g):
        """Load a synth.PoseGraph: vertices, its prior, then the factors in arrival order."""
        self.add_batch(g.keys, g.init, np.zeros(0, np.uint64), np.zeros(0, np.uint64), np.zeros((0, 3)), np.zeros((0, 9)))
        self.add_prior(g.prior_key, g.prior_pose, g.prior_cov)
        self.add_batch(np.zeros(0, np.uint64), np.zeros((0, 3)), g.k1, g.k2, g.z, g.cov)
        return self

    def add_prior_information(self, key: int, pose, info6):
        p, q = _f64(pose, 3), _f64(info6, 6)
        self._ck(self._lib.gs_add_prior_information(self._h, int(key), _dp(p), _dp(q)))

    def add_between_information(self, k1: int, k2: int, delta, info6):
        d, q = _f64(delta, 3), _f64(info6, 6)
        self._ck(self._lib.gs_add_between_information(self._h, int(k1), int(k2), _dp(d), _dp(q)))

    def load_g2o(self, path) -> dict:
        """Append a g2o / TORO 2-D file (all-or-nothing); returns the counts."""
        c = _lib.GsG2oCounts()
        self._ck(self._lib.gs_load_g2o(self._h, str(path).encode(), C.byref(c)))
        return {k: int(getattr(c, k)) for k, _ in c._fields_}

    def save_g2o(self, path):
        """Write the whole graph with the current values (%.17g: save -> load is bit-exact)."""
        self._ck(self._lib.gs_save_g2o(self._h, str(path).encode()))

    def prepare(self):
        self._ck(self._lib.gs_prepare(self._h))

    def structure_timing(self) -> dict:
        """ms of the last structure build: host structure (K6), staging + upload, communicator setup, graph capture."""
        ms = np.zeros(4)
        self._ck(self._lib.gs_get_structure_timing(self._h, _dp(ms)))
        return dict(zip(("host", "upload", "comm", "capture"), ms.tolist()))

    # -- solving
    def optimize(self, allow_not_converged=True) -> dict:
        st = GsStats()
        rc = self._lib.gs_optimize(self._h, C.byref(st))
        self._ck(rc, ok=(0, _lib.GS_ENOTCONVERGED) if allow_not_converged else (0,))
        d = st.as_dict()
        d["status"] = rc
        return d

    def chi2(self) -> float:
        v = C.c_double()
        self._ck(self._lib.gs_chi2(self._h, C.byref(v)))
        return v.value

    # -- reading
    @property
    def num_vertices(self) -> int:
        n = C.c_size_t()
        self._ck(self._lib.gs_num_vertices(self._h, C.byref(n)))
        return n.value

    @property
    def num_factors(self) -> int:
        n = C.c_size_t()
        self._ck(self._lib.gs_num_factors(self._h, C.byref(n)))
        return n.value

    def transfer_bytes(self):
        """(host->device, device->host) bytes the library has copied so far for this graph."""
        a, b = C.c_double(), C.c_double()
        self._ck(self._lib.gs_get_transfer_bytes(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def marginal_covariances(self, keys) -> np.ndarray:
        """(n, 3, 3) marginal covariances of the given poses at the current values (gtsam::Marginals)."""
        keys = np.ascontiguousarray(np.atleast_1d(keys), dtype=np.uint64)
        out = np.zeros((keys.shape[0], 3, 3))
        self._ck(self._lib.gs_marginal_covariances(self._h, keys.shape[0], keys.ctypes.data_as(c_u64_p), _dp(out)))
        return out

    def pose(self, key: int) -> np.ndarray:
        out = np.zeros(3)
        self._ck(self._lib.gs_get_pose(self._h, int(key), _dp(out)))
        return out

    def poses(self, keys_out=None, xyt_out=None):
        """(keys, xyt) of all vertices in sorted-key order.  `keys_out` (uint64[n]) / `xyt_out` (float64[n,3]) are
        caller-owned buffers to fill (a caller that reads the poses every step should not fault in 32 MB of fresh pages
        each time); new arrays otherwise."""
        n = self.num_vertices
        keys = keys_out if keys_out is not None else np.zeros(n, dtype=np.uint64)
        xyt = xyt_out if xyt_out is not None else np.zeros((n, 3))
        assert keys.dtype == np.uint64 and keys.shape == (n,) and keys.flags.c_contiguous
        assert xyt.dtype == np.float64 and xyt.shape == (n, 3) and xyt.flags.c_contiguous
        m = C.c_size_t()
        self._ck(self._lib.gs_get_poses(self._h, n, keys.ctypes.data_as(c_u64_p), _dp(xyt), C.byref(m)))
        return keys, xyt

    def set_poses(self, keys, xyt):
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        xyt = _f64(xyt, (-1, 3))
        self._ck(self._lib.gs_set_poses(self._h, keys.shape[0], keys.ctypes.data_as(c_u64_p), _dp(xyt)))

    # -- rank-local transfers (several GPUs: a rank moves its own rows only; see include/graphslam_cuda.h)
    def local_range(self):
        """(first, count): the rows of the sorted-key order this rank owns (all of them on one GPU)."""
        first, count = C.c_size_t(), C.c_size_t()
        self._ck(self._lib.gs_local_range(self._h, C.byref(first), C.byref(count)))
        return first.value, count.value

    def local_poses(self, keys_out=None, xyt_out=None, want_keys=True):
        """(keys, xyt) of the owned rows; not collective.  Buffers as in poses(); want_keys=False skips the keys (a caller
        that reads the poses every step knows them from its first read) and returns (None, xyt)."""
        _, n = self.local_range()
        keys = None
        if want_keys:
            keys = keys_out if keys_out is not None else np.zeros(n, dtype=np.uint64)
            assert keys.dtype == np.uint64 and keys.shape == (n,) and keys.flags.c_contiguous
        xyt = xyt_out if xyt_out is not None else np.zeros((n, 3))
        assert xyt.dtype == np.float64 and xyt.shape == (n, 3) and xyt.flags.c_contiguous
        m = C.c_size_t()
        self._ck(self._lib.gs_get_local_poses(self._h, n, keys.ctypes.data_as(c_u64_p) if want_keys else None, _dp(xyt),
                                              C.byref(m)))
        return keys, xyt

    def set_local_poses(self, xyt):
        """New values for the owned rows (collective on several GPUs: the ghost rows are exchanged on the device)."""
        xyt = _f64(xyt, (-1, 3))
        self._ck(self._lib.gs_set_local_poses(self._h, xyt.shape[0], _dp(xyt)))

    # -- parity hooks
    def pattern(self):
        nnzb = C.c_size_t()
        self._ck(self._lib.gs_get_pattern(self._h, None, None, C.byref(nnzb)))
        n = self.num_vertices
        rowptr = np.zeros(n + 1, dtype=np.int32)
        colidx = np.zeros(nnzb.value, dtype=np.int32)
        self._ck(self._lib.gs_get_pattern(self._h, rowptr.ctypes.data_as(c_i32_p), colidx.ctypes.data_as(c_i32_p),
                                          C.byref(nnzb)))
        return rowptr, colidx

    def system(self):
        """(rowptr, colidx, blocks[nnzb,3,3], b[N,3], chi2) linearised at the current values."""
        rowptr, colidx = self.pattern()
        blocks = np.zeros((colidx.shape[0], 3, 3))
        b = np.zeros((self.num_vertices, 3))
        chi2 = C.c_double()
        self._ck(self._lib.gs_get_system(self._h, _dp(blocks), _dp(b), C.byref(chi2)))
        return rowptr, colidx, blocks, b, chi2.value

    def solve_linear(self, lam: float):
        d = np.zeros((self.num_vertices, 3))
        it = C.c_int()
        rr = C.c_double()
        self._ck(self._lib.gs_solve_linear(self._h, float(lam), _dp(d), C.byref(it), C.byref(rr)))
        return d, it.value, rr.value

    def trace(self):
        n = C.c_size_t()
        self._ck(self._lib.gs_get_trace(self._h, 0, None, None, None, C.byref(n)))
        m = n.value
        errors = np.zeros(m + 1)
        lambdas = np.zeros(max(m, 1))
        tries = np.zeros(max(m, 1), dtype=np.int32)
        self._ck(self._lib.gs_get_trace(self._h, m, _dp(errors), _dp(lambdas), tries.ctypes.data_as(c_i32_p), C.byref(n)))
        return errors, lambdas[:m], tries[:m]

    def profile_pcg(self, lam: float = 1e-5, iters: int = 50):
        """Average CUDA-event duration (ms) and algorithmic bytes of the three PCG kernels."""
        ms, by = np.zeros(3), np.zeros(3)
        self._ck(self._lib.gs_profile_pcg(self._h, float(lam), int(iters), _dp(ms), _dp(by)))
        return dict(zip(("spmv", "pcg_update", "pcg_dir"), zip(ms.tolist(), by.tolist())))

    def profile_outer(self, reps: int = 5):
        ms, by = np.zeros(3), np.zeros(3)
        self._ck(self._lib.gs_profile_outer(self._h, int(reps), _dp(ms), _dp(by)))
        return dict(zip(("assemble", "retract", "chi2"), zip(ms.tolist(), by.tolist())))

    def closest_keyframe(self, key: int, skip: int = 10):
        out = C.c_uint64()
        dist = C.c_double()
        self._ck(self._lib.gs_closest_keyframe(self._h, int(key), int(skip), C.byref(out), C.byref(dist)))
        return out.value, dist.value

    def closest_keyframe_noalloc(self, key: int, skip: int = 10):
        """gs_closest_keyframe_noalloc: the same query with its scratch in a fixed device array (allocates nothing)."""
        out = C.c_uint64()
        dist = C.c_double()
        self._ck(self._lib.gs_closest_keyframe_noalloc(self._h, int(key), int(skip), C.byref(out), C.byref(dist)))
        return out.value, dist.value


def nccl_unique_id() -> bytes:
    """ncclUniqueId bytes for gs_create_distributed; call on ONE rank and hand the bytes to all."""
    lib = _lib.load()
    n = C.c_size_t()
    lib.gs_nccl_unique_id(None, 0, C.byref(n))
    buf = C.create_string_buffer(n.value)
    st = lib.gs_nccl_unique_id(buf, n.value, C.byref(n))
    if st != _lib.GS_OK:
        raise GraphSlamError(st, "gs_nccl_unique_id failed (libnccl.so.2 not loadable?)")
    return buf.raw


def distributed_graph(**overrides) -> "Graph":
    """One rank of a multi-GPU graph inside an initialised torch.distributed job (torchrun): rank 0 draws the
    NCCL id, torch.distributed broadcasts it, each rank binds to cuda:LOCAL_RANK."""
    import os

    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device("cuda", local)
    raw = nccl_unique_id() if rank == 0 else bytes(128)
    t = torch.tensor(list(raw), dtype=torch.uint8, device=dev if dist.get_backend() == "nccl" else "cpu")
    dist.broadcast(t, 0)
    return Graph(nccl_id=bytes(t.cpu().tolist()), rank=rank, world_size=world, device=local, **overrides)


def compose_reference(a, b) -> np.ndarray:
    """graph.hpp:30-43 bit-for-bit (host)."""
    a, b, out = _f64(a, 3), _f64(b, 3), np.zeros(3)
    _lib.load().gs_compose_reference(_dp(a), _dp(b), _dp(out))
    return out


def compose(a, b) -> np.ndarray:
    a, b, out = _f64(a, 3), _f64(b, 3), np.zeros(3)
    _lib.load().gs_compose(_dp(a), _dp(b), _dp(out))
    return out


class GraphNode:
    """The reference node's bookkeeping (graph.cpp:27-132) over the library, ROS stripped.

    `registration` dicts carry what common/Registration.msg carries that the graph node reads:
    first_frame_flag, keyframe_flag, loop_closure_flag, factor_new {id_1, delta(x,y,theta), covariance[9]},
    factor_loop {id_1, id_2, delta, covariance}.  `faithful_compose=True` seeds new vertices with
    the reference's own `compose` (translation dropped, graph.hpp:37-38); default is the correct one.
    """

    sigma_xy_prior = 0.1          # graph.cpp:13
    sigma_th_prior = 0.1          # graph.cpp:14
    keyframes_to_skip_in_loop_closing = 10   # graph.cpp:15

    def __init__(self, faithful_compose: bool = False, **params):
        self.graph = Graph(**params)
        self.keyframe_ids = 0     # graph.cpp:207
        self.keyframes: list[dict] = []
        self._compose = compose_reference if faithful_compose else compose

    def prior_factor(self, registration=None):
        self.keyframe_ids += 1                                    # graph.cpp:31
        q = np.diag([self.sigma_xy_prior ** 2, self.sigma_xy_prior ** 2, self.sigma_th_prior ** 2]).reshape(9)
        pose = np.zeros(3)
        self.keyframes.append(dict(id=self.keyframe_ids, pose_opti=pose.copy()))
        self.graph.add_vertex(self.keyframe_ids, pose)            # graph.cpp:58
        self.graph.add_prior(self.keyframe_ids, pose, q)          # graph.cpp:57

    def new_factor(self, registration):
        self.keyframe_ids += 1                                    # graph.cpp:68
        f = registration["factor_new"]
        # graph.cpp:72 composes from input.keyframe_last.pose_opti — what the scanner obtained from the last_keyframe
        # service (graph.cpp:134-144) before publishing; a message without it falls back to the node's own last keyframe
        last = registration.get("keyframe_last") or self.keyframes[-1]
        pose_new = self._compose(last["pose_opti"], f["delta"])   # graph.cpp:72
        self.keyframes.append(dict(id=self.keyframe_ids, pose_opti=pose_new.copy()))
        self.graph.add_vertex(self.keyframe_ids, pose_new)        # graph.cpp:86
        self.graph.add_between(f["id_1"], self.keyframe_ids, f["delta"], f["covariance"])   # graph.cpp:87-92

    def loop_factor(self, registration):
        f = registration["factor_loop"]
        self.graph.add_between(f["id_1"], f["id_2"], f["delta"], f["covariance"])           # graph.cpp:106-111

    def solve(self) -> dict:
        stats = self.graph.optimize()                             # graph.cpp:119
        keys, xyt = self.graph.poses()
        by_key = dict(zip(keys.tolist(), xyt))
        for kf in self.keyframes:                                 # graph.cpp:122-125
            kf["pose_opti"] = by_key[kf["id"]].copy()
        return stats                                              # warm start kept in the library, graph.cpp:130

    def registration_callback(self, registration, solve: bool = False):
        """graph.cpp:180-200; `solve` enables the call the reference has commented out (:195)."""
        if registration.get("first_frame_flag"):
            self.prior_factor(registration)
        elif registration.get("keyframe_flag"):
            self.new_factor(registration)
            if registration.get("loop_closure_flag"):
                self.loop_factor(registration)
            if solve:
                self.solve()

    def last_keyframe(self):
        return self.keyframes[-1] if self.keyframes else None     # graph.cpp:134-144

    def publish_keyframes(self):
        """common/Keyframes as the node publishes it (graph.cpp:18-25): (ids, poses) of every keyframe."""
        return (np.array([kf["id"] for kf in self.keyframes], dtype=np.uint64),
                np.array([kf["pose_opti"] for kf in self.keyframes], dtype=np.float64).reshape(-1, 3))

    def closest_keyframe(self, key: int):
        """graph.cpp:146-178 on the device copy of the values."""
        return self.graph.closest_keyframe(key, self.keyframes_to_skip_in_loop_closing)
