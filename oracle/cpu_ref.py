"""ctypes view of oracle/cpu_ref.c (C99 + OpenMP port of the path).  TEST INFRASTRUCTURE ONLY:
imported by tests/, smoke() and bench.py's cpu_baseline / --impl reference legs, never by the product."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "libgs_cpu_ref.so")


class CpuParams(C.Structure):
    _fields_ = [("max_iterations", C.c_int), ("rel_error_tol", C.c_double), ("abs_error_tol", C.c_double),
                ("error_tol", C.c_double), ("lambda_initial", C.c_double), ("lambda_factor", C.c_double),
                ("lambda_upper", C.c_double), ("lambda_lower", C.c_double), ("min_model_fidelity", C.c_double),
                ("diagonal_damping", C.c_int), ("jacobian_mode", C.c_int), ("pcg_rel_tol", C.c_double),
                ("pcg_max_iters", C.c_int), ("pcg_iter_cap_total", C.c_int), ("pcg_time_cap_ms", C.c_double),
                ("solver", C.c_int)]

SOLVER_PCG, SOLVER_CHOLESKY = 0, 1


class CpuStats(C.Structure):
    _fields_ = [("iterations", C.c_int), ("outer_loops", C.c_int), ("inner_iterations", C.c_int), ("converged", C.c_int),
                ("chi2_initial", C.c_double), ("chi2_final", C.c_double), ("lambda_final", C.c_double),
                ("pcg_iterations_total", C.c_int64), ("ms_structure", C.c_double), ("ms_linearize", C.c_double),
                ("ms_solve", C.c_double), ("ms_update", C.c_double), ("ms_total", C.c_double), ("threads", C.c_int),
                ("ms_analyze", C.c_double), ("ms_factor", C.c_double), ("chol_nnz", C.c_double), ("chol_flops", C.c_double), ("capped", C.c_int)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


_lib = None
HOST_TAG = os.path.join(HERE, "_build", "host.txt")


def _host_tag() -> str:
    """CPU model + ISA flags of this machine: the port is built -march=native, so a library built on another kind of
    host (the repo snapshot travels to the GPU box with its built .so files) is rebuilt here rather than trusted."""
    model, flags = "", ""
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name") and not model:
                    model = line.split(":", 1)[1].strip()
                elif line.startswith("flags") and not flags:
                    flags = " ".join(sorted(line.split(":", 1)[1].split()))
                if model and flags:
                    break
    except OSError:
        pass
    return model + "\n" + flags + "\n"


def load():
    global _lib
    if _lib is None:
        srcs = [os.path.join(HERE, f) for f in ("cpu_ref.c", "cpu_chol.c", "cpu_chol.h")]
        tag = _host_tag()
        try:
            same_host = open(HOST_TAG).read() == tag
        except OSError:
            same_host = False
        if not os.path.exists(LIB) or os.path.getmtime(LIB) < max(os.path.getmtime(f) for f in srcs) or not same_host:
            subprocess.run(["make", "-s", "-B", "-C", HERE, "_build/libgs_cpu_ref.so"], check=True)
            with open(HOST_TAG, "w") as f:
                f.write(tag)
        _lib = C.CDLL(LIB)
        _lib.gs_cpu_default_params.argtypes = [C.POINTER(CpuParams)]
        _lib.gs_cpu_threads.restype = C.c_int
        _lib.gs_cpu_set_threads.argtypes = [C.c_int]
        dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
        _lib.gs_cpu_optimize.restype = C.c_int
        _lib.gs_cpu_optimize.argtypes = [C.c_int, C.c_int, ip, ip, dp, dp, C.c_int, ip, dp, dp, dp, C.POINTER(CpuParams),
                                         C.POINTER(CpuStats), dp, ip, C.c_int]
    return _lib


def host_cores() -> int:
    """Cores this process may run on (what a timed CPU baseline should use)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def set_threads(n: int | None = None) -> int:
    """Set the OpenMP thread count explicitly (torchrun exports OMP_NUM_THREADS=1 to its children); returns it."""
    lib = load()
    lib.gs_cpu_set_threads(int(n or host_cores()))
    return int(lib.gs_cpu_threads())


def optimize(g, init=None, **overrides):
    """Run the C port on a synth.PoseGraph.  Returns (poses, stats dict, errors, tries)."""
    lib = load()
    prm = CpuParams()
    lib.gs_cpu_default_params(C.byref(prm))
    for k, v in overrides.items():
        if not hasattr(prm, k):
            raise AttributeError(k)
        setattr(prm, k, v)
    keys = np.asarray(g.keys)
    skeys = np.sort(keys)
    ei = np.ascontiguousarray(np.searchsorted(skeys, g.k1), dtype=np.int32)
    ej = np.ascontiguousarray(np.searchsorted(skeys, g.k2), dtype=np.int32)
    z = np.ascontiguousarray(g.z, dtype=np.float64)
    cov = np.ascontiguousarray(g.cov, dtype=np.float64)
    pidx = np.ascontiguousarray(np.searchsorted(skeys, [g.prior_key]), dtype=np.int32)
    ppose = np.ascontiguousarray(g.prior_pose, dtype=np.float64).reshape(1, 3)
    pcov = np.ascontiguousarray(g.prior_cov, dtype=np.float64).reshape(1, 9)
    order = np.argsort(keys, kind="stable")
    x = np.ascontiguousarray((g.init if init is None else init)[order], dtype=np.float64).copy()
    cap = max(prm.max_iterations + 8, 16)
    errs = np.zeros(cap + 1)
    tries = np.zeros(cap, dtype=np.int32)
    st = CpuStats()
    dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
    rc = lib.gs_cpu_optimize(g.n, g.e, ei.ctypes.data_as(ip), ej.ctypes.data_as(ip), z.ctypes.data_as(dp),
                             cov.ctypes.data_as(dp), 1, pidx.ctypes.data_as(ip), ppose.ctypes.data_as(dp),
                             pcov.ctypes.data_as(dp), x.ctypes.data_as(dp), C.byref(prm), C.byref(st), errs.ctypes.data_as(dp),
                             tries.ctypes.data_as(ip), cap)
    d = st.as_dict()
    d["status"] = rc
    m = d["outer_loops"]
    return x, d, errs[:m + 1], tries[:m]
