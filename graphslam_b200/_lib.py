"""ctypes binding of libgraphslam_cuda.so (include/graphslam_cuda.h).  Fails loudly if the library
is missing: there is no Python / CPU fallback for the optimiser."""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

c_double_p = C.POINTER(C.c_double)
c_u64_p = C.POINTER(C.c_uint64)
c_i32_p = C.POINTER(C.c_int32)

(GS_SOLVER_NO_CUDA_GRAPH, GS_SOLVER_NCCL_PER_OP, GS_SOLVER_NO_PERSISTENT, GS_SOLVER_DIR_KERNEL, GS_SOLVER_DIR_ON_THE_FLY,
 GS_SOLVER_BLOCK_JACOBI_ONLY) = (1, 2, 4, 8, 16, 32)
GS_OK, GS_EINVAL, GS_ENOTSPD, GS_EDUPKEY, GS_ENOKEY, GS_ECUDA, GS_ENCCL, GS_ENOTCONVERGED, GS_ENAN, GS_ENOMEM = range(10)
STATUS_NAMES = ["GS_OK", "GS_EINVAL", "GS_ENOTSPD", "GS_EDUPKEY", "GS_ENOKEY", "GS_ECUDA", "GS_ENCCL",
                "GS_ENOTCONVERGED", "GS_ENAN", "GS_ENOMEM"]


class GsParams(C.Structure):
    _fields_ = [("max_iterations", C.c_int), ("rel_error_tol", C.c_double), ("abs_error_tol", C.c_double),
                ("error_tol", C.c_double), ("lambda_initial", C.c_double), ("lambda_factor", C.c_double),
                ("lambda_upper", C.c_double), ("lambda_lower", C.c_double), ("min_model_fidelity", C.c_double),
                ("diagonal_damping", C.c_int), ("jacobian_mode", C.c_int), ("pcg_rel_tol", C.c_double),
                ("pcg_max_iters", C.c_int), ("device", C.c_int), ("rank", C.c_int), ("world_size", C.c_int),
                ("solver_flags", C.c_int), ("coarse_aggregate_size", C.c_int), ("peer_wait_limit_ms", C.c_int),
                ("reserved", C.c_int * 5)]


class GsStats(C.Structure):
    _fields_ = [("iterations", C.c_int), ("outer_loops", C.c_int), ("inner_iterations", C.c_int),
                ("converged", C.c_int), ("chi2_initial", C.c_double), ("chi2_final", C.c_double),
                ("lambda_final", C.c_double), ("pcg_iterations_total", C.c_int64), ("ms_structure", C.c_double),
                ("ms_linearize", C.c_double), ("ms_solve", C.c_double), ("ms_update", C.c_double),
                ("ms_total", C.c_double), ("edges_per_sec", C.c_double), ("gn_iters_per_sec", C.c_double),
                ("gpu_launches", C.c_double), ("ms_device", C.c_double), ("h2d_bytes", C.c_double),
                ("d2h_bytes", C.c_double), ("p2p_wait_ticks", C.c_double * 3), ("p2p_waits", C.c_double),
                ("ms_structure_parts", C.c_double * 4), ("ms_coarse_setup", C.c_double)]

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_ if k not in ("ms_structure_parts", "p2p_wait_ticks")}
        d["p2p_wait_ticks"] = list(self.p2p_wait_ticks)
        d["ms_structure_parts"] = dict(zip(("host", "upload", "comm", "capture"), self.ms_structure_parts))
        d["gpu_launches"] = int(self.gpu_launches)
        return d


class GsG2oCounts(C.Structure):
    _fields_ = [("vertices", C.c_size_t), ("edges", C.c_size_t), ("priors", C.c_size_t), ("fixed", C.c_size_t),
                ("skipped_lines", C.c_size_t)]


# every symbol include/graphslam_cuda.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "gs_default_params": (None, [C.POINTER(GsParams)]),
    "gs_version": (C.c_int, []),
    "gs_create": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(GsParams)]),
    "gs_nccl_unique_id": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "gs_create_distributed": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(GsParams), C.c_void_p, C.c_size_t]),
    "gs_destroy": (C.c_int, [C.c_void_p]),
    "gs_set_params": (C.c_int, [C.c_void_p, C.POINTER(GsParams)]),
    "gs_add_vertex": (C.c_int, [C.c_void_p, C.c_uint64, c_double_p]),
    "gs_add_prior": (C.c_int, [C.c_void_p, C.c_uint64, c_double_p, c_double_p]),
    "gs_add_between": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint64, c_double_p, c_double_p]),
    "gs_add_batch": (C.c_int, [C.c_void_p, C.c_size_t, c_u64_p, c_double_p, C.c_size_t, c_u64_p, c_u64_p,
                               c_double_p, c_double_p]),
    "gs_add_prior_information": (C.c_int, [C.c_void_p, C.c_uint64, c_double_p, c_double_p]),
    "gs_add_between_information": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint64, c_double_p, c_double_p]),
    "gs_load_g2o": (C.c_int, [C.c_void_p, C.c_char_p, C.POINTER(GsG2oCounts)]),
    "gs_save_g2o": (C.c_int, [C.c_void_p, C.c_char_p]),
    "gs_prepare": (C.c_int, [C.c_void_p]),
    "gs_get_structure_timing": (C.c_int, [C.c_void_p, c_double_p]),
    "gs_optimize": (C.c_int, [C.c_void_p, C.POINTER(GsStats)]),
    "gs_get_pose": (C.c_int, [C.c_void_p, C.c_uint64, c_double_p]),
    "gs_get_poses": (C.c_int, [C.c_void_p, C.c_size_t, c_u64_p, c_double_p, C.POINTER(C.c_size_t)]),
    "gs_set_poses": (C.c_int, [C.c_void_p, C.c_size_t, c_u64_p, c_double_p]),
    "gs_local_range": (C.c_int, [C.c_void_p, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "gs_get_local_poses": (C.c_int, [C.c_void_p, C.c_size_t, c_u64_p, c_double_p, C.POINTER(C.c_size_t)]),
    "gs_set_local_poses": (C.c_int, [C.c_void_p, C.c_size_t, c_double_p]),
    "gs_chi2": (C.c_int, [C.c_void_p, c_double_p]),
    "gs_num_vertices": (C.c_int, [C.c_void_p, C.POINTER(C.c_size_t)]),
    "gs_get_transfer_bytes": (C.c_int, [C.c_void_p, c_double_p, c_double_p]),
    "gs_num_factors": (C.c_int, [C.c_void_p, C.POINTER(C.c_size_t)]),
    "gs_get_pattern": (C.c_int, [C.c_void_p, c_i32_p, c_i32_p, C.POINTER(C.c_size_t)]),
    "gs_get_system": (C.c_int, [C.c_void_p, c_double_p, c_double_p, c_double_p]),
    "gs_solve_linear": (C.c_int, [C.c_void_p, C.c_double, c_double_p, C.POINTER(C.c_int), c_double_p]),
    "gs_marginal_covariance": (C.c_int, [C.c_void_p, C.c_uint64, c_double_p]),
    "gs_marginal_covariances": (C.c_int, [C.c_void_p, C.c_size_t, c_u64_p, c_double_p]),
    "gs_get_trace": (C.c_int, [C.c_void_p, C.c_size_t, c_double_p, c_double_p, c_i32_p, C.POINTER(C.c_size_t)]),
    "gs_profile_pcg": (C.c_int, [C.c_void_p, C.c_double, C.c_int, c_double_p, c_double_p]),
    "gs_profile_outer": (C.c_int, [C.c_void_p, C.c_int, c_double_p, c_double_p]),
    "gs_last_error": (C.c_char_p, [C.c_void_p]),
    "gs_compose_reference": (None, [c_double_p, c_double_p, c_double_p]),
    "gs_compose": (None, [c_double_p, c_double_p, c_double_p]),
    "gs_closest_keyframe": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, c_u64_p, c_double_p]),
    "gs_closest_keyframe_noalloc": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, c_u64_p, c_double_p]),
}

class GsEmuProblem(C.Structure):
    _fields_ = [("n", C.c_int32), ("e", C.c_int32), ("np", C.c_int32), ("ei", c_i32_p), ("ej", c_i32_p), ("z", c_double_p),
                ("om", c_double_p), ("prior_idx", c_i32_p), ("prior_xyt", c_double_p), ("prior_om", c_double_p),
                ("xyt", c_double_p)]


class GsEmuResult(C.Structure):
    _fields_ = [("blocks", c_double_p), ("b", c_double_p), ("delta", c_double_p), ("xyt_new", c_double_p),
                ("chi2", C.c_double), ("chi2_check", C.c_double), ("chi2_new", C.c_double), ("dHd", C.c_double),
                ("db", C.c_double), ("rel_residual", C.c_double), ("pcg_iters", C.c_int32), ("pcg_done", C.c_int32),
                ("n_ghost_total", C.c_int32), ("n_pairs_total", C.c_int32), ("halo_send_total", C.c_int32),
                ("n_factor_copies", C.c_int32), ("n_block_copies", C.c_int32)]


# test-only entries (include/graphslam_cuda_debug.h): the product's own host code, callable on a CPU box
DEBUG_SIGNATURES = {
    "gs_host_partition": (C.c_int, [C.c_int32, C.c_int32, c_i32_p, c_i32_p, C.c_int, C.c_int, c_i32_p, c_i32_p, c_i32_p,
                                    c_i32_p]),
    "gs_emu_run": (C.c_int, [C.POINTER(GsEmuProblem), C.c_int, C.c_int, C.c_double, C.c_int, C.c_double,
                             C.POINTER(GsEmuResult)]),
    "gs_emu_run_form": (C.c_int, [C.POINTER(GsEmuProblem), C.c_int, C.c_int, C.c_double, C.c_int, C.c_double, C.c_int,
                                  C.POINTER(GsEmuResult)]),
    "gs_host_aggregation": (C.c_int, [C.c_int32, C.c_int32, c_i32_p, c_i32_p, C.c_int32, c_i32_p, c_i32_p, c_i32_p, c_i32_p]),
    "gs_host_pattern": (C.c_int, [C.c_int32, C.c_int32, c_i32_p, c_i32_p, c_i32_p, c_i32_p, C.POINTER(C.c_size_t),
                                  c_i32_p]),
    "gs_host_g2o": (C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(C.c_size_t), c_u64_p, c_double_p, c_u64_p, c_u64_p,
                              c_double_p, c_double_p, c_u64_p, c_double_p, c_double_p]),
    "gs_host_between": (C.c_int, [c_double_p] * 4 + [C.c_int] + [c_double_p] * 8),
}

_LIB = None
_DEBUG_LIB = None


def lib_path() -> str:
    """The in-tree product library; GS_LIB (read here, in the Python test/bench harness, not by the library) points
    the harness at a compile-time variant built by build.build_variant for A/B measurements."""
    return os.environ.get("GS_LIB") or _build.LIB


def _bind(path: str, signatures) -> C.CDLL:
    lib = C.CDLL(path)
    for name, (res, args) in signatures.items():
        fn = getattr(lib, name)          # AttributeError = missing export = broken build
        fn.restype = res
        fn.argtypes = args
    return lib


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load libgraphslam_cuda.so; raise if it cannot be found or built."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        if not build_if_missing:
            raise RuntimeError(f"{path} is missing: run `python -m graphslam_b200.build` (no CPU fallback exists)")
        _build.build()
    _LIB = _bind(path, SIGNATURES)
    return _LIB


def load_debug(build_if_missing: bool = True) -> C.CDLL:
    """Load libgraphslam_cuda_debug.so (tests only: CPU emulator + host hooks of include/graphslam_cuda_debug.h)."""
    global _DEBUG_LIB
    if _DEBUG_LIB is not None:
        return _DEBUG_LIB
    path = os.environ.get("GS_DEBUG_LIB") or _build.DEBUG_LIB
    if not os.path.exists(path):
        if not build_if_missing:
            raise RuntimeError(f"{path} is missing: run `python -m graphslam_b200.build`")
        _build.build()
    _DEBUG_LIB = _bind(path, DEBUG_SIGNATURES)
    return _DEBUG_LIB
