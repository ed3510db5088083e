"""ctypes view of oracle/_ref/libgraph_hpp_ref.so — the reference's own graph.hpp (compose, covariance_to_eigen,
eigen_to_covariance; /root/reference/src/graph/include/graph.hpp:30-68) compiled from where it lies by
`make -C oracle ref`.  TEST INFRASTRUCTURE ONLY: the one piece of the path whose reference code runs here; it pins
gs_compose_reference, the oracle's compose_reference and the committed golden table bit for bit.  The .so is built in
this container (where /root/reference exists) and travels to the GPU box; when it is absent `available()` is False and
the tests fall back to the golden table it generated."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libgraph_hpp_ref.so")
_lib = None


def available() -> bool:
    return os.path.exists(LIB)


def _load():
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB)
        dp = C.POINTER(C.c_double)
        _lib.ref_compose.argtypes = [dp, dp, dp]
        _lib.ref_compose.restype = None
        _lib.ref_covariance_roundtrip.argtypes = [dp, dp, dp]
        _lib.ref_covariance_roundtrip.restype = None
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def compose(a, b) -> np.ndarray:
    """compose(input_1, input_2) of graph.hpp:30-43, as compiled."""
    a = np.ascontiguousarray(a, dtype=np.float64).reshape(3)
    b = np.ascontiguousarray(b, dtype=np.float64).reshape(3)
    out = np.zeros(3)
    _load().ref_compose(_dp(a), _dp(b), _dp(out))
    return out


def covariance_roundtrip(cov9):
    """(Q row-major as covariance_to_eigen fills it, covariance[9] as eigen_to_covariance writes it back)."""
    c = np.ascontiguousarray(cov9, dtype=np.float64).reshape(9)
    q, back = np.zeros(9), np.zeros(9)
    _load().ref_covariance_roundtrip(_dp(c), _dp(q), _dp(back))
    return q.reshape(3, 3), back
