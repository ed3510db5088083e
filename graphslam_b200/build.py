"""In-tree build of libgraphslam_cuda.so (nvcc, sm_100a only) — no JIT cache, the .so travels with the repo."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libgraphslam_cuda.so")
DEBUG_LIB = os.path.join(LIBDIR, "libgraphslam_cuda_debug.so")   # test-only: CPU emulator + host hooks, never shipped

NVCC_FLAGS = [
    "-std=c++17", "-O3", "-lineinfo",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "--fmad=true",                       # FP64 FMA contraction on; never --use_fast_math
    "-Xcompiler", "-fPIC,-O3,-Wall,-Wno-unused-function,-Wno-unknown-pragmas",
    "-shared", "-cudart", "static",
    "-lcusolver", "-lcublas",            # dense Cholesky (potrf) and inverse (trsm + syrk) of the two-level preconditioner's
                                         # coarse matrix (setup only)
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libgraphslam_cuda cannot be built (there is no CPU fallback)")


def sources(debug: bool = False):
    d = os.path.join(CSRC, "debug") if debug else CSRC
    return [os.path.join(d, f) for f in sorted(os.listdir(d)) if f.endswith(".cu")]


def deps(debug: bool = False):
    out = sources(debug)
    out += [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".hpp", ".h"))]
    out.append(os.path.join(ROOT, "include", "graphslam_cuda.h"))
    out.append(os.path.join(ROOT, "include", "graphslam_cuda_debug.h"))
    return out


def needs_build(lib: str = LIB, debug: bool = False) -> bool:
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    return any(os.path.getmtime(d) > t for d in deps(debug))


def _compile(lib: str, srcs, extra, verbose: bool) -> str:
    os.makedirs(os.path.dirname(lib), exist_ok=True)
    cmd = [_nvcc(), *NVCC_FLAGS, *extra, "-I", os.path.join(ROOT, "include"), "-o", lib, *srcs]
    if verbose:
        print(" ".join(cmd), flush=True)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError(f"nvcc failed building {os.path.basename(lib)}")
    if verbose and (res.stdout or res.stderr):
        print(res.stdout + res.stderr)
    return lib


def build(force: bool = False, verbose: bool = False, extra=()) -> str:
    """The product library and the test-only debug library (emulator + host hooks), each when out of date."""
    if force or needs_build(LIB):
        _compile(LIB, sources(), extra, verbose)
    if force or needs_build(DEBUG_LIB, debug=True):
        _compile(DEBUG_LIB, sources(debug=True), (), verbose)
    return LIB


def build_variant(name: str, defines, verbose: bool = False, debug: bool = False) -> str:
    """A compile-time variant of the product library (or, debug=True, of the test-only debug library) for A/B
    measurements (tools/variant_probe.py): same sources with -D switches (GS_H_LAYOUT, GS_F_LAYOUT, ...), written to
    lib/variants/libgraphslam_cuda[_debug]_<name>.so."""
    lib = os.path.join(LIBDIR, "variants", f"libgraphslam_cuda_{'debug_' if debug else ''}{name}.so")
    if needs_build(lib, debug):
        _compile(lib, sources(debug), [f"-D{d}" for d in defines], verbose)
    return lib


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True,
          extra=("-Xptxas", "-v") if "--ptxas" in sys.argv else ())
    print(LIB)
