"""Short GPU probe of switchable kernel variants (today: GS_UPDATE_STAGED, see csrc/kernels.cuh; the variants that
lost — evict-first / no-L1 loads of H, alternating sweep direction, evict-first H stores and cp.async slot prefetch
in assembly — are in git history (commits 2bad73e, 2a75c8b) and their numbers in profiles/r1_v7_variant_probe.jsonl).
For each combination, on one config (default C3):

  * CUDA-event time of every kernel inside the real sequence (gs_profile_pcg / gs_profile_outer),
  * wall time per PCG iteration of a capped solve replayed from the captured CUDA graph (how gs_optimize runs it),
  * SHA-1 of the solve's delta: every variant must reproduce the default path's bits,

and the same bit check on a small ragged graph (odd number of rows in the last tile, three-kernel path forced).
Results are appended line by line to gpurun_out/variant_probe.jsonl, so a cut-off run still leaves what it
measured.  No torch import: the library is driven through ctypes only.

  python tools/variant_probe.py [n_poses] [capped_iters] [only: comma-separated variant names] [config]
"""
import ctypes as C
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from graphslam_b200 import _lib, graph, synth  # noqa: E402

FLAGS = ("GS_UPDATE_STAGED", "GS_ASM_NODUP")
VARIANTS = [
    ("default", {}),
    ("unstaged", {"GS_UPDATE_STAGED": "0"}),
    ("asm_nodup", {"GS_ASM_NODUP": "1"}),       # round-2 candidate: k_assemble_nodup, three CTAs per SM
    ("default again", {}),
]
OUT = os.path.join(ROOT, "gpurun_out", "variant_probe.jsonl")


def emit(rec):
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    with open(OUT, "a") as f:
        f.write(json.dumps(rec) + "\n")
    print(json.dumps(rec), flush=True)


def set_env(env):
    for k in FLAGS:
        os.environ.pop(k, None)
    os.environ.update(env)


def capped_solve(G, lam):
    """gs_solve_linear with the iteration cap hit on purpose: (delta, iterations, seconds)."""
    lib = _lib.load()
    d = np.zeros((G.num_vertices, 3))
    it, rr = C.c_int(), C.c_double()
    t0 = time.perf_counter()
    st = lib.gs_solve_linear(G._h, float(lam), d.ctypes.data_as(_lib.c_double_p), C.byref(it), C.byref(rr))
    dt = time.perf_counter() - t0
    if st not in (_lib.GS_OK, _lib.GS_ENOTCONVERGED):
        raise RuntimeError(f"gs_solve_linear: status {st}: {lib.gs_last_error(G._h).decode()}")
    return d, it.value, dt


def run(g, name, env, iters, extra_env=None, profile=True):
    set_env(env)
    os.environ.update(extra_env or {})
    rec = {"variant": name, "n": int(g.n), "env": env}
    with graph.Graph(pcg_max_iters=iters, pcg_rel_tol=1e-300) as G:
        G.add_graph(g)
        G.prepare()
        if profile:
            prof = G.profile_pcg(1e-5, 40)
            outer = G.profile_outer(3)
            rec["us"] = {k: round(v[0] * 1e3, 2) for k, v in {**prof, **outer}.items()}
            rec["us"]["pcg_iteration"] = round(sum(v[0] for v in prof.values()) * 1e3, 2)
        capped_solve(G, 1e-5)                       # warm-up: assemble, graph capture
        d, it, dt = capped_solve(G, 1e-5)
        rec["capped_solve"] = {"iterations": it, "us_per_iteration_wall": round(dt / max(it, 1) * 1e6, 2)}
        rec["sha1"] = hashlib.sha1(d.tobytes()).hexdigest()
    for k in (extra_env or {}):
        os.environ.pop(k, None)
    emit(rec)
    return rec


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 and int(sys.argv[1]) > 0 else None
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 640
    only = sys.argv[3].split(",") if len(sys.argv) > 3 else None
    variants = [v for v in VARIANTS if only is None or v[0] in only]
    t0 = time.time()
    # small ragged graph first (cheap): 4001 rows = 15 full tiles + 161 rows; three-kernel path forced
    gs_ = synth.config("C3", n=4001)
    small = [run(gs_, name, env, 96, {"GS_PCG_FUSED": "0"}, profile=False) for name, env in variants if name != "default again"]
    emit({"check": "small ragged graph bits", "equal": len({r["sha1"] for r in small}) == 1})
    cid = sys.argv[4] if len(sys.argv) > 4 else "C3"
    g = synth.config(cid, n=n)
    emit({"generated": g.name, "n": int(g.n), "e": int(g.e), "s": round(time.time() - t0, 1)})
    big = [run(g, name, env, iters) for name, env in variants]
    emit({"check": "C3 bits", "equal": len({r["sha1"] for r in big}) == 1, "total_s": round(time.time() - t0, 1)})


if __name__ == "__main__":
    main()
