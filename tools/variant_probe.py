"""Short GPU probe of kernel variants: compile-time variants of the library (storage layouts, built by
graphslam_b200.build.build_variant into lib/variants/) x run-time execution paths (gs_params.solver_flags).  For each combination, on one config:

  * CUDA-event time of every kernel inside the real sequence (gs_profile_pcg / gs_profile_outer),
  * wall time per PCG iteration of a capped solve replayed from the captured CUDA graph (how gs_optimize runs it),
  * SHA-1 of the solve's delta: every variant must reproduce the default path's bits,

and the same bit check on a small ragged graph (odd number of rows in the last tile, persistent kernel off).
Each library variant runs in its own child process (the ctypes handle is per process; GS_LIB points the Python
harness at the variant).  Results are appended line by line to gpurun_out/variant_probe.jsonl, so a cut-off run
still leaves what it measured.  No torch import: the library is driven through ctypes only.

  python tools/variant_probe.py build                       # here (no GPU): compile the variant libraries
  python tools/variant_probe.py run [config] [n_poses] [capped_iters] [libs: a,b,..] [flagsets: x,y,..]
"""
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

LIBS = {                       # name -> -D switches ([] = the in-tree default library); decided probes are deleted, their
    "default": [],             # records stay under profiles/ (r2_variant_probe2.jsonl, r2_chi2_pipeline_probe.jsonl)
    "h0f0": ["GS_H_LAYOUT=0", "GS_F_LAYOUT=0"],   # round 1's plane layouts
    "chi2_0": ["GS_CHI2_PIPE=0"],                 # k_chi2 with one slot after the other
}
NP = 4
FLAGSETS = {                   # name -> solver_flags (persistent kernel off everywhere: the probe is about the big-graph path);
    "two_level": NP,           # bits are compared within one flag set: the two preconditioners do not share bits
    "block_jacobi": NP | 32,
}
OUT = os.path.join(ROOT, "gpurun_out", "variant_probe.jsonl")


def emit(rec):
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    with open(OUT, "a") as f:
        f.write(json.dumps(rec) + "\n")
    print(json.dumps(rec), flush=True)


def lib_file(name):
    from graphslam_b200 import build
    return build.LIB if not LIBS[name] else os.path.join(build.LIBDIR, "variants", f"libgraphslam_cuda_{name}.so")


def capped_solve(G, lam):
    """gs_solve_linear with the iteration cap hit on purpose: (delta, iterations, seconds)."""
    from graphslam_b200 import _lib
    lib = _lib.load()
    d = np.zeros((G.num_vertices, 3))
    it, rr = C.c_int(), C.c_double()
    t0 = time.perf_counter()
    st = lib.gs_solve_linear(G._h, float(lam), d.ctypes.data_as(_lib.c_double_p), C.byref(it), C.byref(rr))
    dt = time.perf_counter() - t0
    if st not in (_lib.GS_OK, _lib.GS_ENOTCONVERGED):
        raise RuntimeError(f"gs_solve_linear: status {st}: {lib.gs_last_error(G._h).decode()}")
    return d, it.value, dt


def run(g, libname, fname, iters, profile=True):
    from graphslam_b200 import graph
    rec = {"lib": libname, "flags": fname, "graph": g.name, "n": int(g.n)}
    with graph.Graph(pcg_max_iters=iters, pcg_rel_tol=1e-300, solver_flags=FLAGSETS[fname]) as G:
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
        _, _, blocks, b, chi2 = G.system() if g.n <= 50000 else (0, 0, np.zeros(1), np.zeros(1), 0.0)
        rec["sha1_system"] = hashlib.sha1(blocks.tobytes() + b.tobytes()).hexdigest()
    emit(rec)
    return rec


def child(cid, n, iters, libname, fnames):
    from graphslam_b200 import synth
    t0 = time.time()
    small = synth.config("C3", n=4001)              # 15 full tiles + 161 rows
    for f in fnames:
        run(small, libname, f, 96, profile=False)
    g = synth.config(cid, n=n)
    emit({"lib": libname, "generated": g.name, "n": int(g.n), "e": int(g.e), "s": round(time.time() - t0, 1)})
    for f in fnames:
        run(g, libname, f, iters)


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "run"
    if mode == "build":
        from graphslam_b200 import build
        build.build()
        for name, defs in LIBS.items():
            if defs:
                print(build.build_variant(name, defs), flush=True)
        return
    if mode == "child":
        cid, n, iters, libname = sys.argv[2], int(sys.argv[3]) or None, int(sys.argv[4]), sys.argv[5]
        child(cid, n, iters, libname, sys.argv[6].split(","))
        return
    cid = sys.argv[2] if len(sys.argv) > 2 else "C3"
    n = sys.argv[3] if len(sys.argv) > 3 else "0"
    iters = sys.argv[4] if len(sys.argv) > 4 else "320"
    libs = sys.argv[5].split(",") if len(sys.argv) > 5 else list(LIBS)
    fsets = sys.argv[6] if len(sys.argv) > 6 else ",".join(FLAGSETS)
    for name in libs:
        path = lib_file(name)
        if not os.path.exists(path):
            emit({"lib": name, "error": f"{path} missing: run `python tools/variant_probe.py build` first"})
            continue
        env = dict(os.environ, GS_LIB=path)
        res = subprocess.run([sys.executable, os.path.abspath(__file__), "child", cid, n, iters, name, fsets], env=env,
                             capture_output=True, text=True, timeout=600)
        if res.returncode != 0:
            emit({"lib": name, "error": res.stderr[-1500:]})
    # bit checks over everything recorded for this config in this run
    recs = [json.loads(l) for l in open(OUT)] if os.path.exists(OUT) else []
    for key in sorted({(r["graph"], r["n"], r["flags"]) for r in recs if "sha1" in r}):
        rs = [r for r in recs if (r.get("graph"), r.get("n"), r.get("flags")) == key and "sha1" in r]
        emit({"check": f"bits on {key[0]} n={key[1]} {key[2]}", "variants": len(rs), "solve_equal": len({r["sha1"] for r in rs}) == 1,
              "system_equal": len({r["sha1_system"] for r in rs}) == 1})


if __name__ == "__main__":
    main()
