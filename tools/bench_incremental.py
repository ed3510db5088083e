"""Config C5 (BASELINE.json configs[4]): online incremental mode.

The graph grows from 10,000 to 50,000 poses in 400 batches of 100 keyframes (their odometry and loop-closure factors
arrive with them); after every batch the whole graph is re-optimised, warm-started from the previous optimum exactly
as the reference's solve() does (src/graph/src/graph.cpp:119, 130: a fresh LM per call, `initial = poses_opti`).
New keyframes are initialised by dead reckoning from the last optimised pose (graph.cpp:71-86).

Reported: latency per update (host wall clock around gs_add_batch + gs_optimize + reading the newest pose back,
i.e. what the ROS callback would wait for), p50 / p99 / mean / max, with its split into structure rebuild and solve;
and, on a bounded sample of updates, the same update on the CPU port (oracle/cpu_ref.c, all host threads).

    python tools/bench_incremental.py [--updates 400] [--cpu-updates 3] > profiles/rN_c5_incremental.json
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from graphslam_b200 import graph, synth  # noqa: E402


def dead_reckon(last_pose, z_odo):
    """poses of a batch from the last optimised pose and the batch's odometry measurements (body-frame compose)"""
    out = np.zeros((z_odo.shape[0], 3))
    cur = np.asarray(last_pose, dtype=np.float64)
    for k in range(z_odo.shape[0]):
        cur = synth.compose(cur[None, :], z_odo[k][None, :])[0]
        out[k] = cur
    return out


def run(n0=10_000, batch=100, updates=400, cpu_updates=3, verbose=True):
    """The C5 sequence from n0 poses on; returns the result dict (see the module docstring)."""
    g = synth.config("C5")
    n_final = n0 + batch * updates
    assert n_final <= g.n, "C5 has 50,000 poses"
    k1, k2 = g.k1.astype(np.int64), g.k2.astype(np.int64)
    arrival = np.maximum(k1, k2)
    is_odo = (k2 == k1 + 1)

    G = graph.Graph()
    G.add_graph(g.prefix(n0))
    t0 = time.perf_counter()
    st0 = G.optimize()
    t_first = time.perf_counter() - t0
    if verbose:
        print(f"[c5] initial solve of {n0} poses: {t_first * 1e3:.1f} ms, {st0['iterations']} LM iterations, "
              f"{st0['pcg_iterations_total']} PCG iterations, chi2 {st0['chi2_initial']:.5g} -> {st0['chi2_final']:.5g}",
              file=sys.stderr, flush=True)

    lat, ms_struct, ms_solve, lm_its, pcg_its, chi2, parts = [], [], [], [], [], [], []
    cpu_cases = []
    cpu_at = set(np.linspace(0, updates - 1, cpu_updates).astype(int).tolist()) if cpu_updates else set()
    for u in range(updates):
        lo, hi = n0 + u * batch, n0 + (u + 1) * batch
        m = (arrival > lo) & (arrival <= hi)
        # odometry measurement into each new keyframe (first one in arrival order)
        zo = np.zeros((batch, 3))
        mo = np.flatnonzero(m & is_odo)
        zo[k2[mo] - lo - 1] = g.z[mo]
        keys = g.keys[lo:hi]
        x_before = None
        if u in cpu_at:
            _, x_before = G.poses()
        t0 = time.perf_counter()
        last = G.pose(int(g.keys[lo - 1]))
        poses = dead_reckon(last, zo)
        G.add_batch(keys, poses, g.k1[m], g.k2[m], g.z[m], g.cov[m])
        st = G.optimize()
        newest = G.pose(int(keys[-1]))
        dt = time.perf_counter() - t0
        lat.append(dt * 1e3)
        ms_struct.append(st["ms_structure"])
        parts.append(st["ms_structure_parts"])
        ms_solve.append(st["ms_solve"])
        lm_its.append(st["iterations"])
        pcg_its.append(st["pcg_iterations_total"])
        chi2.append(st["chi2_final"])
        if u in cpu_at:
            cpu_cases.append((hi, x_before.copy(), poses.copy(), lo))
        if verbose and (u + 1) % 50 == 0:
            print(f"[c5] update {u + 1}/{updates}: N={hi} latency {dt * 1e3:.1f} ms ({st['iterations']} LM, "
                  f"{st['pcg_iterations_total']} PCG, structure {st['ms_structure']:.1f} ms) newest pose {newest}",
                  file=sys.stderr, flush=True)
    G.close()

    cpu = None
    if cpu_cases:
        from oracle import cpu_ref   # the checker / CPU baseline, never on the product path
        threads = cpu_ref.set_threads()
        cpu_ref.optimize(synth.config("C3", n=20000), max_iterations=1, pcg_iter_cap_total=20)   # thread pool warm-up
        res = {"cholesky": [], "pcg": []}
        for hi, x_before, poses_new, lo in cpu_cases:
            sub = g.prefix(hi)
            init = np.zeros((hi, 3))           # the same start the GPU update had: previous optimum + dead-reckoned batch
            init[:lo] = x_before[:lo]
            init[lo:hi] = poses_new
            for name, solver in (("cholesky", 1), ("pcg", 0)):
                t0 = time.perf_counter()
                cpu_ref.optimize(sub, init=init, pcg_rel_tol=1e-12, solver=solver)
                res[name].append((time.perf_counter() - t0) * 1e3)
        cpu = dict(kind="port", cores=threads, updates=len(cpu_cases), sizes=[c[0] for c in cpu_cases],
                   latency_ms_cholesky=res["cholesky"], latency_ms_pcg=res["pcg"],
                   mean_ms=float(min(np.mean(res["cholesky"]), np.mean(res["pcg"]))),
                   variant="cholesky" if np.mean(res["cholesky"]) <= np.mean(res["pcg"]) else "pcg",
                   note="full solve() of the same update (structure, ordering and LM to GTSAM's stopping rule) from the "
                        "same start the GPU update had")

    lat = np.array(lat)
    return dict(metric="incremental_update_latency", unit="ms", higher_is_better=False,
                config=dict(workload="C5-incremental: 224-wide lawn-mower SE(2) graph growing %d -> %d poses in %d batches of "
                                     "%d keyframes, full re-optimisation per batch, warm start" %
                                     (n0, n_final, updates, batch)),
                p50=float(np.percentile(lat, 50)), p99=float(np.percentile(lat, 99)), mean=float(lat.mean()),
                max=float(lat.max()), first_solve_ms=t_first * 1e3,
                mean_structure_ms=float(np.mean(ms_struct)),
                mean_structure_parts_ms={k: float(np.mean([p_[k] for p_ in parts])) for k in parts[0]},
                p50_structure_ms=float(np.percentile(ms_struct, 50)), p50_solve_ms=float(np.percentile(ms_solve, 50)),
                mean_solve_ms=float(np.mean(ms_solve)),
                mean_lm_iterations=float(np.mean(lm_its)), mean_pcg_iterations=float(np.mean(pcg_its)),
                final_chi2=chi2[-1], n_final=n_final, cpu_baseline=cpu, dtype="f64", data="synthetic")


def main_from_bench(args):
    """`python bench.py --config C5`: the full C5 sequence (400 updates) as one JSON line."""
    out = run(updates=400, cpu_updates=0 if args.no_cpu_baseline else 3)
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n0", type=int, default=10_000)
    ap.add_argument("--batch", type=int, default=100)
    ap.add_argument("--updates", type=int, default=400)
    ap.add_argument("--cpu-updates", type=int, default=3, help="updates re-run on the CPU port (0 = skip)")
    args = ap.parse_args()
    print(json.dumps(run(args.n0, args.batch, args.updates, args.cpu_updates)))


if __name__ == "__main__":
    main()
