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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n0", type=int, default=10_000)
    ap.add_argument("--batch", type=int, default=100)
    ap.add_argument("--updates", type=int, default=400)
    ap.add_argument("--cpu-updates", type=int, default=3, help="updates re-run on the CPU port (0 = skip)")
    args = ap.parse_args()

    g = synth.config("C5")
    n_final = args.n0 + args.batch * args.updates
    assert n_final <= g.n, "C5 has 50,000 poses"
    k1, k2 = g.k1.astype(np.int64), g.k2.astype(np.int64)
    arrival = np.maximum(k1, k2)
    is_odo = (k2 == k1 + 1)

    G = graph.Graph()
    G.add_graph(g.prefix(args.n0))
    t0 = time.perf_counter()
    st0 = G.optimize()
    t_first = time.perf_counter() - t0
    print(f"[c5] initial solve of {args.n0} poses: {t_first * 1e3:.1f} ms, {st0['iterations']} LM iterations, "
          f"{st0['pcg_iterations_total']} PCG iterations, chi2 {st0['chi2_initial']:.5g} -> {st0['chi2_final']:.5g}",
          file=sys.stderr, flush=True)

    lat, ms_struct, ms_solve, lm_its, pcg_its, chi2 = [], [], [], [], [], []
    cpu_cases = []
    cpu_at = set(np.linspace(0, args.updates - 1, args.cpu_updates).astype(int).tolist()) if args.cpu_updates else set()
    for u in range(args.updates):
        lo, hi = args.n0 + u * args.batch, args.n0 + (u + 1) * args.batch
        m = (arrival > lo) & (arrival <= hi)
        # odometry measurement into each new keyframe (first one in arrival order)
        zo = np.zeros((args.batch, 3))
        mo = np.flatnonzero(m & is_odo)
        zo[k2[mo] - lo - 1] = g.z[mo]
        keys = g.keys[lo:hi]
        t0 = time.perf_counter()
        last = G.pose(int(g.keys[lo - 1]))
        poses = dead_reckon(last, zo)
        G.add_batch(keys, poses, g.k1[m], g.k2[m], g.z[m], g.cov[m])
        st = G.optimize()
        newest = G.pose(int(keys[-1]))
        dt = time.perf_counter() - t0
        lat.append(dt * 1e3)
        ms_struct.append(st["ms_structure"])
        ms_solve.append(st["ms_solve"])
        lm_its.append(st["iterations"])
        pcg_its.append(st["pcg_iterations_total"])
        chi2.append(st["chi2_final"])
        if u in cpu_at:
            _, x_now = G.poses()
            cpu_cases.append((hi, x_now.copy(), poses.copy(), lo))
        if (u + 1) % 50 == 0:
            print(f"[c5] update {u + 1}/{args.updates}: N={hi} latency {dt * 1e3:.1f} ms ({st['iterations']} LM, "
                  f"{st['pcg_iterations_total']} PCG, structure {st['ms_structure']:.1f} ms) newest pose {newest}",
                  file=sys.stderr, flush=True)
    G.close()

    cpu = None
    if cpu_cases:
        from oracle import cpu_ref   # the checker / CPU baseline, never on the product path
        cpu_ref.optimize(synth.config("C3", n=20000), max_iterations=1, pcg_iter_cap_total=20)   # thread pool warm-up
        ms = []
        for hi, x_after, poses_new, lo in cpu_cases:
            sub = g.prefix(hi)
            # warm start as the GPU run had it: previous optimum for the old poses is approximated by the
            # optimum after the update (the CPU gets an easier start than the GPU did), new poses dead-reckoned
            init = x_after.copy()
            init[lo:hi] = poses_new
            t0 = time.perf_counter()
            _, st, _, _ = cpu_ref.optimize(sub, init=init, pcg_rel_tol=1e-12)
            ms.append((time.perf_counter() - t0) * 1e3)
            threads = int(st["threads"])
        cpu = dict(kind="port", cores=threads, updates=len(ms), sizes=[c[0] for c in cpu_cases], latency_ms=ms,
                   mean_ms=float(np.mean(ms)))

    lat = np.array(lat)
    out = dict(metric="incremental_update_latency", unit="ms", higher_is_better=False,
               config=dict(workload="C5-incremental: 224-wide lawn-mower SE(2) graph growing %d -> %d poses in %d batches of "
                                    "%d keyframes, full re-optimisation per batch, warm start" %
                                    (args.n0, n_final, args.updates, args.batch)),
               p50=float(np.percentile(lat, 50)), p99=float(np.percentile(lat, 99)), mean=float(lat.mean()),
               max=float(lat.max()), first_solve_ms=t_first * 1e3,
               mean_structure_ms=float(np.mean(ms_struct)), mean_solve_ms=float(np.mean(ms_solve)),
               mean_lm_iterations=float(np.mean(lm_its)), mean_pcg_iterations=float(np.mean(pcg_its)),
               final_chi2=chi2[-1], n_final=n_final, cpu_baseline=cpu, dtype="f64", data="synthetic")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
