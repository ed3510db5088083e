#!/usr/bin/env python
"""bench.py — GN/LM iterations per second on the 1M-pose SE(2) graph (BASELINE.json metric, config C3).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C3] [--n POSES]

A "step" is one outer Levenberg-Marquardt iteration of the hot path (linearise + assemble, PCG to pcg_rel_tol with
the two-level preconditioner — block-Jacobi plus a rigid-body coarse space — that is the library's default, retract,
chi2, accept/reject) on the resident graph, continuing one optimisation trajectory from the synthetic initial guess;
the line's `block_jacobi_only` block repeats the steps with the plain block-Jacobi preconditioner.  One JSON line on stdout (rank 0); everything else goes to stderr.

  value     device-resident throughput: K steps / device time (CUDA events on the library's stream)
  e2e       the same through the C ABI with HOST buffers, every step: gs_set_local_poses (H2D) + gs_optimize(1) +
            gs_get_local_poses (D2H) — each rank moves the rows it owns
  roofline  dominant kernel (BSR SpMV): SURVEY §8(d) algorithmic bytes / CUDA-event launch duration
  cpu_baseline / --impl reference: oracle/cpu_ref.c (C + OpenMP port of the path with a sparse multifrontal Cholesky, what
            GTSAM's default solver does; all host threads): real un-capped LM steps of the same workload.  The reference's own optimiser is GTSAM (absent, unpinned) — see DESIGN.md.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "gn_iters_per_sec"
UNIT = "GN(LM) iterations/s"
PCG_TOL = 1e-12
_CPU_WARM = False


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0, period_ms=200):
        self.rows, self.proc, self.index, self.period = [], None, index, period_ms

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", str(self.period)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_graph(cid, n):
    from graphslam_b200 import synth
    t = time.time()
    g = synth.config(cid, n=n)
    log(f"[bench] generated {g.name}: N={g.n} E={g.e} in {time.time() - t:.1f}s")
    return g


def workload_name(g):
    return (f"{g.name}: synthetic lawn-mower SE(2) pose graph, {g.n} poses / {g.e} odometry+loop-closure factors, "
            f"FP64 LM, linear solves to relative residual {PCG_TOL:g}")


# ----------------------------------------------------------------------------------------- CPU arm
def config_dict(g):
    """The workload both arms run; printed unchanged by both (run-specific details go under "run")."""
    return {"workload": workload_name(g), "n_poses": int(g.n), "n_factors": int(g.e), "pcg_rel_tol": PCG_TOL,
            "step": "one outer LM iteration (linearise, damped solve, retract, chi2, accept/reject), warm-started, lambda reset"}


def pose_digest(x):
    """SHA-256 of the pose array's bytes (float64, sorted-key order): equal digests = bitwise equal results."""
    return hashlib.sha256(np.ascontiguousarray(x, dtype=np.float64).tobytes()).hexdigest()


def cpu_threads():
    """Set the OpenMP thread count of the CPU port explicitly to the host's cores (torchrun exports OMP_NUM_THREADS=1
    to its children) and fail loudly if that did not take."""
    from oracle import cpu_ref
    want = cpu_ref.host_cores()
    got = cpu_ref.set_threads(want)
    if want > 1 and got <= 1:
        raise SystemExit(f"bench.py: the CPU arm would run on {got} thread of {want} cores")
    return got


def cpu_warm():
    global _CPU_WARM
    if not _CPU_WARM:   # spin up the OpenMP thread pool outside the measurement
        from graphslam_b200 import synth
        from oracle import cpu_ref
        cpu_ref.optimize(synth.config("C3", n=20000), max_iterations=1, pcg_iter_cap_total=20)
        _CPU_WARM = True


def cpu_step(g, x, solver, **kw):
    """One real LM iteration of the C/OpenMP port from values x (lambda reset, like gs_optimize(max_iterations=1)).
    Time = the call minus its one-off structure build and ordering, which the GPU arm also keeps out of a step."""
    from oracle import cpu_ref
    x1, st, errs, _ = cpu_ref.optimize(g, init=x, max_iterations=1, rel_error_tol=0.0, abs_error_tol=0.0,
                                       pcg_rel_tol=PCG_TOL, solver=solver, **kw)
    st["sec"] = (st["ms_total"] - st["ms_structure"] - st["ms_analyze"]) * 1e-3
    return x1, st


def cpu_pcg_sample(g, x, cap_iters, time_cap_s, pcg_per_lm=None):
    """The CPU PCG variant on one LM iteration: run to the tolerance unless it exceeds `time_cap_s` (or `cap_iters`
    iterations when no time cap is given); then the measured ms per PCG iteration is reported, and scaled to
    `pcg_per_lm` iterations only as a labelled extrapolation."""
    kw = dict(pcg_time_cap_ms=time_cap_s * 1e3) if time_cap_s else dict(pcg_iter_cap_total=cap_iters)
    _, st = cpu_step(g, x, 0, **kw)
    its = max(int(st["pcg_iterations_total"]), 1)
    per = st["ms_solve"] / its
    fixed = st["ms_linearize"] + st["ms_update"]
    finished = not st["capped"]
    out = {"ms_per_pcg_iteration": per, "pcg_iterations_run": its, "finished": bool(finished),
           "ms_linearize_update": fixed, "edges_linearized_per_sec": g.e / max(st["ms_linearize"] * 1e-3, 1e-12)}
    if finished:
        out["sec_per_step"] = st["sec"]
    elif pcg_per_lm:
        out["sec_per_step_extrapolated"] = (fixed + per * pcg_per_lm) * 1e-3
        out["extrapolated_to_pcg_iterations"] = int(pcg_per_lm)
    else:
        out["sec_per_step_lower_bound"] = st["sec"]
    return out


def superlu_sample(n=100_000):
    """Informational third figure: wall time of the numpy/SuperLU oracle's damped solve on a bounded
    sample of the workload (the same generator cut to `n` poses)."""
    from graphslam_b200 import synth
    from oracle import gs_oracle as O
    g = synth.config("C3", n=n)
    prob = O.Problem.from_graph(g)
    h, b, _ = prob.linearize(g.init, 0)
    t = time.time()
    O.solve_damped(h, b, 1e-5)
    return {"sec_per_solve": time.time() - t, "sample": f"one damped solve (scipy splu, 1 thread) on C3 cut to {n} poses"}


def run_reference(args):
    """--impl reference: the reference's algorithm for the path on the host cores.  What graph.cpp:119 runs is GTSAM's
    LM with a multifrontal Cholesky; GTSAM is absent (unpinned third-party dependency), so this is the oracle's C
    restatement: every step is a REAL, un-capped LM iteration of oracle/cpu_ref.c with its sparse multifrontal
    Cholesky (oracle/cpu_chol.c), warm-started like the GPU arm's steps.  The CPU PCG variant (the CUDA path's own
    algorithm) is timed beside it on one step, capped by wall time; `value` is the faster variant's."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = cpu_threads()
    g = make_graph(args.config, args.n)
    cpu_warm()
    K, W = args.steps, args.warmup
    x = np.ascontiguousarray(g.init)
    secs, chi2 = [], None
    t_begin = time.time()
    skipped = 0
    for s in range(W + K):
        if time.time() - t_begin > args.ref_budget_s and len(secs) >= 2:
            skipped = W + K - s
            break
        x, st = cpu_step(g, x, 1)
        chi2 = st["chi2_final"]
        if s >= W:
            secs.append(st["sec"])
        log(f"[bench/reference] step {s}: {st['sec']:.2f} s (factor {st['ms_factor'] * 1e-3:.2f} s, "
            f"solve {(st['ms_solve'] - st['ms_factor']) * 1e-3:.2f} s, linearise {st['ms_linearize'] * 1e-3:.2f} s; ordering "
            f"{st['ms_analyze'] * 1e-3:.2f} s not counted) chi2 {st['chi2_initial']:.6g} -> {st['chi2_final']:.6g}")
    if not secs:      # budget hit during warm-up: the steps so far are what there is
        secs = [st["sec"]]
    sec_chol = statistics.mean(secs)
    pcg = None
    if not args.no_cpu_pcg:
        pcg = cpu_pcg_sample(g, np.ascontiguousarray(g.init), args.cpu_pcg_cap, args.cpu_pcg_time_cap_s, args.pcg_per_lm)
        log(f"[bench/reference] CPU PCG variant: {json.dumps(pcg)}")
    sec_pcg = (pcg or {}).get("sec_per_step") or (pcg or {}).get("sec_per_step_lower_bound")
    variant = "cholesky" if (sec_pcg is None or sec_chol <= sec_pcg) else "pcg"
    sec = sec_chol if variant == "cholesky" else sec_pcg
    value = 1.0 / sec
    sample = (f"{len(secs)} real LM iterations of oracle/cpu_ref.c on the full graph, sparse multifrontal Cholesky "
              f"(nnz(L) {st['chol_nnz']:.3g}, {st['chol_flops']:.3g} flop per factorisation), {threads} threads; "
              f"structure build and ordering not counted" + (f"; {skipped} step(s) skipped at the {args.ref_budget_s}-s budget" if skipped else ""))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": K,
            "warmup": W, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config_dict(g),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "variant": variant,
                             "sample": sample, "cholesky": {"sec_per_step": sec_chol, "steps_measured": len(secs)},
                             "pcg": pcg},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "steps_measured": len(secs), "chi2_final": repr(float(chi2)),
            "pose_sha256": pose_digest(x)}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------- our arm
def _lib_flags():
    from graphslam_b200 import _lib
    return {k: getattr(_lib, k) for k in dir(_lib) if k.startswith("GS_SOLVER_")}


def kernel_table(prof, peak):
    return {name: {"ms": round(ms, 5), "algorithmic_MB": round(by / 1e6, 2), "GBps": round(by / (ms * 1e-3) / 1e9, 1),
                   "frac": round(by / (ms * 1e-3) / 1e9 / peak, 4)} for name, (ms, by) in prof.items() if ms > 0}


def roofline_pass(graph_mod, g, device, peak, iters):
    """Per-kernel CUDA-event timings of one config inside the real kernel sequence (gs_profile_pcg / gs_profile_outer)."""
    with graph_mod.Graph(device=device) as G:
        G.add_graph(g)
        G.prepare()
        prof = G.profile_pcg(1e-5, iters)
        outer = G.profile_outer(3)
    return kernel_table({**prof, **outer}, peak)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from graphslam_b200 import graph, synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the optimiser has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def vmax(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def vsum(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    g = make_graph(args.config, args.n)
    K, W = args.steps, args.warmup
    peak, peak_src = load_peaks()
    step_params = dict(max_iterations=1, rel_error_tol=0.0, abs_error_tol=0.0, pcg_rel_tol=PCG_TOL)

    def new_graph():
        if world > 1:
            return graph.distributed_graph(**step_params)
        return graph.Graph(device=local_rank, **step_params)

    # ---- device-resident leg: graph uploaded and structure built before the timed region
    G = new_graph()
    t0 = time.time()
    G.add_graph(g)
    t_add = time.time() - t0
    t0 = time.time()
    G.prepare()
    t_prep = time.time() - t0
    prep_parts = G.structure_timing()
    log(f"[bench r{rank}] host add {t_add:.2f}s, prepare {t_prep:.2f}s: " + json.dumps({k: round(v, 1) for k, v in prep_parts.items()}) + " ms")
    for _ in range(W):
        G.optimize()
    barrier()
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    stats = []
    t0 = time.time()
    for _ in range(K):
        stats.append(G.optimize())
    barrier()
    wall = time.time() - t0
    clocks = sampler.stop() if sampler else None
    dev_s = vmax(sum(s["ms_device"] for s in stats) * 1e-3)     # device time, max over ranks
    value = K / dev_s
    pcg_total = sum(s["pcg_iterations_total"] for s in stats)
    log(f"[bench r{rank}] PCG iterations per step (after {W} warm-up steps): " +
        json.dumps([int(s["pcg_iterations_total"]) for s in stats]))
    solves = sum(s["inner_iterations"] for s in stats)
    launches = vsum(float(sum(s["gpu_launches"] for s in stats)))
    ms_lin = vmax(sum(s["ms_linearize"] for s in stats))
    ms_solve = vmax(sum(s["ms_solve"] for s in stats))
    ms_coarse = vmax(sum(s.get("ms_coarse_setup", 0.0) for s in stats))
    ms_upd = vmax(sum(s["ms_update"] for s in stats))
    waits = None
    if world > 1:   # peer-memory path: what CTA 0 of this rank spent waiting for other GPUs, per PCG iteration (SM clock ticks)
        n_w = max(sum(s["p2p_waits"] for s in stats), 1.0)
        waits = {k: sum(s["p2p_wait_ticks"][i] for s in stats) / n_w for i, k in enumerate(("spmv", "update", "dir"))}
    _, x_res = G.poses()            # collective on several ranks; every rank receives all poses
    chi2_final = stats[-1]["chi2_final"]
    digest = pose_digest(x_res)
    log(f"[bench r{rank}] resident: {K} steps in {dev_s:.3f}s device ({wall:.3f}s wall), {pcg_total} PCG iterations, "
        f"chi2 {stats[0]['chi2_initial']:.17g} -> {chi2_final:.17g}, poses sha256 {digest[:16]}")

    # ---- roofline of the dominant kernel, CUDA events around every launch on the library's stream
    prof = G.profile_pcg(1e-5, args.profile_iters)
    outer = G.profile_outer(3)
    spmv_ms, spmv_bytes = prof["spmv"]
    achieved = spmv_bytes / (spmv_ms * 1e-3) / 1e9
    kernels = kernel_table({**prof, **outer}, peak)
    log(f"[bench r{rank}] kernels: " + json.dumps(kernels))
    traffic = None
    tf = os.path.join(ROOT, "profiles", "spmv_traffic.json")
    if os.path.exists(tf) and world == 1:
        try:
            traffic = json.load(open(tf)).get(f"{g.name}:{g.n}")
        except Exception:
            traffic = None
    G.close()

    # ---- end-to-end leg through the C ABI with host buffers: every step uploads the values and reads them back
    keys = np.ascontiguousarray(g.keys)
    G2 = new_graph()
    G2.add_graph(g)
    G2.prepare()
    x = np.ascontiguousarray(g.init)
    for _ in range(W):                   # same warm-up as the resident leg: the timed steps start from the same state
        G2.set_poses(keys, x)
        G2.optimize()
        _, x = G2.poses()
    x_warm = x.copy()
    # every rank moves the rows it owns and nothing else (gs_set_local_poses / gs_get_local_poses: 32 N / P bytes each way per
    # rank; the ghost rows follow device to device) — with the collective gs_get_poses every rank would receive and convert all
    # N poses every step.  One GPU owns every row: the same bytes as gs_set_poses / gs_get_poses.
    first, count = G2.local_range()
    xl = np.ascontiguousarray(x[first:first + count])   # the caller's host buffers, reused every step
    kl, _ = G2.local_poses(xyt_out=xl.copy())           # the keys of the owned rows: read once
    assert np.array_equal(kl, np.sort(keys)[first:first + count])
    barrier()
    h2d0, d2h0 = G2.transfer_bytes()
    t0 = time.time()
    for _ in range(K):
        G2.set_local_poses(xl)           # host -> device ((x,y,cos,sin) doubles of this rank's rows)
        G2.optimize()
        G2.local_poses(xyt_out=xl, want_keys=False)   # device -> host, into the same host buffer
    barrier()
    e2e_s = vmax(time.time() - t0)
    e2e = K / e2e_s
    h2d1, d2h1 = G2.transfer_bytes()     # counted by the library from the copies it issued
    G2.close()
    h2d_step = vsum(h2d1 - h2d0) / K
    d2h_step = vsum(d2h1 - d2h0) / K

    extra = {}
    cpu = None
    if world == 1 and rank == 0 and not args.no_extras:
        # ---- the same steps with the block-Jacobi preconditioner alone (the north_star's literal algorithm; also what the
        #      multi-GPU path runs): up to 3 steps from the same warmed-up state
        try:
            kb = min(K, 3)
            with graph.Graph(device=local_rank, solver_flags=_lib_flags()["GS_SOLVER_BLOCK_JACOBI_ONLY"], **step_params) as GB:
                GB.add_graph(g)
                GB.set_poses(keys, x_warm)
                sb = [GB.optimize() for _ in range(kb)]
            sec = sum(s_["ms_device"] for s_ in sb) * 1e-3 / kb
            extra["block_jacobi_only"] = {"value": 1.0 / sec, "unit": UNIT, "ms_per_step": sec * 1e3, "steps": kb,
                                          "pcg_iters_per_step": sum(s_["pcg_iterations_total"] for s_ in sb) / kb,
                                          "us_per_pcg_iteration": sec * 1e6 / max(sum(s_["pcg_iterations_total"] for s_ in sb) / kb, 1),
                                          "chi2_final": repr(float(sb[-1]["chi2_final"]))}
            log(f"[bench] block-Jacobi only: {json.dumps(extra['block_jacobi_only'])}")
        except Exception as exc:
            extra["block_jacobi_only"] = {"error": str(exc)[:300]}
        # ---- the north-star gate names C4 as well: per-kernel roofline of the dense loop-closure graph
        try:
            t0 = time.time()
            g4 = synth.config("C4")
            extra["roofline_c4"] = {"workload": f"{g4.name}: {g4.n} poses / {g4.e} factors", "peak": peak, "unit": "GB/s",
                                    "kernels": roofline_pass(graph, g4, local_rank, peak, args.profile_iters)}
            log(f"[bench] C4 roofline pass {time.time() - t0:.1f}s: " + json.dumps(extra["roofline_c4"]["kernels"]))
        except Exception as exc:
            extra["roofline_c4"] = {"error": str(exc)[:300]}
        # ---- BASELINE config C5: latency per incremental update (bounded sample: the last updates of the sequence)
        try:
            from tools import bench_incremental
            t0 = time.time()
            extra["incremental_c5"] = bench_incremental.run(n0=50_000 - 100 * args.c5_updates, updates=args.c5_updates,
                                                            cpu_updates=0 if args.no_cpu_baseline else 1)
            log(f"[bench] C5 incremental sample {time.time() - t0:.1f}s: p50 {extra['incremental_c5']['p50']:.1f} ms")
        except Exception as exc:
            extra["incremental_c5"] = {"error": str(exc)[:300]}

    # ---- CPU baseline on a bounded sample (rank 0, N = 1): both CPU variants + the SuperLU oracle's time
    # the CPU PCG variant is block-Jacobi PCG: its extrapolation uses the iteration count block-Jacobi PCG needed on the
    # GPU for the same steps (extra["block_jacobi_only"]), never the two-level count
    bj = extra.get("block_jacobi_only", {})
    pcg_per_lm = int(round(bj["pcg_iters_per_step"])) if "pcg_iters_per_step" in bj else None
    if not args.no_cpu_baseline and world == 1:
        threads = cpu_threads()
        cpu_warm()
        x0 = np.ascontiguousarray(g.init)
        _, st = cpu_step(g, x0, 1)
        chol = {"sec_per_step": st["sec"], "sec_factor": st["ms_factor"] * 1e-3, "sec_ordering_not_counted": st["ms_analyze"] * 1e-3,
                "nnz_L": st["chol_nnz"], "flop_per_factorisation": st["chol_flops"]}
        pcg = cpu_pcg_sample(g, x0, args.cpu_pcg_cap, 0, pcg_per_lm)
        sec_pcg = pcg.get("sec_per_step") or pcg.get("sec_per_step_extrapolated")
        variant = "cholesky" if (sec_pcg is None or chol["sec_per_step"] <= sec_pcg) else "pcg"
        sec = chol["sec_per_step"] if variant == "cholesky" else sec_pcg
        try:
            slu = superlu_sample()
        except Exception as exc:
            slu = {"error": str(exc)[:200]}
        cpu = {"value": 1.0 / sec, "unit": UNIT, "cores": threads, "kind": "port", "variant": variant,
               "sample": ("1 real LM iteration of oracle/cpu_ref.c from the initial values on the full graph with the sparse "
                          "multifrontal Cholesky (what the reference's GTSAM call does), and the CPU PCG variant capped at "
                          f"{pcg['pcg_iterations_run']} iterations ({pcg['ms_per_pcg_iteration']:.2f} ms each; scaled to the "
                          f"{pcg_per_lm} iterations per solve block-Jacobi PCG needed on the GPU, as a labelled extrapolation); "
                          f"value = the faster"),
               "cholesky": chol, "pcg": pcg, "superlu_oracle": slu,
               "edges_linearized_per_sec": pcg["edges_linearized_per_sec"]}

    two_level = g.n >= 1024
    run = {"preconditioner": ("two-level: 3x3 block-Jacobi + rigid-body coarse space over graph aggregates" if two_level
                              else "3x3 block-Jacobi"),
           "partition": (f"vertex ranges over {world} ranks; halo rows and dot-product segment sums travel by stores into peer "
                         f"memory (NVLink) from inside the PCG kernels; NCCL only for setup and the final pose gather")
           if world > 1 else "single GPU",
           "l2": "inputs larger than L2 (H alone is %.0f MB per PCG iteration on this rank)" % (spmv_bytes / 1e6),
           "pcg_iters_per_step": pcg_total / K, "us_per_pcg_iteration": dev_s / max(pcg_total, 1) * 1e6,
           "structure_build_s": round(prep_parts["host"] * 1e-3, 3), "prepare_s": round(t_prep, 3),
           "prepare_parts_ms": {k: round(v, 1) for k, v in prep_parts.items()}, "host_add_s": round(t_add, 3), "wall_s_timed_region": round(wall, 3),
           "p2p_wait_ticks_per_iteration_rank0": waits}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": dev_s / K * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": config_dict(g), "run": run,
            "chi2_final": repr(float(chi2_final)), "pose_sha256": digest,
            "edges_linearized_per_sec": g.e * len(stats) / (ms_lin * 1e-3),
            "roofline": {"bound": "hbm", "kernel": "k_spmv", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic,
                         "traffic_source": "ncu --set full capture committed under profiles/ (not measured by this run)",
                         "peak_source": peak_src, "kernels": kernels,
                         "note": ("rank 0's kernels on its own rows" if world > 1 else "whole graph") +
                                 ("; two-level iteration: 'pcg_update' = k_pcg_update_staged + k_restrict + k_coarse_mv timed "
                                  "together and 'pcg_dir' = k_dir_coarse, against SURVEY 8(d)'s bytes of the update and the "
                                  "direction alone (per-kernel times: profiles/r2_launches_C3_1M.csv)" if two_level else "")},
            "cpu_baseline": cpu,
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d_step, "d2h_bytes_per_step": d2h_step,
                    "calls": "gs_set_local_poses + gs_optimize + gs_get_local_poses per step on every rank (host buffers; bytes "
                             "summed over ranks, counted by the library)"},
            "gpu_launches": int(launches), "clocks": clocks,
            "breakdown_ms_per_step": {"linearize_assemble": ms_lin / K, "pcg": ms_solve / K, "update": ms_upd / K,
                                      "coarse_setup_inside_pcg": ms_coarse / K}}
    line.update(extra)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C3")
    ap.add_argument("--n", type=int, default=None, help="override the pose count (default: the config's full size)")
    ap.add_argument("--cpu-pcg-cap", type=int, default=120, help="PCG iterations in the GPU arm's bounded CPU PCG sample")
    ap.add_argument("--cpu-pcg-time-cap-s", type=float, default=60.0,
                    help="reference arm: wall-time cap of the CPU PCG variant's one LM iteration")
    ap.add_argument("--pcg-per-lm", type=int, default=None)
    ap.add_argument("--profile-iters", type=int, default=40)
    ap.add_argument("--ref-budget-s", type=float, default=900.0, help="reference arm: stop starting new steps after this")
    ap.add_argument("--c5-updates", type=int, default=30, help="incremental (C5) updates sampled by the default line")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cpu-pcg", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the C4 roofline pass and the C5 incremental sample")
    args = ap.parse_args()
    if args.config == "C5":
        from tools import bench_incremental
        return bench_incremental.main_from_bench(args)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
