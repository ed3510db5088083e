#!/usr/bin/env python
"""bench.py — GN/LM iterations per second on the 1M-pose SE(2) graph (BASELINE.json metric, config C3).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C3] [--n POSES]

A "step" is one outer Levenberg-Marquardt iteration of the hot path (linearise + assemble, block-Jacobi PCG
to pcg_rel_tol, retract, chi2, accept/reject) on the resident graph, continuing one optimisation trajectory
from the synthetic initial guess.  One JSON line on stdout (rank 0); everything else goes to stderr.

  value     device-resident throughput: K steps / device time (CUDA events on the library's stream)
  e2e       the same through the C ABI with HOST buffers, every step: gs_set_poses (H2D) + gs_optimize(1) +
            gs_get_poses (D2H)
  roofline  dominant kernel (BSR SpMV): SURVEY §8(d) algorithmic bytes / CUDA-event launch duration
  cpu_baseline / --impl reference: oracle/cpu_ref.c (C + OpenMP port of the path, all host threads) on a bounded
            sample of the same workload.  The reference's own optimiser is GTSAM (absent, unpinned) — see DESIGN.md.
"""
from __future__ import annotations

import argparse
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
            f"FP64 LM, block-Jacobi PCG to {PCG_TOL:g}")


# ----------------------------------------------------------------------------------------- CPU arm
def cpu_sample(g, pcg_cap, pcg_per_lm=None):
    """One LM iteration of the C/OpenMP port on the full graph with PCG capped at `pcg_cap` iterations.
    Returns seconds per LM iteration, scaled to `pcg_per_lm` PCG iterations when the cap was hit."""
    from oracle import cpu_ref
    global _CPU_WARM
    if not _CPU_WARM:   # spin up the OpenMP thread pool outside the measurement
        from graphslam_b200 import synth
        cpu_ref.optimize(synth.config("C3", n=20000), max_iterations=1, pcg_iter_cap_total=20)
        _CPU_WARM = True
    x, st, errs, tries = cpu_ref.optimize(g, max_iterations=1, rel_error_tol=0.0, abs_error_tol=0.0,
                                          pcg_rel_tol=PCG_TOL, pcg_iter_cap_total=pcg_cap)
    its = max(int(st["pcg_iterations_total"]), 1)
    per_pcg = st["ms_solve"] / its
    fixed = st["ms_linearize"] + st["ms_update"]
    capped = its >= pcg_cap
    n_pcg = pcg_per_lm if (capped and pcg_per_lm) else its
    sec = (fixed + per_pcg * n_pcg) * 1e-3
    return dict(sec_per_lm=sec, ms_per_pcg=per_pcg, ms_fixed=fixed, pcg_run=its, pcg_scaled_to=n_pcg,
                threads=int(st["threads"]), capped=capped, edges_per_sec=g.e / max(st["ms_linearize"] * 1e-3, 1e-12))


def run_reference(args):
    """--impl reference: the CPU implementation of the path on the host cores (oracle port; the reference's own
    optimiser is an absent third-party library).  Each step = one bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    g = make_graph(args.config, args.n)
    pcg_per_lm = args.pcg_per_lm or estimate_pcg_per_lm(g, args.warmup, args.steps)
    cap = args.cpu_pcg_cap
    if pcg_per_lm is None and g.n <= 250_000:
        cap = 2_000_000_000      # no iteration count on record and the graph is small enough: solve to the tolerance
    secs = []
    last = None
    for s in range(args.warmup + args.steps):
        last = cpu_sample(g, cap, pcg_per_lm)
        if s >= args.warmup:
            secs.append(last["sec_per_lm"])
        log(f"[bench/reference] step {s}: {last['sec_per_lm']:.2f} s per LM iteration "
            f"({last['ms_per_pcg']:.2f} ms per PCG iteration x {last['pcg_scaled_to']})")
    sec = statistics.mean(secs)
    value = 1.0 / sec
    if last["capped"] and pcg_per_lm:
        sample = (f"1 LM iteration on the full graph, PCG capped at {cap} iterations and scaled to "
                  f"{pcg_per_lm} iterations/LM-iteration (count the same algorithm needs at tol {PCG_TOL:g})")
    elif last["capped"]:
        sample = (f"1 LM iteration on the full graph, PCG capped at {cap} iterations and NOT scaled (no iteration count "
                  f"on record for this graph): an upper bound on the CPU rate")
    else:
        sample = f"1 LM iteration on the full graph, PCG to tol {PCG_TOL:g} ({last['pcg_run']} iterations)"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(g), "n_poses": g.n, "n_factors": g.e, "pcg_rel_tol": PCG_TOL},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": last["threads"], "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def estimate_pcg_per_lm(g, warmup=0, steps=1):
    """PCG iterations one LM iteration needs on this graph, averaged over the LM steps the GPU arm times (steps
    warmup+1 .. warmup+steps of the warm-started sequence): from the recorded per-step counts
    (profiles/pcg_iters.json, logged by the GPU arm: same algorithm, same tolerance, and the count is a property
    of the problem, not of the machine), else measured on the GPU when one is present, else the sample's own count."""
    rec = os.path.join(ROOT, "profiles", "pcg_iters.json")
    if os.path.exists(rec):
        try:
            per_step = json.load(open(rec)).get(f"{g.name}:{g.n}")
        except Exception:
            per_step = None
        if isinstance(per_step, (int, float)) and per_step > 0:     # older record: one average
            return int(round(per_step))
        if per_step:
            take = [per_step[min(k, len(per_step) - 1)] for k in range(warmup, warmup + max(steps, 1))]
            return int(round(sum(take) / len(take)))
    try:
        import torch
        if torch.cuda.is_available():
            from graphslam_b200 import graph
            with graph.Graph(max_iterations=1, rel_error_tol=0.0, abs_error_tol=0.0, pcg_rel_tol=PCG_TOL) as G:
                G.add_graph(g)
                counts = [G.optimize() for _ in range(warmup + max(steps, 1))][warmup:]
            return int(round(sum(c["pcg_iterations_total"] / max(c["inner_iterations"], 1) for c in counts) / len(counts)))
    except Exception as exc:  # no GPU for the reference arm is fine
        log(f"[bench/reference] no GPU estimate of PCG iterations ({exc})")
    return None


# ----------------------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    from graphslam_b200 import graph

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
    log(f"[bench r{rank}] host add {t_add:.2f}s, structure build + upload {t_prep:.2f}s")
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
    ms_upd = vmax(sum(s["ms_update"] for s in stats))
    log(f"[bench r{rank}] resident: {K} steps in {dev_s:.3f}s device ({wall:.3f}s wall), {pcg_total} PCG iterations, "
        f"chi2 {stats[0]['chi2_initial']:.6g} -> {stats[-1]['chi2_final']:.6g}")

    # ---- roofline of the dominant kernel, CUDA events around every launch on the library's stream
    prof = G.profile_pcg(1e-5, args.profile_iters)
    outer = G.profile_outer(3)
    spmv_ms, spmv_bytes = prof["spmv"]
    achieved = spmv_bytes / (spmv_ms * 1e-3) / 1e9
    kernels = {}
    for name, (ms, by) in {**prof, **outer}.items():
        kernels[name] = {"ms": round(ms, 5), "algorithmic_MB": round(by / 1e6, 2), "GBps": round(by / (ms * 1e-3) / 1e9, 1),
                         "frac": round(by / (ms * 1e-3) / 1e9 / peak, 4)}
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
    barrier()
    h2d0, d2h0 = G2.transfer_bytes()
    t0 = time.time()
    for _ in range(K):
        G2.set_poses(keys, x)            # host -> device (this rank's rows + halo, (x,y,cos,sin) doubles)
        G2.optimize()
        _, x = G2.poses()                # device -> host (all N poses on every rank)
    barrier()
    e2e_s = vmax(time.time() - t0)
    e2e = K / e2e_s
    h2d1, d2h1 = G2.transfer_bytes()     # counted by the library from the copies it issued
    G2.close()
    h2d_step = vsum(h2d1 - h2d0) / K
    d2h_step = vsum(d2h1 - d2h0) / K

    # ---- CPU baseline on a bounded sample (rank 0, N = 1)
    pcg_per_lm = int(round(pcg_total / max(solves, 1)))
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        c = cpu_sample(g, args.cpu_pcg_cap, pcg_per_lm)
        cpu = {"value": 1.0 / c["sec_per_lm"], "unit": UNIT, "cores": c["threads"], "kind": "port",
               "sample": (f"1 LM iteration of oracle/cpu_ref.c (C+OpenMP) on the full graph, PCG capped at {c['pcg_run']} "
                          f"iterations ({c['ms_per_pcg']:.2f} ms each) and scaled to the {pcg_per_lm} iterations per solve "
                          f"the same algorithm needed on the GPU; linearise+update {c['ms_fixed']:.0f} ms"),
               "edges_linearized_per_sec": c["edges_per_sec"]}
    if rank == 0:
        try:
            os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
            rec = os.path.join(ROOT, "profiles", "pcg_iters.json")
            d = json.load(open(rec)) if os.path.exists(rec) else {}
            key = f"{g.name}:{g.n}"
            per_step = d.get(key) if isinstance(d.get(key), list) else []
            per_step += [per_step[-1] if per_step else 0] * max(W + K - len(per_step), 0)
            for k, s in enumerate(stats):          # LM steps W+1 .. W+K of the warm-started sequence
                per_step[W + k] = int(round(s["pcg_iterations_total"] / max(s["inner_iterations"], 1)))
            if W and not any(per_step[:W]):        # warm-up steps not recorded yet: best guess = the first timed one
                per_step[:W] = [per_step[W]] * W
            d[key] = per_step
            json.dump(d, open(rec, "w"), sort_keys=True)
        except Exception:
            pass

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": dev_s / K * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(g), "n_poses": g.n, "n_factors": g.e, "pcg_rel_tol": PCG_TOL,
                       "partition": f"vertex ranges over {world} rank(s), NCCL halo exchange + segment-sum all-gathers"
                       if world > 1 else "single GPU",
                       "l2": "inputs larger than L2 (H alone is %.0f MB per PCG iteration on this rank)" % (spmv_bytes / 1e6),
                       "pcg_iters_per_step": pcg_total / K, "structure_build_s": round(t_prep, 3),
                       "host_add_s": round(t_add, 3), "wall_s_timed_region": round(wall, 3)},
            "edges_linearized_per_sec": g.e * len(stats) / (ms_lin * 1e-3),
            "roofline": {"bound": "hbm", "kernel": "k_spmv", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src, "kernels": kernels,
                         "note": "rank 0's kernels on its own rows" if world > 1 else "whole graph"},
            "cpu_baseline": cpu,
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d_step, "d2h_bytes_per_step": d2h_step},
            "gpu_launches": int(launches), "clocks": clocks,
            "breakdown_ms_per_step": {"linearize_assemble": ms_lin / K, "pcg": ms_solve / K, "update": ms_upd / K}}
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
    ap.add_argument("--cpu-pcg-cap", type=int, default=120, help="PCG iterations in the bounded CPU sample")
    ap.add_argument("--pcg-per-lm", type=int, default=None)
    ap.add_argument("--profile-iters", type=int, default=40)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
