"""bench.py's output contract, as far as it can be exercised without a GPU: the reference arm (`--impl reference`,
the CPU port on the host cores) prints ONE JSON line with the keys the driver reads, on the same metric / unit /
config as our arm; and our arm refuses to run without a CUDA device instead of falling back to the CPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, env=None):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=600,
                          cwd=ROOT, env=dict(os.environ, **(env or {})))


def test_reference_arm_prints_one_json_line_with_the_contract_keys():
    res = _run("--impl", "reference", "--steps", "1", "--warmup", "0", "--n", "6000")
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [ln for ln in res.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "gn_iters_per_sec" and d["unit"] == "GN(LM) iterations/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 0
    assert d["value"] > 0 and abs(d["ms_per_step"] * 1e-3 * d["value"] - 1.0) < 1e-9
    assert d["dtype"] == "f64" and d["data"] == "synthetic" and d["vs_baseline"] is None and d["gpu_launches"] == 0
    assert "workload" in d["config"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    # both CPU variants are there: real Cholesky steps and the PCG variant beside them; the line's value is the faster one's
    assert cb["variant"] in ("cholesky", "pcg") and cb["cholesky"]["sec_per_step"] > 0 and cb["pcg"]["ms_per_pcg_iteration"] > 0
    assert d["steps_measured"] == 1 and len(d["pose_sha256"]) == 64 and float(d["chi2_final"]) > 0
    assert set(d["config"]) == {"workload", "n_poses", "n_factors", "pcg_rel_tol", "step"}
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly():
    res = _run("--impl", "reference", "--steps", "1", "--warmup", "0", "--n", "6000", "--gpus", "2", env={"RANK": "1"})
    assert res.returncode == 0 and res.stdout.strip() == ""


def test_reference_arm_sets_its_thread_count_itself():
    """torchrun exports OMP_NUM_THREADS=1 to its children; the CPU arm must not inherit that."""
    res = _run("--impl", "reference", "--steps", "1", "--warmup", "0", "--n", "6000", "--no-cpu-pcg", env={"OMP_NUM_THREADS": "1"})
    assert res.returncode == 0, res.stderr[-2000:]
    d = json.loads(res.stdout.strip().splitlines()[-1])
    assert d["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))


def test_our_arm_needs_a_cuda_device():
    try:
        import torch
        if torch.cuda.is_available():
            return
    except Exception:
        pass
    res = _run("--steps", "1", "--warmup", "0", "--n", "6000")
    assert res.returncode != 0 and "no CPU fallback" in (res.stderr + res.stdout)
    assert not [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
