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
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly():
    res = _run("--impl", "reference", "--steps", "1", "--warmup", "0", "--n", "6000", "--gpus", "2", env={"RANK": "1"})
    assert res.returncode == 0 and res.stdout.strip() == ""


def test_recorded_iteration_counts_cover_the_benchmark_steps():
    """The reference arm scales its bounded sample to the PCG iteration counts recorded for the very LM steps our arm
    times (default --warmup 3 --steps 3 on C3): the record must hold them."""
    rec = json.load(open(os.path.join(ROOT, "profiles", "pcg_iters.json")))
    per_step = rec["C3-grid1M:1000000"]
    assert isinstance(per_step, list) and len(per_step) >= 6 and all(isinstance(v, int) and v > 1000 for v in per_step)


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
