"""Multi-GPU parity (needs >= 2 GPUs on the box; skipped otherwise): the NCCL path over 2 (and 4, 8 when present)
ranks must reproduce the single-GPU result bit for bit — poses, chi2 trace, accept/reject sequence and PCG
iteration counts."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


# execution paths of the multi-GPU PCG iteration (gs_params.solver_flags; the worker takes the flag word as argv[3]).
_NCCL, _DIRK, _OTF, _BJ = 2, 8, 16, 32
# block-Jacobi PCG: peer-memory kernels (two or three per iteration) or one NCCL call per exchange; two-level
# preconditioner: peer-memory kernels (default) or NCCL per exchange.  Each is compared with ONE GPU running the same
# preconditioner.
_FORMS = [("peer-memory-2-kernels", _BJ | _OTF), ("peer-memory-3-kernels", _BJ | _DIRK), ("nccl-per-op", _BJ | _NCCL),
          ("two-level-default", 0), ("two-level-nccl-per-op", _NCCL)]


def _run_worker(world, flags, cid, n, timeout=900):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "mgpu_worker.py"), cid, str(n),
           str(flags)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    assert res.returncode == 0 and "MGPU_OK" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]


@pytest.mark.parametrize("form", _FORMS, ids=[f[0] for f in _FORMS])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_multi_gpu_path_bitwise_equals_single_gpu(world, form):
    """peer memory: halo rows and dot-product segment sums travel by stores into the peers' windows from inside the PCG
    kernels, with two (direction formed on the fly in the product) or three kernels per iteration; nccl-per-op: one
    NCCL call per exchange.  All must reproduce the single-GPU bits."""
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    _run_worker(world, form[1], "C3", 30000)


@pytest.mark.parametrize("flags", [0, _BJ], ids=["two-level", "block-jacobi-peer-memory"])
@pytest.mark.parametrize("cid,n", [("C3", 250000), ("C4", 40000)])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_multi_gpu_bitwise_at_scale(world, cid, n, flags):
    """The same bit-for-bit check on graphs whose ranks are large enough for the two-kernel iteration (C3, 250k poses)
    and on the dense loop-closure graph, whose halo is scattered over many tiles (C4)."""
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    _run_worker(world, flags, cid, n)
