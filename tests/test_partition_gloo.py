"""world_size-2 test of the N>1 host path on CPU (gloo): each process builds ITS rank's vertex-range partition
with the library's own host code, the NCCL id travels over torch.distributed exactly as on the GPU box, and
the two ranks exchange their halo send lists over gloo to check that they pair up with the peer's ghost slots.
Without a GPU gs_create_distributed must refuse to run (no CPU fallback)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    try:
        sys.path.insert(0, ROOT)
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                          LOCAL_RANK=str(rank))
        import torch
        import torch.distributed as dist
        from graphslam_b200 import _lib, graph, synth
        dist.init_process_group("gloo", rank=rank, world_size=world)
        lib = _lib.load_debug()
        g = synth.config("C3", n=6000)
        ei, ej = g.dense_edges()
        ip = _lib.c_i32_p
        info = np.zeros(8, np.int32)
        ghosts, peers, send = np.zeros(g.n, np.int32), np.zeros(5 * world, np.int32), np.zeros(g.n, np.int32)
        rc = lib.gs_host_partition(g.n, g.e, ei.ctypes.data_as(ip), ej.ctypes.data_as(ip), world, rank, info.ctypes.data_as(ip),
                                   ghosts.ctypes.data_as(ip), peers.ctypes.data_as(ip), send.ctypes.data_as(ip))
        assert rc == 0
        v0, v1, n_ghost, n_peers, n_send = (int(v) for v in info[:5])
        assert n_peers == 1 and n_send > 0 and n_ghost > 0
        prank, soff, scnt, goff, gcnt = (int(v) for v in peers[:5])
        assert prank == 1 - rank
        # ownership ranges tile [0, N) on tile boundaries
        rng = torch.tensor([v0, v1], dtype=torch.int64)
        allr = [torch.zeros(2, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(allr, rng)
        assert allr[0][0] == 0 and allr[0][1] == allr[1][0] and allr[1][1] == g.n and int(allr[0][1]) % 256 == 0
        # my send list must be exactly the peer's ghost list (same order): exchange over gloo
        mine = torch.from_numpy(send[soff:soff + scnt].astype(np.int64))
        cnt = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(cnt, torch.tensor([scnt], dtype=torch.int64))
        theirs = torch.zeros(int(cnt[prank]), dtype=torch.int64)
        if rank == 0:
            dist.send(mine, 1)
            dist.recv(theirs, 1)
        else:
            dist.recv(theirs, 0)
            dist.send(mine, 0)
        assert int(cnt[prank]) == gcnt
        assert np.array_equal(theirs.numpy(), ghosts[goff:goff + gcnt].astype(np.int64))
        # every ghost really is an endpoint of a cut factor
        own = lambda v: (v >= v0) & (v < v1)
        cut = own(ei) != own(ej)
        want = np.unique(np.r_[ei[cut & ~own(ei)], ej[cut & ~own(ej)]])
        assert np.array_equal(want, ghosts[:n_ghost])
        # the NCCL id travels over torch.distributed; without a GPU the library refuses (no CPU fallback)
        if not torch.cuda.is_available():
            raw = bytes(range(128)) if rank == 0 else bytes(128)
            t = torch.tensor(list(raw), dtype=torch.uint8)
            dist.broadcast(t, 0)
            assert bytes(t.tolist()) == bytes(range(128))
            with pytest.raises(graph.GraphSlamError) as e:
                graph.Graph(nccl_id=bytes(t.tolist()), rank=rank, world_size=world)
            assert e.value.status == _lib.GS_ECUDA
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except BaseException as exc:  # noqa: BLE001
        import traceback
        q.put((rank, "FAIL: " + "".join(traceback.format_exception(exc))))


def test_two_rank_partition_and_halo_lists_over_gloo(lib):
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=240) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == {0: "ok", 1: "ok"}, res
