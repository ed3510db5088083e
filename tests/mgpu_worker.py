"""Worker of tests/test_multi_gpu.py (launched by torchrun, one process per GPU): optimise the same graph on
WORLD_SIZE ranks over NCCL and, on rank 0, on one GPU; the two must agree bitwise (fixed reduction tree)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from graphslam_b200 import graph, synth  # noqa: E402


def main():
    cid = sys.argv[1] if len(sys.argv) > 1 else "C3"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 30000
    flags = int(sys.argv[3]) if len(sys.argv) > 3 else 0     # gs_params.solver_flags of the multi-GPU graph
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    g = synth.config(cid, n=n, full_cov=True)
    G = graph.distributed_graph(solver_flags=flags)
    G.add_graph(g)
    chi0 = G.chi2()
    st = G.optimize()
    keys, x = G.poses()
    errors, lambdas, tries = G.trace()
    # warm-started second solve after appending nothing: exercises values staying on the devices
    st2 = G.optimize()
    # marginals: the owner of each pose broadcasts its rows; same bits as one GPU (same PCG)
    pick = np.array([g.keys[0], g.keys[n // 2], g.keys[-1]], dtype=np.uint64)
    cov = G.marginal_covariances(pick)
    # closest_keyframe on the partitioned poses: every rank scans its own rows, same answer everywhere
    near = [G.closest_keyframe(int(g.keys[-1]), 10), G.closest_keyframe(int(g.keys[n // 3]), n // 2)]
    # switching the damping on a live graph (the captured PCG graph is rebuilt: its product reads another scaling)
    G.set_poses(g.keys, g.init)
    G.set_params(diagonal_damping=1)
    st3 = G.optimize()
    # rank-local transfers: a rank reads its own rows (same bits as its slice of the collective read), writes them back
    # displaced (the ghost rows follow on the device), and the next solve must be the one a single GPU runs after
    # gs_set_poses with the same values
    _, x3 = G.poses()
    first, count = G.local_range()
    lk, lx = G.local_poses()
    local_ok = np.array_equal(lk, np.sort(g.keys)[first:first + count]) and np.array_equal(lx, x3[first:first + count])
    shift = np.array([0.01, -0.02, 0.003])
    G.set_local_poses(lx + shift)
    _, x4 = G.poses()                        # collective: every rank sees everybody's new rows
    d4 = x4 - (x3 + shift)
    d4[:, 2] = (d4[:, 2] + np.pi) % (2 * np.pi) - np.pi          # theta comes back in (-pi, pi]
    local_ok = local_ok and float(np.abs(d4).max()) <= 1e-12
    st4 = G.optimize()
    G.close()
    cnt = torch.tensor([count, 1 if local_ok else 0], device="cuda")
    dist.all_reduce(cnt)
    local_ok = int(cnt[0].item()) == g.n and int(cnt[1].item()) == dist.get_world_size()   # the ranges tile the rows
    ok = True
    if rank == 0:
        # the single-GPU reference runs the same preconditioner as the ranks (flag 32 = block-Jacobi only): the two
        # preconditioners solve the same systems to the same tolerance, but not with the same bits
        with graph.Graph(device=local, solver_flags=flags & 32) as G1:
            G1.add_graph(g)
            chi0_1 = G1.chi2()
            st1 = G1.optimize()
            _, x1 = G1.poses()
            e1, l1, t1 = G1.trace()
            st1b = G1.optimize()
            cov1 = G1.marginal_covariances(pick)
            near1 = [G1.closest_keyframe(int(g.keys[-1]), 10), G1.closest_keyframe(int(g.keys[n // 3]), n // 2)]
            G1.set_poses(g.keys, g.init)
            G1.set_params(diagonal_damping=1)
            st1c = G1.optimize()
            _, y3 = G1.poses()
            G1.set_poses(np.sort(g.keys), y3 + shift)
            st1d = G1.optimize()
        ok = (local_ok and st4["chi2_final"] == st1d["chi2_final"] and st4["pcg_iterations_total"] == st1d["pcg_iterations_total"] and
              st3["chi2_final"] == st1c["chi2_final"] and st3["pcg_iterations_total"] == st1c["pcg_iterations_total"] and
              near == near1 and np.array_equal(cov, cov1) and chi0 == chi0_1 and np.array_equal(x, x1) and np.array_equal(errors, e1) and tries.tolist() == t1.tolist()
              and st["pcg_iterations_total"] == st1["pcg_iterations_total"] and st2["chi2_final"] == st1b["chi2_final"])
        print(f"[mgpu] local transfers ok={local_ok} st4 {st4['chi2_final']!r}/{st4['pcg_iterations_total']} 1 GPU {st1d['chi2_final']!r}/"
              f"{st1d['pcg_iterations_total']}", flush=True)
        print(f"[mgpu] world={dist.get_world_size()} N={g.n} iterations={st['iterations']} pcg={st['pcg_iterations_total']} "
              f"chi2 {st['chi2_initial']:.6g}->{st['chi2_final']:.6g} bitwise_equal_to_single_gpu={ok} "
              f"ms_solve {st['ms_solve']:.1f} (1 GPU: {st1['ms_solve']:.1f})", flush=True)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.barrier()
    dist.destroy_process_group()
    if int(flag.item()) != 1:
        raise SystemExit("multi-GPU result differs from the single-GPU result")
    if rank == 0:
        print("MGPU_OK", flush=True)


if __name__ == "__main__":
    main()
