"""Per-stage CUDA-event timing of the PCG iteration on WORLD_SIZE GPUs (launch with torchrun), nothing else:
  torchrun --nproc-per-node 2 tools/profile_mgpu.py [C3] [n] [pcg_iters]
Every rank prints its own averages.  Used for the peer-memory path experiments (GS_P2P=0/1 and debug switches)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from graphslam_b200 import graph, synth  # noqa: E402

cid = sys.argv[1] if len(sys.argv) > 1 else "C3"
n = int(sys.argv[2]) if len(sys.argv) > 2 and int(sys.argv[2]) > 0 else None
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 200
rank, local = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
g = synth.config(cid, n=n)
G = graph.distributed_graph()
G.add_graph(g)
G.prepare()
for rep in range(2):
    dist.barrier()
    p = G.profile_pcg(1e-5, iters)
    print(f"[r{rank}] rep {rep} " + " ".join(f"{k} {v[0] * 1e3:.1f}us" for k, v in p.items()) +
          f" sum {sum(v[0] for v in p.values()) * 1e3:.1f}us", flush=True)
G.close()
dist.barrier()
dist.destroy_process_group()
