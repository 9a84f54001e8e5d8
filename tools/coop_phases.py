"""Phase timeline of the persistent cooperative kernel on a coupled scene (piadmm_debug_coop_trace):

  python tools/coop_phases.py c5            # one GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 tools/coop_phases.py c5

Prints, per outer iteration, the microseconds between the phase boundaries on every rank (x incl. its exchange barrier,
slots, count, edges, z, the z exchange barrier, decide).  A diagnostic, not a bench value."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from pi_admm_b200.parallel import VehicleShardedSolver  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c5"
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    (xt, spd, refwin), V, N, D = bench.coupled_scene(name)
    p = bench.coupled_params(N)
    vs = VehicleShardedSolver(p, V, D, device=local)
    vs.setup(xt, spd, refwin)
    vs.sol.coop_trace()                      # enable
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local}")
    for _ in range(4):
        flush.fill_(1)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        vs.iterate()
        torch.cuda.synchronize()
    t = vs.sol.coop_trace().astype(np.float64) / 1e3          # us since the kernel start
    names = ["x+xbar", "slots", "count", "edges", "z", "zbar", "decide"]
    rows = [f"rank {rank}: exchange={vs.exchange} begin {t[1]:.1f} us, total {t[-1]:.1f} us"]
    body = np.diff(t[1:])
    for it in range(len(body) // 7):
        seg = body[it * 7:(it + 1) * 7]
        rows.append(f"  iter {it + 1}: " + "  ".join(f"{n} {v:6.1f}" for n, v in zip(names, seg)) + f"   = {seg.sum():.1f} us")
    out = "\n".join(rows)
    if world > 1:
        outs = [None] * world
        dist.all_gather_object(outs, out)
        if rank == 0:
            print("\n".join(outs))
        dist.barrier()
        vs.close()
        dist.destroy_process_group()
    else:
        print(out)


if __name__ == "__main__":
    main()
