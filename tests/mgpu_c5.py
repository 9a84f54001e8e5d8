"""Multi-GPU check (run under torchrun on N GPUs, NCCL): config-5-style vehicle-sharded consensus
over NVLink against the oracle's C twin, and config-4 scenario sharding.  Not collected by pytest
(needs >1 GPU):  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/mgpu_c5.py"""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from oracle import ctwin
    from util import make_params, rel_err
    from pi_admm_b200.parallel import VehicleShardedSolver
    from pi_admm_b200.scenarios import c3_intersection, c5_grid
    V, N, D = 64, 30, 8
    p = make_params("matlab_pi", "default", num_ho=N, beta=10.0, Pnorm=1.0, rho=2.0, eps_pri=0.5, eps_dual=0.5,
                    al_all=0, iter_num=12)
    xt, spd, rw = c3_intersection(V=V, N=N)
    vs = VehicleShardedSolver(p, V, D, device=local)
    vs.setup(xt, spd, rw)
    vs.iterate()
    torch.cuda.synchronize()
    got = vs.sol.to_numpy()
    ref = ctwin.step(p, xt, spd, rw, D=D)
    ok = np.array_equal(got["n_outer"], ref["n_outer"]) and np.array_equal(got["curr_iter"], ref["curr_iter"])
    errs = {k: rel_err(got[k], ref[k]) for k in ("u", "pos", "err_rk", "err_sk")}
    errs["hat_owned"] = rel_err(got["hat"][:, vs.lo:vs.hi], ref["hat"][:, vs.lo:vs.hi])
    ok = ok and max(errs.values()) <= 1e-6
    print(f"[rank {rank}/{world}] c3-sharded V={V}: ok={ok} n_outer={got['n_outer']} edges={int(ref['n_edge_solves'][0])} errs={errs}", flush=True)
    # config 5 at full size: timing of the coupled step (device events, max over ranks)
    xt, spd, rw = c5_grid(N=30)
    V5 = xt.shape[1]
    vs5 = VehicleShardedSolver(p, V5, 8, device=local)
    vs5.setup(xt, spd, rw)
    vs5.iterate()                       # warm-up
    torch.cuda.synchronize(); dist.barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); vs5.iterate(); e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    g5 = vs5.sol.to_numpy()
    solves = torch.tensor([float(g5["n_veh_solves"][0])], device="cuda", dtype=torch.float64)
    dist.all_reduce(solves)
    if rank == 0:
        print(f"c5 V={V5} N=30 on {world} GPUs: {ms.item():.2f} ms per coupled MPC step, n_outer={int(g5['n_outer'][0])}, "
              f"agent-QP solves={int(solves.item())} -> {solves.item() / ms.item() * 1e3:.3e} solves/s, "
              f"edges last iter={int(g5['edge_n'][0])}, status={int(g5['status'][0])}", flush=True)
    allok = torch.tensor([1.0 if ok else 0.0], device="cuda")
    dist.all_reduce(allok, op=dist.ReduceOp.MIN)
    dist.barrier(); dist.destroy_process_group()
    sys.exit(0 if allok.item() == 1.0 else 1)


if __name__ == "__main__":
    main()
