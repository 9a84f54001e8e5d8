"""world_size-2 gloo tests (CPU) of the host-side multi-GPU logic in pi_admm_b200/parallel.py:
scenario sharding of config 4 (no data-path collective; only timings/counters are reduced) and the
row all-gather / residual all-reduce used by the vehicle-sharded config 5 loop."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pi_admm_b200.parallel import shard_range, reduce_timing_ms, reduce_sum, all_gather_rows


def test_shard_range_partitions():
    for n in (1, 7, 4096, 4097):
        for world in (1, 2, 3, 8):
            parts = [shard_range(n, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import sys
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        sys.path.insert(0, root); sys.path.insert(0, os.path.join(root, "tests"))
        from oracle import ctwin
        from util import make_params
        from pi_admm_b200.scenarios import c4_scenarios
        # --- config 4: scenario sharding, the oracle's C twin stands in for the GPU step
        S = 10
        p = make_params("matlab_pi", "default", num_ho=15, eps_pri=1.0, eps_dual=1.0)
        xt, spd, rw = c4_scenarios(S=S, N=15, seed=11)
        lo, hi = shard_range(S, rank, world)
        mine = ctwin.step(p, xt[lo:hi], spd[lo:hi], rw[lo:hi], nthreads=1)
        parts = [None] * world
        dist.all_gather_object(parts, (lo, hi, mine["u"], mine["n_outer"]))
        full = ctwin.step(p, xt, spd, rw, nthreads=1)
        u = np.concatenate([x[2] for x in sorted(parts)]); n_outer = np.concatenate([x[3] for x in sorted(parts)])
        ok_c4 = bool(np.array_equal(u, full["u"]) and np.array_equal(n_outer, full["n_outer"]))
        tmax = reduce_timing_ms(10.0 + rank)
        tot = reduce_sum([float(mine["n_veh_solves"].sum()), 1.0])
        ok_red = (tmax == 10.0 + world - 1) and tot[0] == float(full["n_veh_solves"].sum()) and tot[1] == world
        # --- config 5 collectives: rows all-gather (in place) and partial-sum all-reduce
        V, F = 8, 6
        truth = torch.arange(V * F, dtype=torch.float64).view(1, V, F)
        loc = torch.zeros(1, V, F, dtype=torch.float64)
        lo, hi = shard_range(V, rank, world)
        loc[:, lo:hi] = truth[:, lo:hi]
        all_gather_rows(loc, lo, hi)
        part = torch.tensor([[1.5 * (rank + 1), 2.0]], dtype=torch.float64)
        dist.all_reduce(part, op=dist.ReduceOp.SUM)
        ok_c5 = bool(torch.equal(loc, truth)) and part.tolist() == [[1.5 * sum(range(1, world + 1)), 2.0 * world]]
        q.put((rank, ok_c4, ok_red, ok_c5))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_gloo_sharding():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok_c4, ok_red, ok_c5 in res:
        assert ok_c4, f"rank {rank}: sharded config-4 results differ from the unsharded run"
        assert ok_red, f"rank {rank}: timing / counter reductions wrong"
        assert ok_c5, f"rank {rank}: row all-gather / partial all-reduce wrong"
