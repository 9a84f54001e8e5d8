"""Multi-GPU check + timing of the vehicle-sharded coupled step (BASELINE config 5), run under torchrun:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port 29511 tests/mgpu_c5.py

For both exchanges (peer-mapped persistent kernel, NCCL baseline): the full 4096-vehicle scene on G ranks against the
oracle's C twin (rank 0 computes it, everybody compares), results bit-identical across ranks, then device-timed steps
(L2 flushed, barrier on both sides, max over ranks).  One JSON line per exchange on rank 0.  Not collected by pytest
(no test_ prefix): pytest's multi-rank case is tests/test_gpu_coupled.py::test_vehicle_sharded_two_ranks_equal_twin."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import bench  # noqa: E402
from pi_admm_b200.parallel import VehicleShardedSolver, cut_edges  # noqa: E402
from util import rel_err  # noqa: E402


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    name = os.environ.get("MGPU_SCENE", "c5")
    steps = int(os.environ.get("MGPU_STEPS", "10"))
    (xt, spd, refwin), V, N, D = bench.coupled_scene(name)
    p = bench.coupled_params(N)
    ref = None
    if rank == 0:
        # the twin in the kernels' (spatial) vehicle numbering: the residual's veh1 side depends on the numbering
        from oracle import ctwin
        from pi_admm_b200.parallel import spatial_order
        pm = spatial_order(xt)
        inv = np.empty_like(pm); inv[pm] = np.arange(V)
        r = ctwin.step(p, np.ascontiguousarray(xt[:, pm]), np.ascontiguousarray(np.broadcast_to(spd, (1, V))[:, pm]),
                       np.ascontiguousarray(refwin[:, pm]), D=D, nthreads=len(os.sched_getaffinity(0)))
        ref = {k: r[k][:, inv] for k in ("u", "pos", "xt_next")}
        ref.update({k: r[k] for k in ("n_outer", "edge_n", "err_rk", "err_sk")})
        ref["edge_ij"] = pm[r["edge_ij"]]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local}")
    stream = torch.cuda.current_stream()
    for exchange in ("peer", "nccl"):
        vs = VehicleShardedSolver(p, V, D, device=local, exchange=exchange)
        vs.setup(xt, spd, refwin)
        for _ in range(3):
            vs.iterate()
        vs.plant_step()
        out = vs.results()
        # bit-identical across ranks: compare a checksum of u / pos / histories
        h = torch.tensor([float(np.frombuffer(out[k].tobytes(), dtype=np.uint8).astype(np.float64).sum()) for k in ("u", "pos", "err_rk")],
                         dtype=torch.float64, device=f"cuda:{local}")
        hs = [torch.zeros_like(h) for _ in range(world)]
        dist.all_gather(hs, h)
        same = all(bool(torch.equal(hs[0], x)) for x in hs)
        ms = []
        for _ in range(steps):
            flush.fill_(1)
            dist.barrier(); torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream); vs.iterate(); b.record(stream)
            torch.cuda.synchronize()
            ms.append(a.elapsed_time(b))
        t = torch.tensor([float(np.mean(ms))], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            err = {k: rel_err(out[k], ref[k]) for k in ("u", "pos", "xt_next", "err_rk", "err_sk")}
            ok = (np.array_equal(out["n_outer"], ref["n_outer"]) and np.array_equal(out["edge_n"], ref["edge_n"])
                  and max(err.values()) <= 1e-6 and same)
            n = int(ref["edge_n"][0])
            print(json.dumps({"scene": name, "world": world, "exchange_requested": exchange, "exchange": vs.exchange,
                              "note": vs.exchange_note, "ms_per_step": float(t.item()), "outer_iters": int(out["n_outer"][0]),
                              "status": int(out["status"][0]), "parity_vs_cpu_port": ok, "rel_err": err,
                              "identical_across_ranks": same,
                              "cut_edges_last_iter": cut_edges(vs.perm, ref["edge_ij"][0, :n], world)}), flush=True)
            assert ok, (exchange, err, same)
        dist.barrier()
        vs.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
