"""Multi-GPU plumbing (one process per GPU, torch.distributed over NCCL / NVLink).

Two ways the path shards:

* independent scenarios (config 4): ``shard_range`` splits the batch, every rank runs the fused
  persistent kernel on its slice, NO data-path collective (bench.py reduces only timings / counters);
* one coupled scene (config 5): vehicles are partitioned contiguously; per outer iteration the ranks
  all-gather the ``pos_old`` / ``primal_u`` rows they x-updated, detect edges redundantly (identical
  on every rank), solve every edge that has an owned endpoint (a cross-GPU edge is solved on both
  owners: same fp64 inputs, same result, so z-hat / lambda need no second exchange) and all-reduce
  the two residual sums.  ``VehicleShardedSolver`` drives that loop over the phase entry points.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .solver import PIADMMSolver


def shard_range(n: int, rank: int, world: int):
    """Contiguous, balanced split of range(n): first n % world ranks get one extra item."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def reduce_timing_ms(local_ms: float, device=None) -> float:
    """Max over ranks of a device-timed duration (the contract of bench.py)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(local_ms)
    t = torch.tensor([local_ms], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def reduce_sum(values, device=None):
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(v) for v in t.tolist()]


def all_gather_rows(full: torch.Tensor, lo: int, hi: int, dim: int = 1):
    """In-place all-gather of the rows [lo,hi) (equal-sized shards) of `full` along `dim`."""
    world = dist.get_world_size()
    n = full.shape[dim]
    if n % world != 0 or hi - lo != n // world:
        raise ValueError("all_gather_rows needs equal contiguous shards")
    if dim != 1 or full.shape[0] != 1:
        raise ValueError("expected a [1, V, ...] tensor sharded along V")
    flat = full.view(n, -1)
    dist.all_gather_into_tensor(flat, flat[lo:hi].contiguous())


class VehicleShardedSolver:
    """Config 5: ONE coupled scene, vehicles sharded over the ranks of the default process group."""

    def __init__(self, params, V, D, rank=None, world=None, device=None):
        self.rank = dist.get_rank() if rank is None else rank
        self.world = dist.get_world_size() if world is None else world
        if V % self.world:
            raise ValueError("V must be divisible by the number of ranks")
        self.sol = PIADMMSolver(params, S=1, V=V, D=D, device=device)
        self.lo, self.hi = shard_range(V, self.rank, self.world)
        self.sol.set_ownership(self.lo, self.hi)

    def setup(self, xt, spd, refwin):
        self.sol.setup(xt, spd, refwin)

    def iterate(self):
        s = self.sol
        s.begin_step()
        for _ in range(s.I):
            s.x_update()
            all_gather_rows(s.buf["pos"], self.lo, self.hi)          # coupling variables over NVLink
            all_gather_rows(s.buf["u"], self.lo, self.hi)
            s.detect_edges()
            s.z_update()
            s.residual_partials()
            dist.all_reduce(s.buf["res_part"], op=dist.ReduceOp.SUM)  # global primal / dual residuals
            s.residual_decide()
            if s.num_active() == 0:
                break


def iterate_lockstep(solvers):
    """Single-process emulation of VehicleShardedSolver.iterate() for `solvers` = the per-rank
    PIADMMSolver instances (ownership already set, same inputs): the collectives become copies.
    Used on one GPU (tests) -- the ranks run as sequential launches, never as kernels that wait on
    one another."""
    I = solvers[0].I
    for s in solvers:
        s.begin_step()
    for _ in range(I):
        for s in solvers:
            s.x_update()
        for name in ("pos", "u"):                                   # all-gather
            for src in solvers:
                lo, hi = src.own
                for dst in solvers:
                    if dst is not src:
                        dst.buf[name][:, lo:hi].copy_(src.buf[name][:, lo:hi])
        for s in solvers:
            s.detect_edges(); s.z_update(); s.residual_partials()
        tot = sum(s.buf["res_part"].clone() for s in solvers)       # all-reduce (sum)
        for s in solvers:
            s.buf["res_part"].copy_(tot)
            s.residual_decide()
        if all(s.num_active() == 0 for s in solvers):
            break
