"""Multi-GPU plumbing (one process per GPU, torch.distributed for the rendezvous).

Two ways the path shards:

* independent scenarios (config 4): ``shard_range`` splits the batch, every rank runs the fused
  persistent kernel on its slice, NO data-path collective (bench.py reduces only timings / counters);
* one coupled scene (config 5): ``VehicleShardedSolver``.  Vehicles are ordered SPATIALLY
  (``spatial_order``: strips along the axis of larger extent) and every rank owns one contiguous strip of that
  order, so most collision edges have both endpoints on one rank; an edge that crosses a cut is solved on both
  owners (same fp64 inputs, same result: z-hat / lambda need no second exchange).  Two exchanges:

  - ``exchange="peer"`` (default): the persistent cooperative kernel of csrc/piadmm_coop.cuh runs on every rank and
    stores its owned ``pos_old`` / ``primal_u`` rows and per-edge residual terms straight into the peers' exchange
    blocks over NVLink (CUDA IPC mapped memory); the grid barriers behind the x and z phases span all ranks.
    torch.distributed only carries the 64-byte IPC handles once.  No NCCL call and no host round trip in the step.
  - ``exchange="nccl"`` (baseline): the phase kernels, one ``all_gather`` of the owned ``pos_old`` rows and one of the
    two residual partial sums per outer iteration (summed in rank order on every rank: deterministic), stop decision
    on the device, the host looks at the done flag one iteration late (never waits for the GPU inside the loop).

  Results are bit-identical for every world size and both exchanges (tests/test_gpu_coupled.py, tests/mgpu_c5.py).
"""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.distributed as dist

from . import _abi
from .solver import PIADMMSolver


def shard_range(n: int, rank: int, world: int):
    """Contiguous, balanced split of range(n): first n % world ranks get one extra item."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def spatial_order(xt) -> np.ndarray:
    """Permutation that sorts the vehicles of ONE scene into strips: primary key = the coordinate with the larger
    extent (ties: y), secondary = the other coordinate, then the vehicle index.  Independent of the world size, so the
    vehicle numbering the kernels see -- and with it slot order, edge order and every floating-point sum -- is the same
    on 1, 2, 4 and 8 GPUs."""
    xy = np.asarray(xt, dtype=np.float64).reshape(-1, 3)[:, :2]
    ext = xy.max(axis=0) - xy.min(axis=0)
    a = 0 if ext[0] > ext[1] else 1
    return np.lexsort((np.arange(len(xy)), xy[:, 1 - a], xy[:, a]))


def cut_edges(perm, edge_ij, world):
    """How many of the edges (original numbering) cross an ownership cut of the spatial order (diagnostic)."""
    inv = np.empty_like(perm); inv[perm] = np.arange(len(perm))
    owner = np.zeros(len(perm), dtype=np.int64)
    for r in range(world):
        lo, hi = shard_range(len(perm), r, world)
        owner[lo:hi] = r
    e = np.asarray(edge_ij).reshape(-1, 2)
    return int((owner[inv[e[:, 0]]] != owner[inv[e[:, 1]]]).sum())


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


def gather_peer_handles(mine: bytes, world: int, device) -> bytes:
    """all-gather of the ranks' IPC handles (the only thing torch.distributed carries for the peer exchange)."""
    t = torch.frombuffer(bytearray(mine), dtype=torch.uint8).to(device)
    out = torch.empty(world * len(mine), dtype=torch.uint8, device=device)
    dist.all_gather_into_tensor(out, t)
    return bytes(out.cpu().numpy().tobytes())


class VehicleShardedSolver:
    """Config 5: ONE coupled scene, vehicles sharded over the ranks of the default process group (world == 1: the same
    spatial numbering on one GPU, so that results can be compared bit for bit across world sizes)."""

    def __init__(self, params, V, D, rank=None, world=None, device=None, exchange=None):
        ini = dist.is_available() and dist.is_initialized()
        self.rank = (dist.get_rank() if ini else 0) if rank is None else rank
        self.world = (dist.get_world_size() if ini else 1) if world is None else world
        self.V = V
        self.sol = PIADMMSolver(params, S=1, V=V, D=D, device=device)
        self.lo, self.hi = shard_range(V, self.rank, self.world)
        self.sol.set_ownership(self.lo, self.hi)
        self.exchange = "none"
        self.exchange_note = ""
        self.perm = np.arange(V)
        self._have_perm = False
        self._perm_dev = None
        self._stage = None
        if self.world > 1:
            want = exchange or os.environ.get("PIADMM_EXCHANGE", "peer")
            if want not in ("peer", "nccl"):
                raise ValueError("exchange must be 'peer' or 'nccl'")
            if want == "peer":
                ok, why = self._connect_peers()
                if not ok:                       # every rank takes the same branch (the outcome is all-reduced)
                    want = "nccl"
                    self.exchange_note = f"peer exchange unavailable ({why}); NCCL baseline path"
            if want == "nccl" and V % self.world:
                raise ValueError("the NCCL exchange needs V divisible by the number of ranks")
            self.exchange = want
            if want == "nccl":
                self._done_host = torch.zeros(2, dtype=torch.int32).pin_memory()
                self._done_ev = [torch.cuda.Event(), torch.cuda.Event()]
                self._parts = torch.zeros(self.world, 2, dtype=torch.float64, device=self.sol.device)

    def _connect_peers(self):
        err = ""
        handles = b""
        try:
            mine = self.sol.peer_export()
        except _abi.PIADMMError as e:
            mine, err = bytes(_abi.PEER_HANDLE_BYTES), str(e)
        handles = gather_peer_handles(mine, self.world, self.sol.device)
        if not err:
            try:
                self.sol.peer_connect(self.rank, self.world, handles)
            except _abi.PIADMMError as e:
                err = str(e)
        bad = torch.tensor([1 if err else 0], dtype=torch.int32, device=self.sol.device)
        dist.all_reduce(bad, op=dist.ReduceOp.MAX)
        if int(bad.item()):
            if not err:
                self.sol.peer_disconnect()
            return False, err or "a peer failed to map the exchange blocks"
        dist.barrier()                           # every rank has mapped every block before the first step
        return True, ""

    # ------------------------------------------------------------------ per-step API (original vehicle numbering outside)
    def setup(self, xt, spd, refwin, reorder=None):
        """Inputs of one MPC step in the caller's vehicle numbering (host arrays).  The spatial order is computed at the
        first call and kept afterwards (reorder=True recomputes it): any fixed numbering is valid, vehicles move a few
        metres per step, and a stable numbering keeps the per-vehicle K^-1 cache warm.  The rows are permuted into pinned
        staging buffers and go to the device as three asynchronous copies."""
        V = self.V
        if torch.is_tensor(xt) and torch.is_tensor(spd) and torch.is_tensor(refwin) and self._have_perm and not reorder:
            # (pinned) host tensors in the caller's numbering: three asynchronous H2D copies, the permutation is a gather on
            # the device (no host pass over the rows at all)
            if self._perm_dev is None:
                self._perm_dev = torch.from_numpy(self.perm.copy()).to(self.sol.device)
                self._tmp = {k: torch.empty_like(v) for k, v in ((n, self.sol.buf[n]) for n in ("xt", "spd", "refwin"))}
            for name, src in (("xt", xt), ("spd", spd), ("refwin", refwin)):
                self._tmp[name].copy_(src.view_as(self._tmp[name]), non_blocking=True)
                torch.index_select(self._tmp[name], 1, self._perm_dev, out=self.sol.buf[name])
            return
        xt, spd, refwin = (x.numpy() if torch.is_tensor(x) else x for x in (xt, spd, refwin))
        xt = np.asarray(xt, dtype=np.float64).reshape(V, 3)
        spd = np.broadcast_to(np.asarray(spd, dtype=np.float64).reshape(-1), (V,))
        refwin = np.asarray(refwin, dtype=np.float64).reshape(V, 2, -1)
        if reorder or (reorder is None and not self._have_perm):
            self.perm = spatial_order(xt)
            self._have_perm = True
            self._perm_dev = None
        if self._stage is None:
            mk = lambda *sh: torch.zeros(sh, dtype=torch.float64).pin_memory()
            self._stage = dict(xt=mk(1, V, 3), spd=mk(1, V), refwin=mk(1, V, 2, self.sol.N + 1))
            self._stage_np = {k: v.numpy() for k, v in self._stage.items()}
        p = self.perm
        np.take(xt, p, axis=0, out=self._stage_np["xt"][0], mode="clip")          # (mode="clip": no bounds-check buffering, 2.6x faster)
        np.take(spd, p, axis=0, out=self._stage_np["spd"][0], mode="clip")
        np.take(refwin.reshape(V, -1), p, axis=0, out=self._stage_np["refwin"][0].reshape(V, -1), mode="clip")
        for k, v in self._stage.items():
            self.sol.buf[k].copy_(v, non_blocking=True)

    def iterate(self):
        s = self.sol
        if self.world == 1 or self.exchange == "peer":
            s.iterate()                          # ONE cooperative launch per rank: all outer iterations, exchange fused in
            return
        s.begin_step()
        dh, ev = self._done_host, self._done_ev
        for it in range(s.I):
            s.x_update()
            all_gather_rows(s.buf["pos"], self.lo, self.hi)          # coupling variables over NVLink
            s.detect_edges()
            s.z_update()
            s.residual_partials()
            dist.all_gather_into_tensor(self._parts.view(-1), s.buf["res_part"].view(-1))
            s.buf["res_part"].copy_(self._parts.sum(dim=0, keepdim=True))   # same data, same order on every rank
            s.residual_decide()
            dh[it & 1:(it & 1) + 1].copy_(s.buf["done"], non_blocking=True)
            ev[it & 1].record(torch.cuda.current_stream(s.device))
            if it >= 1:                          # the flag of the PREVIOUS iteration: its copy finished long ago
                ev[(it - 1) & 1].synchronize()
                if int(dh[(it - 1) & 1]):
                    break
        all_gather_rows(s.buf["u"], self.lo, self.hi)

    def plant_step(self):
        self.sol.plant_step()

    def close(self):
        """Collective: every rank unmaps its peers' exchange blocks before anybody frees its own."""
        if self.world > 1 and self.exchange == "peer":
            torch.cuda.synchronize(self.sol.device)
            self.sol.peer_disconnect()
            dist.barrier()
        self.sol.close()

    def results(self):
        """Outputs in the ORIGINAL vehicle numbering (u, pos, xt_next per vehicle; scene-wide scalars and histories)."""
        o = self.sol.to_numpy()
        inv = np.empty_like(self.perm); inv[self.perm] = np.arange(self.V)
        out = {k: o[k][:, inv] for k in ("u", "pos", "xt_next")}
        for k in ("n_outer", "curr_iter", "edge_n", "status", "err_rk", "err_sk", "flag"):
            out[k] = o[k]
        n = int(o["edge_n"][0])
        out["edge_ij"] = self.perm[o["edge_ij"][0, :n]]
        return out


def iterate_lockstep(solvers):
    """Single-process emulation of the NCCL form of VehicleShardedSolver.iterate() for `solvers` = the per-rank
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
        tot = torch.stack([s.buf["res_part"].clone() for s in solvers]).sum(dim=0)   # all-gather + rank-order sum
        for s in solvers:
            s.buf["res_part"].copy_(tot)
            s.residual_decide()
        if all(s.num_active() == 0 for s in solvers):
            break
