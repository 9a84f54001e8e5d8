#!/usr/bin/env python
"""bench.py -- agent-QP solves/s of one PI-ADMM MPC step over a batch of synthetic scenarios.

  python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload c4|c4n30|c3|c5]

Workload (config 4 of BASELINE.json): S = 4096 independent two-vehicle intersection
scenarios PER GPU (weak scaling, no data-path collective), each one MPC step run to PI-ADMM
convergence by ONE launch of the fused persistent kernel.  One "step" = that MPC step over the
whole batch; `value` = agent-QP (vehicle x-update) solves per second, all GPUs together, inputs
resident in HBM; `e2e` = the same through the host-buffer C-ABI entry point piadmm_step_host
(pinned host inputs -> H2D -> step -> D2H of u/pos/xt_next/iteration counts inside the timed region).

--workload c3 / c5 (not the default, not the driver's line): ONE coupled scene (64-vehicle intersection / 4096-vehicle
grid, N=30) through the generic-V phase kernels; under torchrun the vehicles are sharded over the ranks (NCCL all-gather
of pos/u rows + all-reduce of the residual sums per outer iteration), "scaling": "strong".  `ms_per_step` is then the
BASELINE metric's second half, ms per coupled MPC step.

--impl reference: the reference's CasADi+OSQP stack is not installable here (no wheel, no network),
so this arm times the oracle's C twin (oracle/piadmm_ref.c, kind "port") on all host cores.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (S per GPU, horizon N)
    "c4": (4096, 15),
    "c4n30": (4096, 30),
}


COUPLED = {
    # name: (vehicles, horizon N, neighbour slots D)   -- BASELINE configs 3 and 5, SURVEY.md 8(d)
    "c3": (64, 30, 8),
    "c5": (4096, 30, 8),
}


def coupled_params(N):
    """Synthetic coupled scenes: dual="pi", AL term over activated edges, convex edge problem."""
    import pi_admm_b200 as pa
    p = pa.preset("matlab_pi")
    p.update(num_ho=N, beta=10.0, Pnorm=1.0, rho=2.0, eps_pri=0.5, eps_dual=0.5, al_all=0, iter_num=12)
    return p


def coupled_scene(name):
    from pi_admm_b200.scenarios import c3_intersection, c5_grid
    V, N, D = COUPLED[name]
    return (c3_intersection(V=V, N=N) if name == "c3" else c5_grid(N=N)), V, N, D


def coupled_config(name, V, N, D, world):
    what = {"c3": "c3: ONE 64-vehicle 4-arm intersection (BASELINE config 3)",
            "c5": "c5: ONE 4096-vehicle coupled grid scene, 64+64 lanes x 32 vehicles (BASELINE config 5)"}[name]
    return {"workload": f"{what}, horizon N={N}, one coupled MPC step (outer-iteration cap 12) through the generic-V "
                        "phase kernels", "vehicles": V, "horizon": N, "neighbour_slots": D,
            "semantics": "dual=pi (MATLAB PI+anti-windup), AL over activated edges, edge=convex, eps_pri=eps_dual=0.5, "
                         "inner=osqp_default",
            "parallelism": "single GPU" if world == 1 else
                           f"vehicles sharded x{world}, NCCL all-gather of pos/u rows + all-reduce of residual sums per outer iteration",
            "l2": "L2 flushed (256 MiB write) before every timed step"}


def bench_params(N):
    """C1 problem constants (casadi/PI_ADMM_class.py:15-28) with the PI + anti-windup dual law and
    the stop test of the MATLAB CVX script (dual="pi"), OSQP-default inner tolerances."""
    import pi_admm_b200 as pa
    p = pa.preset("matlab_pi")
    p.update(num_ho=N, eps_pri=0.1, eps_dual=0.1, round_u=4, round_edge=4)
    return p


def flops_per_inner_iter(N, edge):
    """F_iter = 2n^2 + 12m + 6n per inner ADMM iteration (n = N, m = 4N-2);
    the edge problem here is two halves of n = N unknowns with m = 5N-2 rows each."""
    if not edge:
        return 2 * N * N + 12 * (4 * N - 2) + 6 * N
    return 2 * (2 * N * N + 12 * (5 * N - 2) + 6 * N)


def algorithmic_bytes(N, n_veh_solves, n_edge_solves, S):
    """fp64 bytes one launch has to move: per agent-QP 8*[3 + N + 4(N+1)(1+deg)] with deg = 1
    (SURVEY.md 8(d)); per edge problem both sides' pos/dual in, hat/dual out."""
    per_qp = 8 * (3 + N + 4 * (N + 1) * 2)
    per_edge = 8 * (2 * 4 * (N + 1) + 2 * 4 * (N + 1))
    return n_veh_solves * per_qp + n_edge_solves * per_edge


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-i", str(self.index), "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush(); self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in self.f.read().splitlines():
            t = [x.strip() for x in line.split(",")]
            if len(t) < 9:
                continue
            try:
                sm.append(float(t[1])); mx.append(float(t[2]))
            except ValueError:
                continue
            for nm, v in zip(names, t[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args, rank, world):
    """CPU arm: oracle C twin on all host cores; rank 0 only."""
    if rank != 0:
        return
    import __graft_entry__ as ge
    ge.build()
    from oracle import ctwin
    from pi_admm_b200.scenarios import c4_scenarios
    if args.workload in COUPLED:
        return run_reference_coupled(args)
    S, N = WORKLOADS[args.workload]
    p = bench_params(N)
    xt, spd, refwin = c4_scenarios(S=S, N=N)
    arrs = ctwin.alloc(S, 2, 1, N, p.iter_num)
    cores = host_cores()         # all host cores, whatever OMP_NUM_THREADS says (torchrun sets it to 1)
    for _ in range(args.warmup):
        ctwin.step(p, xt, spd, refwin, arrs=arrs, nthreads=cores)
    t0 = time.perf_counter()
    solves = 0
    for _ in range(args.steps):
        r = ctwin.step(p, xt, spd, refwin, arrs=arrs, nthreads=cores)
        solves += int(r["n_veh_solves"].sum())
    dt = time.perf_counter() - t0
    val = solves / dt
    line = {
        "impl": "reference", "metric": "agent-QP solves/s", "value": val, "unit": "solves/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args.workload, S, N, 1),
        "cpu_baseline": {"value": val, "unit": "solves/s", "cores": cores, "kind": "port",
                         "sample": f"the full {S}-scenario MPC step x {args.steps} steps on {cores} OpenMP threads "
                                   "(oracle/piadmm_ref.c; CasADi+OSQP not installable)"},
        "e2e": {"value": val, "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def run_reference_coupled(args):
    """CPU arm of the coupled workloads: the C twin's generic-V step on all host cores (a c5 step takes seconds:
    the run is bounded to ~60 s whatever --steps says, and says so)."""
    from oracle import ctwin
    (xt, spd, refwin), V, N, D = coupled_scene(args.workload)
    p = coupled_params(N)
    cores = host_cores()
    t0 = time.perf_counter()
    r = ctwin.step(p, xt, spd, refwin, D=D, nthreads=cores)              # warm-up, also sizes the run
    one = time.perf_counter() - t0
    steps = max(1, min(args.steps, int(60.0 / max(one, 1e-3))))
    t0 = time.perf_counter()
    solves = 0
    for _ in range(steps):
        r = ctwin.step(p, xt, spd, refwin, D=D, nthreads=cores)
        solves += int(r["n_veh_solves"].sum())
    dt = time.perf_counter() - t0
    val = solves / dt
    print(json.dumps({
        "impl": "reference", "metric": "agent-QP solves/s", "value": val, "unit": "solves/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": 1, "ms_per_step": 1e3 * dt / steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": coupled_config(args.workload, V, N, D, 1),
        "cpu_baseline": {"value": val, "unit": "solves/s", "cores": cores, "kind": "port",
                         "sample": f"{steps} coupled MPC steps of the whole scene on {cores} OpenMP threads "
                                   "(oracle/piadmm_ref.c; CasADi+OSQP not installable)"},
        "e2e": {"value": val, "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def workload_config(name, S, N, world):
    return {"workload": f"{name}: {S} independent two-vehicle intersection scenarios per GPU (BASELINE config 4), "
                        f"horizon N={N}, one MPC step to PI-ADMM convergence",
            "scenarios_per_gpu": S, "vehicles_per_scenario": 2, "horizon": N,
            "semantics": "C1 constants (casadi/PI_ADMM_class.py:15-28), dual=pi (MATLAB PI+anti-windup), "
                         "edge=convex, eps_pri=eps_dual=0.1, inner=osqp_default (eps 1e-3, adaptive rho)",
            "parallelism": f"scenario-sharded x{world}, no collective",
            "l2": "L2 flushed (256 MiB write) before every timed step"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS) + sorted(COUPLED))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    from pi_admm_b200.scenarios import c4_scenarios
    from pi_admm_b200.solver import PIADMMSolver
    from pi_admm_b200 import _abi
    import ctypes as C

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; there is no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # stdout carries exactly one JSON line: NCCL prints its version banner to stdout when the communicator is
        # created (NCCL_DEBUG=VERSION/WARN in the environment), so fd 1 points at stderr until that has happened
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    if args.workload in COUPLED:
        return run_coupled(args, rank, local_rank, world)
    S, N = WORKLOADS[args.workload]
    p = bench_params(N)
    # every rank gets its own S scenarios (different seed per rank: weak scaling, no exchange)
    xt, spd, refwin = c4_scenarios(S=S, N=N, seed=4096 + rank)
    sol = PIADMMSolver(p, S=S, V=2, device=local_rank)
    sol.setup(xt, spd, refwin)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=sol.device)
    stream = torch.cuda.current_stream()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing -----------------------------------------------------------
    for _ in range(args.warmup):
        sol.iterate()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    l0 = sol.launch_count
    barrier()
    for a, b in evs:
        flush.fill_(1)                 # L2 flush, outside the event pair
        a.record(stream)
        sol.iterate()                  # ONE fused persistent kernel: all outer iterations of the step
        b.record(stream)
    barrier()
    launches = sol.launch_count - l0
    step_ms = [a.elapsed_time(b) for a, b in evs]
    total_ms = sum(step_ms)
    out = sol.to_numpy()
    n_veh = int(out["n_veh_solves"].sum()); n_edge = int(out["n_edge_solves"].sum())
    veh_inner = int(out["veh_inner"].sum()); edge_inner = int(out["edge_inner"].sum())
    n_outer = out["n_outer"]

    # ---- end to end through the host-buffer C-ABI entry point ------------------------------
    io = sol.alloc_host_io()
    io["xt"].copy_(torch.from_numpy(xt)); io["spd"].copy_(torch.from_numpy(spd)); io["refwin"].copy_(torch.from_numpy(refwin))
    # (a) every step uploads xt, spd and the reference window (piadmm_step_host): reported as e2e.full_upload
    h2d_full, d2h = sol.host_io_bytes(io)
    for _ in range(args.warmup):
        sol.step_host_tensors(io)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sol.step_host_tensors(io)      # synchronous: returns after the D2H copies completed
    torch.cuda.synchronize()
    e2e_full_s = time.perf_counter() - t0
    barrier()
    # (b) the reference's own data flow: ref_traj is built once (casadi/PI_ADMM_class.py:31-37) and stays
    # resident; per MPC step only the plant state xt goes in (piadmm_step_host_at) -> the headline e2e
    from pi_admm_b200.scenarios import c4_reference
    sol.set_reference(c4_reference(S=S, seed=4096 + rank))
    h2d, _ = sol.host_io_bytes(io, resident_reference=True)
    for _ in range(args.warmup):
        sol.step_host_at(12, io)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sol.step_host_at(12, io)       # synchronous: returns after the D2H copies completed
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    barrier()
    for k in ("u", "pos", "curr_iter", "n_outer"):      # same step, same results as the device-resident run
        assert np.array_equal(io[k].numpy(), out[k]), f"e2e output {k} differs from the device-resident run"
    clocks = sampler.stop() if rank == 0 else None

    # ---- max over ranks, sums of work ------------------------------------------------------
    stats = torch.tensor([total_ms, e2e_s * 1e3, e2e_full_s * 1e3], dtype=torch.float64, device=sol.device)
    work = torch.tensor([n_veh, n_edge, veh_inner, edge_inner], dtype=torch.float64, device=sol.device)
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
        dist.all_reduce(work, op=dist.ReduceOp.SUM)
    total_ms, e2e_ms, e2e_full_ms = [float(x) for x in stats.tolist()]
    g_veh, g_edge, g_vi, g_ei = [float(x) for x in work.tolist()]

    if rank == 0:
        ms_per_step = total_ms / args.steps
        value = g_veh / (ms_per_step * 1e-3)
        e2e_value = g_veh / (e2e_ms / args.steps * 1e-3)
        # roofline of the dominant (only) kernel k_fused_v2, per launch on this rank
        flops = veh_inner * flops_per_inner_iter(N, False) + edge_inner * flops_per_inner_iter(N, True)
        kern_s = (sum(step_ms) / args.steps) * 1e-3
        tf = C.c_double(0); pm = C.c_double(0)
        rc = sol.lib.piadmm_fp64_peak(local_rank, 200000, C.byref(tf), C.byref(pm))
        fp64_peak = float(tf.value) if rc == 0 else None
        achieved = flops / kern_s / 1e12
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        abytes = algorithmic_bytes(N, n_veh, n_edge, S)
        traffic = None
        try:   # dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` capture of this kernel
            tj = json.load(open(os.path.join(ROOT, "profiles", "r1_traffic.json")))
            if args.workload == "c4":
                traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
        except Exception:
            pass
        roofline = {"bound": "fp64", "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s",
                    "frac": (achieved / fp64_peak) if fp64_peak else None, "traffic": traffic,
                    "kernel": "k_fused_v2", "peak_source": "DFMA micro-benchmark of this repo (piadmm_fp64_peak), "
                    "MEASURED_PEAKS.json has no fp64 figure",
                    "flops_per_launch": flops, "inner_iters_per_launch": veh_inner + edge_inner,
                    "hbm": {"achieved": abytes / kern_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                            "frac": abytes / kern_s / 1e9 / hbm_peak, "algorithmic_bytes": abytes,
                            "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"}}
        line = {
            "metric": "agent-QP solves/s", "value": value, "unit": "solves/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args.workload, S, N, world),
            "e2e": {"value": e2e_value, "unit": "solves/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms / args.steps,
                    "api": "piadmm_step_host_at (C ABI, pinned host buffers, copies inside; reference trajectories "
                           "resident on the device as in the reference's PI_ADMM_CASADI.ref_traj, per step H2D = plant "
                           "state xt + step index, D2H = u, pos, xt_next, curr_iter, n_outer)",
                    "full_upload": {"value": g_veh / (e2e_full_ms / args.steps * 1e-3), "ms_per_step": e2e_full_ms / args.steps,
                                    "h2d_bytes_per_step": h2d_full, "d2h_bytes_per_step": d2h,
                                    "api": "piadmm_step_host (xt, spd and the reference window uploaded every step)"}},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roofline,
            "work_per_step": {"agent_qp_solves": g_veh, "edge_solves": g_edge, "veh_inner_iters": g_vi,
                              "edge_inner_iters": g_ei, "outer_iters_mean": float(n_outer.mean()),
                              "outer_iters_max": int(n_outer.max()),
                              "ms_per_mpc_step": ms_per_step},
        }
        if not args.no_cpu_baseline and world == 1:       # reported at N=1 only (rank 0)
            line["cpu_baseline"] = cpu_baseline(p, xt, spd, refwin, S, N)
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_coupled(args, rank, local_rank, world):
    """--workload c3 / c5: one coupled scene through the phase kernels (generic V).  world > 1: vehicles sharded over the
    ranks (parallel.VehicleShardedSolver).  Timing: CUDA events around the whole coupled step, max over ranks."""
    import ctypes as C
    import torch
    import torch.distributed as dist
    from pi_admm_b200.parallel import VehicleShardedSolver
    from pi_admm_b200.solver import PIADMMSolver
    (xt, spd, refwin), V, N, D = coupled_scene(args.workload)
    p = coupled_params(N)
    if world > 1:
        vs = VehicleShardedSolver(p, V, D, device=local_rank)
        sol, step = vs.sol, vs.iterate
    else:
        sol = PIADMMSolver(p, S=1, V=V, D=D, device=local_rank)
        step = sol.iterate
    sol.setup(xt, spd, refwin)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=sol.device)
    stream = torch.cuda.current_stream()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = sol.launch_count
    step_ms = []
    for _ in range(args.steps):
        flush.fill_(1)                                   # L2 flush, outside the event pair
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        step()                                           # all outer iterations of the coupled MPC step
        b.record(stream)
        torch.cuda.synchronize()
        step_ms.append(a.elapsed_time(b))
    launches = sol.launch_count - l0
    out = sol.to_numpy()
    # ---- end to end (world == 1): host arrays in, step, plant step, results back on the host
    e2e = None
    if world == 1:
        pin = [torch.from_numpy(np.ascontiguousarray(x)).pin_memory() for x in (xt, spd, refwin)]
        keys = ("u", "pos", "xt_next", "n_outer", "curr_iter")
        host = {k: torch.empty_like(sol.buf[k], device="cpu").pin_memory() for k in keys}

        def e2e_step():
            sol.setup(*pin)
            sol.iterate()
            sol.plant_step()
            for k in keys:
                host[k].copy_(sol.buf[k], non_blocking=True)
            torch.cuda.synchronize()
        for _ in range(args.warmup):
            e2e_step()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
        assert np.array_equal(host["u"].numpy(), out["u"]) and np.array_equal(host["n_outer"].numpy(), out["n_outer"])
        e2e = {"ms_per_step": e2e_ms, "h2d": sum(t.numel() * t.element_size() for t in pin),
               "d2h": sum(t.numel() * t.element_size() for t in host.values())}
    clocks = sampler.stop() if rank == 0 else None
    stats = torch.tensor([sum(step_ms)], dtype=torch.float64, device=sol.device)
    # every rank books the solves of the vehicles / edges it owns... x-updates: owned vehicles only
    work = torch.tensor([float(out["n_veh_solves"][0]), float(out["veh_inner"][0])], dtype=torch.float64, device=sol.device)
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
        dist.all_reduce(work, op=dist.ReduceOp.SUM)
    if rank == 0:
        ms = float(stats[0]) / args.steps
        n_veh, veh_inner = float(work[0]), float(work[1])
        n_edge, edge_inner = float(out["n_edge_solves"][0]), float(out["edge_inner"][0])   # world > 1: rank 0's share
        flops = veh_inner * flops_per_inner_iter(N, False) + edge_inner * flops_per_inner_iter(N, True)
        tf = C.c_double(0); pm = C.c_double(0)
        rc = sol.lib.piadmm_fp64_peak(local_rank, 200000, C.byref(tf), C.byref(pm))
        peak = float(tf.value) if rc == 0 else None
        ach = flops / (ms * 1e-3) / 1e12
        line = {
            "metric": "agent-QP solves/s", "value": n_veh / (ms * 1e-3), "unit": "solves/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": coupled_config(args.workload, V, N, D, world),
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "fp64", "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                         "frac": ach / peak if peak else None, "traffic": None,
                         "kernel": "whole coupled step (k_x_update + detection + k_z_update + residual kernels)",
                         "note": "latency bound by construction: every outer iteration waits for the slowest vehicle "
                                 "QP and the slowest edge problem of the scene (one dependent chain of inner "
                                 "iterations each); the fraction says how little of the GPU one scene can fill",
                         "peak_source": "DFMA micro-benchmark of this repo (piadmm_fp64_peak)",
                         "flops_per_launch": flops},
            "work_per_step": {"agent_qp_solves": n_veh, "edge_solves": n_edge, "veh_inner_iters": veh_inner,
                              "edge_inner_iters": edge_inner, "outer_iters": int(out["n_outer"][0]),
                              "status": int(out["status"][0]), "ms_per_mpc_step": ms}}
        if e2e:
            line["e2e"] = {"value": n_veh / (e2e["ms_per_step"] * 1e-3), "unit": "solves/s", "ms_per_step": e2e["ms_per_step"],
                           "h2d_bytes_per_step": e2e["h2d"], "d2h_bytes_per_step": e2e["d2h"],
                           "api": "PIADMMSolver.setup (pinned xt, spd, reference window -> device) + iterate + plant_step + "
                                  "D2H of u, pos, xt_next, n_outer, curr_iter into pinned buffers"}
        if not args.no_cpu_baseline and world == 1:
            import __graft_entry__ as ge  # noqa: F401
            from oracle import ctwin
            cores = host_cores()
            t0 = time.perf_counter(); reps = 0; solves = 0
            while reps == 0 or (time.perf_counter() - t0 < 10.0 and reps < 200):
                r = ctwin.step(p, xt, spd, refwin, D=D, nthreads=cores)
                solves += int(r["n_veh_solves"].sum()); reps += 1
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": solves / dt, "unit": "solves/s", "cores": cores, "kind": "port",
                                    "sample": f"{reps} coupled MPC steps of the same scene (~{dt:.1f} s) with {cores} "
                                              "OpenMP threads, oracle/piadmm_ref.c"}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def cpu_baseline(p, xt, spd, refwin, S, N):
    """Oracle C twin on the box's host cores, ~10 s of CPU work on the same workload."""
    import __graft_entry__ as ge  # noqa: F401
    from oracle import ctwin
    arrs = ctwin.alloc(S, 2, 1, N, p.iter_num)
    cores = host_cores()
    ctwin.step(p, xt, spd, refwin, arrs=arrs, nthreads=cores)
    t0 = time.perf_counter(); reps = 0; solves = 0
    while time.perf_counter() - t0 < 10.0 and reps < 2000:
        r = ctwin.step(p, xt, spd, refwin, arrs=arrs, nthreads=cores)
        solves += int(r["n_veh_solves"].sum()); reps += 1
    dt = time.perf_counter() - t0
    return {"value": solves / dt, "unit": "solves/s", "cores": cores, "kind": "port",
            "sample": f"{reps} repetitions of the same {S}-scenario MPC step (~{dt:.1f} s) with {cores} OpenMP threads, "
                      "oracle/piadmm_ref.c"}


if __name__ == "__main__":
    main()
