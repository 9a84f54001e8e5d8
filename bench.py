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
    # capacity scenes (not BASELINE configs): the config-5 generator with 4x / 8x the vehicles, for the multi-GPU strong-scaling
    # measurement of a scene whose first x phase is many waves of one GPU (profiles/r2_scaling.md)
    "c5x4": (16384, 30, 8),
    "c5x8": (32768, 30, 8),
}
C5_GRID = {"c5": (64, 32), "c5x4": (128, 64), "c5x8": (128, 128)}      # (lanes per direction, vehicles per lane)
COUPLED_WHAT = {
    "c3": "c3: ONE 64-vehicle 4-arm intersection (BASELINE config 3)",
    "c5": "c5: ONE 4096-vehicle coupled grid scene, 64+64 lanes x 32 vehicles (BASELINE config 5; generator with the "
          "rejection rule clear=3 m: no pair starts inside dis_thres)",
    "c5x4": "c5x4: ONE 16384-vehicle coupled grid scene, 128+128 lanes x 64 vehicles (the config-5 generator at 4x the vehicles; "
            "a capacity scene, not a BASELINE config)",
    "c5x8": "c5x8: ONE 32768-vehicle coupled grid scene, 128+128 lanes x 128 vehicles (the config-5 generator at 8x the vehicles; "
            "a capacity scene, not a BASELINE config)",
}


def coupled_params(N):
    """Synthetic coupled scenes: dual="pi" (per-directed-edge integrators, clamp +-30 with
    back-calculation), AL term over activated edges, convex edge problem, the reference's outer cap of 100
    (casadi/PI_ADMM_class.py:25).  Gains: kI = 1.5, kP = 0 -- with the reference's own gains (kI = rho = 2, sigmoid
    kP in [2, 3.5], MATLAB :21-23,160) the outer loop does not reach consensus on many-edge scenes (profiles/r2_convergence.md,
    tools/coupled_gain_scan.py): every kP > 0 we scanned slows it down or destabilises it."""
    import pi_admm_b200 as pa
    p = pa.preset("matlab_pi")
    p.update(num_ho=N, beta=10.0, Pnorm=1.0, rho=2.0, eps_pri=0.5, eps_dual=0.5, al_all=0, iter_num=100,
             kI=1.5, theta1=0.0, theta2=0.0)
    return p


def coupled_scene(name):
    from pi_admm_b200.scenarios import c3_intersection, c5_grid
    V, N, D = COUPLED[name]
    if name == "c3":
        return c3_intersection(V=V, N=N), V, N, D
    lanes, per_lane = C5_GRID[name]
    return c5_grid(lanes=lanes, per_lane=per_lane, N=N), V, N, D


def coupled_config(name, V, N, D, world):
    what = COUPLED_WHAT[name]
    return {"workload": f"{what}, horizon N={N}, one coupled MPC step to PI-ADMM convergence (outer cap 100)",
            "vehicles": V, "horizon": N, "neighbour_slots": D,
            "semantics": "dual=pi (per-edge integrators, clamp +-30 + back-calculation; kI=1.5, kP=0), AL over activated edges, "
                         "edge=convex, eps_pri=eps_dual=0.5, inner=osqp_default",
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
    from oracle import ctwin          # builds oracle/libpiadmm_ref.so if missing; the CUDA library is NOT loaded in this arm
    ctwin.build()
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
        # the GPU arm's config at this N (the metric is a throughput: the CPU arm times a bounded SAMPLE of that workload,
        # one rank's 4096 scenarios per step, whatever N is -- said in cpu_baseline.sample)
        "config": workload_config(args.workload, S, N, max(1, args.gpus)),
        "cpu_baseline": {"value": val, "unit": "solves/s", "cores": cores, "kind": "port",
                         "sample": f"one rank's share of the workload ({S} scenarios = one full MPC step) x {args.steps} steps on "
                                   f"{cores} OpenMP threads (oracle/piadmm_ref.c; CasADi+OSQP not installable)"},
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
                        f"horizon N={N}, one MPC step to PI-ADMM convergence (on these inputs: <= 2-3 outer iterations, every "
                        "scenario leaves the loop through the reference's empty-edge rule; see work_per_step.rank0)",
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
    ap.add_argument("--no-capacity", action="store_true", help="skip the 16384-vehicle capacity scene of the `coupled` records")
    ap.add_argument("--no-coupled", action="store_true", help="skip the coupled (config 5 / 3) records of the default line")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    # one slice of the host cores per rank (eight ranks sharing all cores of one NUMA node lost 15 % end to
    # end to each other's graph launches and zero-copy drains); set before torch / CUDA create their threads
    pinned_cores = None
    if world > 1 and os.environ.get("PIADMM_PIN", "1") != "0":
        try:
            cores = sorted(os.sched_getaffinity(0))
            n = len(cores) // world
            if n >= 1:
                pinned_cores = cores[local_rank * n:(local_rank + 1) * n]
                os.sched_setaffinity(0, pinned_cores)
        except (AttributeError, OSError):
            pinned_cores = None

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

    # ---- the coupled scenes (BASELINE configs 5 and 3): config 5 on ALL ranks of this run (vehicle-sharded, strong
    # scaling), config 3 at N=1.  Reported beside the config-4 headline as `coupled`.
    coupled, config1, variants = None, None, None
    if not args.no_coupled:
        coupled = {"c5": measure_coupled("c5", args, rank, local_rank, world)}
        if not args.no_capacity:
            # capacity scene (16384 coupled vehicles, not a BASELINE config): the strong-scaling curve of ONE coupled scene
            # that is large enough to have waves to shard, on ALL ranks of this run (profiles/r2_scaling.md)
            coupled["c5x4"] = measure_coupled("c5x4", args, rank, local_rank, world, with_cpu=False, steps=min(args.steps, 30))
        if world == 1:
            coupled["c3"] = measure_coupled("c3", args, rank, local_rank, world)
            few = max(3, min(args.steps, 5))
            coupled["c5_inner_tight"] = measure_coupled("c5", args, rank, local_rank, world, with_cpu=False, inner="tight", steps=few)
            coupled["c3_inner_tight"] = measure_coupled("c3", args, rank, local_rank, world, with_cpu=False, inner="tight", steps=few)
            coupled["c3_joint_edge"] = measure_joint(args, local_rank)
            config1 = measure_config1(args, local_rank)
            variants = measure_c4_variants(args, local_rank, S, N)

    # ---- max over ranks, sums of work ------------------------------------------------------
    stats = torch.tensor([total_ms, e2e_s * 1e3, e2e_full_s * 1e3], dtype=torch.float64, device=sol.device)
    work = torch.tensor([n_veh, n_edge, veh_inner, edge_inner], dtype=torch.float64, device=sol.device)
    per_rank = None
    if world > 1:
        gathered = [torch.zeros_like(stats) for _ in range(world)]
        dist.all_gather(gathered, stats)                  # per-rank times: how much of the max-over-ranks is spread
        per_rank = [[float(x) / args.steps for x in g.tolist()] for g in gathered]
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
        traffic, traffic_source = None, None
        try:   # dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` capture of this kernel
            tpath = [f for f in ("r2_traffic.json", "r1_traffic.json") if os.path.exists(os.path.join(ROOT, "profiles", f))][0]
            tj = json.load(open(os.path.join(ROOT, "profiles", tpath)))
            if args.workload == "c4":
                traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
                traffic_source = f"profiles/{tpath} (one ncu --set full capture of this kernel), NOT measured in this run"
        except Exception:
            pass
        roofline = {"bound": "fp64", "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s",
                    "frac": (achieved / fp64_peak) if fp64_peak else None, "traffic": traffic,
                    "traffic_source": traffic_source, "kernel": "k_fused_v2", "peak_source": "DFMA micro-benchmark of this repo (piadmm_fp64_peak), "
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
                    "ms_per_step_by_rank": [round(r[1], 5) for r in per_rank] if per_rank else None,
                    "api": "piadmm_step_host_at (C ABI, pinned host buffers, copies inside; reference trajectories "
                           "resident on the device as in the reference's PI_ADMM_CASADI.ref_traj, per step H2D = plant "
                           "state xt + step index, D2H = u, pos, xt_next, curr_iter, n_outer)",
                    "full_upload": {"value": g_veh / (e2e_full_ms / args.steps * 1e-3), "ms_per_step": e2e_full_ms / args.steps,
                                    "h2d_bytes_per_step": h2d_full, "d2h_bytes_per_step": d2h,
                                    "api": "piadmm_step_host (xt, spd and the reference window uploaded every step)"}},
            "gpu_launches": int(launches),
            "host_cores_per_rank": len(pinned_cores) if pinned_cores else host_cores(),
            "clocks": clocks,
            "roofline": roofline,
            "edge_qp_solves_per_s": g_edge / (ms_per_step * 1e-3),
            "work_per_step": {"agent_qp_solves": g_veh, "edge_solves": g_edge, "veh_inner_iters": g_vi,
                              "edge_inner_iters": g_ei, "outer_iters_mean": float(n_outer.mean()),
                              "outer_iters_max": int(n_outer.max()),
                              "ms_per_mpc_step": ms_per_step, "rank0": exit_summary(out, p)},
            "coupled": coupled,
            "config1": config1,
            "c4_variants": variants,
        }
        if not args.no_cpu_baseline and world == 1:       # reported at N=1 only (rank 0)
            line["cpu_baseline"] = cpu_baseline(p, xt, spd, refwin, S, N)
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def exit_summary(out, p):
    """How the scenarios left the outer loop (casadi/main.py:115-118 empty edge list / :174-178 stop test / the cap)."""
    n_outer = np.asarray(out["n_outer"]).reshape(-1)
    status = np.asarray(out["status"]).reshape(-1)
    edge_n = np.asarray(out["edge_n"]).reshape(-1)
    flag = np.asarray(out["flag"]).reshape(-1)
    rk = np.asarray(out["err_rk"]).reshape(len(n_outer), -1)
    sk = np.asarray(out["err_sk"]).reshape(len(n_outer), -1)
    last = np.clip(n_outer - 1, 0, rk.shape[1] - 1)
    rk_l = rk[np.arange(len(n_outer)), last]; sk_l = sk[np.arange(len(n_outer)), last]
    cap = (status & 8) != 0
    never = (flag == 0) & ~cap                                  # no collision ever seen: break in the first iteration
    empty = (edge_n == 0) & (flag != 0) & ~cap                  # the edge list emptied: residuals exactly 0
    stop = ~cap & ~never & ~empty
    hist = {int(k): int(v) for k, v in zip(*np.unique(n_outer, return_counts=True))}
    return {"exit": {"stop_test_with_active_edges": int(stop.sum()), "empty_edge_list": int(empty.sum()),
                     "no_collision_at_all": int(never.sum()), "iteration_cap": int(cap.sum())},
            "converged": bool(not cap.any()), "outer_iters_hist": hist,
            "residual_at_exit": {"err_rk_max": float(rk_l.max()), "err_sk_max": float(sk_l.max()),
                                 "eps_pri": float(p.eps_pri), "eps_dual": float(p.eps_dual)},
            "outer_cap": int(p.iter_num)}


def measure_coupled(name, args, rank, local_rank, world, with_cpu=True, inner="osqp_default", steps=None):
    """One coupled scene (c3 / c5) on `world` ranks: vehicles in the spatial order, sharded by strips
    (parallel.VehicleShardedSolver; world == 1: the same numbering on one GPU).  Device-timed steps (L2 flushed, barrier
    on both sides, max over ranks), end-to-end steps from host arrays, exit / convergence summary, a checksum of the
    results (bit-identical for every world size), and at world == 1 the CPU port timed and used as the checker."""
    import ctypes as C
    import hashlib
    import torch
    import torch.distributed as dist
    from pi_admm_b200.parallel import VehicleShardedSolver
    (xt, spd, refwin), V, N, D = coupled_scene(name)
    p = coupled_params(N)
    if inner == "tight":
        import pi_admm_b200 as pa
        pa.inner_tight(p)
    vs = VehicleShardedSolver(p, V, D, device=local_rank)
    sol = vs.sol
    vs.setup(xt, spd, refwin)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=sol.device)
    stream = torch.cuda.current_stream()
    steps, warmup = max(1, steps or args.steps), max(3, args.warmup)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        vs.iterate()
    barrier()
    l0 = sol.launch_count
    step_ms = []
    for _ in range(steps):
        flush.fill_(1)                                   # L2 flush, outside the event pair
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        vs.iterate()                                     # all outer iterations of the coupled MPC step
        b.record(stream)
        torch.cuda.synchronize()
        step_ms.append(a.elapsed_time(b))
    launches = sol.launch_count - l0
    vs.plant_step()
    out = vs.results()
    raw = sol.to_numpy()
    # ---- end to end: host arrays in (H2D inside setup), step, plant step, results back on the host
    h2d = sum(np.asarray(x).nbytes for x in (xt, spd, refwin))
    keys = ("u", "pos", "xt_next", "n_outer", "curr_iter")
    host = {k: torch.empty_like(sol.buf[k], device="cpu").pin_memory() for k in keys}

    pin_in = [torch.from_numpy(np.ascontiguousarray(np.broadcast_to(x, s_))).pin_memory()
              for x, s_ in ((xt, (1, V, 3)), (spd, (1, V)), (refwin, (1, V, 2, N + 1)))]

    def e2e_step():
        vs.setup(*pin_in)                                # pinned host inputs -> device (3 async copies + device gather)
        vs.iterate()
        vs.plant_step()
        for k in keys:
            host[k].copy_(sol.buf[k], non_blocking=True)
        torch.cuda.synchronize()
    for _ in range(warmup):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        e2e_step()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / steps
    assert np.array_equal(host["u"].numpy(), raw["u"]), "coupled e2e results differ from the device-timed run"
    stats = torch.tensor([sum(step_ms) / steps, e2e_ms], dtype=torch.float64, device=sol.device)
    work = torch.tensor([float(raw["n_veh_solves"][0]), float(raw["veh_inner"][0]), float(raw["n_edge_solves"][0]),
                         float(raw["edge_inner"][0])], dtype=torch.float64, device=sol.device)
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
        dist.all_reduce(work, op=dist.ReduceOp.SUM)
    ms, e2e_ms = float(stats[0]), float(stats[1])
    n_veh, veh_inner, n_edge, edge_inner = [float(x) for x in work.tolist()]
    rec = None
    if rank == 0:
        flops = veh_inner * flops_per_inner_iter(N, False) + edge_inner * flops_per_inner_iter(N, True)
        tf = C.c_double(0); pm = C.c_double(0)
        rc = sol.lib.piadmm_fp64_peak(local_rank, 200000, C.byref(tf), C.byref(pm))
        peak = float(tf.value) if rc == 0 else None
        ach = flops / (ms * 1e-3) / 1e12
        sha = hashlib.sha256(np.ascontiguousarray(out["u"]).tobytes() + np.ascontiguousarray(out["pos"]).tobytes()).hexdigest()
        stored = None
        try:
            stored = json.load(open(os.path.join(ROOT, "profiles", "r2_coupled_checksums.json"))).get(name)
        except Exception:
            pass
        what = COUPLED_WHAT[name]
        rec = {"workload": f"{what}, horizon N={N}, one coupled MPC step to PI-ADMM convergence (outer cap {p.iter_num})",
               "vehicles": V, "horizon": N, "neighbour_slots": D, "n_gpus": world, "scaling": "strong",
               "semantics": "dual=pi (per-edge integrators, clamp +-30 + back-calculation; kI=1.5, kP=0: profiles/r2_convergence.md), "
                            "AL over activated edges, edge=convex, eps_pri=eps_dual=0.5, inner=" + inner,
               "parallelism": "single GPU, persistent cooperative kernel" if world == 1 else
                              f"vehicles sharded x{world} by spatial strips; exchange={vs.exchange}" +
                              (f" [{vs.exchange_note}]" if vs.exchange_note else ""),
               "exchange": vs.exchange, "ms_per_step": ms, "steps": steps, "warmup": warmup,
               "agent_qp_solves_per_s": n_veh / (ms * 1e-3), "edge_qp_solves_per_s": n_edge / (ms * 1e-3),
               "gpu_launches_per_step": launches / steps,
               "l2": "L2 flushed (256 MiB write) before every timed step",
               "e2e": {"ms_per_step": e2e_ms, "value": n_veh / (e2e_ms * 1e-3), "unit": "solves/s", "h2d_bytes_per_step": int(h2d),
                       "d2h_bytes_per_step": int(sum(t.numel() * t.element_size() for t in host.values())),
                       "api": "VehicleShardedSolver.setup (pinned host xt, spd, reference window -> device) + iterate + plant_step + D2H "
                              "of u, pos, xt_next, n_outer, curr_iter into pinned buffers"},
               "work_per_step": {"agent_qp_solves": n_veh, "agent_qp_solves_reference_loop": float(V * int(out["n_outer"][0])),
                                 "x_memo": "vehicle QPs whose inputs did not change since their last solve are not re-solved "
                                           "(exact; the reference loop re-solves all V in every outer iteration)",
                                 "edge_solves_incl_redundant_cut_edges": n_edge,
                                 "veh_inner_iters": veh_inner, "edge_inner_iters": edge_inner,
                                 "outer_iters": int(out["n_outer"][0]), "status": int(out["status"][0])},
               "roofline": {"bound": "fp64", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak if peak else None,
                            "flops_per_step": flops, "kernel": "k_coop (whole coupled step)",
                            "note": "one scene = a few hundred microseconds of dependent phases; the fraction says how little of "
                                    "the GPU one coupled scene can fill, the batched config-4 line is the throughput figure",
                            "peak_source": "DFMA micro-benchmark of this repo (piadmm_fp64_peak)"},
               "result_sha256": sha, "matches_stored_checksum": (sha == stored) if stored else None}
        rec.update(exit_summary(out, p))
        if with_cpu and world == 1 and not args.no_cpu_baseline:
            from oracle import ctwin
            cores = host_cores()
            t0 = time.perf_counter(); reps = 0; solves = 0
            pm_ = vs.perm                  # the same vehicle numbering as the kernels (the residual's veh1 side depends on it)
            xt_p, spd_p, rw_p = (np.ascontiguousarray(np.asarray(x_)[:, pm_]) for x_ in (xt, np.broadcast_to(spd, (1, V)), refwin))
            while reps == 0 or (time.perf_counter() - t0 < 8.0 and reps < 200):
                r = ctwin.step(p, xt_p, spd_p, rw_p, D=D, nthreads=cores)
                solves += int(r["n_veh_solves"].sum()); reps += 1
            dt = time.perf_counter() - t0
            rel = lambda a_, b_: float(np.abs(a_ - b_).max() / max(1.0, np.abs(b_).max()))
            rec["cpu_baseline"] = {"value": solves / dt, "unit": "solves/s", "ms_per_step": 1e3 * dt / reps, "cores": cores,
                                   "kind": "port", "sample": f"{reps} coupled MPC steps of the same scene (~{dt:.1f} s) with "
                                                             f"{cores} OpenMP threads, oracle/piadmm_ref.c"}
            rec["parity_vs_cpu_port"] = {"outer_iters_equal": bool(np.array_equal(raw["n_outer"], r["n_outer"])),
                                         "edge_lists_equal": bool(np.array_equal(raw["edge_ij"], r["edge_ij"]) and np.array_equal(raw["nbr"], r["nbr"])),
                                         "solve_counts_equal": bool(np.array_equal(raw["n_veh_solves"], r["n_veh_solves"]) and np.array_equal(raw["veh_inner"], r["veh_inner"]) and np.array_equal(raw["edge_inner"], r["edge_inner"])),
                                         "u_rel_err": rel(raw["u"], r["u"]), "pos_rel_err": rel(raw["pos"], r["pos"]),
                                         "err_rk_rel_err": rel(raw["err_rk"], r["err_rk"]), "err_sk_rel_err": rel(raw["err_sk"], r["err_sk"])}
    barrier()
    vs.close()
    return rec


def measure_config1(args, local_rank):
    """BASELINE config 1, the reference's own case: ONE two-vehicle scenario, all 50 - num_ho = 35 receding-horizon steps
    of casadi/main.py:43-201, in the script's semantics (python_current) and in dual=pi, both inner modes.  Per MPC step
    the latency of the host-buffer call (piadmm_step_host_at: plant state in, u / pos / xt_next / counts out, synchronous)
    -- mean and max over the steps with an active edge, as SURVEY 8(d) asks -- and the whole loop on the device."""
    import torch
    from pi_admm_b200 import main_intersection
    from pi_admm_b200.PI_ADMM_class import PI_ADMM_CASADI
    out = {}
    for sem in ("python_current", "matlab_pi"):
        for inner in ("osqp_default", "tight"):
            P = PI_ADMM_CASADI(semantics=sem, inner=inner)
            N, V = int(P.param.num_ho), int(P.param.num_veh)
            sol = P.solver(1)
            spd = np.asarray(P.param.spd, dtype=np.float64)[None]
            sol.setup(np.zeros((1, V, 3)), spd, np.zeros((1, V, 2, N + 1)), rho=float(P.param.rho))
            sol.set_reference(np.ascontiguousarray(P.ref_traj).reshape(1, V, 2, -1))
            io = sol.alloc_host_io()
            n_steps = P.ref_traj.shape[1] - N
            best = None
            for rep in range(3):                     # the third repetition is reported (clocks, caches, graph warm)
                io["xt"].copy_(torch.tensor([[[-10.0, 0.0, 0.0], [0.0, 20.0, -np.pi / 2]]], dtype=torch.float64))
                ms, active, outer = [], [], []
                for step in range(n_steps):
                    t0 = time.perf_counter()
                    sol.step_host_at(step, io)
                    ms.append((time.perf_counter() - t0) * 1e3)
                    outer.append(int(io["n_outer"][0]))
                    active.append(outer[-1] > 1)
                    io["xt"].copy_(io["xt_next"])
                best = (ms, active, outer)
            ms, active, outer = best
            act = [m for m, a in zip(ms, active) if a] or [0.0]
            # the same loop without leaving the device
            t_dev = None
            for rep in range(3):
                sol.setup(np.array([[[-10.0, 0.0, 0.0], [0.0, 20.0, -np.pi / 2]]]), spd, np.zeros((1, V, 2, N + 1)), rho=float(P.param.rho))
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); sol.run_closed_loop(0, n_steps); e1.record(); torch.cuda.synchronize()
                t_dev = e0.elapsed_time(e1)
            out[f"{sem}/{inner}"] = {"steps": n_steps, "steps_with_active_edge": int(sum(active)),
                                     "ms_per_step_mean_active": float(np.mean(act)), "ms_per_step_max_active": float(np.max(act)),
                                     "ms_per_step_mean_all": float(np.mean(ms)), "ms_total": float(np.sum(ms)),
                                     "outer_iters_total": int(sum(outer)), "outer_iters_max": int(max(outer)),
                                     "device_loop_ms_total": float(t_dev), "device_loop_ms_per_step": float(t_dev / n_steps)}
            P._solver.close()
    return {"workload": "c1: casadi/main.py two-vehicle intersection, 35 MPC steps, ONE scenario (BASELINE config 1); latency of the "
                        "synchronous host-buffer call per step (wall clock) and of the on-device loop (CUDA events)",
            "runs": out}


def measure_c4_variants(args, local_rank, S, N):
    """The headline workload in inner="tight" (parity mode: results are solver independent) and as a CLOSED LOOP (20
    consecutive MPC steps on the device: inputs, active sets and the scheduling order change from step to step, the
    K^-1 cache stays valid because the speeds do not change) -- ADVICE r1: the replayed step is the best case."""
    import torch
    import pi_admm_b200 as pa
    from pi_admm_b200.scenarios import c4_scenarios, c4_reference
    from pi_admm_b200.solver import PIADMMSolver
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local_rank}")
    steps = max(3, min(args.steps, 10))
    out = {}
    p = pa.inner_tight(bench_params(N))
    xt, spd, refwin = c4_scenarios(S=S, N=N)
    sol = PIADMMSolver(p, S=S, V=2, device=local_rank)
    sol.setup(xt, spd, refwin)
    for _ in range(3):
        sol.iterate()
    ms = []
    for _ in range(steps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); sol.iterate(); b.record(); torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    o = sol.to_numpy()
    out["inner_tight"] = {"ms_per_step": float(np.mean(ms)), "agent_qp_solves_per_s": float(o["n_veh_solves"].sum() / (np.mean(ms) * 1e-3)),
                          "edge_qp_solves_per_s": float(o["n_edge_solves"].sum() / (np.mean(ms) * 1e-3)),
                          "veh_inner_iters": int(o["veh_inner"].sum()), "edge_inner_iters": int(o["edge_inner"].sum()),
                          "inner": "eps 1e-9, max_iter 20000", "steps": steps}
    sol.close()
    p = bench_params(N)
    first, n = 5, 20
    xt, spd, _ = c4_scenarios(S=S, N=N, step=first)
    sol = PIADMMSolver(p, S=S, V=2, device=local_rank)
    sol.set_reference(c4_reference(S=S, T=max(50, first + n + N + 1)))   # every window [step, step + N] of the run inside it
    t = None
    for rep in range(3):
        sol.setup(xt, spd, np.zeros((S, 2, 2, N + 1)))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); h = sol.run_closed_loop(first, n); e1.record(); torch.cuda.synchronize()
        t = e0.elapsed_time(e1)
    no = h["n_outer_his"].cpu().numpy()
    out["closed_loop"] = {"mpc_steps": n, "first_step": first, "ms_per_mpc_step": float(t / n), "ms_total": float(t),
                          "outer_iters_mean": float(no.mean()), "outer_iters_max": int(no.max()),
                          "outer_iters_hist": {int(k): int(v) for k, v in zip(*np.unique(no, return_counts=True))},
                          "note": "piadmm_run_closed_loop: window gather, outer loop, plant step and histories on the device, "
                                  "no host round trip per MPC step; no L2 flush between steps (one asynchronous call)"}
    sol.close()
    return out


def measure_joint(args, local_rank):
    """The joint N-vehicle RSU edge problem (piadmm_solve_joint: one thread-block cluster per scene) on the
    config-3 scene: the pairs that collide after the first vehicle QPs (at most 8 per vehicle), one joint solve per timed
    step; also a batch of 16 copies of the scene in one launch (one cluster each).  Device-timed, L2 flushed."""
    import ctypes as C
    import torch
    from pi_admm_b200.solver import PIADMMSolver
    from pi_admm_b200.scenarios import c3_intersection
    V, N = 64, 30
    p = coupled_params(N).update(x_memo=0)
    xt, spd, refwin = c3_intersection(V=V, N=N)
    out = {"workload": "ONE joint edge problem over the 64 vehicles of the config-3 intersection (N = 30): hinge over the pairs that "
                       "collide after the first vehicle QPs, <= 8 per vehicle; matlab_old_files/ADMM_fmincon_multi_veh_intesection_"
                       "static.m:250-291 in its convex statement",
           "kernel": "k_joint<32>: 8 CTAs x 256 threads per scene (one cluster), t~ exchanged through distributed shared memory, "
                     "one cluster barrier per inner iteration"}
    for S in (1, 16):
        sol = PIADMMSolver(p, S=S, V=V, D=1, device=local_rank)
        sol.setup(np.broadcast_to(xt, (S, V, 3)).copy(), np.broadcast_to(spd, (S, V)).copy(), np.broadcast_to(refwin, (S, V, 2, N + 1)).copy())
        sol.begin_step(); sol.x_update()
        e_cap = V * (V - 1) // 2
        en = torch.zeros(S, dtype=torch.int32, device=sol.device); eij = torch.zeros((S, e_cap, 2), dtype=torch.int32, device=sol.device)
        rc = sol.lib.piadmm_edge_list_static(C.c_void_p(sol.buf["pos"].data_ptr()), S, V, N, float(p.dis_thres), int(p.thr_squared), e_cap,
                                             C.c_void_p(en.data_ptr()), C.c_void_p(eij.data_ptr()), sol._stream())
        assert rc == 0
        n0 = int(en[0]); lst = sorted(map(tuple, eij[0, :n0].cpu().numpy().tolist()))
        deg = [0] * V; pairs = []
        for (i, j) in lst:
            if deg[i] < 8 and deg[j] < 8:
                pairs.append((i, j)); deg[i] += 1; deg[j] += 1
        pij = np.broadcast_to(np.array(pairs, np.int32), (S, len(pairs), 2)).copy(); pn = np.full(S, len(pairs), np.int32)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=sol.device)
        for _ in range(3):
            u, it, st = sol.solve_joint(pn, pij)
        reps = max(5, min(args.steps, 20)); ms = []
        for _ in range(reps):
            flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); u, it, st = sol.solve_joint(pn, pij); b.record(); torch.cuda.synchronize()
            ms.append(a.elapsed_time(b))
        iters = int(it[0]); t = float(np.mean(ms))
        # per inner iteration and vehicle: the N x N mat-vec, the 4N-2 bound rows and one position-row copy per incident pair
        flops = iters * S * (V * (2 * N * N + 12 * (4 * N - 2) + 6 * N) + 2 * len(pairs) * 12 * N)
        out[f"S{S}"] = {"scenes_per_launch": S, "pairs": len(pairs), "colliding_pairs_found": n0, "inner_iters": iters, "status": int(st[0]),
                        "ms_per_joint_solve": t, "us_per_inner_iter": t * 1e3 / max(iters, 1), "joint_solves_per_s": S / (t * 1e-3),
                        "algorithmic_tflops": flops / (t * 1e-3) / 1e12, "reps": reps}
        sol.close()
    return out


def run_coupled(args, rank, local_rank, world):
    """--workload c3 / c5: the coupled record as its own line (`ms_per_step` = the BASELINE metric's second half)."""
    import torch.distributed as dist
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    rec = measure_coupled(args.workload, args, rank, local_rank, world)
    clocks = sampler.stop() if rank == 0 else None
    if rank == 0:
        line = {"metric": "agent-QP solves/s", "value": rec["agent_qp_solves_per_s"], "unit": "solves/s", "n_gpus": world,
                "steps": rec["steps"], "warmup": rec["warmup"], "ms_per_step": rec["ms_per_step"], "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {k: rec[k] for k in ("workload", "vehicles", "horizon", "neighbour_slots", "semantics", "parallelism", "l2")},
                "gpu_launches": int(round(rec["gpu_launches_per_step"] * rec["steps"])), "clocks": clocks}
        for k in ("e2e", "roofline", "work_per_step", "exit", "converged", "outer_iters_hist", "residual_at_exit", "outer_cap",
                  "result_sha256", "matches_stored_checksum", "cpu_baseline", "parity_vs_cpu_port", "exchange"):
            if k in rec:
                line[k] = rec[k]
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
