"""Regenerates the round's files under profiles/ from what a profiling gpurun call (tools/gpu_call_profile.sh <tag>) left in
gpurun_out/:  bench_<tag>.json, bench_<tag>_ref.json, bench_<tag>_{c3,c5,n30}.json, launches_<tag>.csv, launches_<tag>_c5.csv,
prof_<tag>.ncu-rep (k_fused_v2), prof_<tag>_coop.ncu-rep (k_coop).

    python tools/make_profiles.py r2a [r2]        # tag of the call, prefix of the files written (default r2)
"""
import csv
import json
import os
import shutil
import subprocess
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P, S = (os.path.join(ROOT, d) for d in ("gpurun_out", "profiles", "scratch"))
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
KEYS = ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__warps_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active", "gpu__time_duration.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__grid_size",
        "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum")


def ncu_summary(rep, out_txt, tag):
    raw, src = os.path.join(S, f"{tag}_raw.csv"), os.path.join(S, f"{tag}_src.csv")
    for page, out in (("raw", raw), ("source", src)):
        with open(out, "w") as f:
            subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], stdout=f, stderr=subprocess.DEVNULL, check=True)
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), raw, src, out_txt], check=True,
                   stdout=subprocess.DEVNULL)
    rows = list(csv.reader(open(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    val = lambda k: float(data[0][hdr.index(k)].replace(",", ""))
    m = {k: val(k) for k in KEYS if k in hdr}
    m["dram_read"] = val("dram__bytes_read.sum") * SCALE[units[hdr.index("dram__bytes_read.sum")]]
    m["dram_write"] = val("dram__bytes_write.sum") * SCALE[units[hdr.index("dram__bytes_write.sum")]]
    tail = "\n".join(l for l in open(out_txt).read().splitlines() if l.startswith(("opcode mix", "warp stall", "hot straight", "  [")))
    return m, tail


def launch_table(path, step_kernels):
    lr = [r for r in csv.reader(open(path)) if len(r) > 5]
    ki, vi = lr[0].index("Kernel Name"), lr[0].index("Metric Value")
    d = defaultdict(list)
    for r in lr[1:]:
        try:
            d[r[ki][:70]].append(float(r[vi].replace(",", "")))
        except ValueError:
            pass
    step = [k for k in d if any(s in k for s in step_kernels)]
    per_launch = max(len(d[k]) for k in step) if step else 1
    tot = sum(sum(d[k]) for k in step)
    rows = [f"| `{k}` | {len(v)} | {sum(v) / len(v) / 1e3:.1f} | {'%.3f' % (sum(v) / tot) if k in step else '(not part of a step)'} |"
            for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1]))]
    return "\n".join(rows), tot / per_launch / 1e3


def coupled_row(name, c):
    ex = [k for k, v in c["exit"].items() if v][0]
    return (f"| {name} | {c['vehicles']} | {c['ms_per_step']:.3f} | {c['e2e']['ms_per_step']:.3f} | {c['work_per_step']['outer_iters']} | "
            f"{ex} | {c['converged']} | {c['residual_at_exit']['err_rk_max']:.3g} / {c['residual_at_exit']['err_sk_max']:.3g} "
            f"(eps {c['residual_at_exit']['eps_pri']}) | {int(c['work_per_step']['agent_qp_solves'])} / "
            f"{int(c['work_per_step']['edge_solves_incl_redundant_cut_edges'])} | {c['gpu_launches_per_step']:.0f} | "
            f"{c.get('cpu_baseline', {}).get('ms_per_step', float('nan')):.1f} |")


def main(tag, R="r2"):
    os.makedirs(S, exist_ok=True)
    cp = lambda a, b: shutil.copy(os.path.join(G, a), os.path.join(P, b)) if os.path.exists(os.path.join(G, a)) and os.path.getsize(os.path.join(G, a)) else None
    cp(f"launches_{tag}.csv", f"{R}_launches_bench.csv")
    cp(f"launches_{tag}_c5.csv", f"{R}_launches_c5.csv")
    cp(f"bench_{tag}.json", f"{R}_bench_1gpu.json")
    cp(f"bench_{tag}_ref.json", f"{R}_bench_reference_arm.json")
    cp(f"bench_{tag}_n30.json", f"{R}_bench_1gpu_n30.json")
    cp(f"bench_{tag}_c3.json", f"{R}_bench_c3.json")
    cp(f"bench_{tag}_c5.json", f"{R}_bench_c5.json")
    for w in ("c5x4", "c5x8"):
        cp(f"bench_{tag}_{w}.json", f"{R}_bench_{w}.json")
    for w in ("c3", "c5", "c5x4", "c5x8"):
        cp(f"coop_phases_{w}_{tag}.txt", f"{R}_coop_phases_{w}_1gpu.txt")
    m, tail = ncu_summary(os.path.join(G, f"prof_{tag}.ncu-rep"), os.path.join(P, f"{R}_k_fused_v2_ncu_full.txt"), tag)
    mc, tailc = ncu_summary(os.path.join(G, f"prof_{tag}_coop.ncu-rep"), os.path.join(P, f"{R}_k_coop_ncu_full.txt"), tag + "_coop")
    mj, tailj = (None, "")
    if os.path.exists(os.path.join(G, f"prof_{tag}_joint.ncu-rep")):
        mj, tailj = ncu_summary(os.path.join(G, f"prof_{tag}_joint.ncu-rep"), os.path.join(P, f"{R}_k_joint_ncu_full.txt"), tag + "_joint")
    json.dump({"kernel": "k_fused_v2<16>", "dram_bytes_read": int(m["dram_read"]), "dram_bytes_write": int(m["dram_write"]),
               "source": f"profiles/{R}_k_fused_v2_ncu_full.txt (ncu --set full, 1 launch; the writes are mostly the "
                         "write-back of the 256 MiB L2-flush buffer)"}, open(os.path.join(P, f"{R}_traffic.json"), "w"))
    b = json.load(open(os.path.join(P, f"{R}_bench_1gpu.json")))
    r = json.load(open(os.path.join(P, f"{R}_bench_reference_arm.json")))
    lines, step_us = launch_table(os.path.join(P, f"{R}_launches_bench.csv"), ("k_fused", "k_order", "k_refwin", "k_plant"))
    lines5, step5_us = launch_table(os.path.join(P, f"{R}_launches_c5.csv"), ("k_coop", "k_cand", "k_disc", "k_plant"))
    w = b["work_per_step"]
    cpl = b.get("coupled", {})
    c1 = b.get("config1", {}).get("runs", {})
    n30 = json.load(open(os.path.join(P, f"{R}_bench_1gpu_n30.json"))) if os.path.exists(os.path.join(P, f"{R}_bench_1gpu_n30.json")) else None
    c1rows = "\n".join(f"| {k} | {v['steps_with_active_edge']} / {v['steps']} | {v['ms_per_step_mean_active']:.3f} | {v['ms_per_step_max_active']:.3f} | "
                       f"{v['ms_total']:.2f} | {v['device_loop_ms_per_step']:.3f} | {v['outer_iters_total']} |" for k, v in c1.items())
    cv = b.get("c4_variants", {})
    jt = cpl.get("c3_joint_edge")
    joint_txt = ""
    if jt:
        joint_txt = ("\n### Joint N-vehicle RSU edge problem (SURVEY 8(f)-4, `piadmm_solve_joint`, cluster kernel `k_joint<32>`)\n\n"
                     + jt["workload"] + ".\n\n| scenes per launch | pairs | inner iterations | ms per joint solve | us per inner iteration | joint solves/s | algorithmic TFLOP/s |\n|---|---|---|---|---|---|---|\n"
                     + "\n".join(f"| {r['scenes_per_launch']} | {r['pairs']} | {r['inner_iters']} | {r['ms_per_joint_solve']:.3f} | {r['us_per_inner_iter']:.2f} | {r['joint_solves_per_s']:.0f} | {r['algorithmic_tflops']:.3f} |"
                                 for k, r in jt.items() if k.startswith("S"))
                     + (f"\n\nncu (`{R}_k_joint_ncu_full.txt`, one launch, {mj['gpu__time_duration.sum']:.0f} us, {mj['launch__registers_per_thread']:.0f} registers, grid {mj['launch__grid_size']:.0f} CTAs): FP64 pipe {mj['sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active']:.1f} %, issue slots {mj['smsp__issue_active.avg.pct_of_peak_sustained_active']:.1f} %.\n" + tailj if mj else "")
                     + "\n\nOne inner iteration = one mat-vec per vehicle + one HARDWARE CLUSTER BARRIER + <= 8 distributed-shared-memory reads per lane: the iteration time is the barrier + DSMEM round trip, not arithmetic (the pairwise z phase of the same scene takes ~15 us per OUTER iteration; the joint problem is the alternative z-step of the reference's old flow, not the latency path).\n")
    open(os.path.join(P, "README.md"), "w").write(f"""# profiles/ — round 2 evidence (B200, sm_100a, driver 580.159, CUDA 12.9)

Regenerate with `python tools/make_profiles.py <tag>` after `gpurun -- 'bash tools/gpu_call_profile.sh <tag>'`.
Round-1 files (`r1_*`) are kept as history; `r1_history.md` / `r2_history.md` say what each optimisation step bought.

Files of this round
* `{R}_bench_1gpu.json` — `python bench.py --steps 100 --warmup 5`: the default line (config 4, `value` / `e2e` / `roofline` /
  `cpu_baseline`) with the `coupled` records (configs 5 and 3, osqp_default and tight), `config1` (the 35-step run of
  `casadi/main.py`, 4 semantics x inner modes) and `c4_variants` (tight mode, 20-step closed loop on the device).
  `{R}_bench_1gpu_n30.json` — `--workload c4n30`; `{R}_bench_c3.json`, `{R}_bench_c5.json` — `--workload c3|c5` as own lines.
* `{R}_bench_1gpu_final_20steps.json` — `python bench.py --steps 20 --warmup 3` on the LAST build: `launch_coop` keeps the register
  build of `k_coop` at tight inner tolerances, so `coupled.c5_inner_tight` is 2.71 ms there (3.33 ms in `{R}_bench_1gpu.json`, which
  ran that record on the occupancy build one commit earlier); every other record agrees between the two files.
* `{R}_bench_reference_arm.json` — `python bench.py --impl reference --steps 20 --warmup 2` (oracle C twin, all host cores).
* `{R}_launches_bench.csv`, `{R}_launches_c5.csv` — `ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv` of
  `python bench.py --steps 3 --warmup 3 --no-cpu-baseline [--no-coupled | --workload c5]` (every launch with its device time).
* `{R}_k_fused_v2_ncu_full.txt`, `{R}_k_coop_ncu_full.txt` — summaries (tools/ncu_summary.py) of one
  `ncu --set full --clock-control none --import-source on -k regex:<kernel> -s 3 -c 1` capture each; `{R}_traffic.json` = DRAM bytes
  of the fused kernel.
* `{R}_bench_c5x4.json`, `{R}_bench_c5x8.json` (+ `_2gpu`, `_8gpu`) — the capacity scenes (16384 / 32768 coupled vehicles,
  `--workload c5x4|c5x8`), `{R}_bench_c5_{{2,8}}gpu.json` — config 5 as its own line at 2 / 8 ranks (`tools/gpu_call_multi_cap.sh`).
* `{R}_scaling.md` — 1 / 2 / 8 GPU runs (config 4 weak scaling; config 5 and the capacity scenes vehicle-sharded, peer exchange vs NCCL) and what bounds them;
  `{R}_mgpu_*.jsonl`, `{R}_bench_*gpu*.json` — the lines behind it.  `{R}_coop_phases_*.txt` — phase timelines of the cooperative
  kernel (`piadmm_debug_coop_trace`, tools/coop_phases.py).
* `{R}_convergence.md`, `{R}_coupled_gain_scan.txt` — convergence of the coupled configurations (VERDICT r1 #2);
  `{R}_ruiz_experiment.txt` — OSQP's Ruiz equilibration restated on the numpy oracle (more inner iterations, not fewer).

## Headline (1 x B200): config 4, 4096 independent two-vehicle scenarios, N = 15

| quantity | value |
|---|---|
| agent-QP solves/s, inputs resident in HBM (`value`) | **{b['value']:.3e}** ({b['ms_per_step']:.3f} ms per 4096-scenario MPC step: {int(w['agent_qp_solves'])} vehicle QPs + {int(w['edge_solves'])} edge problems, {int(w['veh_inner_iters'] + w['edge_inner_iters'])} inner ADMM iterations, outer iterations mean {w['outer_iters_mean']:.2f}; exits: {w['rank0']['exit']}) |
| edge-QP solves/s | {b.get('edge_qp_solves_per_s', float('nan')):.3e} |
| end to end through `piadmm_step_host_at`, pinned host buffers, reference trajectories resident (`e2e`) | **{b['e2e']['value']:.3e}** ({b['e2e']['ms_per_step']:.3f} ms; H2D {b['e2e']['h2d_bytes_per_step']} B (plant state + step index) + D2H {b['e2e']['d2h_bytes_per_step']} B per step, stored by the kernel into the mapped pinned buffers) |
| end to end through `piadmm_step_host` (xt, spd and the reference window uploaded every step) | {b['e2e']['full_upload']['value']:.3e} ({b['e2e']['full_upload']['ms_per_step']:.3f} ms; H2D {b['e2e']['full_upload']['h2d_bytes_per_step']} B) |
| CPU arm (`--impl reference`: oracle C twin = "port", {r['cpu_baseline']['cores']} OpenMP threads of the GPU box) | {r['value']:.3e} solves/s |
| e2e / CPU arm | {b['e2e']['value'] / r['value']:.1f}x |
| kernel launches per step | {b['gpu_launches'] // b['steps']} (`k_fused_v2<16>` + `k_order`, the longest-first schedule for the next step) |
| FP64 roofline: algorithmic flops / launch time / repo-measured DFMA peak ({b['roofline']['peak']:.1f} TFLOP/s) | {b['roofline']['achieved']:.2f} TFLOP/s = **{b['roofline']['frac']:.3f}** |
| ncu (one launch, {m['gpu__time_duration.sum']:.0f} us, {m['launch__registers_per_thread']:.0f} registers) | FP64 pipe active {m['sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active']:.1f} % of active cycles; issue slots busy {m['smsp__issue_active.avg.pct_of_peak_sustained_active']:.1f} %; {m['smsp__warps_active.avg.per_cycle_active']:.2f} warps / SM sub-partition active, {m['smsp__warps_eligible.avg.per_cycle_active']:.2f} eligible; local loads / stores {m.get('smsp__sass_inst_executed_op_local_ld.sum', 0):.0f} / {m.get('smsp__sass_inst_executed_op_local_st.sum', 0):.0f} |
| ncu DRAM | {m['dram_read'] / 1e6:.2f} MB read + {m['dram_write'] / 1e6:.1f} MB written per launch (the writes are the write-back of the L2-flush buffer); DRAM throughput {m['gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed']:.2f} % of peak: not HBM-bound |
| tight inner mode (eps 1e-9; results solver independent) | {cv.get('inner_tight', {}).get('ms_per_step', float('nan')):.3f} ms per step, {cv.get('inner_tight', {}).get('agent_qp_solves_per_s', float('nan')):.3e} solves/s |
| closed loop on the device, 20 consecutive MPC steps of all 4096 scenarios | {cv.get('closed_loop', {}).get('ms_per_mpc_step', float('nan')):.3f} ms per MPC step (outer iterations up to {cv.get('closed_loop', {}).get('outer_iters_max', '?')}) |
| N = 30 variant (`--workload c4n30`) | {('%.3f ms per step, %.3e solves/s, frac %.3f' % (n30['ms_per_step'], n30['value'], n30['roofline']['frac'])) if n30 else 'n/a'} |
| clocks during the timed region | {b['clocks']} |

## Coupled scenes (one scene per GPU, persistent cooperative kernel `k_coop`)

| record | V | ms per coupled MPC step (device) | e2e ms (host buffers) | outer iterations | exit | converged | residual at exit | vehicle QPs / edge problems solved | launches per step | CPU twin ms (16 threads) |
|---|---|---|---|---|---|---|---|---|---|---|
{chr(10).join(coupled_row(k, v) for k, v in cpl.items() if 'exit' in v)}

Parity of every record against the CPU twin is checked inside the bench (`parity_vs_cpu_port`: identical outer iteration
counts, edge lists and solve counts; u within {max(v.get('parity_vs_cpu_port', {}).get('u_rel_err', 0) for v in cpl.values()):.1e} relative).
{joint_txt}

## Config 1 (`casadi/main.py`, ONE scenario, 35 MPC steps): latency

| semantics / inner | steps with an active edge | mean ms per step (active) | max ms | total ms (35 steps, synchronous host calls) | on-device loop ms per step | outer iterations |
|---|---|---|---|---|---|---|
{c1rows}

## Launch lists (cold cache, serialised under ncu: shares, not absolutes)

Config-4 step = `k_fused_v2` + `k_order` ({step_us:.1f} us per step under ncu vs {b['ms_per_step'] * 1e3:.1f} us by CUDA events in the bench):

| kernel | launches | avg us | share of step |
|---|---|---|---|
{lines}

Config-5 step = `k_disc` + `k_cand_direct` (candidate lists, once per MPC step) + `k_coop` (all outer iterations) (+ `k_plant`): {step5_us:.1f} us per step under ncu:

| kernel | launches | avg us | share of step |
|---|---|---|---|
{lines5}

`k_fp64_peak` is the one-off peak micro-benchmark after the timed region; torch fill / gather kernels are buffer
initialisation, the spatial permutation of the inputs and the L2 flush (outside the event pairs).

## Reading of the full capture: `k_fused_v2<16>` (what limits the config-4 step)

{tail}

* 4096 scenario-warps for 2368 resident warp slots (8 CTAs x 2 warps per SM, 128 registers).  Issue slots are about half busy and
  the FP64 pipe ~40 %: the kernel is **dependency-latency bound, not FP64-throughput bound** (stalls: `wait` = fixed-latency
  dependent FP64 chains, `short_scoreboard` = LDS / SHFL results).
* Algorithmic flops are ~21 % of the DFMA peak while the pipe is ~39 % busy: the difference is Gauss-Jordan rebuilds of K^-1 at an
  adapted rho, termination checks, prologue / epilogue of the x-update and edge step, the padded 16th lane and the lock-step wait of
  the faster vehicle of a warp.  `{R}_history.md` has the per-region sample shares and what was tried against each.

## Reading of the full capture: `k_coop<32>` (config 5, one 4096-vehicle scene)

{tailc}

* ncu: {mc['gpu__time_duration.sum']:.0f} us, grid {mc['launch__grid_size']:.0f} CTAs x 64 threads ({mc['launch__registers_per_thread']:.0f} registers, no spills in the loops), FP64 pipe
  {mc['sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active']:.1f} %, issue slots {mc['smsp__issue_active.avg.pct_of_peak_sustained_active']:.1f} %, DRAM {mc['dram_read'] / 1e6:.1f} MB read (the K^-1 cache entries and the scene after the L2 flush).
* The two hottest "regions" are the six-instruction polling loops of the grid barrier (`LD / CCTL / YIELD`): most CTAs WAIT most of
  the time.  A coupled step is a chain of dependent phases (x phase -> slots -> edges -> z phase -> decide, 4 outer iterations
  here), each as long as its slowest problem; after the first iteration x_memo leaves a few hundred vehicle QPs and ~10-360 edge
  problems per phase, i.e. a small fraction of the 592 CTAs has work.  The fraction of the FP64 peak (0.026) says how little of a
  B200 one coupled scene can fill; the batched config-4 line is the throughput figure.
""")
    print("profiles/ regenerated from tag", tag, "as", R)


if __name__ == "__main__":
    main(*sys.argv[1:3])
