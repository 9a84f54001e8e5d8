"""Latency of ONE scenario-warp running alone on the GPU (no contention): fused-kernel time and inner iteration
counts of selected config-4 scenarios, and of the same scenario replicated over every warp slot.
    python tools/single_scenario_latency.py [scenario ...]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from pi_admm_b200.scenarios import c4_scenarios  # noqa: E402
from pi_admm_b200.solver import PIADMMSolver  # noqa: E402


def run(xt, spd, rw, N, reps=5):
    S = xt.shape[0]
    sol = PIADMMSolver(bench.bench_params(N), S=S, V=2)
    sol.setup(xt, spd, rw)
    best = 1e9
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); sol.iterate(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    o = sol.to_numpy()
    r = (best * 1e3, int(o["veh_inner"].max()), int(o["edge_inner"].max()), int(o["n_outer"].max()))
    sol.close()
    return r


def main(which, N=15):
    xt, spd, rw = c4_scenarios(S=4096, N=N)
    for s in which:
        one = run(xt[s:s + 1], spd[s:s + 1], rw[s:s + 1], N)
        rep = lambda a, n: np.repeat(a[s:s + 1], n, axis=0)
        full = run(rep(xt, 2368), rep(spd, 2368), rep(rw, 2368), N)
        print(f"scenario {s}: alone {one[0]:.1f} us; x2368 (every warp slot busy with it) {full[0]:.1f} us; "
              f"veh_inner {one[1]} edge_inner {one[2]} outer {one[3]}")


if __name__ == "__main__":
    main([int(a) for a in sys.argv[1:]] or [1220, 3712, 29, 0, 1, 2])
