"""ms per MPC step of the on-device receding-horizon run (piadmm_run_closed_loop) for config 4:
4096 perturbed two-vehicle scenarios advancing `n` consecutive MPC steps without a host round trip.
    python tools/closed_loop_bench.py [S] [first] [n]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from pi_admm_b200.scenarios import c4_scenarios, c4_reference  # noqa: E402
from pi_admm_b200.solver import PIADMMSolver  # noqa: E402


def main(S=4096, first=5, n=25, N=15):
    p = bench.bench_params(N)
    xt, spd, _ = c4_scenarios(S=S, N=N, step=first)
    sol = PIADMMSolver(p, S=S, V=2)
    sol.set_reference(c4_reference(S=S))
    for rep in range(3):
        sol.setup(xt, spd, np.zeros((S, 2, 2, N + 1)))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = sol.launch_count
        e0.record(); h = sol.run_closed_loop(first, n); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    no = h["n_outer_his"].cpu().numpy()
    veh = 2 * no.sum()
    print(f"closed loop on device: S={S} scenarios x {n} MPC steps (steps {first}..{first + n - 1}): {ms:.2f} ms total, "
          f"{ms / n:.3f} ms per MPC step, {veh / ms * 1e3:.3e} agent-QP solves/s, outer iterations per step mean "
          f"{no.mean():.2f} max {no.max()}, kernel launches {sol.launch_count - l0} ({(sol.launch_count - l0) / n:.0f} per step)")


if __name__ == "__main__":
    a = [int(x) for x in sys.argv[1:]]
    main(*a)
