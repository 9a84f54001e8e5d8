"""Receding-horizon run of a COUPLED scene (config 3: 64-vehicle intersection) on one GPU: per MPC step the persistent
cooperative kernel (all outer iterations), the plant step, xt <- xt_next, the reference window slides along each vehicle's
straight line.  Reports per step: device ms, outer iterations, how the loop was left, edges solved, and the smallest
distance between any two vehicles' ACTUAL positions (a domain property: the planner is there to keep that above the
collision distance).      python tools/coupled_closed_loop.py [n_steps] [c3|c5] [stop_dist_test 0|1]"""
import json
import os
import sys

import numpy as np
import torch

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import bench  # noqa: E402
from pi_admm_b200.scenarios import c3_intersection, c5_grid  # noqa: E402
from pi_admm_b200.solver import PIADMMSolver  # noqa: E402


def main():
    n_steps = int(sys.argv[1]) if len(sys.argv) > 1 else 25
    scene = sys.argv[2] if len(sys.argv) > 2 else "c3"
    N = 30
    p = bench.coupled_params(N)
    if len(sys.argv) > 3:
        p.update(stop_dist_test=int(sys.argv[3]))     # MATLAB :202 extra stop condition (casadi/main.py:174 has it commented out)
    xt, spd, refwin = c3_intersection(V=64, N=N) if scene == "c3" else c5_grid(N=N)
    V = xt.shape[1]
    x0 = xt.copy()
    sol = PIADMMSolver(p, S=1, V=V, D=8)
    k = np.arange(N + 1)
    rows = []
    dmin_run = np.inf
    for step in range(n_steps):
        # reference of vehicle i at MPC step t: its straight constant-speed line through the origin, window [t, t + N]
        th = x0[0, :, 2]
        for c, f in ((0, np.cos), (1, np.sin)):
            refwin[0, :, c] = x0[0, :, c][:, None] + (step + k)[None, :] * p.dt * spd[0][:, None] * f(th)[:, None]
        sol.setup(xt, spd, refwin)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); sol.iterate(); sol.plant_step(); b.record(); torch.cuda.synchronize()
        o = sol.to_numpy()
        n = int(o["n_outer"][0]); en = int(o["edge_n"][0]); st = int(o["status"][0])
        rk = float(o["err_rk"][0, n - 1]) if n else 0.0
        exit_ = "cap" if st & 8 else ("no edge" if en == 0 else "stop test")
        xt = o["xt_next"].copy()
        from scipy.spatial import cKDTree
        dd, _ = cKDTree(xt[0, :, :2]).query(xt[0, :, :2], k=2)
        d = dd[:, 1]
        dmin_run = min(dmin_run, d.min())
        rows.append(dict(step=step, ms=a.elapsed_time(b), outer=n, exit=exit_, edges_last=en, edge_solves=int(o["n_edge_solves"][0]),
                         rk=rk, status=st, min_dist=float(d.min()), max_abs_u=float(np.abs(o["u"]).max())))
        print(json.dumps(rows[-1]))
    ms = np.array([r["ms"] for r in rows[2:]])
    print(json.dumps(dict(summary=True, stop_dist_test=int(p.stop_dist_test), scene=scene, V=V, N=N, steps=n_steps, ms_mean=float(ms.mean()), ms_max=float(ms.max()),
                          outer_max=max(r["outer"] for r in rows), capped_steps=sum(r["exit"] == "cap" for r in rows),
                          min_distance_over_run=float(dmin_run), dis_thres=float(p.dis_thres))))


if __name__ == "__main__":
    main()
