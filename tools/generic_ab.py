"""A/B timing of the generic-V step on four scenes (PIADMM_LIB selects the build)."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import torch
import bench
from pi_admm_b200.scenarios import c3_intersection, c5_grid
from pi_admm_b200.solver import PIADMMSolver
scenes = {"c3 N=30": (c3_intersection(V=64, N=30), 30), "c5 N=30": (c5_grid(N=30), 30),
          "c3 N=15": (c3_intersection(V=64, N=15, r0=7.0, gap=5.0), 15), "c5 N=15": (c5_grid(N=15), 15)}
for name, ((xt, spd, rw), N) in scenes.items():
    p = bench.coupled_params(N)
    s = PIADMMSolver(p, S=1, V=xt.shape[1], D=8); s.setup(xt, spd, rw)
    for _ in range(3):
        s.iterate()
    ts = []
    for _ in range(5):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); s.iterate(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    o = s.to_numpy()
    print(f"{os.environ.get('PIADMM_LIB', 'default')[-12:]:>12s} {name}: {min(ts):.3f} ms (median {sorted(ts)[2]:.3f}) outer {int(o['n_outer'][0])} "
          f"edges {int(o['n_edge_solves'][0])} veh_inner {int(o['veh_inner'][0])} edge_inner {int(o['edge_inner'][0])}", flush=True)
    s.close()
