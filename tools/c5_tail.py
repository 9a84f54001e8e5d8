"""How much of a coupled step is the tail of its slowest inner solves?  Runs config 3 / 5 with the inner iteration cap
in_max_iter lowered (a diagnostic: capped runs are NOT the reference's results) and prints step time and work."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import torch
import bench
from pi_admm_b200.solver import PIADMMSolver

for name in ("c3", "c5"):
    (xt, spd, rw), V, N, D = bench.coupled_scene(name)
    for cap in (4000, 800, 400, 200, 100, 50):
        p = bench.coupled_params(N)
        p.update(in_max_iter=cap)
        s = PIADMMSolver(p, S=1, V=V, D=D); s.setup(xt, spd, rw)
        for _ in range(3):
            s.iterate()
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); s.iterate(); b.record(); torch.cuda.synchronize()
        o = s.to_numpy()
        print(f"{name} in_max_iter {cap:5d}: {a.elapsed_time(b):6.2f} ms  outer {int(o['n_outer'][0])} veh_inner {int(o['veh_inner'][0])} "
              f"edge_inner {int(o['edge_inner'][0])} edges {int(o['n_edge_solves'][0])} status {int(o['status'][0])}", flush=True)
        s.close()
