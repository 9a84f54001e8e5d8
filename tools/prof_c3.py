"""Small driver for profiling the generic-V phase kernels on the config-3 scene (64 vehicles, N=30)."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import torch
from util import make_params
from pi_admm_b200.solver import PIADMMSolver
from pi_admm_b200.scenarios import c3_intersection
p = make_params("matlab_pi", "default", num_ho=30, beta=10.0, Pnorm=1.0, rho=2.0, eps_pri=0.5, eps_dual=0.5, al_all=0, iter_num=12)
xt, spd, rw = c3_intersection(V=64, N=30)
s = PIADMMSolver(p, S=1, V=64, D=8); s.setup(xt, spd, rw)
for _ in range(2):
    s.iterate()
torch.cuda.synchronize()
print("n_outer", int(s.buf["n_outer"][0]), "edges", int(s.buf["n_edge_solves"][0]))
