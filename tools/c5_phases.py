import sys, time; import os; R=os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0,R); sys.path.insert(0,os.path.join(R,'tests'))
import numpy as np, torch
from util import make_params
from pi_admm_b200.solver import PIADMMSolver
from pi_admm_b200.scenarios import c5_grid, c3_intersection
p = make_params("matlab_pi", "default", num_ho=30, beta=10.0, Pnorm=1.0, rho=2.0, eps_pri=0.5, eps_dual=0.5, al_all=0, iter_num=12)
for name,(xt,spd,rw),D in (("c5",c5_grid(N=30),8),("c3",c3_intersection(V=64,N=30),8)):
    V=xt.shape[1]
    s=PIADMMSolver(p,S=1,V=V,D=D); s.setup(xt,spd,rw)
    for rep in range(2):
        tm={k:0.0 for k in ('x','detect','z','res','active')}
        s.begin_step(); torch.cuda.synchronize(); t0=time.perf_counter()
        for it in range(s.I):
            for k,fn in (('x',s.x_update),('detect',s.detect_edges),('z',s.z_update),('res',s.update_residuals)):
                a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
                a.record(); fn(); b.record(); torch.cuda.synchronize(); tm[k]+=a.elapsed_time(b)
            t=time.perf_counter(); na=s.num_active(); tm['active']+=(time.perf_counter()-t)*1e3
            if na==0: break
        print(name,'V',V,'iters',it+1,'wall %.2f ms'%((time.perf_counter()-t0)*1e3),{k:round(v,3) for k,v in tm.items()}, 'edges', int(s.buf['n_edge_solves'][0]), 'inner veh', int(s.buf['veh_inner'][0]), 'edge', int(s.buf['edge_inner'][0]))
    a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True); a.record(); s.iterate(); b.record(); torch.cuda.synchronize(); print(name,'iterate() total %.2f ms'%a.elapsed_time(b))
