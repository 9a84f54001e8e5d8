"""Per-scenario start / end times of one fused step.  Needs a debug build of the library:
    nvcc ... -DPIADMM_TRACE -o /tmp/libpiadmm_trace.so piadmm.cu obca_consensus.cu;  PIADMM_LIB=/tmp/libpiadmm_trace.so python tools/trace_step.py"""
import ctypes as C, os, sys
import numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import bench
from pi_admm_b200 import _abi
from pi_admm_b200.scenarios import c4_scenarios
from pi_admm_b200.solver import PIADMMSolver
S, N = 4096, 15
xt, spd, rw = c4_scenarios(S=S, N=N)
sol = PIADMMSolver(bench.bench_params(N), S=S, V=2); sol.setup(xt, spd, rw)
for _ in range(4): sol.iterate()
torch.cuda.synchronize()
lib = _abi.load_library()
tr = np.zeros((S, 4), dtype=np.int64)
lib.piadmm_debug_read_trace.argtypes = [C.c_void_p, C.c_int]
print('rc', lib.piadmm_debug_read_trace(tr.ctypes.data_as(C.c_void_p), S))
t0 = tr[:, 0].min(); st = (tr[:, 0] - t0) / 1e3; en = (tr[:, 1] - t0) / 1e3; sm = tr[:, 2]; mode = tr[:, 3]
o = sol.to_numpy(); cost = o['veh_inner'] + 4 * o['edge_inner']
print('kernel span %.1f us; modes' % en.max(), np.bincount(mode.astype(int)), 'SMs seen', len(np.unique(sm)), 'max smid', sm.max())
idx = np.argsort(-en)[:12]
for i in idx: print('late: s %d start %.1f end %.1f dur %.1f cost %d mode %d sm %d' % (i, st[i], en[i], en[i]-st[i], cost[i], mode[i], sm[i]))
hv = np.argsort(-cost)[:10]
for i in hv: print('heavy: s %d start %.1f end %.1f dur %.1f cost %d mode %d sm %d' % (i, st[i], en[i], en[i]-st[i], cost[i], mode[i], sm[i]))
print('end-time percentiles', np.percentile(en, [50, 90, 99, 99.9, 100]).round(1))
print('starts after 100us:', (st > 100).sum(), ' after 150:', (st > 150).sum(), 'last start %.1f' % st.max())
print('scenarios per SM: min %d max %d' % (np.bincount(sm.astype(int)).min(), np.bincount(sm.astype(int)).max()))
