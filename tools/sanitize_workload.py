"""Small workload that launches every kernel of libpiadmm.so once or twice, for compute-sanitizer:

    compute-sanitizer --tool memcheck|racecheck|synccheck|initcheck python tools/sanitize_workload.py

Sizes are tiny (the tools slow kernels down 10-100x); results are checked for finiteness only -- parity is the
job of tests/, this script is about memory and synchronisation hazards."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
from util import make_params  # noqa: E402
from pi_admm_b200.scenarios import c3_intersection, c4_reference, c4_scenarios  # noqa: E402
from pi_admm_b200.solver import PIADMMSolver  # noqa: E402


def fused(N, S=6):
    """k_fused_v2<16|32>, k_order, k_refwin, k_advance, host-buffer step (graph + zero copy), closed loop."""
    p = bench.bench_params(N)
    xt, spd, rw = c4_scenarios(S=S, N=N, step=12)
    sol = PIADMMSolver(p, S=S, V=2)
    sol.setup(xt, spd, rw)
    for _ in range(2):                       # second pass: K^-1 cache hits + longest-first order
        sol.iterate()
    torch.cuda.synchronize()
    out = sol.to_numpy()
    assert np.isfinite(out["u"]).all() and np.isfinite(out["pos"]).all()
    sol.set_reference(c4_reference(S=S))
    io = sol.alloc_host_io()
    io["xt"].copy_(torch.as_tensor(xt))
    for _ in range(2):
        sol.step_host_at(12, io)
    torch.cuda.synchronize()
    assert torch.isfinite(io["u"]).all()
    h = sol.run_closed_loop(10, 3)
    torch.cuda.synchronize()
    assert torch.isfinite(h["u_vec"]).all()
    sol.step_host(xt, spd, rw)
    sol.close()


def phases(N, V=12):
    """k_x_update, k_bbox / k_hits / k_slots / k_edges, k_z_update, k_res_*, k_plant for a coupled scene."""
    p = make_params("matlab_pi", "default", num_ho=N, beta=10.0, Pnorm=1.0, rho=2.0, eps_pri=0.5, eps_dual=0.5,
                    al_all=0, iter_num=3)
    xt, spd, rw = c3_intersection(V=V, N=N)
    sol = PIADMMSolver(p, S=1, V=V, D=8)
    sol.setup(xt, spd, rw)
    sol.iterate()
    sol.plant_step()
    torch.cuda.synchronize()
    out = sol.to_numpy()
    assert np.isfinite(out["u"]).all()
    sol.set_ownership(0, V // 2)             # vehicle-sharded variant of the same kernels
    sol.begin_step()
    sol.x_update(); sol.detect_edges(); sol.z_update(); sol.residual_partials(); sol.residual_decide()
    torch.cuda.synchronize()
    sol.close()


def ragged(N=15, V=5):
    """Warps that mix a live and a masked group: odd V, a batch with a finished scenario, an odd ownership cut."""
    p = make_params("matlab_pi", "default", num_ho=N, beta=10.0, Pnorm=1.0, rho=2.0, eps_pri=0.5, eps_dual=0.5,
                    al_all=0, iter_num=3)
    parts = [c3_intersection(V=8, N=N, seed=s, r0=r0, gap=g) for s, r0, g in ((3, 7.0, 5.0), (4, 60.0, 25.0), (5, 7.0, 5.0))]
    xt, spd, rw = (np.ascontiguousarray(np.concatenate([q[i][:, :V] for q in parts], axis=0)) for i in range(3))
    sol = PIADMMSolver(p, S=3, V=V, D=V - 1)
    sol.setup(xt, spd, rw)
    sol.iterate()
    torch.cuda.synchronize()
    assert np.isfinite(sol.to_numpy()["u"]).all()
    sol.set_ownership(0, 3)
    sol.begin_step()
    for _ in range(2):
        sol.x_update(); sol.detect_edges(); sol.z_update(); sol.residual_partials(); sol.residual_decide()
    torch.cuda.synchronize()
    sol.close()


def obca(S=5):
    import obca_synth
    from pi_admm_b200.obca_consensus import OBCAConsensus
    o = OBCAConsensus(S=S)
    prev = o.create_bar_state()
    ctrl_prev = torch.zeros((S, 2, 14), dtype=torch.float64, device="cuda")
    o.begin_time_step()
    for step in obca_synth.sequence(S, 3, seed=3):
        for f in ("local_x", "lamb_ij", "A", "b", "Z_bar"):
            getattr(o.bar_state, f).copy_(torch.as_tensor(step[f]).cuda())
        o.check_converge(0.1, 0.1)
        ctrl = torch.as_tensor(step["ctrl"]).cuda().contiguous()
        o.residuals(ctrl, ctrl_prev, prev)
        o.consensus_tail(prev, ctrl, ctrl_prev, rho=1.0)
    o.lambda_update(0.5)
    o.iterate_next_state(o.bar_state)
    torch.cuda.synchronize()


def joint():
    """k_joint<16|32> (cluster + distributed shared memory; 1 and 2 CTAs per cluster, a rho change) and the joint flow."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import test_joint_edge as T
    for V, N, cap in ((5, 8, 4), (10, 30, 3), (20, 15, 3)):
        sc = T.scene(V, N, seed=V + N, spread=4.0 + 0.35 * V)
        pairs = T.near_pairs(sc[0], sc[1], sc[2], cap)
        u, it, st = T._gpu_joint([sc, sc], [pairs, pairs[:1]], N, "default")
        assert np.isfinite(u).all()
    from pi_admm_b200.joint import JointRSUSolver
    ap = make_params("matlab_pi", "default", num_ho=10, beta=10.0, Pnorm=1.0, rho=2.0, eps_pri=0.5, eps_dual=0.5, kI=1.5, iter_num=2)
    xt, spd, rw = c3_intersection(V=6, N=10, r0=5.0, gap=4.0)
    js = JointRSUSolver(ap, S=1, V=6)
    js.setup(xt, spd, rw)
    assert np.isfinite(js.iterate()["u"]).all()
    js.close()


def static():
    """piadmm_edge_list_static (k_static_bbox / k_static_hits / k_static_emit) on two shapes back to back."""
    import ctypes as C
    from pi_admm_b200 import _abi
    lib = _abi.load_library()
    for S, V, N in ((1, 70, 30), (9, 2, 15), (1, 70, 30)):
        pos = torch.randn(S, V, 2, N + 1, dtype=torch.float64, device="cuda") * 5.0
        en = torch.zeros(S, dtype=torch.int32, device="cuda"); eij = torch.zeros(S, 64, 2, dtype=torch.int32, device="cuda")
        rc = lib.piadmm_edge_list_static(C.c_void_p(pos.data_ptr()), S, V, N, 2.0, 1, 64, C.c_void_p(en.data_ptr()),
                                         C.c_void_p(eij.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0
    torch.cuda.synchronize()


if __name__ == "__main__":
    which = sys.argv[1:] or ["fused15", "fused30", "phases15", "phases30", "ragged", "obca", "joint", "static"]
    for w in which:
        {"fused15": lambda: fused(15), "fused30": lambda: fused(30), "phases15": lambda: phases(15),
         "phases30": lambda: phases(30), "ragged": ragged, "obca": obca, "joint": joint, "static": static}[w]()
        print("ok", w, flush=True)
