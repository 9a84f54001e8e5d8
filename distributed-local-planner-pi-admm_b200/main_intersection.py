"""Drop-in form of the reference driver script casadi/main.py (two-vehicle intersection, config 1).

``run(...)`` keeps the script's loop shape and variable names (casadi/main.py:22-201); the blocks
that built CasADi problems and called ``ca.qpsol("osqp")`` are replaced by calls into the CUDA solver:

    casadi/main.py:81-106   vehicle loop      ->  sol.x_update()
    casadi/main.py:110-118  edge_mat / flag   ->  sol.detect_edges()
    casadi/main.py:121-162  edge loop + dual  ->  sol.z_update()
    casadi/main.py:165-180  residuals / stop  ->  sol.update_residuals()

``fast=True`` runs each MPC step as ONE fused persistent kernel instead (same results);
``device_loop=True`` runs the whole receding-horizon loop on the device (piadmm_run_closed_loop: reference
trajectories resident, plant step and histories in kernels, no host round trip per MPC step).
Run as a script:  python -m pi_admm_b200.main_intersection
"""
from __future__ import annotations

import time

import numpy as np

from .PI_ADMM_class import PI_ADMM_CASADI, check_status


def run(semantics="python_current", inner="osqp_default", fast=False, verbose=False, n_steps=None,
        png=None, device_loop=False, **overrides):
    # PI_ADMM structure
    PI_ADMM = PI_ADMM_CASADI(semantics=semantics, inner=inner, **overrides)
    # initialize the parameter (casadi/main.py:25)
    xt = np.array([[-10, 0, 0], [0, 20, -np.pi / 2]], dtype=np.float64)
    total_steps = int(PI_ADMM.param.Nt / PI_ADMM.param.dt - PI_ADMM.param.num_ho)
    n_steps = total_steps if n_steps is None else n_steps
    iter_his = np.zeros(total_steps)
    primal_u = np.zeros([int(PI_ADMM.param.num_veh), int(PI_ADMM.param.num_ho)])
    x_vec, theta_vec, u_vec, err_rk_his, n_outer_his = [], [], [], [], []
    sol = PI_ADMM.solver(1)
    spd = np.asarray(PI_ADMM.param.spd, dtype=np.float64)[None]
    rho = float(PI_ADMM.param.rho)
    t = time.time()
    total_iter = 0
    if device_loop:
        V, N = int(PI_ADMM.param.num_veh), int(PI_ADMM.param.num_ho)
        sol.setup(xt[None], spd, np.zeros((1, V, 2, N + 1)), rho=rho)
        sol.set_reference(np.ascontiguousarray(PI_ADMM.ref_traj).reshape(1, V, 2, -1))
        h = {k: v.cpu().numpy() for k, v in sol.run_closed_loop(0, n_steps).items()}
        for i in range(n_steps):          # failures are reported after the run (the loop never leaves the device)
            check_status(int(h["status_his"][i, 0]), getattr(PI_ADMM, "error_on_fail", True), f"t_step {i + 1}")
        iter_his[:n_steps] = h["iter_his"][:, 0]
        out = dict(x_vec=[h["x_vec"][i, 0].T.copy() for i in range(n_steps)],
                   theta_vec=[h["theta_vec"][i, 0].copy() for i in range(n_steps)],
                   u_vec=[h["u_vec"][i, 0].copy() for i in range(n_steps)], iter_his=iter_his,
                   err_rk=[h["err_rk_his"][i] for i in range(n_steps)], n_outer=h["n_outer_his"][:, 0].tolist(),
                   elapsed_time=time.time() - t, launches=sol.launch_count)
        if png:
            _plot(out["x_vec"], png)
        return out
    for num_step in range(n_steps):
        # per-step inputs; hat_pos_old / dual_var_old / flag are re-initialised by begin_step (main.py:52-72)
        sol.setup(xt[None], spd, PI_ADMM.ref_window(num_step)[None], rho=rho)
        if fast:
            sol.iterate()
        else:
            sol.begin_step()
            for i_iter in range(PI_ADMM.param.iter_num):
                total_iter += 1
                sol.x_update()            # perform decentralized optimization on each vehicle
                sol.detect_edges()        # judge if collision
                sol.z_update()            # perform optimization for the edge side + dual variable update
                sol.update_residuals()    # summary
                if sol.num_active() == 0:
                    break
        # finish the iterations of optimization in one time step (main.py:185-201)
        sol.plant_step()
        o = sol.to_numpy()
        # casadi/main.py:96 vs :145: a failed vehicle QP raises (error_on_fail default), a failed edge solve does not
        check_status(int(o["status"][0]), getattr(PI_ADMM, "error_on_fail", True), f"t_step {num_step + 1}")
        curr_iter, n_outer = int(o["curr_iter"][0]), int(o["n_outer"][0])
        primal_u = o["u"][0]
        iter_his[num_step] = curr_iter
        ut = primal_u[:, 0]
        xt = o["xt_next"][0].copy()
        rho = float(o["rho_s"][0])        # only dual_mode "pi_old" mutates it (old main.py:137)
        err_rk_his.append(o["err_rk"][0, :n_outer].copy())
        n_outer_his.append(n_outer)
        if verbose:
            print("t_step: {}, iter: {}, max dual: {}, min dual: {}, rho: {}, veh_x: {}, veh_y: {}".format(
                num_step + 1, curr_iter, o["dual"].max(), o["dual"].min(), rho, xt[:, 0], xt[:, 1]))
        x_vec.append(np.copy(np.transpose(xt[:, :2])))
        theta_vec.append(np.copy(np.transpose(xt[:, -1])))
        u_vec.append(np.copy(ut))
    elapsed_time = time.time() - t
    if png:
        _plot(x_vec, png)
    return dict(x_vec=x_vec, theta_vec=theta_vec, u_vec=u_vec, iter_his=iter_his, err_rk=err_rk_his,
                n_outer=n_outer_his, elapsed_time=elapsed_time, launches=sol.launch_count)


def _plot(x_vec, path):
    """casadi/main.py:205-220 headless (Agg) -- skipped when matplotlib is not installed."""
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except ImportError:
        return
    v1 = np.stack([a[:, 0] for a in x_vec]); v2 = np.stack([a[:, 1] for a in x_vec])
    fig, ax = plt.subplots()
    ax.scatter(v1[:, 0], v1[:, 1], color="blue"); ax.scatter(v2[:, 0], v2[:, 1], color="red")
    ax.set_title("2D Scatter of Two Lines"); ax.set_xlabel("X Axis"); ax.set_ylabel("Y Axis")
    fig.savefig(path)


if __name__ == "__main__":
    out = run(verbose=True)
    print("iterations:", int(out["iter_his"].sum()), "elapsed: %.3f s" % out["elapsed_time"])
