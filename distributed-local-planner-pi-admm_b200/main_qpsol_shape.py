"""The two-vehicle intersection run (config 1) written the way casadi/main.py is written: one symbolic variable, one
cost / constraint call and one ``ca.qpsol("solver", "osqp", ...)`` per vehicle and per edge, with the script's own numpy
statements for the collision test, the roll-out of z-hat, the dual update, the residuals and the plant step around them.
The only difference to a script against the reference class is the two import lines: ``ca`` is
``pi_admm_b200.ca_shim`` and every ``qpsol`` call is solved by the CUDA kernels.

This is the compatibility route (a handful of small launches and copies per QP, as slow as its shape demands);
``main_intersection.run(fast=True)`` / ``PI_ADMM_CASADI.step`` run the same MPC step as ONE persistent kernel.
Statement blocks follow casadi/main.py:43-201; tests/test_gpu_parity.py::test_qpsol_shaped_script_equals_fused_step
checks that both routes give the same closed loop.
"""
from __future__ import annotations

import numpy as np

from . import ca_shim as ca
from .PI_ADMM_class import PI_ADMM_CASADI


def run(n_steps=None, **overrides):
    PI_ADMM = PI_ADMM_CASADI(semantics="python_current", **overrides)
    par = PI_ADMM.param
    V, N = int(par.num_veh), int(par.num_ho)
    xt = np.array([[-10, 0, 0], [0, 20, -np.pi / 2]], dtype=np.float64)            # main.py:25
    total = int(par.Nt / par.dt - N)
    n_steps = total if n_steps is None else n_steps
    iter_his, x_vec, theta_vec, u_vec, err_rk_his, n_outer_his = np.zeros(total), [], [], [], [], []
    primal_u = np.zeros((V, N))
    opts_veh = {"verbose": False, "print_problem": False, "print_time": False}
    opts_edge = dict(opts_veh, error_on_fail=False)
    for num_step in range(n_steps):
        # one-step-ahead points, the linearisation point of the collision constraint (main.py:48-49)
        x_seed = np.around(xt[:, 0] + par.dt * par.spd * np.cos(xt[:, 2]), 4)
        y_seed = np.around(xt[:, 1] + par.dt * par.spd * np.sin(xt[:, 2]), 4)
        # local / edge / dual variables start from zero in every MPC step (main.py:52-72)
        pos_old = np.zeros((2 * V, N + 1))
        hat_pos_old = np.empty((V, V), dtype=object)
        dual_var_old = np.empty((V, V), dtype=object)
        for i in range(V):
            for j in range(V):
                hat_pos_old[i, j] = np.zeros((2, N + 1))
                dual_var_old[i, j] = np.zeros((2, N + 1))
        last_iter_hat_pos = hat_pos_old.copy()
        edge_mat = np.zeros((V, V))
        flag, curr_iter, n_outer, err_rk_vec = 0, 1, 0, []
        for i_iter in range(int(par.iter_num)):
            n_outer += 1
            # ---- vehicle side (main.py:81-106)
            for i_veh in range(V):
                ui_t = ca.SX.sym("ui_t", 1, N)
                cost_veh = PI_ADMM.cost_function_primal(num_step, ui_t, xt, hat_pos_old, dual_var_old, i_veh)
                _, cieq_veh = PI_ADMM.nonlcon_function(ui_t)
                F_prob_veh = ca.qpsol("solver", "osqp", {"x": ui_t, "f": cost_veh, "g": ca.reshape(cieq_veh, -1, 1)}, opts_veh)
                sol = F_prob_veh(lbg=0)
                veh_u = np.around(np.array(sol["x"]), 4)
                pos_old[2 * i_veh], pos_old[2 * i_veh + 1], _ = PI_ADMM.dynamic_update_local(xt[i_veh], veh_u, i_veh, if_SX=0)
                primal_u[i_veh] = veh_u
            # ---- collision links (main.py:110-118; squared distance against the un-squared threshold, as written)
            for i in range(V):
                for j in range(i + 1, V):
                    d2 = np.square(pos_old[2 * i] - pos_old[2 * j]) + np.square(pos_old[2 * i + 1] - pos_old[2 * j + 1])
                    edge_mat[i, j] = int(max(d2 < par.dis_thres))
            if np.sum(edge_mat) == 0 and flag == 0:
                break
            flag = 1
            # ---- edge side (main.py:121-162)
            edge_row, edge_col = np.where(edge_mat == 1)
            for veh1, veh2 in zip(edge_row, edge_col):
                prev_pred_pos = np.array([[x_seed[veh1], y_seed[veh1]], [x_seed[veh2], y_seed[veh2]]])
                xt_edge = np.vstack((xt[veh1], xt[veh2]))
                pos_old_edge = np.vstack((pos_old[2 * veh1:2 * veh1 + 2], pos_old[2 * veh2:2 * veh2 + 2]))
                dual_edge = np.vstack((dual_var_old[veh1, veh2], dual_var_old[veh2, veh1]))
                hat_u = ca.SX.sym("hat_u", 1, 2 * N)
                halves = ca.vcat([hat_u[:N], hat_u[N:]])
                cost_edge = PI_ADMM.cost_function_edge(halves, xt_edge, pos_old_edge, dual_edge, prev_pred_pos)
                _, cieq_edge = PI_ADMM.nonlcon_function(halves)
                F_prob_edge = ca.qpsol("solver", "osqp", {"x": hat_u, "f": cost_edge, "g": cieq_edge}, opts_edge)
                result = F_prob_edge(lbg=0)
                optimal_u_edge = np.around(np.array(result["x"]).reshape(2, N), 4)
                hx, hy, _ = PI_ADMM.dynamic_update_edge(xt_edge, optimal_u_edge, if_SX=0)
                hat_pos_old[veh1, veh2] = np.vstack((hx[0], hy[0]))
                hat_pos_old[veh2, veh1] = np.vstack((hx[1], hy[1]))
                # plain dual update (main.py:161-162)
                dual_var_old[veh1, veh2] = dual_var_old[veh1, veh2] + par.rho * (pos_old[2 * veh1:2 * veh1 + 2] - hat_pos_old[veh1, veh2])
                dual_var_old[veh2, veh1] = dual_var_old[veh2, veh1] + par.rho * (pos_old[2 * veh2:2 * veh2 + 2] - hat_pos_old[veh2, veh1])
            # ---- residuals and stop test (main.py:165-180)
            error_sk = error_rk = 0.0
            for veh1, veh2 in zip(edge_row, edge_col):
                error_sk += 2 * np.sqrt(np.sum((par.rho * (last_iter_hat_pos[veh1, veh2] - hat_pos_old[veh1, veh2])) ** 2))
                error_rk += 2 * np.sqrt(np.sum((pos_old[2 * veh1:2 * veh1 + 2] - hat_pos_old[veh1, veh2]) ** 2))
            err_rk_vec.append(error_rk)
            if error_rk <= par.eps_pri and error_sk <= par.eps_dual:
                curr_iter = i_iter + 1
                break
            last_iter_hat_pos = hat_pos_old        # an alias, as in the reference (:180): the dual residual is 0 from now on
        # ---- plant step and histories (main.py:185-201)
        iter_his[num_step] = curr_iter
        x_new, y_new, th_new = PI_ADMM.dynamic_update_edge(xt, primal_u, if_SX=0)
        xt = np.stack((x_new[:, 1], y_new[:, 1], th_new[:, 1]), axis=1)
        x_vec.append(np.copy(xt[:, :2].T)); theta_vec.append(np.copy(xt[:, 2])); u_vec.append(np.copy(primal_u[:, 0]))
        err_rk_his.append(np.array(err_rk_vec)); n_outer_his.append(n_outer)
    return dict(x_vec=x_vec, theta_vec=theta_vec, u_vec=u_vec, iter_his=iter_his, err_rk=err_rk_his, n_outer=n_outer_his)


if __name__ == "__main__":
    out = run()
    print("iterations:", int(out["iter_his"].sum()))
