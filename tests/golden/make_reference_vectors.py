"""Golden vectors produced by the REFERENCE'S OWN CODE, run in the build container (where /root/reference exists):

    python tests/golden/make_reference_vectors.py        # writes tests/golden/reference_vectors.json

The reference's classes import `casadi` and `bunch`, which are not installed here; tests/golden/_ref_import_stubs.py puts
numpy-backed stand-ins for exactly the container functions the class bodies call into sys.modules, the reference modules are
then imported UNMODIFIED from /root/reference and their methods are called on numeric arrays.  What is recorded:

  casadi/PI_ADMM_class.py      param, ref_traj (:13-37); dynamic_update_local numeric branch (:45-70); dynamic_update_edge numeric
                               branch (:77-105); cost_function_primal (:114-135) evaluated at 0, +-e_i and e_i+e_j -> the exact
                               (H, q) of the quadratic it is, and at random u; nonlcon_function (:172-192) at random u;
                               cost_function_edge (:145-169, the Python non-convex statement) at random hat_u.
  casadi_old_PI_ADMM/main.py   the dual / PI update block :127-164 exec'd verbatim over several outer iterations (the PI_OLD law).
  matlab_old_files/ADMM_CVX_two_veh_intesection_PI_antiwindup.m
                               param :6-25, collision test :105-121, dual / PI update + saturation + back-calculation :152-189,
                               residual sums and stop test :191-209; the bodies of dynamic_update_local / _edge (:312-397),
                               cost_function_primal (:254-273), cost_function_edge (:287-302, the CONVEX edge statement)
                               -- the .m text executed by tests/golden/_matlab_subset.py.
  Distributed_planner/decentralized/optimizer.py, util.py
                               mid_state (:351-373), lambda_update (:330-335), check_converge (:225-235),
                               iterate_next_state (:337-344), bar_state_update (:205-222) with
                               compute_square_halfspaces_ca_prob (util.py:76-101).

The solvers (OSQP / IPOPT inside the CasADi wheel) are NOT available, so no solver output is recorded: these vectors pin the
problem construction, the roll-outs, the constraints and the consensus arithmetic of the oracle and of the CUDA path to the
reference's own statements; iterates of the inner solver stay unpinned."""
import importlib
import importlib.util
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import _ref_import_stubs as stubs  # noqa: E402

REF = "/root/reference"


def load_class():
    stubs.install()
    spec = importlib.util.spec_from_file_location("ref_PI_ADMM_class", os.path.join(REF, "casadi", "PI_ADMM_class.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m.PI_ADMM_CASADI


def scal(x):
    return float(np.asarray(x).reshape(-1)[0])


def vehicle_case(PI, rng, num_step, veh, xt, N, spd):
    PI.param.num_ho = N
    PI.param.spd = np.array(spd, dtype=float)
    V = PI.param.num_veh
    hat = np.empty((V, V), dtype=object); dual = np.empty((V, V), dtype=object)
    for i in range(V):
        for j in range(V):
            hat[i, j] = rng.normal(0, 3, (2, N + 1)); dual[i, j] = rng.normal(0, 1, (2, N + 1))
    f = lambda u: scal(PI.cost_function_primal(num_step, stubs.Mat(np.asarray(u).reshape(1, N)), xt, hat, dual, veh))
    E = np.eye(N)
    f0 = f(np.zeros(N))
    fp = np.array([f(E[i]) for i in range(N)]); fm = np.array([f(-E[i]) for i in range(N)])
    q = (fp - fm) / 2.0
    H = np.zeros((N, N))
    for i in range(N):
        H[i, i] = fp[i] + fm[i] - 2.0 * f0
        for j in range(i + 1, N):
            H[i, j] = H[j, i] = f(E[i] + E[j]) - fp[i] - fp[j] + f0
    us = rng.normal(0, 0.3, (4, N))
    vals = [f(u) for u in us]
    cons = [np.asarray(PI.nonlcon_function(stubs.Mat(u.reshape(1, N)))[1]).ravel().tolist() for u in us]
    roll = []
    for u in us:
        x, y, th = PI.dynamic_update_local(xt[veh], np.around(u.reshape(1, N), 4), veh, if_SX=0)
        roll.append([np.asarray(x).tolist(), np.asarray(y).tolist(), np.asarray(th).tolist()])
    other = 1 - veh
    return dict(num_step=num_step, veh=veh, N=N, spd=list(map(float, spd)), xt=xt.tolist(),
                hat=hat[veh, other].tolist(), dual=dual[veh, other].tolist(), f0=f0, H=H.tolist(), q=q.tolist(),
                u=us.tolist(), cost=vals, cieq=cons, rollout_local_of_rounded_u=roll)


def edge_case(PI, rng, xt, N, spd):
    PI.param.num_ho = N
    PI.param.spd = np.array(spd, dtype=float)
    pos_old = rng.normal(0, 5, (4, N + 1)); dual = rng.normal(0, 1, (4, N + 1))
    seed = np.stack((np.around(xt[:, 0] + PI.param.dt * PI.param.spd * np.cos(xt[:, 2]), 4),
                     np.around(xt[:, 1] + PI.param.dt * PI.param.spd * np.sin(xt[:, 2]), 4)), 1)      # casadi/main.py:48-49
    us = rng.normal(0, 0.25, (4, 2, N))
    cost = [scal(PI.cost_function_edge(stubs.Mat(u), xt, pos_old, dual, seed)) for u in us]
    roll = []
    for u in us:
        x, y, th = PI.dynamic_update_edge(xt, np.around(u, 4), if_SX=0)
        roll.append([np.asarray(x).tolist(), np.asarray(y).tolist(), np.asarray(th).tolist()])
    # the CONVEX statement (MATLAB ...PI_antiwindup.m:276-304 with the edge roll-out linearised about xt(:,3), :383-396) composed
    # from the reference's OWN two Python functions: cost_function_edge (:145-169) with its roll-out replaced by the reference's
    # linearised dynamic_update_local (:45-70) of each vehicle.  Both bodies are the reference's; only the composition is ours.
    def lin_edge(xt_, u_, if_SX):
        rows = [PI.dynamic_update_local(xt_[i], stubs.Mat(np.asarray(u_)[i].reshape(1, N)), i, if_SX=1) for i in range(2)]
        st = lambda c: stubs.Mat(np.vstack([np.asarray(r[c]).reshape(1, -1) for r in rows]))
        return st(0), st(1), st(2)
    nonlin = PI.dynamic_update_edge
    PI.dynamic_update_edge = lin_edge
    # seeds that put the pair at the collision threshold, so that the hinge is active for some steps and some hat_u
    pos_c = np.vstack((xt[0, 0] + np.arange(N + 1) * 0.1 * spd[0] * np.cos(xt[0, 2]), xt[0, 1] + np.arange(N + 1) * 0.1 * spd[0] * np.sin(xt[0, 2]),
                       xt[1, 0] + np.arange(N + 1) * 0.1 * spd[1] * np.cos(xt[1, 2]), xt[1, 1] + np.arange(N + 1) * 0.1 * spd[1] * np.sin(xt[1, 2])))
    us_c = np.vstack((np.zeros((1, 2, N)), rng.normal(0, 0.3, (7, 2, N))))
    cost_c = [scal(PI.cost_function_edge(stubs.Mat(u), xt, pos_c, dual, seed)) for u in us_c]
    PI.dynamic_update_edge = nonlin
    return dict(N=N, spd=list(map(float, spd)), xt=xt.tolist(), pos_old=pos_old.tolist(), dual=dual.tolist(), seed=seed.tolist(),
                hat_u=us.tolist(), cost_python_nonconvex=cost, rollout_edge_of_rounded_u=roll,
                convex=dict(pos_old=pos_c.tolist(), hat_u=us_c.tolist(), cost=cost_c))


def obca_cases(rng):
    stubs.install()
    sys.path.insert(0, os.path.join(REF, "Distributed_planner"))
    dec = importlib.import_module("decentralized")
    out = []
    for case in range(4):
        opt = dec.OBCAOptimizer(min_dis=0.1, prob=1)
        bar = opt.create_bar_state()
        opt.bar_state = bar
        V, H = opt.num_veh, opt.N_horz - 1
        init = dict(lamb_ij0=bar.lamb_ij.tolist(), lamb_bar0=bar.lamb_bar.tolist())
        # local solutions of a fictitious iteration: states near the reference trajectories, multipliers near the warm start
        fullx = []
        for v in range(V):
            st = opt.ref_traj[v][3 * case + 1: 3 * case + 1 + H].copy()
            st += rng.normal(0, 1, st.shape) * np.array([0.5, 0.3, 0.5, 0.05, 0.02])
            lam = np.abs(bar.lamb_ij[v] + rng.normal(0, 0.05, (H, 4)))
            fullx.append(np.hstack((st, lam)))
        opt.bar_state_update(fullx, None)
        A, b, local_x = bar.A.copy(), bar.b.copy(), bar.local_x.copy()
        bar.Z_bar = np.round(np.stack([fx + rng.normal(0, 0.05 * (case + 1), fx.shape) for fx in fullx]), 4)
        if case == 3:          # a state in which check_converge is true: opposite half-spaces, equal multipliers
            bar.A[1] = -bar.A[0]; bar.lamb_ij[1] = bar.lamb_ij[0]; bar.b[1] = -bar.b[0] - 0.2 / np.maximum(bar.lamb_ij[0].sum(-1, keepdims=True), 1e-9)
        conv = bool(opt.check_converge(0.1))
        A_cc, b_cc, lij_cc = bar.A.copy(), bar.b.copy(), bar.lamb_ij.copy()
        rho = [1.0, 0.5, 2.0, 1.0][case]
        lb0 = bar.lamb_bar.copy()
        opt.lambda_update(rho)
        lb1 = bar.lamb_bar.copy()
        before = {f: getattr(bar, f).copy() for f in ("Z_bar", "A", "b", "lamb_bar", "lamb_ij", "local_x")}
        nxt = opt.iterate_next_state(bar)
        out.append(dict(fullx=[fx.tolist() for fx in fullx], A=A.tolist(), b=b.tolist(), local_x=local_x.tolist(),
                        Z_bar=before["Z_bar"].tolist(), check_A=A_cc.tolist(), check_b=b_cc.tolist(), check_lamb_ij=lij_cc.tolist(),
                        check_converge=conv, thres=0.1, min_dis=0.1, rho=rho, lamb_bar_before=lb0.tolist(), lamb_bar_after=lb1.tolist(),
                        shift_before={k: v.tolist() for k, v in before.items()},
                        shift_after={f: getattr(nxt, f).tolist() for f in before}, **init))
    return out


def old_pi_block_cases(rng):
    """The dual / PI update block of the OLD Python script, casadi_old_PI_ADMM/main.py:127-164 (gains from the minimum
    distance, output-then-integrate, saturation +-20, back-calculation 2*diff, global Frobenius residuals, stop test with
    the distance condition, the last_iter_hat_pos alias): those source lines are EXEC'd verbatim, several outer iterations in
    a row, in a namespace that holds what the script holds at that point (the old class instance imported unmodified)."""
    stubs.install()
    spec = importlib.util.spec_from_file_location("ref_old_PI_ADMM_class", os.path.join(REF, "casadi_old_PI_ADMM", "PI_ADMM_class.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    lines = open(os.path.join(REF, "casadi_old_PI_ADMM", "main.py")).read().split("\n")[126:164]
    assert "dual variable update" in lines[0] and "sum_iter_num += 1" in lines[-1]
    src = "for num_step in [num_step_]:\n    for i_iter in [i_iter_]:\n" + "\n".join(lines) + "\n"
    code = compile(src, "casadi_old_PI_ADMM/main.py:127-164", "exec")
    out = []
    for case, (trad, scale) in enumerate(((0, 1.0), (0, 6.0), (1, 3.0))):
        PI = m.PI_ADMM_CASADI()
        N, V = PI.param.num_ho, PI.param.num_veh
        xt = np.array([[-3.0, 0.2, 0.05], [0.3, 6.0, -1.5]]) + rng.normal(0, 0.2, (2, 3))
        hat_pos_old = np.zeros((2 * V, N + 1))
        ns = dict(PI_ADMM=PI, np=np, xt=xt, trad=trad, windup_sat=20, sum_err=0, diff_val=0, err_rk_vec=[], sum_iter_num=0,
                  dual_var_old=np.zeros((2 * V, N + 1)), hat_pos_old=hat_pos_old, last_iter_hat_pos=hat_pos_old.copy(),
                  curr_iter=-1, num_step_=3)
        its = []
        for it in range(6):
            primal_u = rng.normal(0, 0.2, (V, N))
            pos_old = rng.normal(0, 1.0, (2 * V, N + 1)) * scale
            new_hat = pos_old + rng.normal(0, 0.5, pos_old.shape) * scale * (0.02 if (case == 0 and it >= 4) else 1.0)
            for r in range(2 * V):
                hat_pos_old[r] = new_hat[r]             # the script writes hat_pos_old row by row, in place (:121-125)
            ns.update(primal_u=primal_u, pos_old=pos_old, i_iter_=it, curr_iter=-1)
            exec(code, ns)
            dv = ns["diff_val"]
            its.append(dict(primal_u=primal_u.tolist(), pos_old=pos_old.tolist(), hat_pos_old=new_hat.tolist(),
                            dual_var_old=np.asarray(ns["dual_var_old"]).tolist(), sum_err=np.broadcast_to(np.asarray(ns["sum_err"], dtype=float), pos_old.shape).tolist(),
                            diff_val=np.broadcast_to(np.asarray(dv, dtype=float), pos_old.shape).tolist(), rho=float(PI.param.rho),
                            error_sk=float(ns["error_sk"]), error_rk=float(ns["error_rk"]), dis_vec=np.asarray(ns["dis_vec"]).tolist(),
                            stopped=bool(ns["curr_iter"] == it + 1)))
        out.append(dict(trad=trad, xt=xt.tolist(), spd=np.asarray(PI.param.spd, dtype=float).tolist(), N=int(N), windup_sat=20,
                        eps_pri=float(PI.param.eps_pri), eps_dual=float(PI.param.eps_dual), dis_thres=float(PI.param.dis_thres),
                        dt=float(PI.param.dt), L=float(PI.param.L), iterations=its))
    return out


def matlab_pi_block_cases(rng):
    """The default dual law (dual="pi") comes from MATLAB source, which cannot run here.  Its statements are therefore
    EXECUTED FROM THE .m TEXT by the small interpreter of tests/golden/_matlab_subset.py:
    matlab_old_files/ADMM_CVX_two_veh_intesection_PI_antiwindup.m  :6-25 (param), :43 (windup_sat), :105-121 (collision test,
    early exit, find), :122-124 + :152-189 (per-edge dual / PI update, saturation, back-calculation), :190-209 (residual
    sums, stop test with the distance condition, last_iter_hat_pos, curr_iter).  The CVX solve and the roll-out between
    :126 and :151 are replaced by the harness writing pos_old / hat_pos_old for the iteration."""
    import _matlab_subset as ms
    path = os.path.join(REF, "matlab_old_files", "ADMM_CVX_two_veh_intesection_PI_antiwindup.m")
    L = open(path).read().split("\n")
    seg = lambda a, b: "\n".join(L[a - 1:b])
    assert L[5].startswith("param.dt") and L[24].startswith("param.windup") and "windup_sat" in L[42] and "edge_mat(i, j)" in L[108]
    assert L[121].strip().startswith("for i_edge") and "veh2 = edge_col" in L[123] and "dual variable update" in L[151]
    assert L[188].strip() == "end" and "%% summary" in L[190] and "curr_iter = i_iter;" in L[208]
    body = ("for i_iter = i_iter_: i_iter_\n" + seg(105, 121) + "\n" + seg(122, 124) + "\n" + seg(152, 189) + "\n" + seg(191, 209) + "\nend\n")
    out = []
    for case, (PI, windup, scale, close) in enumerate(((1, 1, 1.0, 1), (1, 1, 8.0, 1), (0, 1, 2.0, 1), (1, 0, 3.0, 1), (1, 1, 1.0, 0))):
        it = ms.Interp()
        it.run(seg(6, 25) + "\n" + seg(43, 43))
        P = it.ns["param"]
        P["PI"], P["windup"] = float(PI), float(windup)
        N, V = int(P["num_ho"]), int(P["num_veh"])
        z = lambda: np.zeros((2, N + 1))
        cell = lambda: ms.Cell({(i, j): z() for i in (1, 2) for j in (1, 2)})
        it.ns.update(hat_pos_old=cell(), dual_var_old=cell(), diff_val_veh=cell(), edge_mat=np.zeros((V, V)),
                     kP_mat=P["kP"] * np.ones((V, V)), sum_err_veh1=0.0, sum_err_veh2=0.0, err_rk_vec=np.zeros((0, 0)), flag=0.0,
                     curr_iter=1.0, sum_iter_num=0.0)
        it.ns["last_iter_hat_pos"] = ms.Interp.value_copy(it.ns["hat_pos_old"])
        its = []
        for k in range(6):
            base = rng.normal(0, 1.0, (2, N + 1)) * scale
            sep = (0.3 if close else 40.0) if k < 5 else 40.0            # last iteration: far apart (the edge drops out)
            pos_old = np.vstack((base, base + sep + rng.normal(0, 0.3, (2, N + 1))))
            shrink = 0.002 if (case == 0 and k >= 3) else 1.0
            hat = {(1, 2): pos_old[:2] + rng.normal(0, 0.5, (2, N + 1)) * scale * shrink,
                   (2, 1): pos_old[2:] + rng.normal(0, 0.5, (2, N + 1)) * scale * shrink}
            if case == 0 and k >= 3:
                pos_old[2:, 1] = pos_old[:2, 1] + 3.0              # dis_vec(2) > dis_thres, so that the stop test can pass
            it.ns["pos_old"] = pos_old.copy()
            for key, v in hat.items():
                it.ns["hat_pos_old"][key] = v.copy()               # what :149-150 would have written for the colliding pair
            it.ns["i_iter_"] = float(k + 1)
            before = it.ns["sum_iter_num"]
            it.run(body)
            ns = it.ns
            edges = int(np.sum(mat_(ns["edge_mat"]) == 1))
            its.append(dict(pos_old=pos_old.tolist(), hat12=hat[(1, 2)].tolist(), hat21=hat[(2, 1)].tolist(), edges=edges,
                            flag=float(ns["flag"]), went_on=bool(ns["sum_iter_num"] > before),
                            dual12=mat_(ns["dual_var_old"][(1, 2)]).tolist(), dual21=mat_(ns["dual_var_old"][(2, 1)]).tolist(),
                            diff12=mat_(ns["diff_val_veh"][(1, 2)]).tolist(), diff21=mat_(ns["diff_val_veh"][(2, 1)]).tolist(),
                            sum_err1=np.broadcast_to(mat_(ns["sum_err_veh1"]), (2, N + 1)).tolist(),
                            sum_err2=np.broadcast_to(mat_(ns["sum_err_veh2"]), (2, N + 1)).tolist(),
                            kP=float(mat_(ns["kP_mat"])[0, 1]), error_rk=float(ns.get("error_rk", 0.0)), error_sk=float(ns.get("error_sk", 0.0)),
                            curr_iter=float(ns["curr_iter"])))
        out.append(dict(param={k: (np.asarray(v).tolist() if isinstance(v, np.ndarray) else v) for k, v in P.items()},
                        windup_sat=float(it.ns["windup_sat"]), iterations=its))
    return out


def matlab_function_cases(rng):
    """Function bodies of the same MATLAB script, executed from the .m text with the arguments preset in the namespace:
    dynamic_update_local (:312-348, both branches: numeric non-linear roll-out / linearised about xt(3)),
    dynamic_update_edge (:357-397, both branches), cost_function_primal (:254-273, with the linearised roll-out of :332-347
    run first, as its line :258 does), cost_function_edge (:287-302 after the linearised edge roll-out :379-396 of its line
    :286) -- the CONVEX edge statement this repository implements, evaluated by the reference's own MATLAB statements."""
    import _matlab_subset as ms
    path = os.path.join(REF, "matlab_old_files", "ADMM_CVX_two_veh_intesection_PI_antiwindup.m")
    L = open(path).read().split("\n")
    seg = lambda a, b: "\n".join(L[a - 1:b])
    assert L[311].strip().startswith("if nargin  == 4") and L[347].strip() == "end" and L[356].strip().startswith("if nargin  == 3")
    assert L[396].strip() == "end" and "ref_val_x = ref_traj" in L[253] and "cost = cost_norm" in L[272]
    assert "last_dis = prev_pred_pos" in L[286] and "cost = cost_punish" in L[301]
    out = []
    for case, (N, spd, close) in enumerate(((8, (4.0, 8.0), 1), (15, (6.0, 5.0), 1), (5, (8.0, 4.0), 0))):
        it0 = ms.Interp(); it0.run(seg(6, 25))
        P = it0.ns["param"]; P["num_ho"] = float(N); P["spd"] = np.array(spd).reshape(2, 1)
        xt = np.array([[-3.0, 0.2, 0.1], [0.4, 5.0, -1.45]]) + rng.normal(0, 0.2, (2, 3))
        if not close:
            xt[1, :2] += 40.0
        T = 50
        ref_traj = rng.normal(0, 3, (4, T))
        rec = dict(N=N, spd=list(spd), xt=xt.tolist(), ref_traj=ref_traj.tolist(), param={k: (np.asarray(v).tolist() if isinstance(v, np.ndarray) else v) for k, v in P.items()})
        zer = lambda r: np.zeros((r, N + 1))
        # --- roll-outs of one vehicle (numeric: nargin == 4; symbolic branch: nargin == 7) and of the pair (3 / 6)
        rolls = []
        for v in (1, 2):
            u = rng.normal(0, 0.25, (1, N))
            r_ = dict(veh=v, u=u.tolist())
            for nargin, key in ((4, "nonlinear"), (7, "linear")):
                ns = dict(param=P, xt=xt[v - 1:v].copy(), u=u, veh_index=float(v), nargin=float(nargin), x_pred=zer(1), y_pred=zer(1), theta_pred=zer(1))
                ms.Interp(ns).run(seg(312, 348))
                r_[key] = [ns["x_pred"].tolist(), ns["y_pred"].tolist(), ns["theta_pred"].tolist()]
            rolls.append(r_)
        rec["local_rollouts"] = rolls
        ue = rng.normal(0, 0.25, (2, N))
        er = dict(u=ue.tolist())
        for nargin, key in ((3, "nonlinear"), (6, "linear")):
            ns = dict(param=P, xt=xt.copy(), u=ue, nargin=float(nargin), x_pred=zer(2), y_pred=zer(2), theta_pred=zer(2))
            ms.Interp(ns).run(seg(357, 397))
            er[key] = [ns["x_pred"].tolist(), ns["y_pred"].tolist(), ns["theta_pred"].tolist()]
        rec["edge_rollouts"] = er
        # --- cost_function_primal at 0, +-e_i, e_i+e_j (exact H, q) and at random u
        veh, num_step = 1 + case % 2, 3 + case
        hat = ms.Cell({(i, j): rng.normal(0, 3, (2, N + 1)) for i in (1, 2) for j in (1, 2)})
        dual = ms.Cell({(i, j): rng.normal(0, 1, (2, N + 1)) for i in (1, 2) for j in (1, 2)})

        def f_primal(u):
            ns = dict(param=P, xt=xt[veh - 1:veh].copy(), u=np.asarray(u, dtype=float).reshape(1, N), veh_index=float(veh), nargin=7.0,
                      x_pred=zer(1), y_pred=zer(1), theta_pred=zer(1), ref_traj=ref_traj, num_step=float(num_step), hat_pos=hat, dual_var=dual)
            I = ms.Interp(ns)
            I.run(seg(312, 348))                     # = the call at :258
            I.run(seg(254, 257) + "\n" + seg(259, 273))
            return float(ns["cost"])
        E = np.eye(N)
        f0 = f_primal(np.zeros(N)); fp = np.array([f_primal(E[i]) for i in range(N)]); fm = np.array([f_primal(-E[i]) for i in range(N)])
        H = np.zeros((N, N))
        for i in range(N):
            H[i, i] = fp[i] + fm[i] - 2 * f0
            for j in range(i + 1, N):
                H[i, j] = H[j, i] = f_primal(E[i] + E[j]) - fp[i] - fp[j] + f0
        us = rng.normal(0, 0.3, (3, N))
        other = 3 - veh
        rec["primal"] = dict(veh=veh, num_step=num_step, hat=hat[(veh, other)].tolist(), dual=dual[(veh, other)].tolist(), f0=f0,
                             q=((fp - fm) / 2).tolist(), H=H.tolist(), u=us.tolist(), cost=[f_primal(u) for u in us])
        # --- cost_function_edge: the convex statement
        k = np.arange(N + 1)
        pos_old = np.vstack([xt[v, c] + k * 0.1 * spd[v] * (np.cos, np.sin)[c](xt[v, 2]) for v in (0, 1) for c in (0, 1)]) + rng.normal(0, 0.1, (4, N + 1))
        dual_e = rng.normal(0, 0.5, (4, N + 1))
        prev = np.stack((xt[:, 0] + 0.1 * np.array(spd) * np.cos(xt[:, 2]), xt[:, 1] + 0.1 * np.array(spd) * np.sin(xt[:, 2])), 1)   # :54-55 (no rounding in MATLAB)

        def f_edge(hu):
            ns = dict(param=P, xt=xt.copy(), u=np.asarray(hu, dtype=float), hat_u=np.asarray(hu, dtype=float), nargin=6.0, x_pred=zer(2), y_pred=zer(2), theta_pred=zer(2),
                      edge_pos=zer(4), pos_old=pos_old, dual_var_old=dual_e, prev_pred_pos=prev)
            I = ms.Interp(ns)
            I.run(seg(357, 397))                     # = the call at :286 (symbolic branch: linearised about xt(:,3))
            I.run(seg(287, 302))
            return float(ns["cost"])
        hus = np.vstack((np.zeros((1, 2, N)), rng.normal(0, 0.3, (7, 2, N))))
        rec["edge_cost"] = dict(pos_old=pos_old.tolist(), dual=dual_e.tolist(), prev_pred_pos=prev.tolist(), hat_u=hus.tolist(), cost=[f_edge(h) for h in hus])
        out.append(rec)
    return out


def matlab_joint_cost_cases(rng):
    """the body of cost_function_edge of matlab_old_files/ADMM_fmincon_multi_veh_intesection_static.m
    (:262-289 -- ONE edge problem over all num_veh vehicles), executed from the .m text on the LINEARISED roll-out of
    ...PI_antiwindup.m:357-397 (nargin == 6 branch, which loops over param.num_veh) -- i.e. the convex statement of the joint
    problem that oracle/joint.py and k_joint solve.  V = 2, 3, 4.  For V > 2 the script's `last_dis(1)`, `last_dis(2)` (:270)
    are LINEAR indices into a (V-1) x 2 matrix; the value recorded is what the text computes (oracle: as_written=True)."""
    import _matlab_subset as ms
    L = open(os.path.join(REF, "matlab_old_files", "ADMM_fmincon_multi_veh_intesection_static.m")).read().split("\n")
    Lp = open(os.path.join(REF, "matlab_old_files", "ADMM_CVX_two_veh_intesection_PI_antiwindup.m")).read().split("\n")
    assert L[249].startswith("function cost = cost_function_edge") and L[261].strip() == "num_veh = param.num_veh;"
    assert L[288].strip().startswith("cost = cost_punish + cost_u + cost_AL") and "last_dis(1)" in L[269]
    assert Lp[356].strip().startswith("if nargin  == 3") and Lp[396].strip() == "end"
    body, roll = "\n".join(L[261:289]), "\n".join(Lp[356:397])
    base = np.array([[-3.0, 0.2, 0.1], [0.4, 5.0, -1.45], [3.0, 1.0, 2.9], [-0.5, -4.0, 1.5]])
    speeds = np.array([4.0, 8.0, 6.0, 5.0])
    out = []
    for V, N in ((2, 5), (2, 15), (3, 8), (4, 6)):
        P = ms.Struct(dt=0.1, L=1.0, num_ho=float(N), num_veh=float(V), dis_thres=2.0, spd=speeds[:V].reshape(V, 1), beta=1000.0,
                      Pnorm=5.0, Pcost=1.0, rho=2.0)
        xt = base[:V] + rng.normal(0, 0.2, (V, 3))
        spd = speeds[:V]
        prev = np.stack((xt[:, 0] + 0.1 * spd * np.cos(xt[:, 2]), xt[:, 1] + 0.1 * spd * np.sin(xt[:, 2])), 1)
        k = np.arange(N + 1)
        pos_old = np.vstack([xt[v, c] + k * 0.1 * spd[v] * (np.cos, np.sin)[c](xt[v, 2]) for v in range(V) for c in (0, 1)]) + rng.normal(0, 0.1, (2 * V, N + 1))
        dual = rng.normal(0, 0.5, (2 * V, N + 1))
        hus = np.vstack((np.zeros((1, V, N)), rng.normal(0, 0.3, (5, V, N))))
        costs = []
        for hu in hus:
            z = lambda: np.zeros((V, N + 1))
            ns = dict(param=P, xt=xt.copy(), u=hu, hat_u=hu, nargin=6.0, x_pred=z(), y_pred=z(), theta_pred=z(),
                      pos_old=pos_old, dual_var_old=dual, prev_pred_pos=prev)
            I = ms.Interp(ns)
            I.run(roll)
            I.run(body)
            costs.append(float(ns["cost"]))
        out.append(dict(V=V, N=N, spd=spd.tolist(), xt=xt.tolist(), prev_pred_pos=prev.tolist(), pos_old=pos_old.tolist(), dual=dual.tolist(),
                        hat_u=hus.tolist(), cost=costs))
    return out


def mat_(x):
    x = np.asarray(x, dtype=np.float64)
    return x.reshape(1, 1) if x.ndim == 0 else x


def main():
    rng = np.random.default_rng(20261105)
    PI_cls = load_class()
    PI = PI_cls()
    out = {"generator": "tests/golden/make_reference_vectors.py (reference code imported unmodified from /root/reference; numpy stand-ins "
                        "for the casadi / bunch containers, no solver)",
           "param": {k: (v.tolist() if isinstance(v, np.ndarray) else v) for k, v in PI.param.items()},
           "ref_traj": PI.ref_traj.tolist()}
    xt0 = np.array([[-10.0, 0.0, 0.0], [0.0, 20.0, -np.pi / 2]])
    veh = []
    for (step, v, N, spd, xt) in [(0, 0, 15, (4, 8), xt0), (3, 1, 15, (4, 8), xt0 + rng.normal(0, 0.3, (2, 3))),
                                  (12, 0, 15, (4.3, 7.4), np.array([[-5.2, 0.1, 0.05], [0.2, 10.4, -1.5]])),
                                  (5, 1, 30, (6.5, 5.0), np.array([[-7.0, 0.3, 0.7], [0.4, 12.0, -1.2]])),
                                  (20, 0, 8, (8, 4), np.array([[-1.0, 0.0, 2.1], [0.0, 3.0, -0.4]]))]:
        veh.append(vehicle_case(PI_cls(), rng, step, v, np.asarray(xt, dtype=float), N, spd))
    out["vehicle"] = veh
    near = np.array([[-4.0, 0.0, 0.0], [0.0, 7.0, -np.pi / 2]])        # close enough for the hinge of the edge cost to be active
    out["edge"] = [edge_case(PI_cls(), rng, near + rng.normal(0, 0.3, (2, 3)), N, spd) for N, spd in ((15, (4, 8)), (30, (6, 7)), (5, (8, 4)))]
    out["obca"] = obca_cases(rng)
    out["old_pi_block"] = old_pi_block_cases(np.random.default_rng(77))
    out["matlab_pi_block"] = matlab_pi_block_cases(np.random.default_rng(78))
    out["matlab_functions"] = matlab_function_cases(np.random.default_rng(79))
    out["matlab_joint_cost"] = matlab_joint_cost_cases(np.random.default_rng(80))
    path = os.path.join(HERE, "reference_vectors.json")
    json.dump(out, open(path, "w"))
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
