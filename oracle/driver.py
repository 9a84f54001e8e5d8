"""CPU ORACLE (test infrastructure, never a product path) -- outer PI-ADMM loop.

Literal restatement of one receding-horizon step of
/root/reference/casadi/main.py:46-192 (Python "current": plain dual update and
the quirks listed in SURVEY.md appendix D), with the PI + anti-windup dual
updates of
  * matlab_old_files/ADMM_CVX_two_veh_intesection_PI_antiwindup.m:152-188 (dual="pi")
  * casadi_old_PI_ADMM/main.py:127-151                                   (dual="pi_old")
selectable through ``Modes``.  The edge step is the convex statement
(MATLAB :276-304 with :383-396); the CasADi Taylor-at-0 model of the non-convex
Python edge cost is NOT reproduced (ill-posed).

OUTER LOOP PINNED to the reference's own script: tests/golden/reference_main_run.json is /root/reference/casadi/main.py run
unmodified (QP solves stood in by oracle/solver.py, tests/golden/run_reference_main.py); run_closed_loop in python_current
semantics reproduces its histories exactly (tests/test_reference_vectors.py).  The two PI laws are pinned statement by statement:
pi_old against lines 127-164 of casadi_old_PI_ADMM/main.py exec'd verbatim, pi against the MATLAB script's own text executed by
tests/golden/_matlab_subset.py (:6-25,105-124,152-209) -- and the WHOLE MATLAB script run from its text for its 42 MPC steps
(tests/golden/run_reference_matlab.py): run_closed_loop in matlab_pi semantics (tight) reproduces its iter_his and trajectories.  The inner solver's iterates stay unpinned (oracle/problem.py header).
"""
from __future__ import annotations

import numpy as np

from .problem import (Params, affine_rollout, edge_qp_blocks, literal_dynamic_update_edge,
                      vehicle_qp)
from .solver import InnerSettings, solve_box_qp, solve_edge_pair


class Modes:
    """Semantics switches; defaults = casadi/main.py as written (config 1)."""

    def __init__(self, **kw):
        self.dual = "plain"            # "plain" | "pi" | "pi_old"
        self.thr_squared = 0           # 0: d^2 < dis_thres (main.py:113 quirk); 1: d^2 < dis_thres^2 (MATLAB :109)
        self.dual_res_alias = 1        # main.py:180 aliasing -> s == 0 from outer iteration 2
        self.stop_dist_test = 0        # MATLAB :202 / old main.py:157 extra "dis_vec[1] > dis_thres"
        self.pos_nonlinear = 0         # MATLAB :312-330 numeric pos_old roll-out is non-linear
        self.curr_iter_latest = 0      # MATLAB :209 curr_iter = i_iter every iteration
        self.round_u = 4               # main.py:103 ; -1 = no rounding (MATLAB)
        self.round_edge = 4            # main.py:153 ; old main.py:119 rounds to 0 decimals
        self.al_all = 1                # AL over every j != i (PI_ADMM_class.py:126-129) vs only activated edges
        self.always_edge = 0           # old main.py: edge step every iteration, no collision test
        self.residual_global = 0       # old main.py:154-155 both-sides Frobenius norms (no 2x)
        # dual="pi" (MATLAB :20-25,43)
        self.PI = 1
        self.theta1 = 5.0
        self.theta2 = 3.0
        self.kI = None                 # None -> param.rho (MATLAB :21)
        self.windup = 1
        self.windup_sat = 30.0
        self.backcalc = 1.0            # MATLAB adds diff once; old main.py:142 adds 2*diff
        # dual="pi_old" (old main.py:24,135-137)
        self.K_I = 3.0
        self.KP_num = 5.0
        self.KP_max = 2.5
        self.rho_num = 4.0
        self.rho_min = 1.0
        self.rho_max = 5.0
        self.trad = 0
        for k, v in kw.items():
            setattr(self, k, v)

    @classmethod
    def python_current(cls):
        return cls()

    @classmethod
    def matlab_pi(cls, **kw):
        d = dict(dual="pi", thr_squared=1, dual_res_alias=0, stop_dist_test=1, pos_nonlinear=1,
                 curr_iter_latest=1, round_u=-1, round_edge=-1, windup_sat=30.0, backcalc=1.0)
        d.update(kw)
        return cls(**d)

    @classmethod
    def python_old_pi(cls, **kw):
        d = dict(dual="pi_old", always_edge=1, stop_dist_test=1, residual_global=1, windup_sat=20.0,
                 backcalc=2.0, round_edge=4)
        d.update(kw)
        return cls(**d)


def _round(a, dec):
    return np.around(a, dec) if dec >= 0 else np.asarray(a, dtype=np.float64)


def pi_gain(md: Modes, dmin):
    """MATLAB ...PI_antiwindup.m:160: kP = theta1 - theta2 / (1 + exp(-min(dis_vec)))."""
    return md.theta1 - md.theta2 / (1.0 + np.exp(-dmin))


def pi_integrate_and_output(md: Modes, rho, dmin, e, integ, diffv):
    """MATLAB ...PI_antiwindup.m:160-166 for one directed edge: integrate (kI = rho unless given, :21) THEN output."""
    kI = rho if md.kI is None else md.kI
    integ = integ + kI * e + md.backcalc * diffv
    return integ + pi_gain(md, dmin) * e, integ


def pi_old_gains(md: Modes, dmin):
    """casadi_old_PI_ADMM/main.py:135-137: K_P = min(5/d_min, 2.5), rho = max(1, min(5, 4/d_min)) (K_I = 3 is md.K_I)."""
    return min(md.KP_num / dmin, md.KP_max), max(md.rho_min, min(md.rho_max, md.rho_num / dmin))


def pi_old_output_and_integrate(md: Modes, dmin, e, dual, integ, diffv):
    """casadi_old_PI_ADMM/main.py:138-142 for one block of rows: output THEN integrate (trad: plain update + diff)."""
    K_P, rho = pi_old_gains(md, dmin)
    if md.trad:
        return dual + rho * e + diffv, integ
    return integ + K_P * e, integ + md.K_I * e + md.backcalc * diffv


def clamp_with_backcalc(md: Modes, raw):
    """Saturation + back-calculation term (old main.py:144-151, MATLAB :173-188): diff = clamped - raw, 0 where unclipped."""
    cl = np.minimum(md.windup_sat, np.maximum(raw, -md.windup_sat))
    return cl, cl - raw


def run_step(p: Params, md: Modes, inner: InnerSettings, xt, spd, refwin, rho_state=None):
    """One MPC step (all outer iterations).

    xt (V,3), spd (V,), refwin (V,2,N+1).  Returns dict with primal_u (V,N), pos (V,2,N+1),
    hat/dual dicts keyed (i,j), curr_iter (reference's iter_his entry), n_outer (iterations
    executed), err_rk / err_sk histories, per-solve inner iteration counts, xt_next, rho.
    """
    V = xt.shape[0]
    N = refwin.shape[2] - 1
    rho = float(p.rho if rho_state is None else rho_state)
    zeros = lambda: np.zeros((2, N + 1))
    seed = np.stack((_round(xt[:, 0] + p.dt * spd * np.cos(xt[:, 2]), 4),
                     _round(xt[:, 1] + p.dt * spd * np.sin(xt[:, 2]), 4)), axis=1)   # main.py:48-49
    seed_exact = np.stack((xt[:, 0] + p.dt * spd * np.cos(xt[:, 2]),
                           xt[:, 1] + p.dt * spd * np.sin(xt[:, 2])), axis=1)
    pos = np.zeros((V, 2, N + 1))
    hat, dual, integ, diffv, last_hat = {}, {}, {}, {}, {}
    active = set()                       # ordered pairs (i,j) present in the AL term
    if md.al_all:
        for i in range(V):
            for j in range(V):
                if i != j:
                    active.add((i, j))
    for (i, j) in [(i, j) for i in range(V) for j in range(V) if i != j]:
        hat[i, j] = zeros(); dual[i, j] = zeros(); integ[i, j] = zeros(); diffv[i, j] = zeros()
        last_hat[i, j] = zeros()
    primal_u = np.zeros((V, N))
    flag = 0
    curr_iter = 1
    n_outer = 0
    err_rk, err_sk, veh_iters, edge_iters = [], [], [], []
    ever = set()                         # undirected edges ever activated this step
    aliased = False
    for i_iter in range(p.iter_num):
        n_outer += 1
        # ---- x-update (main.py:81-106) ----
        for i in range(V):
            nb = sorted(j for (a, j) in active if a == i)
            H, q = vehicle_qp(p, xt[i], spd[i], refwin[i, 0], refwin[i, 1],
                              [hat[i, j] for j in nb], [dual[i, j] for j in nb], rho=rho)
            u, info = solve_box_qp(H, q, inner)
            veh_iters.append(info["iters"])
            u = _round(u, md.round_u)
            primal_u[i] = u
            if md.pos_nonlinear:
                x, y, _ = literal_dynamic_update_edge(p, xt[i:i + 1], u[None, :], spd[i:i + 1])
                pos[i, 0], pos[i, 1] = x[0], y[0]
            else:
                pos[i, 0], pos[i, 1] = affine_rollout(p, xt[i], spd[i], u)
        # ---- collision test (main.py:110-118) ----
        thr = p.dis_thres ** 2 if md.thr_squared else p.dis_thres
        edges = []
        for i in range(V):
            for j in range(i + 1, V):
                d2 = (pos[i, 0] - pos[j, 0]) ** 2 + (pos[i, 1] - pos[j, 1]) ** 2
                if md.always_edge or np.any(d2 < thr):
                    edges.append((i, j))
        if len(edges) == 0 and flag == 0:
            break
        flag = 1
        # ---- z-update + dual update per edge (main.py:121-162) ----
        rho_it = rho                     # every edge of this outer iteration sees the same penalty
        for (v1, v2) in edges:
            ever.add((v1, v2))
            active.add((v1, v2)); active.add((v2, v1))
            xt_e = np.vstack((xt[v1], xt[v2])); spd_e = np.array([spd[v1], spd[v2]])
            pos_e = np.vstack((pos[v1], pos[v2]))
            dual_e = np.vstack((dual[v1, v2], dual[v2, v1]))
            blk = edge_qp_blocks(p, xt_e, spd_e, pos_e, dual_e, np.vstack((seed[v1], seed[v2])), rho=rho_it)
            uh, info = solve_edge_pair(blk["H"], blk["q"], blk["a"], blk["c"], p.beta, inner)
            edge_iters.append(info["iters"])
            uh = _round(uh, md.round_edge)
            hx, hy, _ = literal_dynamic_update_edge(p, xt_e, uh, spd_e)       # main.py:156-158 (non-linear)
            new1 = np.vstack((hx[0], hy[0])); new2 = np.vstack((hx[1], hy[1]))
            if aliased and md.dual_res_alias:
                last_hat[v1, v2] = new1; last_hat[v2, v1] = new2              # alias: same objects
            hat[v1, v2] = new1; hat[v2, v1] = new2
            e1 = pos[v1] - hat[v1, v2]; e2 = pos[v2] - hat[v2, v1]
            if md.dual == "plain" or (md.dual == "pi" and not md.PI):
                dual[v1, v2] = dual[v1, v2] + rho_it * e1
                dual[v2, v1] = dual[v2, v1] + rho_it * e2
            elif md.dual == "pi":
                dmin = np.sqrt(((pos[v1] - pos[v2]) ** 2).sum(axis=0)).min()
                for key, e in (((v1, v2), e1), ((v2, v1), e2)):
                    dual[key], integ[key] = pi_integrate_and_output(md, rho_it, dmin, e, integ[key], diffv[key])
            elif md.dual == "pi_old":
                px, py, _ = literal_dynamic_update_edge(p, xt_e, np.vstack((primal_u[v1], primal_u[v2])), spd_e)
                dmin = np.sqrt((px[0] - px[1]) ** 2 + (py[0] - py[1]) ** 2).min()
                rho = pi_old_gains(md, dmin)[1]
                for key, e in (((v1, v2), e1), ((v2, v1), e2)):
                    dual[key], integ[key] = pi_old_output_and_integrate(md, dmin, e, dual[key], integ[key], diffv[key])
            else:
                raise ValueError(md.dual)
            if md.dual != "plain" and md.windup:
                for key in ((v1, v2), (v2, v1)):
                    dual[key], diffv[key] = clamp_with_backcalc(md, dual[key])
        # ---- residuals (main.py:165-178) ----
        sk = 0.0; rk = 0.0
        for (v1, v2) in edges:
            if md.residual_global:
                sk += np.sqrt(np.sum((rho * (last_hat[v1, v2] - hat[v1, v2])) ** 2)
                              + np.sum((rho * (last_hat[v2, v1] - hat[v2, v1])) ** 2))
                rk += np.sqrt(np.sum((pos[v1] - hat[v1, v2]) ** 2) + np.sum((pos[v2] - hat[v2, v1]) ** 2))
            else:
                sk += 2 * np.sqrt(np.sum((rho * (last_hat[v1, v2] - hat[v1, v2])) ** 2))
                rk += 2 * np.sqrt(np.sum((pos[v1] - hat[v1, v2]) ** 2))
        err_rk.append(rk); err_sk.append(sk)
        dist_ok = True
        if md.stop_dist_test:
            for (v1, v2) in sorted(k for k in active if k[0] < k[1]):
                if not np.sqrt(np.sum((seed_exact[v1] - seed_exact[v2]) ** 2)) > p.dis_thres:
                    dist_ok = False
        if rk <= p.eps_pri and sk <= p.eps_dual and dist_ok:
            curr_iter = i_iter + 1
            break
        if md.dual_res_alias:
            aliased = True
            for key in hat:
                last_hat[key] = hat[key]
        else:
            for key in hat:
                last_hat[key] = hat[key].copy()
        if md.curr_iter_latest:
            curr_iter = i_iter + 1
    # ---- plant step (main.py:185-192) ----
    px, py, pth = literal_dynamic_update_edge(p, xt, primal_u, spd)
    xt_next = np.stack((px[:, 1], py[:, 1], pth[:, 1]), axis=1)
    return dict(primal_u=primal_u, pos=pos, hat=hat, dual=dual, integ=integ, diff=diffv,
                curr_iter=curr_iter, n_outer=n_outer, err_rk=np.array(err_rk), err_sk=np.array(err_sk),
                veh_iters=veh_iters, edge_iters=edge_iters, xt_next=xt_next, rho=rho, flag=flag,
                edges_ever=sorted(ever))


def run_closed_loop(p: Params, md: Modes, inner: InnerSettings, xt0, spd, ref_traj, n_steps=None):
    """casadi/main.py:43-201: the receding-horizon loop.  ref_traj (2V, T)."""
    xt = np.array(xt0, dtype=np.float64)
    V = xt.shape[0]; N = p.num_ho
    T = ref_traj.shape[1]
    n_steps = T - N if n_steps is None else n_steps
    out = dict(x_vec=[], theta_vec=[], u_vec=[], iter_his=[], n_outer=[], err_rk=[], err_sk=[], steps=[])
    rho = p.rho
    for num_step in range(n_steps):
        refwin = np.stack([ref_traj[2 * i:2 * i + 2, num_step:num_step + N + 1] for i in range(V)])
        r = run_step(p, md, inner, xt, spd, refwin, rho_state=rho)
        rho = r["rho"] if md.dual == "pi_old" else rho
        xt = r["xt_next"]
        out["x_vec"].append(xt[:, :2].T.copy()); out["theta_vec"].append(xt[:, 2].copy())
        out["u_vec"].append(r["primal_u"][:, 0].copy())
        out["iter_his"].append(r["curr_iter"]); out["n_outer"].append(r["n_outer"])
        out["err_rk"].append(r["err_rk"]); out["err_sk"].append(r["err_sk"])
        out["steps"].append(r)
    return out
