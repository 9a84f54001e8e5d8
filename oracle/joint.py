"""CPU ORACLE (test infrastructure, never a product path) -- the joint N-vehicle RSU edge problem, SURVEY.md 8(f)-4.

Restates ``cost_function_edge`` of
/root/reference/matlab_old_files/ADMM_fmincon_multi_veh_intesection_static.m:250-291 (= casadi_old_PI_ADMM/
PI_ADMM_class.py:141-178): ONE edge problem over all vehicles of a scene,

    cost = beta * sum_i sum_{j != i} sum_{k=1..N} max(0, d^2 - (2 l_ij'(P_j,k - P_i,k) - |l_ij|^2))      (:265-276)
         + Pcost * |hat_u|^2  +  rho/2 * | pos_old - edge_pos + dual_var_old |^2                          (:285-287)
    s.t.  nonlcon_function(hat_u) on every row                                                            (:240-248)

with ONE consensus copy and ONE multiplier per vehicle (``pos_old`` / ``dual_var_old`` are 2V x (N+1)), in its convex
statement: the roll-out linearised about each vehicle's heading (MATLAB ...PI_antiwindup.m:383-396), l_ij = seed_j - seed_i.
The ordered-pair sum counts every unordered pair twice (both terms are identical), i.e. the hinge weight per pair is 2 beta.

Deviations, stated:
* the reference builds ``last_dis`` from LINEAR indices (``last_dis(1)``, ``last_dis(2)``, :268), which is the pair's own
  (dx, dy) only for V = 2 -- for V > 2 it reads the dx of the first two other vehicles; restated here per pair (the
  evident intent, and what every two-vehicle statement does);
* the pair set is a caller-given list (default: every pair i < j); the product kernel takes at most 8 pairs per vehicle.

Solver: the OSQP-style ADMM of oracle/solver.py on the split  z_i = [A u_i ; one copy of (M u_i)_{1..N} per incident pair],
so the linear system stays block diagonal (one N x N matrix per vehicle, K_i = H_i + sigma I + rho (A'A + deg_i M'M)) and
the pairs couple only through the closed-form 2-D prox of their hinge.  One rho for the whole problem, termination and
adaptive rho on the residuals of the whole problem.  Pinning: the COST (literal_cost_joint, and through it joint_blocks) equals
the MATLAB function body executed from its text, V = 2 exactly and V = 3, 4 in its as-written form
(tests/golden/reference_vectors.json key matlab_joint_cost); the MINIMISER is pinned against scipy SLSQP on the slack form
and, for V = 2, against solve_edge_pair with beta doubled (tests/test_joint_edge.py); fmincon's own iterates stay unpinned.
"""
from __future__ import annotations

import numpy as np

from .problem import Params, constraint_lower, constraint_matrix, roll_matrix
from .solver import InnerSettings, _new_rho


def all_pairs(V):
    return [(i, j) for i in range(V) for j in range(i + 1, V)]


def seeds(p: Params, xt, spd):
    """casadi/main.py:48-49 / ...static.m:52-53: one-step-ahead points, rounded to 4 decimals as the Python script does."""
    xt = np.asarray(xt, dtype=np.float64)
    return np.stack((np.round(xt[:, 0] + p.dt * spd * np.cos(xt[:, 2]), 4),
                     np.round(xt[:, 1] + p.dt * spd * np.sin(xt[:, 2]), 4)), 1)


def joint_blocks(p: Params, xt, spd, pos_old, dual, pairs, rho=None, prev_pred_pos=None):
    """Per vehicle (H_i, q_i) and per pair (a_i, a_j, c[N]) of the joint problem.
    pos_old, dual: (V, 2, N+1).  ``prev_pred_pos`` (V, 2) replaces the rounded seed points when given.
    Returns dict(H, q, pairs=[(i, j, ai, aj, c)])."""
    rho = p.rho if rho is None else rho
    V, N = len(spd), pos_old.shape[2] - 1
    M = roll_matrix(N)
    kk = np.arange(N + 1)
    sd = seeds(p, xt, spd) if prev_pred_pos is None else np.asarray(prev_pred_pos, dtype=np.float64)
    H, q, sn, cs, g, xb, yb = [], [], [], [], [], [], []
    for i in range(V):
        s, c = np.sin(xt[i, 2]), np.cos(xt[i, 2])
        gi = spd[i] * spd[i] * p.dt * p.dt / p.L
        xbar = xt[i, 0] + kk * p.dt * spd[i] * c
        ybar = xt[i, 1] + kk * p.dt * spd[i] * s
        ex = pos_old[i, 0] + dual[i, 0] - xbar
        ey = pos_old[i, 1] + dual[i, 1] - ybar
        H.append(2.0 * p.Pcost * np.eye(N) + rho * gi * gi * (M.T @ M))
        q.append(rho * gi * (M.T @ (s * ex - c * ey)))
        sn.append(s); cs.append(c); g.append(gi); xb.append(xbar); yb.append(ybar)
    out = []
    for (i, j) in pairs:
        ell = sd[j] - sd[i]
        ai = -2.0 * g[i] * (-ell[0] * sn[i] + ell[1] * cs[i])
        aj = 2.0 * g[j] * (-ell[0] * sn[j] + ell[1] * cs[j])
        dis0 = -2.0 * (ell[0] * xb[i] + ell[1] * yb[i]) + 2.0 * (ell[0] * xb[j] + ell[1] * yb[j]) - np.sum(ell ** 2)
        out.append((i, j, ai, aj, (p.dis_thres ** 2 - dis0)[1:]))
    return dict(H=H, q=q, pairs=out)


def literal_cost_joint(p: Params, hat_u, xt, spd, pos_old, dual, pairs, rho=None, prev_pred_pos=None, as_written=False):
    """The cost of ...static.m:250-291 evaluated literally (ordered pairs, linearised roll-out): pins joint_blocks.
    ``prev_pred_pos`` (V, 2) replaces the rounded seed points when given.  ``as_written=True`` keeps the script's
    linear indices ``last_dis(1)``, ``last_dis(2)`` (:270) -- column-major over the (V-1) x 2 difference matrix, so for
    V > 2 every other vehicle of row i gets the direction (dx to the first other, dx to the second other); ``pairs`` is
    then ignored (the script sums over every ordered pair).  Both forms are checked against the MATLAB text executed by
    tests/golden/_matlab_subset.py (tests/test_reference_vectors.py)."""
    from .problem import affine_rollout
    rho = p.rho if rho is None else rho
    V = len(spd)
    sd = seeds(p, xt, spd) if prev_pred_pos is None else np.asarray(prev_pred_pos, dtype=np.float64)
    P = [np.stack(affine_rollout(p, xt[i], spd[i], hat_u[i])) for i in range(V)]
    cost = 0.0
    if as_written:
        for i in range(V):
            others = [j for j in range(V) if j != i]
            flat = (sd[others] - sd[i]).ravel(order="F")
            ell = flat[:2]
            for j in others:
                cur = P[j][:, 1:] - P[i][:, 1:]
                dis = 2.0 * (ell[0] * cur[0] + ell[1] * cur[1]) - np.sum(ell ** 2)
                cost += p.beta * np.sum(np.maximum(0.0, p.dis_thres ** 2 - dis))
    else:
        for (a, b) in pairs:
            for (i, j) in ((a, b), (b, a)):
                ell = sd[j] - sd[i]
                cur = P[j][:, 1:] - P[i][:, 1:]
                dis = 2.0 * (ell[0] * cur[0] + ell[1] * cur[1]) - np.sum(ell ** 2)
                cost += p.beta * np.sum(np.maximum(0.0, p.dis_thres ** 2 - dis))
    cost += p.Pcost * np.sum(np.asarray(hat_u) ** 2)
    for i in range(V):
        cost += rho / 2.0 * np.sum((pos_old[i] - P[i] + dual[i]) ** 2)
    return cost


def solve_joint(H, q, pairs, hinge, s: InnerSettings):
    """min sum_i [1/2 u_i'H_i u_i + q_i'u_i] + hinge * sum_pairs sum_k max(0, c_k - ai (M u_i)_k - aj (M u_j)_k)
    s.t. A u_i >= l.  Returns u (V,N), info."""
    V, N = len(q), len(q[0])
    Ac = constraint_matrix(N); lc = constraint_lower(N); m = len(lc)
    Ms = roll_matrix(N)[1:]
    inc = [[] for _ in range(V)]                      # incident pairs of every vehicle, in list order
    for pi, (i, j, ai, aj, c) in enumerate(pairs):
        inc[i].append(pi); inc[j].append(pi)
    AcAc, MsMs = Ac.T @ Ac, Ms.T @ Ms
    x = [np.zeros(N) for _ in range(V)]
    zb = [np.zeros(m) for _ in range(V)]; yb = [np.zeros(m) for _ in range(V)]
    zs = {}; ys = {}                                   # (pair, side) -> position-row copies
    for pi in range(len(pairs)):
        for side in (0, 1):
            zs[pi, side] = np.zeros(N); ys[pi, side] = np.zeros(N)
    rho = s.rho0

    def factor(rho):
        return [np.linalg.inv(H[i] + s.sigma * np.eye(N) + rho * (AcAc + len(inc[i]) * MsMs)) for i in range(V)]

    def side_of(pi, i):
        return 0 if pairs[pi][0] == i else 1

    K = factor(rho)
    nref = 0; status = 1; it = 0
    for it in range(1, s.max_iter + 1):
        xt_ = [None] * V
        for i in range(V):
            rhs = s.sigma * x[i] - q[i] + Ac.T @ (rho * zb[i] - yb[i])
            for pi in inc[i]:
                sd_ = side_of(pi, i)
                rhs = rhs + Ms.T @ (rho * zs[pi, sd_] - ys[pi, sd_])
            xt_[i] = K[i] @ rhs
        for i in range(V):
            zt = Ac @ xt_[i]
            x[i] = s.alpha * xt_[i] + (1 - s.alpha) * x[i]
            zr = s.alpha * zt + (1 - s.alpha) * zb[i]
            zn = np.maximum(zr + yb[i] / rho, lc)
            yb[i] = yb[i] + rho * (zr - zn)
            zb[i] = zn
        for pi, (i, j, ai, aj, c) in enumerate(pairs):
            v = []
            zr = []
            for side, veh in ((0, i), (1, j)):
                zt = Ms @ xt_[veh]
                r_ = s.alpha * zt + (1 - s.alpha) * zs[pi, side]
                zr.append(r_); v.append(r_ + ys[pi, side] / rho)
            na = ai * ai + aj * aj
            tau = ai * v[0] + aj * v[1]
            t = hinge / rho
            step = np.where(tau >= c, 0.0, np.where(tau + t * na <= c, t, (c - tau) / na)) if na > 0.0 else np.zeros(N)
            for side, a_ in ((0, ai), (1, aj)):
                zn = v[side] + step * a_
                ys[pi, side] = ys[pi, side] + rho * (zr[side] - zn)
                zs[pi, side] = zn
        if it % s.check == 0:
            rp = rd = n_p = n_d = 0.0
            for i in range(V):
                Ax = Ac @ x[i]; Mx = Ms @ x[i]; Px = H[i] @ x[i]
                Aty = Ac.T @ yb[i]
                rp = max(rp, np.abs(Ax - zb[i]).max()); n_p = max(n_p, np.abs(Ax).max(), np.abs(zb[i]).max())
                for pi in inc[i]:
                    sd_ = side_of(pi, i)
                    Aty = Aty + Ms.T @ ys[pi, sd_]
                    rp = max(rp, np.abs(Mx - zs[pi, sd_]).max())
                    n_p = max(n_p, np.abs(Mx).max(), np.abs(zs[pi, sd_]).max())
                rd = max(rd, np.abs(Px + q[i] + Aty).max())
                n_d = max(n_d, np.abs(Px).max(), np.abs(Aty).max(), np.abs(q[i]).max())
            if rp <= s.eps_abs + s.eps_rel * n_p and rd <= s.eps_abs + s.eps_rel * n_d:
                status = 0
                break
            if s.adaptive:
                rn = _new_rho(rho, rp, rd, n_p, n_d)
                if rn > s.adapt_tol * rho or rn < rho / s.adapt_tol:
                    rho = rn
                    K = factor(rho)
                    nref += 1
    return np.stack(x), dict(iters=it, status=status, refactors=nref, rho=rho)


# ------------------------------------------------------------------ the joint flow: one MPC step
def capped_pairs(ever, V, cap=8):
    """Sorted pair list with at most `cap` pairs per vehicle (the kernel's per-vehicle row state); returns (pairs, dropped)."""
    deg = [0] * V
    out, dropped = [], 0
    for (i, j) in sorted(ever):
        if deg[i] < cap and deg[j] < cap:
            out.append((i, j)); deg[i] += 1; deg[j] += 1
        else:
            dropped += 1
    return out, dropped


def run_joint_step(p: Params, inner: InnerSettings, xt, spd, refwin, thr_squared=1, round_u=-1, round_edge=-1, cap=8, kI=None, pos_nonlinear=1):
    """One MPC step of the JOINT flow -- the loop of casadi_old_PI_ADMM/main.py:60-164 / ...static.m:57-164 with the
    edge side solved as ONE problem over all vehicles (cost_function_edge, ...static.m:250-291):

      x-update per vehicle against its ONE consensus copy (AL term only once the vehicle is in a pair),
      collision test (casadi/main.py:110-113), pairs = every pair that collided so far in this step (with the pairs that
      collide NOW only, the pair set oscillates and the loop does not settle: measured),
      joint edge problem -> hat_u, z-hat_i = non-linear roll-out of hat_u_i (casadi/main.py:156-158),
      lambda_i += kI (pos_i - z-hat_i)  (plain law, ...static.m:131-132; kI = rho unless given),
      global Frobenius residuals (old main.py:154-155) over the vehicles that are in a pair.
    """
    from .problem import vehicle_qp, affine_rollout, literal_dynamic_update_edge
    from .solver import solve_box_qp
    V, N = xt.shape[0], refwin.shape[2] - 1
    rho = float(p.rho)
    kI = rho if kI is None else float(kI)
    pos = np.zeros((V, 2, N + 1)); hat = np.zeros((V, 2, N + 1)); dual = np.zeros((V, 2, N + 1)); last_hat = np.zeros((V, 2, N + 1))
    u = np.zeros((V, N))
    ever, in_pair = set(), np.zeros(V, bool)
    err_rk, err_sk, joint_iters, n_outer, converged, dropped = [], [], [], 0, False, 0
    pairs = []
    thr = p.dis_thres ** 2 if thr_squared else p.dis_thres
    for it in range(p.iter_num):
        n_outer += 1
        for i in range(V):
            hl, dl = ([hat[i]], [dual[i]]) if in_pair[i] else ([], [])
            H, q = vehicle_qp(p, xt[i], spd[i], refwin[i, 0], refwin[i, 1], hl, dl, rho=rho)
            ui, _ = solve_box_qp(H, q, inner)
            u[i] = np.around(ui, round_u) if round_u >= 0 else ui
            if pos_nonlinear:                     # MATLAB ...PI_antiwindup.m:312-330: numeric pos_old is the non-linear roll-out
                x_, y_, _ = literal_dynamic_update_edge(p, xt[i:i + 1], u[i][None, :], spd[i:i + 1])
                pos[i, 0], pos[i, 1] = x_[0], y_[0]
            else:
                pos[i, 0], pos[i, 1] = affine_rollout(p, xt[i], spd[i], u[i])
        for i in range(V):
            for j in range(i + 1, V):
                d2 = (pos[i, 0] - pos[j, 0]) ** 2 + (pos[i, 1] - pos[j, 1]) ** 2
                if np.any(d2 < thr):
                    ever.add((i, j))
        if not ever:
            converged = True
            break
        pairs, dropped = capped_pairs(ever, V, cap)
        for (i, j) in pairs:
            in_pair[i] = in_pair[j] = True
        B = joint_blocks(p, xt, spd, pos, dual, pairs, rho=rho)
        uh, info = solve_joint(B["H"], B["q"], B["pairs"], 2.0 * p.beta, inner)
        joint_iters.append(info["iters"])
        uh = np.around(uh, round_edge) if round_edge >= 0 else uh
        hx, hy, _ = literal_dynamic_update_edge(p, xt, uh, spd)
        new = np.stack((hx, hy), 1)
        act = in_pair[:, None, None]
        hat = np.where(act, new, hat)
        dual = np.where(act, dual + kI * (pos - hat), dual)
        rk = np.sqrt(np.sum(np.where(act, (pos - hat) ** 2, 0.0)))
        sk = np.sqrt(np.sum(np.where(act, (rho * (last_hat - hat)) ** 2, 0.0)))
        err_rk.append(rk); err_sk.append(sk)
        if rk <= p.eps_pri and sk <= p.eps_dual:
            converged = True
            break
        last_hat = hat.copy()
    return dict(u=u, pos=pos, hat=hat, dual=dual, n_outer=n_outer, err_rk=np.array(err_rk), err_sk=np.array(err_sk),
                joint_iters=joint_iters, pairs=pairs, dropped=dropped, converged=converged)
