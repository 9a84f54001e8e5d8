"""CPU ORACLE (test infrastructure, never a product path) -- problem definition.

Plain numpy fp64 restatement of the reference's PI-ADMM problem class
(/root/reference/casadi/PI_ADMM_class.py) and of the well-posed (convex) edge
step of the MATLAB CVX script
(/root/reference/matlab_old_files/ADMM_CVX_two_veh_intesection_PI_antiwindup.m).

Two layers live here:

* ``literal_*``: line-by-line numeric transcriptions of the reference methods
  (loops kept).  They are slow and exist to pin the closed forms below.
* closed forms (``roll_matrix``, ``vehicle_qp``, ``edge_qp_blocks``): what the
  C twin (oracle/piadmm_ref.c) and the CUDA kernels implement.

PINNED, for everything in THIS file, against outputs of the reference's own code:
tests/golden/reference_vectors.json is produced by importing casadi/PI_ADMM_class.py
unmodified from /root/reference (numpy stand-ins for the `casadi` / `bunch` containers,
tests/golden/make_reference_vectors.py) and calling its methods on numeric arrays --
param, ref_traj, both roll-outs, cost_function_primal (its exact (H, q)),
nonlcon_function, cost_function_edge, the convex edge statement composed from the
reference's two functions -- and the MATLAB script's own function bodies (roll-outs,
cost_function_primal, the convex cost_function_edge) executed from the .m text by
tests/golden/_matlab_subset.py; tests/test_reference_vectors.py compares the closed forms
below, the C twin's inputs and the CUDA path with them.
PARITY UNPINNED for solver OUTPUTS: the reference ships no tests / golden vectors and
its solver stack (CasADi 3.5.6.pre2 + vendored OSQP) is not installable here, so no
iterate or iteration count of the inner solver could be checked against the reference
(oracle/solver.py, oracle/driver.py).  DESIGN.md section "Oracle" has the full list.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference arm
may import this package.
"""
from __future__ import annotations

import numpy as np


class Params:
    """Attribute bag with the constants of casadi/PI_ADMM_class.py:15-28."""

    def __init__(self, **kw):
        self.dt = 0.1
        self.Nt = 5
        self.L = 1.0
        self.num_ho = 15
        self.num_veh = 2
        self.dis_thres = 2.0
        self.spd = np.array([4.0, 8.0])
        self.beta = 10.0
        self.Pnorm = 1.0
        self.Pcost = 1.0
        self.iter_num = 100
        self.rho = 2.0
        self.eps_pri = 20.0
        self.eps_dual = 1.0
        for k, v in kw.items():
            setattr(self, k, v)


def reference_trajectories(p: Params) -> np.ndarray:
    """casadi/PI_ADMM_class.py:31-37 -> (4, Nt/dt) array [x_A; 0; 0; y_B]."""
    T = int(p.Nt / p.dt)
    x_A = np.linspace(-10, 10, T)
    y_B = np.linspace(20, -20, T)
    return np.vstack((x_A, np.zeros_like(x_A), np.zeros_like(y_B), y_B))


# --------------------------------------------------------------------------
# literal transcriptions (numeric branch, if_SX=0, and numeric evaluation of
# the symbolic branch)
# --------------------------------------------------------------------------
def literal_dynamic_update_local(p: Params, xt, u, spd):
    """casadi/PI_ADMM_class.py:45-70 (linearised about xt[2]).  u is 1-D."""
    N = len(u)
    x = np.zeros(N + 1); y = np.zeros(N + 1); th = np.zeros(N + 1)
    x[0], y[0], th[0] = xt[0], xt[1], xt[2]
    for k in range(N):
        x_dot = -spd * np.sin(xt[2]) * th[k] + (spd * np.cos(xt[2]) + spd * xt[2] * np.sin(xt[2]))
        x[k + 1] = x[k] + x_dot * p.dt
        y_dot = spd * np.cos(xt[2]) * th[k] + (spd * np.sin(xt[2]) - spd * xt[2] * np.cos(xt[2]))
        y[k + 1] = y[k] + y_dot * p.dt
        th[k + 1] = th[k] + spd / p.L * u[k] * p.dt
    return x, y, th


def literal_dynamic_update_edge(p: Params, xt, u, spd):
    """casadi/PI_ADMM_class.py:77-105 numeric branch (NON-linear roll-out).

    xt (R,3), u (R,N), spd (R,) -- the reference indexes param.spd by the local
    row (``:93``); callers here pass the speeds of the rows explicitly.
    """
    R, N = u.shape
    x = np.zeros((R, N + 1)); y = np.zeros((R, N + 1)); th = np.zeros((R, N + 1))
    x[:, 0], y[:, 0], th[:, 0] = xt[:, 0], xt[:, 1], xt[:, 2]
    for k in range(N):
        for i in range(R):
            v = spd[i]
            x_dot = -v * np.sin(th[i, k]) * th[i, k] + (v * np.cos(th[i, k]) + v * th[i, k] * np.sin(th[i, k]))
            x[i, k + 1] = x[i, k] + x_dot * p.dt
            y_dot = v * np.cos(th[i, k]) * th[i, k] + (v * np.sin(th[i, k]) - v * th[i, k] * np.cos(th[i, k]))
            y[i, k + 1] = y[i, k] + y_dot * p.dt
            th[i, k + 1] = th[i, k] + v / p.L * u[i, k] * p.dt
    return x, y, th


def literal_edge_linear_rollout(p: Params, xt, u, spd):
    """MATLAB ...PI_antiwindup.m:379-396 (symbolic branch: linearised about xt(:,3))."""
    R, N = u.shape
    x = np.zeros((R, N + 1)); y = np.zeros((R, N + 1))
    for i in range(R):
        x[i], y[i], _ = literal_dynamic_update_local(p, xt[i], u[i], spd[i])
    return x, y


def literal_cost_primal(p: Params, u, xt_i, spd_i, ref_x, ref_y, hat_list, dual_list, rho=None):
    """casadi/PI_ADMM_class.py:114-135 evaluated numerically.

    hat_list / dual_list: the 2x(N+1) blocks hat_pos[i,j], dual_var[i,j] of every
    AL neighbour j (the reference sums over ALL j != i).
    """
    rho = p.rho if rho is None else rho
    x, y, _ = literal_dynamic_update_local(p, xt_i, u, spd_i)
    cost_norm = p.Pnorm * np.sum((ref_x - x) ** 2 + (ref_y - y) ** 2)
    cost_smooth = np.sum((u[2:] - 2 * u[1:-1] + u[:-2]) ** 2)
    cost_AL = 0.0
    for hat, dual in zip(hat_list, dual_list):
        cost_AL += rho / 2 * np.sum((np.vstack((x, y)) - hat + dual) ** 2)
    cost_u = p.Pcost * np.sum(u ** 2)
    return cost_norm + cost_smooth + cost_AL + cost_u


def literal_cost_edge_convex(p: Params, hat_u, xt_edge, spd_edge, pos_old_edge, dual_edge, prev_pred_pos, rho=None):
    """MATLAB ...PI_antiwindup.m:276-304 with the linearised roll-out :383-396
    (= casadi/PI_ADMM_class.py:145-169 with the roll-out made affine).

    hat_u (2,N); pos_old_edge/dual_edge (4,N+1) rows [x1;y1;x2;y2]; prev_pred_pos (2,2).
    """
    rho = p.rho if rho is None else rho
    x, y = literal_edge_linear_rollout(p, xt_edge, hat_u, spd_edge)
    last_dis = prev_pred_pos[1] - prev_pred_pos[0]
    curr = np.vstack((x[1, 1:] - x[0, 1:], y[1, 1:] - y[0, 1:]))
    dis_temp = 2 * (last_dis[0] * curr[0] + last_dis[1] * curr[1]) - np.sum(last_dis ** 2)
    cost_punish = p.beta * np.sum(np.maximum(0.0, p.dis_thres ** 2 - dis_temp))
    edge_pos = np.vstack((x[0], y[0], x[1], y[1]))
    cost_u = p.Pcost * np.sum(hat_u ** 2)
    cost_AL = rho / 2 * np.sum((pos_old_edge - edge_pos + dual_edge) ** 2)
    return cost_punish + cost_u + cost_AL


# --------------------------------------------------------------------------
# closed forms
# --------------------------------------------------------------------------
def roll_matrix(N: int) -> np.ndarray:
    """M (N+1,N): M[k,j] = max(k-1-j,0).  Double summation of the steering
    sequence implied by casadi/PI_ADMM_class.py:59-69 (theta integrates u, the
    linearised position integrates theta)."""
    k = np.arange(N + 1)[:, None]
    j = np.arange(N)[None, :]
    return np.maximum(k - 1 - j, 0).astype(np.float64)


def diff1(N: int) -> np.ndarray:
    D = np.zeros((N - 1, N))
    for k in range(N - 1):
        D[k, k] = -1.0; D[k, k + 1] = 1.0
    return D


def diff2(N: int) -> np.ndarray:
    D = np.zeros((N - 2, N))
    for k in range(N - 2):
        D[k, k] = 1.0; D[k, k + 1] = -2.0; D[k, k + 2] = 1.0
    return D


def affine_rollout(p: Params, xt_i, spd_i, u):
    """Closed form of the linearised roll-out: returns (x, y) each (N+1,)."""
    N = len(u)
    s, c = np.sin(xt_i[2]), np.cos(xt_i[2])
    g = spd_i * spd_i * p.dt * p.dt / p.L
    k = np.arange(N + 1)
    m = roll_matrix(N) @ u
    x = xt_i[0] + k * p.dt * spd_i * c - g * s * m
    y = xt_i[1] + k * p.dt * spd_i * s + g * c * m
    return x, y


def constraint_matrix(N: int) -> np.ndarray:
    """casadi/PI_ADMM_class.py:172-192: g(u) = A u + b >= 0 with rows
    [I; -I; -D1; D1] (m = 4N-2), b = [pi/6; pi/6; pi/9; pi/9]."""
    D1 = diff1(N)
    return np.vstack((np.eye(N), -np.eye(N), -D1, D1))


def constraint_lower(N: int) -> np.ndarray:
    return -np.concatenate((np.full(2 * N, np.pi / 6), np.full(2 * N - 2, np.pi / 9)))


def vehicle_qp(p: Params, xt_i, spd_i, ref_x, ref_y, hat_list, dual_list, rho=None):
    """Vehicle QP  min 1/2 u'Hu + q'u  s.t. A u >= l  (SURVEY B.2).

    Returns H (N,N), q (N,).  Constant term dropped.
    """
    rho = p.rho if rho is None else rho
    N = len(ref_x) - 1
    M = roll_matrix(N)
    D2 = diff2(N)
    s, c = np.sin(xt_i[2]), np.cos(xt_i[2])
    g = spd_i * spd_i * p.dt * p.dt / p.L
    d = len(hat_list)
    w = p.Pnorm + rho * d / 2.0
    tx = p.Pnorm * np.asarray(ref_x, dtype=np.float64).copy()
    ty = p.Pnorm * np.asarray(ref_y, dtype=np.float64).copy()
    for hat, dual in zip(hat_list, dual_list):
        tx += rho / 2 * (hat[0] - dual[0])
        ty += rho / 2 * (hat[1] - dual[1])
    e0 = -s * xt_i[0] + c * xt_i[1]
    H = 2.0 * (w * g * g * (M.T @ M) + D2.T @ D2 + p.Pcost * np.eye(N))
    q = 2.0 * g * (M.T @ (w * e0 + s * tx - c * ty))
    return H, q


def edge_qp_blocks(p: Params, xt_edge, spd_edge, pos_old_edge, dual_edge, prev_pred_pos, rho=None):
    """Convex edge problem in the split form the kernels use:

        min  sum_a [1/2 u_a' H_a u_a + q_a' u_a] + beta * sum_k max(0, c_k - a1*(M u_1)_k - a2*(M u_2)_k)
        s.t. A u_a >= l  (a = 1,2),    k = 1..N

    Returns dict(H=[H1,H2], q=[q1,q2], a=(a1,a2), c=(N,) hinge offsets).
    """
    rho = p.rho if rho is None else rho
    N = pos_old_edge.shape[1] - 1
    M = roll_matrix(N)
    kk = np.arange(N + 1)
    ell = prev_pred_pos[1] - prev_pred_pos[0]
    Hs, qs, a, base = [], [], [], []
    for r in range(2):
        th0 = xt_edge[r, 2]
        s, c = np.sin(th0), np.cos(th0)
        v = spd_edge[r]
        g = v * v * p.dt * p.dt / p.L
        xbar = xt_edge[r, 0] + kk * p.dt * v * c
        ybar = xt_edge[r, 1] + kk * p.dt * v * s
        # AL: rho/2 || pos_old - P + lambda ||^2, P = [xbar - g s m ; ybar + g c m]
        ex = pos_old_edge[2 * r] + dual_edge[2 * r] - xbar
        ey = pos_old_edge[2 * r + 1] + dual_edge[2 * r + 1] - ybar
        Hs.append(2.0 * p.Pcost * np.eye(N) + rho * g * g * (M.T @ M))
        qs.append(rho * g * (M.T @ (s * ex - c * ey)))
        sign = -1.0 if r == 0 else 1.0
        # dis_k = 2 ell.(P2_k - P1_k) - |ell|^2 ; coefficient of (M u_r)_k
        a.append(sign * 2.0 * g * (-ell[0] * s + ell[1] * c))
        base.append(sign * 2.0 * (ell[0] * xbar + ell[1] * ybar))
    dis0 = base[0] + base[1] - np.sum(ell ** 2)          # dis_k at u = 0, k = 0..N
    cvec = (p.dis_thres ** 2 - dis0)[1:]                  # hinge active when c_k - a.(Mu)_k > 0
    return dict(H=Hs, q=qs, a=(a[0], a[1]), c=cvec)
