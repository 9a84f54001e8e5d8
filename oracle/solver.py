"""CPU ORACLE (test infrastructure, never a product path) -- inner QP solver.

Restatement of the OSQP ADMM iteration (Stellato et al., "OSQP: an operator
splitting solver for quadratic programs", Alg. 1 + sec. 5.2 adaptive rho) which
the reference reaches through ``ca.qpsol("solver","osqp",...)`` at
/root/reference/casadi/main.py:96,101,146,151.  OSQP itself is a third-party
dependency that is NOT vendored under /root/reference (it lives inside the
CasADi 3.5.6.pre2 wheel pinned at Distributed_planner/environment.yml:230), so
this file restates the published algorithm; documented deviations:

* no Ruiz equilibration (scaling=10 in OSQP) -- iterates differ, minimiser does not;
* adaptive-rho is evaluated at every termination check (every ``check`` its)
  instead of OSQP's wall-clock-derived interval (non-deterministic by design);
* the hinge term of the edge problem is handled by its proximal operator on
  split position rows instead of slack variables (same minimiser; see
  ``solve_edge_pair``).

PARITY UNPINNED (see oracle/problem.py header).
"""
from __future__ import annotations

import numpy as np

from .problem import constraint_matrix, constraint_lower, roll_matrix


class InnerSettings:
    def __init__(self, **kw):
        self.rho0 = 0.1
        self.sigma = 1e-6
        self.alpha = 1.6
        self.eps_abs = 1e-3
        self.eps_rel = 1e-3
        self.max_iter = 4000
        self.check = 25
        self.adaptive = 1
        self.adapt_tol = 5.0
        for k, v in kw.items():
            setattr(self, k, v)

    @classmethod
    def osqp_default(cls):
        return cls()

    @classmethod
    def tight(cls):
        return cls(eps_abs=1e-9, eps_rel=1e-9, max_iter=20000)


def _new_rho(rho, rp, rd, n_p, n_d):
    rn = rho * np.sqrt((rp / max(n_p, 1e-30)) / max(rd / max(n_d, 1e-30), 1e-30))
    return min(max(rn, 1e-6), 1e6)


def solve_dense_qp(P, q, A, l, u, s: InnerSettings):
    """min 1/2 x'Px + q'x  s.t.  l <= A x <= u  (dense OSQP iteration, uniform rho).  Returns x, y, info."""
    n = len(q)
    AtA = A.T @ A
    x = np.zeros(n); z = np.zeros(len(l)); y = np.zeros(len(l))
    rho = s.rho0
    K = np.linalg.inv(P + s.sigma * np.eye(n) + rho * AtA)
    nref = 0; status = 1
    it = 0
    for it in range(1, s.max_iter + 1):
        rhs = s.sigma * x - q + A.T @ (rho * z - y)
        xt = K @ rhs
        zt = A @ xt
        x = s.alpha * xt + (1 - s.alpha) * x
        zr = s.alpha * zt + (1 - s.alpha) * z
        zn = np.minimum(np.maximum(zr + y / rho, l), u)
        y = y + rho * (zr - zn)
        z = zn
        if it % s.check == 0:
            Ax = A @ x; Px = P @ x; Aty = A.T @ y
            rp = np.abs(Ax - z).max(); rd = np.abs(Px + q + Aty).max()
            n_p = max(np.abs(Ax).max(), np.abs(z).max())
            n_d = max(np.abs(Px).max(), np.abs(Aty).max(), np.abs(q).max())
            if rp <= s.eps_abs + s.eps_rel * n_p and rd <= s.eps_abs + s.eps_rel * n_d:
                status = 0
                break
            if s.adaptive:
                rn = _new_rho(rho, rp, rd, n_p, n_d)
                if rn > s.adapt_tol * rho or rn < rho / s.adapt_tol:
                    rho = rn
                    K = np.linalg.inv(P + s.sigma * np.eye(n) + rho * AtA)
                    nref += 1
    return x, y, dict(iters=it, status=status, refactors=nref, rho=rho)


def solve_box_qp(P, q, s: InnerSettings):
    """min 1/2 x'Px + q'x  s.t.  A x >= l  with the reference's 4N-2 rows
    (casadi/PI_ADMM_class.py:172-192, passed to OSQP as l = -b, u = +inf).  Returns x, info(dict)."""
    n = len(q)
    l = constraint_lower(n)
    x, y, info = solve_dense_qp(P, q, constraint_matrix(n), l, np.full(len(l), np.inf), s)
    info["y"] = y
    return x, info


def solve_edge_pair(H, q, a, c, beta, s: InnerSettings):
    """Convex edge problem (oracle.problem.edge_qp_blocks) by ADMM on the split

        z_r = [A u_r ; M_{1..N} u_r]   (r = 1,2)
        g(z) = box indicator on the first 4N-2 rows of each half
             + beta * sum_k max(0, c_k - a1 z_1[s_k] - a2 z_2[s_k])

    so the linear system stays block diagonal (one N x N matrix per vehicle
    half) and the two halves couple only through the closed-form prox of the
    hinge on the pairs (s1_k, s2_k).
    """
    N = len(q[0])
    Ac = constraint_matrix(N); lc = constraint_lower(N); m = len(lc)
    Ms = roll_matrix(N)[1:]
    A = np.vstack((Ac, Ms))
    AtA = A.T @ A
    a = np.asarray(a, dtype=np.float64)
    na = float(a @ a)
    x = [np.zeros(N), np.zeros(N)]
    z = [np.zeros(m + N), np.zeros(m + N)]
    y = [np.zeros(m + N), np.zeros(m + N)]
    rho = s.rho0

    def factor(rho):
        return [np.linalg.inv(H[r] + s.sigma * np.eye(N) + rho * AtA) for r in range(2)]

    K = factor(rho)
    nref = 0; status = 1; it = 0
    for it in range(1, s.max_iter + 1):
        v = [None, None]; zr = [None, None]; zn = [None, None]
        for r in range(2):
            rhs = s.sigma * x[r] - q[r] + A.T @ (rho * z[r] - y[r])
            xt = K[r] @ rhs
            zt = A @ xt
            x[r] = s.alpha * xt + (1 - s.alpha) * x[r]
            zr[r] = s.alpha * zt + (1 - s.alpha) * z[r]
            v[r] = zr[r] + y[r] / rho
            zn[r] = v[r].copy()
            zn[r][:m] = np.maximum(v[r][:m], lc)
        s1 = v[0][m:]; s2 = v[1][m:]
        tau = a[0] * s1 + a[1] * s2
        t = beta / rho
        if na > 0.0:
            step = np.where(tau >= c, 0.0, np.where(tau + t * na <= c, t, (c - tau) / na))
        else:
            step = np.zeros(N)
        zn[0][m:] = s1 + step * a[0]
        zn[1][m:] = s2 + step * a[1]
        for r in range(2):
            y[r] = y[r] + rho * (zr[r] - zn[r])
            z[r] = zn[r]
        if it % s.check == 0:
            rp = rd = n_p = n_d = 0.0
            for r in range(2):
                Ax = A @ x[r]; Px = H[r] @ x[r]; Aty = A.T @ y[r]
                rp = max(rp, np.abs(Ax - z[r]).max())
                rd = max(rd, np.abs(Px + q[r] + Aty).max())
                n_p = max(n_p, np.abs(Ax).max(), np.abs(z[r]).max())
                n_d = max(n_d, np.abs(Px).max(), np.abs(Aty).max(), np.abs(q[r]).max())
            if rp <= s.eps_abs + s.eps_rel * n_p and rd <= s.eps_abs + s.eps_rel * n_d:
                status = 0
                break
            if s.adaptive:
                rn = _new_rho(rho, rp, rd, n_p, n_d)
                if rn > s.adapt_tol * rho or rn < rho / s.adapt_tol:
                    rho = rn
                    K = factor(rho)
                    nref += 1
    return np.vstack((x[0], x[1])), dict(iters=it, status=status, refactors=nref, rho=rho)
