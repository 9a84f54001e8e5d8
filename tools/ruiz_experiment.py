"""Would OSQP's default preconditioning (scaling = 10 Ruiz iterations + cost scaling) shorten the inner solves of this
path?  CPU experiment on the numpy oracle (tools only; nothing here is a product path).

The oracle's dense OSQP iteration (oracle/solver.py: solve_dense_qp) is run on the vehicle QPs of the bench workloads
(a) as the kernels run it (unscaled) and (b) on the Ruiz-equilibrated problem, with OSQP's termination on UNSCALED
residuals, and the inner iteration counts are compared.  OSQP's scaling restated from the OSQP paper, sec. 5.1, and
osqp/src/scaling.c of the 0.6 series (third party, not under /root/reference).

  python tools/ruiz_experiment.py            # config 4 (N=15) and config 5 (N=30) vehicle QPs, outer iterations 1 and 2
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import problem as op, solver as osol  # noqa: E402

MIN_S, MAX_S = 1e-4, 1e4


def _limit(v):
    v = np.where(v < MIN_S, 1.0, v)
    return np.minimum(v, MAX_S)


def ruiz(P, q, A, iters=10):
    n, m = P.shape[0], A.shape[0]
    D, E, c = np.ones(n), np.ones(m), 1.0
    P, q, A = P.copy(), q.copy(), A.copy()
    for _ in range(iters):
        dn = np.maximum(np.abs(P).max(axis=0), np.abs(A).max(axis=0))
        en = np.abs(A).max(axis=1)
        dt, et = 1.0 / np.sqrt(_limit(dn)), 1.0 / np.sqrt(_limit(en))
        P = dt[:, None] * P * dt[None, :]; A = et[:, None] * A * dt[None, :]; q = dt * q
        D *= dt; E *= et
        ct = max(float(_limit(np.array([np.abs(P).max(axis=0).mean()]))[0]), float(_limit(np.array([np.abs(q).max()]))[0]))
        P *= 1.0 / ct; q *= 1.0 / ct; c *= 1.0 / ct
    return P, q, A, D, E, c


def solve_scaled(P, q, A, l, s):
    """OSQP iteration on the equilibrated problem; termination / adaptive rho on unscaled residuals (scaled_termination=0)."""
    Pb, qb, Ab, D, E, c = ruiz(P, q, A)
    lb = E * l
    n, m = len(q), len(l)
    AtA = Ab.T @ Ab
    x = np.zeros(n); z = np.zeros(m); y = np.zeros(m)
    rho = s.rho0
    K = np.linalg.inv(Pb + s.sigma * np.eye(n) + rho * AtA)
    nref = 0
    for it in range(1, s.max_iter + 1):
        xt = K @ (s.sigma * x - qb + Ab.T @ (rho * z - y))
        zt = Ab @ xt
        x = s.alpha * xt + (1 - s.alpha) * x
        zr = s.alpha * zt + (1 - s.alpha) * z
        zn = np.maximum(zr + y / rho, lb)
        y = y + rho * (zr - zn)
        z = zn
        if it % s.check == 0:
            Ax = (Ab @ x) / E; zu = z / E
            Px = (Pb @ x) / D / c; Aty = (Ab.T @ y) / D / c; qu = qb / D / c
            rp = np.abs(Ax - zu).max(); rd = np.abs(Px + qu + Aty).max()
            n_p = max(np.abs(Ax).max(), np.abs(zu).max()); n_d = max(np.abs(Px).max(), np.abs(Aty).max(), np.abs(qu).max())
            if rp <= s.eps_abs + s.eps_rel * n_p and rd <= s.eps_abs + s.eps_rel * n_d:
                return D * x, it, nref
            if s.adaptive:
                rn = osol._new_rho(rho, rp, rd, n_p, n_d)
                if rn > s.adapt_tol * rho or rn < rho / s.adapt_tol:
                    rho = rn; K = np.linalg.inv(Pb + s.sigma * np.eye(n) + rho * AtA); nref += 1
    return D * x, s.max_iter, nref


def sample_qps(N, n, seed, coupled):
    """Vehicle QPs as the outer loop meets them: iteration 1 (no neighbour) and a later one (one neighbour, duals ~ O(10))."""
    rng = np.random.default_rng(seed)
    par = op.Params(num_ho=N)
    out = []
    for t in range(n):
        v = rng.choice([4.0, 6.0, 8.0]) * (1 + rng.uniform(-0.1, 0.1))
        th = rng.choice([0.0, -np.pi / 2])
        xt = np.array([rng.uniform(-10, 10), rng.uniform(-10, 10), th])
        k = np.arange(N + 1)
        rx, ry = xt[0] + k * 0.1 * v * np.cos(th), xt[1] + k * 0.1 * v * np.sin(th)
        if t % 2 == 0:
            hats, duals = [], []
        else:
            off = rng.uniform(-3, 3, size=(2, 1)) * np.linspace(0, 1, N + 1)[None, :]
            hats = [np.vstack((rx, ry)) + off]
            duals = [rng.uniform(-1, 1) * (20.0 if coupled else 5.0) * np.linspace(0, 1, N + 1)[None, :] * np.ones((2, 1))]
        out.append(op.vehicle_qp(par, xt, v, rx, ry, hats, duals, 2.0))
    return out


def main():
    s = osol.InnerSettings.osqp_default()
    for N, coupled in ((15, False), (30, True)):
        A, l = op.constraint_matrix(N), op.constraint_lower(N)
        a, b, err = [], [], []
        for H, q in sample_qps(N, 200, 7 + N, coupled):
            x0, _, i0 = osol.solve_dense_qp(H, q, A, l, np.full(len(l), np.inf), s)
            x1, it1, _ = solve_scaled(H, q, A, l, s)
            a.append(i0["iters"]); b.append(it1); err.append(np.abs(x0 - x1).max())
        a, b = np.array(a), np.array(b)
        hist = lambda v: {int(k): int(c) for k, c in zip(*np.unique(v, return_counts=True))}
        print(f"N={N}: 200 vehicle QPs, inner iterations  unscaled {hist(a)} (mean {a.mean():.1f})   Ruiz-scaled {hist(b)} (mean {b.mean():.1f})"
              f"   max |x_unscaled - x_scaled| = {max(err):.2e} (both within eps 1e-3 of the minimiser)")


if __name__ == "__main__":
    main()
