"""What pins the oracle (the reference has no tests and cannot run here):
literal transcriptions of the reference's cost code vs the closed forms, solver-independent KKT
checks, scipy cross-solves, the reference's own 2-variable QP (casadi/casadi_osqp_test.py:12-14)."""
import numpy as np
import pytest
import scipy.optimize as so

from oracle.problem import (Params, reference_trajectories, vehicle_qp, edge_qp_blocks, roll_matrix,
                            constraint_matrix, constraint_lower, literal_cost_primal,
                            literal_cost_edge_convex, literal_dynamic_update_local, affine_rollout,
                            literal_dynamic_update_edge)
from oracle.solver import InnerSettings, solve_box_qp, solve_dense_qp, solve_edge_pair

RNG = np.random.default_rng(0)


def _vehicle_case(N, spd, th0, deg):
    p = Params(num_ho=N)
    xt = np.array([RNG.uniform(-5, 5), RNG.uniform(-5, 5), th0])
    k = np.arange(N + 1)
    rx = xt[0] + k * p.dt * spd * np.cos(th0) + RNG.normal(0, 0.1, N + 1)
    ry = xt[1] + k * p.dt * spd * np.sin(th0) + RNG.normal(0, 0.1, N + 1)
    hats = [RNG.normal(0, 3, (2, N + 1)) for _ in range(deg)]
    duals = [RNG.normal(0, 1, (2, N + 1)) for _ in range(deg)]
    return p, xt, rx, ry, hats, duals


@pytest.mark.parametrize("N,spd,th0,deg", [(15, 4.0, 0.0, 1), (15, 8.0, -np.pi / 2, 1), (30, 6.5, 0.7, 3),
                                           (8, 8.0, 2.1, 0), (5, 4.0, -0.3, 2)])
def test_vehicle_qp_closed_form_equals_literal_cost(N, spd, th0, deg):
    """cost_function_primal (casadi/PI_ADMM_class.py:114-135) is exactly 1/2 u'Hu + q'u + const."""
    p, xt, rx, ry, hats, duals = _vehicle_case(N, spd, th0, deg)
    H, q = vehicle_qp(p, xt, spd, rx, ry, hats, duals)
    f = lambda u: literal_cost_primal(p, u, xt, spd, rx, ry, hats, duals)
    f0 = f(np.zeros(N))
    for _ in range(5):
        u = RNG.normal(0, 0.3, N)
        quad = 0.5 * u @ H @ u + q @ u
        assert abs((f(u) - f0) - quad) <= 1e-9 * max(1.0, abs(quad))
    assert np.allclose(H, H.T) and np.linalg.eigvalsh(H).min() >= 2 * p.Pcost - 1e-9
    assert H[-1, -1] == pytest.approx(2 * (p.Pcost + 1))     # last column of M is zero


def test_hessian_independent_of_heading():
    p = Params()
    hs = []
    for th in (0.0, -np.pi / 2, 0.7):
        xt = np.array([1.0, 2.0, th])
        hs.append(vehicle_qp(p, xt, 8.0, np.zeros(16), np.zeros(16), [np.zeros((2, 16))], [np.zeros((2, 16))])[0])
    assert np.allclose(hs[0], hs[1], atol=1e-12) and np.allclose(hs[0], hs[2], atol=1e-12)


def test_affine_rollout_equals_literal_loop():
    p = Params(num_ho=15)
    xt = np.array([0.3, -1.0, 0.4]); u = RNG.normal(0, 0.2, 15)
    x, y, _ = literal_dynamic_update_local(p, xt, u, 8.0)
    xa, ya = affine_rollout(p, xt, 8.0, u)
    assert np.abs(x - xa).max() < 1e-12 and np.abs(y - ya).max() < 1e-12
    M = roll_matrix(15)
    assert (M[:, -1] == 0).all() and (M[:2] == 0).all()


def _edge_case(N, seed=1):
    rng = np.random.default_rng(seed)
    p = Params(num_ho=N)
    xt = np.array([[-10 + 4.8 + rng.uniform(-0.5, 0.5), 0.0, 0.0], [0.0, 20 - 9.6 + rng.uniform(-1, 1), -np.pi / 2]])
    spd = np.array([4.0, 8.0]) * (1 + rng.uniform(-0.1, 0.1, 2))
    pos = np.zeros((4, N + 1))
    for i in range(2):
        pos[2 * i], pos[2 * i + 1] = affine_rollout(p, xt[i], spd[i], rng.normal(0, 0.05, N))
    dual = rng.normal(0, 0.3, (4, N + 1))
    seed_pt = np.around(xt[:, :2] + p.dt * spd[:, None] * np.stack((np.cos(xt[:, 2]), np.sin(xt[:, 2])), 1), 4)
    return p, xt, spd, pos, dual, seed_pt


@pytest.mark.parametrize("N", [8, 15, 30])
def test_edge_blocks_equal_literal_convex_cost(N):
    """MATLAB cost_function_edge (:276-304) with the linearised roll-out (:383-396)."""
    p, xt, spd, pos, dual, seed_pt = _edge_case(N)
    blk = edge_qp_blocks(p, xt, spd, pos, dual, seed_pt)
    Ms = roll_matrix(N)[1:]

    def closed(uu):
        u1, u2 = uu[:N], uu[N:]
        quad = sum(0.5 * u @ H @ u + q @ u for u, H, q in zip((u1, u2), blk["H"], blk["q"]))
        return quad + p.beta * np.maximum(0, blk["c"] - blk["a"][0] * (Ms @ u1) - blk["a"][1] * (Ms @ u2)).sum()

    lit = lambda uu: literal_cost_edge_convex(p, uu.reshape(2, N), xt, spd, pos, dual, seed_pt)
    d0 = lit(np.zeros(2 * N)) - closed(np.zeros(2 * N))
    for _ in range(5):
        uu = RNG.normal(0, 0.2, 2 * N)
        assert abs(lit(uu) - closed(uu) - d0) <= 1e-8 * max(1.0, abs(lit(uu)))


def test_reference_two_variable_qp_known_answer():
    """casadi/casadi_osqp_test.py:12-14: min x0^2 + x1^2  s.t. x0 + x1 >= 1, x >= 0  ->  (0.5, 0.5)."""
    P = 2 * np.eye(2); q = np.zeros(2)
    A = np.array([[1.0, 1.0], [1.0, 0.0], [0.0, 1.0]])
    l = np.array([1.0, 0.0, 0.0]); u = np.full(3, np.inf)
    x, y, info = solve_dense_qp(P, q, A, l, u, InnerSettings.tight())
    assert info["status"] == 0 and np.abs(x - 0.5).max() < 1e-7
    x3, _, info3 = solve_dense_qp(P, q, A, l, u, InnerSettings.osqp_default())
    assert info3["status"] == 0 and np.abs(x3 - 0.5).max() < 2e-3


@pytest.mark.parametrize("N,spd", [(15, 8.0), (30, 8.0), (15, 4.0)])
def test_vehicle_qp_kkt_and_scipy(N, spd):
    """tight mode: KKT residuals (solver independent) and agreement with scipy SLSQP."""
    p, xt, rx, ry, hats, duals = _vehicle_case(N, spd, -np.pi / 2, 1)
    ry = ry + 0.05 * np.arange(N + 1) ** 2          # make the steering constraints bind
    H, q = vehicle_qp(p, xt, spd, rx, ry, hats, duals)
    x, info = solve_box_qp(H, q, InnerSettings.tight())
    A = constraint_matrix(N); l = constraint_lower(N)
    y = info["y"]
    assert info["status"] == 0
    assert (A @ x - l).min() > -1e-7                                  # primal feasibility
    assert np.abs(H @ x + q + A.T @ y).max() < 1e-6 * max(1, np.abs(q).max())   # stationarity
    assert y.max() <= 1e-9 and np.abs(y * (A @ x - l)).max() < 1e-6   # y <= 0 for lower bounds, complementarity
    res = so.minimize(lambda u: 0.5 * u @ H @ u + q @ u, np.zeros(N), jac=lambda u: H @ u + q, method="SLSQP",
                      constraints=[dict(type="ineq", fun=lambda u: A @ u - l, jac=lambda u: A)],
                      options=dict(ftol=1e-14, maxiter=500))
    assert np.abs(res.x - x).max() < 1e-5


@pytest.mark.parametrize("N", [8, 15])
def test_edge_problem_against_scipy_slack_qp(N):
    """The prox-split ADMM reaches the minimiser of the slack-variable QP."""
    p, xt, spd, pos, dual, seed_pt = _edge_case(N, seed=3)
    p.beta = 10.0
    blk = edge_qp_blocks(p, xt, spd, pos, dual, seed_pt)
    uh, info = solve_edge_pair(blk["H"], blk["q"], blk["a"], blk["c"], p.beta, InnerSettings.tight())
    assert info["status"] == 0
    Ms = roll_matrix(N)[1:]; A = constraint_matrix(N); l = constraint_lower(N)
    G = np.hstack((blk["a"][0] * Ms, blk["a"][1] * Ms))
    n = 2 * N

    def f(v):
        u1, u2, t = v[:N], v[N:n], v[n:]
        return sum(0.5 * u @ H @ u + q @ u for u, H, q in zip((u1, u2), blk["H"], blk["q"])) + p.beta * t.sum()

    def g(v):
        return np.concatenate((A @ v[:N] - l, A @ v[N:n] - l, v[n:], v[n:] - (blk["c"] - G @ v[:n])))

    fo = f(np.concatenate((uh.ravel(), np.maximum(0, blk["c"] - G @ uh.ravel()))))
    # (1) the SURVEY B.3 slack-variable QP through the generic dense OSQP iteration (different splitting)
    P = np.zeros((3 * N, 3 * N)); P[:N, :N] = blk["H"][0]; P[N:n, N:n] = blk["H"][1]
    qq = np.concatenate((blk["q"][0], blk["q"][1], np.full(N, p.beta)))
    m = len(l)
    AA = np.zeros((2 * m + 2 * N, 3 * N))
    AA[:m, :N] = A; AA[m:2 * m, N:n] = A
    AA[2 * m:2 * m + N, n:] = np.eye(N)
    AA[2 * m + N:, :n] = G; AA[2 * m + N:, n:] = np.eye(N)
    ll = np.concatenate((l, l, np.zeros(N), blk["c"]))
    xs, _, inf2 = solve_dense_qp(P, qq, AA, ll, np.full(len(ll), np.inf), InnerSettings(eps_abs=1e-10, eps_rel=1e-10, max_iter=200000))
    assert inf2["status"] == 0
    assert abs(f(xs) - fo) <= 1e-6 * max(1.0, abs(fo))
    assert np.abs(xs[:n] - uh.ravel()).max() < 1e-5
    # (2) scipy SLSQP from the ADMM point must not find a better feasible point
    v0 = np.concatenate((uh.ravel(), np.maximum(0, blk["c"] - G @ uh.ravel()) + 1e-9))
    res = so.minimize(f, v0, method="SLSQP", constraints=[dict(type="ineq", fun=g)],
                      options=dict(ftol=1e-13, maxiter=500))
    if g(res.x).min() > -1e-9:
        assert fo <= res.fun + 1e-6 * max(1.0, abs(res.fun))


def test_nonlinear_rollout_is_v_cos_theta():
    """casadi/PI_ADMM_class.py:95-101: the three-term x_dot equals v*cos(theta_k)."""
    p = Params(num_ho=15)
    xt = np.array([[0.0, 0.0, 0.3]]); u = RNG.normal(0, 0.3, (1, 15))
    x, y, th = literal_dynamic_update_edge(p, xt, u, np.array([8.0]))
    assert np.allclose(np.diff(x[0]), 8.0 * np.cos(th[0, :-1]) * p.dt, atol=1e-12)
    assert np.allclose(np.diff(y[0]), 8.0 * np.sin(th[0, :-1]) * p.dt, atol=1e-12)


def test_reference_constants():
    p = Params(); ref = reference_trajectories(p)
    assert ref.shape == (4, 50) and ref[0, 0] == -10 and ref[3, -1] == -20 and p.num_ho == 15 and p.rho == 2
