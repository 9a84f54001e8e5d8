"""Probe of SURVEY 8(f)-1: the local and RSU NLPs of BASELINE config 2 (Distributed_planner/decentralized/optimizer.py:40-201,
239-328, util.py:76-101, driver decentralized_overtaking_ADMM.py:30-103), restated in numpy and handed to scipy SLSQP in
place of IPOPT, with the consensus statements of oracle/obca_consensus.py between the solves (RSU variables in the intended
10 + 8 layout; the reference's ca.reshape of a 2x9 block into 10x1 / 8x1, optimizer.py:288,293, cannot run as written).

    python tools/obca_nlp_probe.py 1      # first MPC step, up to 12 ADMM iterations, then the first local solves in detail

What it shows (profiles/r2_obca_nlp_probe.txt): the first local problem of the reference is INFEASIBLE by construction --
bar_state.A = b = 0 at start (optimizer.py:353-355), so (5a) reads -b(X)'L >= min_dis with L >= 0 and, through (5b)
A(X)'L = 0, b(X)'L = b0'L >= 0 -- and the problems of the following iterations stay infeasible (constraint violations of 2-9
at the returned points, dual residual growing to 10^2).  What the reference computes from there on is whatever IPOPT's
restoration phase returns for an infeasible problem; no other solver reproduces that, so there is no solver-independent
answer for a GPU replacement to match.  Test infrastructure / evidence only; nothing imports this."""
import numpy as np, copy, time, sys
from scipy.optimize import minimize
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import obca_consensus as oc

L, W = 3.5, 2.0; lf, lr = 1.5, 1.0; dt = 0.1; T = 5
NH = 8; H = NH - 1
B0 = np.array([L/2, W/2, L/2, W/2])
AVG, VAR, PROB = 0.05, 0.025, 0.95
KQ = np.sqrt(PROB/(1-PROB))

def ref_traj():
    n = int(T/dt)+1
    x1 = np.linspace(0, 100, n); x2 = np.linspace(20, 70, n)
    z = np.zeros(n)
    return [np.stack((x1, z, 20+z, z, z), 1), np.stack((x2, z, 10+z, z, z), 1)]

def halfspace(st):
    x, y, v, th = st[0], st[1], st[2], st[3]
    c, s = np.cos(th), np.sin(th)
    R = np.array([[c, -s], [s, c]])
    A = np.vstack((R.T, -R.T))
    p = np.array([x, y]) + AVG*v*np.array([c, s]) + KQ*np.array([(VAR*v*c)**2, (VAR*v*s)**2])
    return A, B0 + A @ p

def f_dyn(st, u):
    x, y, v, th, de = st
    beta = np.arctan(lr*np.tan(de)/(lr+lf))
    return np.array([v*np.cos(th+beta), v*np.sin(th+beta), u[0], v/lr*np.sin(beta), u[1]])

def unpack(w):
    X = w[:5*NH].reshape(NH, 5); U = w[5*NH:5*NH+2*H].reshape(H, 2); Lm = w[5*NH+2*H:].reshape(H, 4)
    return X, U, Lm

def local_solve(v, t_step, init, bar, rho, ref, min_dis=0.1):
    xr = ref[v][t_step:t_step+NH]
    w0 = np.concatenate((xr.ravel(), np.zeros(2*H), np.ones(4*H)))
    Rw, Qw = 1e4, 1e5
    o = 1 - v
    cst_b = np.array([bar.b[o, i] @ bar.lamb_ij[o, i] for i in range(H)])
    cst_A = np.array([bar.A[o, i].T @ bar.lamb_ij[o, i] for i in range(H)])
    def obj(w):
        X, U, Lm = unpack(w)
        S = np.concatenate((X[1:], Lm), 1)
        return Rw*np.sum(U*U) + Qw*np.sum((X[1:]-xr[1:])**2) + np.sum(bar.lamb_bar[v]*S) + rho/2*np.sum((S-bar.Z_bar[v])**2)
    def ceq(w):
        X, U, Lm = unpack(w)
        out = [X[0]-init]
        for i in range(H):
            out.append(X[i+1]-(X[i]+dt*f_dyn(X[i], U[i])))
        for i in range(1, NH):
            A, b = halfspace(X[i])
            out.append(A.T @ Lm[i-1] + cst_A[i-1])
        return np.concatenate(out)
    def cin(w):
        X, U, Lm = unpack(w)
        out = []
        for i in range(1, NH):
            A, b = halfspace(X[i])
            g = -b @ Lm[i-1] - cst_b[i-1]
            out += [g - min_dis, 1000 - g]
        for i in range(1, NH):
            d = Lm[i-1][:2]-Lm[i-1][2:]
            out += [1 - d @ d]
        return np.array(out)
    lb = np.concatenate((np.tile([0, -20, -20, -2*np.pi, -0.6], NH), np.tile([-5, -20], H), np.zeros(4*H)))
    ub = np.concatenate((np.tile([150, 20, 20, 2*np.pi, 0.6], NH), np.tile([5, 20], H), 1e5*np.ones(4*H)))
    r = minimize(obj, w0, method="SLSQP", bounds=list(zip(lb, ub)),
                 constraints=[{"type": "eq", "fun": ceq}, {"type": "ineq", "fun": cin}], options={"maxiter": 300, "ftol": 1e-9})
    X, U, Lm = unpack(r.x)
    viol = max(np.abs(ceq(r.x)).max(), -min(cin(r.x).min(), 0))
    return X, U, Lm, r, viol

def edge_solve(bar, rho, min_dis=0.1):
    Z = np.zeros((2, H, 9)); info = []
    for i in range(H):
        c = np.stack([np.concatenate((bar.local_x[v, i], bar.lamb_ij[v, i])) + bar.lamb_bar[v, i]/rho for v in range(2)])
        def obj(z):
            return rho/2*np.sum((z.reshape(2, 9)-c)**2)
        def jac(z): return rho*(z-c.ravel())
        def ceq(z):
            z = z.reshape(2, 9)
            A1, _ = halfspace(z[0, :5]); A2, _ = halfspace(z[1, :5])
            return A1.T @ z[0, 5:] + A2.T @ z[1, 5:]
        def cin(z):
            z = z.reshape(2, 9)
            _, b1 = halfspace(z[0, :5]); _, b2 = halfspace(z[1, :5])
            g = -b1 @ z[0, 5:] - b2 @ z[1, 5:]
            return np.array([g-min_dis, 1000-g])
        lb = np.tile(np.concatenate((-1000*np.ones(5), np.zeros(4))), 2); ub = np.tile(np.concatenate((1000*np.ones(5), 1e5*np.ones(4))), 2)
        r = minimize(obj, np.zeros(18), jac=jac, method="SLSQP", bounds=list(zip(lb, ub)),
                     constraints=[{"type": "eq", "fun": ceq}, {"type": "ineq", "fun": cin}], options={"maxiter": 200, "ftol": 1e-10})
        Z[:, i] = np.round(r.x.reshape(2, 9), 4)
        info.append((r.success, r.nit, max(np.abs(ceq(r.x)).max(), -min(cin(r.x).min(), 0))))
    return Z, info

ref = ref_traj()
init = np.stack([r[0] for r in ref])
prev = oc.MidState()
nsteps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
for t_step in range(nsteps):
    bar = copy.deepcopy(prev)
    ctrl_prev = np.zeros((2, H*2)); i_iter = 0
    while True:
        i_iter += 1; rho = 1.0
        t0 = time.time()
        sols = [local_solve(v, t_step, init[v], bar, rho, ref) for v in range(2)]
        for v in range(2):
            X = sols[v][0]
            for i in range(H):
                A, b = halfspace(X[i+1]); bar.A[v, i] = A; bar.b[v, i] = b; bar.local_x[v, i] = X[i+1]
        conv = oc.check_converge(bar, 0.1, 0.1)
        bar.Z_bar, einfo = edge_solve(bar, rho)
        ctrl = [np.append(sols[v][1][:, 0], sols[v][1][:, 1]) for v in range(2)]
        prev, ctrl_prev, p, d, stop = oc.consensus_tail(bar, prev, np.array(ctrl), ctrl_prev, i_iter)
        print(f"step {t_step} iter {i_iter}: local ok {[s[3].success for s in sols]} nit {[s[3].nit for s in sols]} viol {[float('%.2e' % s[4]) for s in sols]} "
              f"edge ok {all(e[0] for e in einfo)} viol {max(e[2] for e in einfo):.1e} conv {conv} primal {p:.4f} dual {d:.4f} ({time.time()-t0:.1f}s)", flush=True)
        if stop: break
        bar = copy.deepcopy(prev) if False else bar
        if i_iter >= 12: break
    init = np.stack([sols[v][0][1] for v in range(2)])

# diagnostics: which constraints are violated at the first local solve of step 0
prev = oc.MidState(); bar = copy.deepcopy(prev)
init = np.stack([r[0] for r in ref])
for v in range(2):
    X, U, Lm, r, viol = local_solve(v, 0, init[v], bar, 1.0, ref)
    print("veh", v, r.message, "viol", viol)
    dyn = np.array([X[i+1]-(X[i]+dt*f_dyn(X[i], U[i])) for i in range(H)])
    print(" dyn viol", np.abs(dyn).max(), "init viol", np.abs(X[0]-init[v]).max())
    for i in range(1, NH):
        A, b = halfspace(X[i])
        print("  i", i, "5a g=", float(-b @ Lm[i-1]), "5b", A.T @ Lm[i-1], "Lm", np.round(Lm[i-1], 3), "X", np.round(X[i], 3))
