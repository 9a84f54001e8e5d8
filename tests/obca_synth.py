"""Synthetic stand-ins for the IPOPT solves of config 2 (they are out of scope): seeded sequences of
local solutions (local_x, lamb_ij, controls), half-spaces (A, b) and RSU solutions (Z_bar) that contract
geometrically, scenario by scenario at a different rate, so that the consensus loop of
Distributed_planner/decentralized_overtaking_ADMM.py:40-97 stops after a scenario-dependent number of
iterations (some never within max_iter)."""
import numpy as np

V, H, NS, NL, NC = 2, 7, 5, 4, 2


def sequence(S, n_iter, seed=2, V=V, H=H):
    """Returns a list (one entry per ADMM iteration) of dicts of [S,...] arrays."""
    rng = np.random.default_rng(seed)
    gamma = rng.uniform(0.25, 0.97, size=(S, 1, 1, 1))            # contraction rate per scenario
    x_star = rng.normal(size=(S, V, H, NS)) * np.array([30.0, 3.0, 8.0, 0.3, 0.1])
    l_star = rng.uniform(0.2, 1.5, size=(S, V, H, NL))
    c_star = rng.normal(size=(S, V, H * NC))
    th = rng.uniform(-0.2, 0.2, size=(S, V, H, 1))
    out = []
    for t in range(n_iter):
        g = gamma ** t
        local_x = x_star + g * rng.normal(size=x_star.shape)
        lamb_ij = l_star + g * 0.1 * rng.normal(size=l_star.shape)
        # RSU solution rounded to 4 decimals as in optimizer.py:322, offset from the local one by ~g
        Z_bar = np.round(np.concatenate((local_x, lamb_ij), -1) + g * 0.05 * rng.normal(size=(S, V, H, NS + NL)), 4)
        ctrl = c_star + g[:, :, :, 0] * rng.normal(size=c_star.shape)
        ang = th + np.arange(4).reshape(1, 1, 1, 4) * (np.pi / 2) + g[..., 0:1] * 0.01 * rng.normal(size=(S, V, H, 4))
        A = np.stack((np.cos(ang), np.sin(ang)), -1)               # [S,V,H,4,2] outward normals of a rectangle
        b = rng.uniform(1.0, 2.5, size=(S, V, H, 4)) * np.where(np.arange(V).reshape(1, V, 1, 1) == 0, 1.0, -1.2)
        # even scenarios: opposing normals and nearly equal multipliers, so that check_converge
        # (optimizer.py:225-235) turns true once the sequence has contracted
        tie = g[::2, 0] * 0.02 * rng.normal(size=lamb_ij[::2, 0].shape)
        A[::2, 1] = -A[::2, 0]
        lamb_ij[::2, 1] = lamb_ij[::2, 0] + tie
        Z_bar[::2, 1, :, NS:] = np.round(lamb_ij[::2, 1] + g[::2, 0] * 0.05 * rng.normal(size=tie.shape), 4)
        out.append(dict(local_x=local_x, lamb_ij=lamb_ij, Z_bar=Z_bar, ctrl=ctrl, A=A, b=b))
    return out


def run_oracle(seq, S, max_iter=50, rho=1.0):
    """The reference loop, scenario by scenario, with the oracle's literal functions."""
    from oracle import obca_consensus as oc
    res = dict(n_iter=np.zeros(S, np.int32), stop=np.zeros(S, np.int32), primal=[], dual=[],
               prev={f: [] for f in oc.MidState.FIELDS}, cur_lamb_bar=[], check=[])
    for s in range(S):
        prev = oc.MidState(V, H)
        cur = None
        ctrl_prev = np.zeros((V, H * NC))
        pr, du, ck = [], [], []
        i_iter = 0
        stopped = False
        for step in seq:
            if stopped:
                break
            i_iter += 1
            if cur is None:
                cur = oc.MidState(V, H)          # optimizer.bar_state = deepcopy(bar_state_prev)
                for f in oc.MidState.FIELDS:
                    setattr(cur, f, getattr(prev, f).copy())
            cur.local_x = step["local_x"][s].copy(); cur.lamb_ij = step["lamb_ij"][s].copy()
            cur.A = step["A"][s].copy(); cur.b = step["b"][s].copy()
            ck.append(oc.check_converge(cur, 0.1, 0.1))
            cur.Z_bar = step["Z_bar"][s].copy()
            prev, ctrl_prev, p, d, stopped = oc.consensus_tail(cur, prev, step["ctrl"][s], ctrl_prev, i_iter,
                                                               rho=rho, max_iter=max_iter)
            pr.append(p); du.append(d)
        res["n_iter"][s] = i_iter; res["stop"][s] = int(stopped)
        res["primal"].append(pr); res["dual"].append(du); res["check"].append(ck)
        for f in oc.MidState.FIELDS:
            res["prev"][f].append(getattr(prev, f))
        res["cur_lamb_bar"].append(cur.lamb_bar)
    for f in oc.MidState.FIELDS:
        res["prev"][f] = np.stack(res["prev"][f])
    res["cur_lamb_bar"] = np.stack(res["cur_lamb_bar"])
    return res
