"""TEST INFRASTRUCTURE ONLY -- CPU oracle of the consensus pieces of the OBCA overtaking ADMM
(SURVEY.md 8(a) row a13).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import this module; the product path never does.

numpy restatement, loop for loop, of
  * ``mid_state``            Distributed_planner/decentralized/optimizer.py:351-373
  * ``lambda_update``        Distributed_planner/decentralized/optimizer.py:330-335
  * ``check_converge``       Distributed_planner/decentralized/optimizer.py:225-235
  * ``iterate_next_state``   Distributed_planner/decentralized/optimizer.py:337-344
  * residuals / stop / carry Distributed_planner/decentralized_overtaking_ADMM.py:79-97
Parity unpinned: the reference has no tests or stored outputs for these functions and its module cannot
be imported here (it imports casadi at the top); the statements are short enough to be restated literally.
"""
from __future__ import annotations

import copy

import numpy as np

# constants of Distributed_planner/decentralized/optimizer.py:10-37 (N_horz=8 -> 7 predicted steps)
NUM_VEH, N_HORZ, N_STATES, N_CONTROLS, N_LOC_LAMBDA, N_DUAL = 2, 8, 5, 2, 4, 4
PRIMAL_THRES = DUAL_THRES = 0.01     # decentralized_overtaking_ADMM.py:86 via optimizer.primal_thres / dual_thres
MAX_ITER = 50                        # `i_iter > 50`, decentralized_overtaking_ADMM.py:86

# warm-start multipliers hard-coded at optimizer.py:357-371 (the only numeric data the reference carries)
LAMB_IJ0 = np.array(
    [[[1.49, 0.566, 0.566, 1.49], [1.438, 0.514, 0.514, 1.438], [1.387, 0.462, 0.462, 1.387],
      [1.336, 0.411, 0.411, 1.336], [1.287, 0.361, 0.361, 1.287], [1.238, 0.312, 0.312, 1.238],
      [1.191, 0.263, 0.263, 1.191]],
     [[1.436] * 4, [1.325] * 4, [1.213] * 4, [1.1] * 4, [0.986] * 4, [0.871] * 4, [0.755] * 4]])


class MidState:
    """optimizer.py:351-373"""

    def __init__(self, num_veh=NUM_VEH, H=N_HORZ - 1, ns=N_STATES, nl=N_LOC_LAMBDA, nd=N_DUAL):
        self.Z_bar = np.zeros((num_veh, H, ns + nl))
        self.A = np.zeros((num_veh, H, nd, 2))
        self.b = np.zeros((num_veh, H, nd))
        self.lamb_bar = 1e-3 * np.ones((num_veh, H, ns + nl))
        if (num_veh, H, nd) == LAMB_IJ0.shape:
            self.lamb_ij = LAMB_IJ0.copy()
        else:
            self.lamb_ij = 1e-3 * np.ones((num_veh, H, nd))     # the commented-out alternative, optimizer.py:372
        self.local_x = np.zeros((num_veh, H, ns))

    FIELDS = ("Z_bar", "A", "b", "lamb_bar", "lamb_ij", "local_x")


def lambda_update(st: MidState, rho):
    """optimizer.py:330-335"""
    V, H = st.local_x.shape[:2]
    for i_veh in range(V):
        for i in range(H):
            local_z = np.concatenate((st.local_x[i_veh, i, :], st.lamb_ij[i_veh, i, :]))
            edge_z = st.Z_bar[i_veh, i, :]
            st.lamb_bar[i_veh, i, :] = st.lamb_bar[i_veh, i, :] + rho * (local_z - edge_z)


def check_converge(st: MidState, thres, min_dis):
    """optimizer.py:225-235, summed over every vehicle (the reference writes the V=2 sum out)."""
    V, H = st.local_x.shape[:2]
    eq_con1 = np.zeros([H, 2])
    ueq_con1 = np.zeros(H)
    for i_tstep in range(H):
        for v in range(V):
            eq_con1[i_tstep, :] += st.A[v, i_tstep, :, :].T @ st.lamb_ij[v, i_tstep, :]
            ueq_con1[i_tstep] += -st.b[v, i_tstep, :].T @ st.lamb_ij[v, i_tstep, :]
    return bool((abs(eq_con1) <= thres).all() and (ueq_con1 >= min_dis).all())


def iterate_next_state(st: MidState):
    """optimizer.py:337-344"""
    for f in MidState.FIELDS:
        a = getattr(st, f)
        setattr(st, f, np.concatenate((a[:, 1:], a[:, -1:]), axis=1))
    return st


def residuals(bar_ctrl, bar_ctrl_prev, lamb_bar, lamb_bar_prev):
    """decentralized_overtaking_ADMM.py:82-85 (np.sqrt(x*x) kept literally), all vehicles summed."""
    primal = 0.0
    for v in range(len(bar_ctrl)):
        dlt = bar_ctrl[v] - bar_ctrl_prev[v]
        primal = primal + np.sum(np.sqrt(dlt * dlt))
    dl = lamb_bar - lamb_bar_prev
    dual = np.sum(np.sqrt(dl * dl))
    return float(primal), float(dual)


def consensus_tail(cur: MidState, prev: MidState, bar_ctrl, bar_ctrl_prev, i_iter, rho=1.0,
                   primal_thres=PRIMAL_THRES, dual_thres=DUAL_THRES, max_iter=MAX_ITER):
    """decentralized_overtaking_ADMM.py:79-97 for one scenario.  ``i_iter`` is the value AFTER the
    ``i_iter += 1`` of :41.  Returns (prev', bar_ctrl_prev', primal, dual, stop)."""
    lambda_update(cur, rho)
    primal, dual = residuals(bar_ctrl, bar_ctrl_prev, cur.lamb_bar, prev.lamb_bar)
    if (primal <= primal_thres and dual <= dual_thres) or i_iter > max_iter:
        return iterate_next_state(prev), bar_ctrl_prev, primal, dual, True
    return copy.deepcopy(cur), copy.deepcopy(bar_ctrl), primal, dual, False


# ---- batched (leading scenario axis) vectorised forms, checked against the loops in tests ----
def lambda_update_batched(local_x, lamb_ij, Z_bar, lamb_bar, rho):
    return lamb_bar + rho * (np.concatenate((local_x, lamb_ij), axis=-1) - Z_bar)


def shift_batched(a):
    return np.concatenate((a[:, :, 1:], a[:, :, -1:]), axis=2)
