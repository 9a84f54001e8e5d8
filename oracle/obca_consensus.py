"""TEST INFRASTRUCTURE ONLY -- CPU oracle of the consensus pieces of the OBCA overtaking ADMM
(SURVEY.md 8(a) row a13).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import this module; the product path never does.

numpy restatement, loop for loop, of
  * ``mid_state``            Distributed_planner/decentralized/optimizer.py:351-373
  * ``bar_state_update``     Distributed_planner/decentralized/optimizer.py:205-222 (+ util.py:12-101 half-spaces)
  * ``lambda_update``        Distributed_planner/decentralized/optimizer.py:330-335
  * ``check_converge``       Distributed_planner/decentralized/optimizer.py:225-235
  * ``iterate_next_state``   Distributed_planner/decentralized/optimizer.py:337-344
  * residuals / stop / carry Distributed_planner/decentralized_overtaking_ADMM.py:79-97
PINNED against the reference's own methods: tests/golden/reference_vectors.json holds what
OBCAOptimizer.bar_state_update / check_converge / lambda_update / iterate_next_state / mid_state returned when the
reference package was imported unmodified from /root/reference (numpy stand-ins for the casadi containers it imports,
tests/golden/make_reference_vectors.py); tests/test_reference_vectors.py compares this module and the CUDA kernels with
them (multipliers and shifts bit-exact).  The driver-inline statements (residuals / stop / carry) have no function in
the reference to call and stay literal restatements.
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


# ---- half-spaces of the vehicle rectangle (Distributed_planner/decentralized/util.py, veh_config.py:9-27) ----
VEH_LENGTH, VEH_WIDTH, AVG_DELAY, VAR_DELAY, DELAY_PROB = 3.5, 2.0, 0.05, 0.025, 0.95


def halfspaces_prob(state):
    """compute_square_halfspaces_ca_prob (util.py:76-101), numeric: A [4,2], b [4] of the delay-tightened rectangle."""
    x, y, v, th = state[0], state[1], state[2], state[3]
    c, s = np.cos(th), np.sin(th)
    R = np.array([[c, -s], [s, c]])
    A = np.vstack((R.T, -R.T))
    b0 = np.array([VEH_LENGTH / 2, VEH_WIDTH / 2, VEH_LENGTH / 2, VEH_WIDTH / 2])          # `if(~base_link)` is always true (:94)
    d_avg = np.array([AVG_DELAY * v * c, AVG_DELAY * v * s])
    d_var = np.array([(VAR_DELAY * v * c) ** 2, (VAR_DELAY * v * s) ** 2])
    return A, b0 + A @ (np.array([x, y]) + d_avg + np.sqrt(DELAY_PROB / (1 - DELAY_PROB)) * d_var)


def halfspaces_vertices(state):
    """generate_vehicle_vertices (util.py:12-47, base_link=False) + compute_square_halfspaces_ca (:49-74), numeric."""
    x, y, th = state[0], state[1], state[3]
    c, s, L, W = np.cos(th), np.sin(th), VEH_LENGTH, VEH_WIDTH
    P = np.array([[x + L / 2 * c - W / 2 * s, y + L / 2 * s + W / 2 * c], [x + L / 2 * c + W / 2 * s, y + L / 2 * s - W / 2 * c],
                  [x - L / 2 * c + W / 2 * s, y - L / 2 * s - W / 2 * c], [x - L / 2 * c - W / 2 * s, y - L / 2 * s + W / 2 * c]])
    A = np.zeros((4, 2)); b = np.zeros(4)
    for i in range(4):
        p1, p2 = P[i], P[(i + 1) % 4]
        nrm = np.sqrt((p2[1] - p1[1]) ** 2 + (p1[0] - p2[0]) ** 2)
        A[i] = np.array([p1[1] - p2[1], p2[0] - p1[0]]) / nrm
        b[i] = (p2[0] * p1[1] - p2[1] * p1[0]) / nrm
    return A, b


def bar_state_update(st: MidState, bar_fullx_vehs, prob=1):
    """optimizer.py:205-222 (lamb_ij is not updated: the reference's line is commented out, :219-220)."""
    V, H = st.local_x.shape[:2]
    for i_tstep in range(H):
        for i_veh in range(V):
            bar_x = np.asarray(bar_fullx_vehs[i_veh])[:, :5]
            A, b = (halfspaces_prob if prob else halfspaces_vertices)(bar_x[i_tstep])
            st.A[i_veh, i_tstep] = A
            st.b[i_veh, i_tstep] = b
            st.local_x[i_veh, i_tstep] = bar_x[i_tstep]
