"""Synthetic inputs of the BASELINE.json configs; numpy only, no solver code."""
from __future__ import annotations

import numpy as np


def reference_trajectories(Nt=5.0, dt=0.1):
    """casadi/PI_ADMM_class.py:31-37: rows [x_A; 0; 0; y_B], int(Nt/dt) columns."""
    T = int(Nt / dt)
    x_A = np.linspace(-10, 10, T)
    y_B = np.linspace(20, -20, T)
    return np.vstack((x_A, np.zeros(T), np.zeros(T), y_B))


def c1_initial_state():
    """casadi/main.py:25."""
    return np.array([[-10.0, 0.0, 0.0], [0.0, 20.0, -np.pi / 2]])


def ref_window(ref_traj, num_step, N):
    """ref_traj (2V,T) -> (V,2,N+1) window used by cost_function_primal (PI_ADMM_class.py:116-117)."""
    V = ref_traj.shape[0] // 2
    return np.stack([ref_traj[2 * i:2 * i + 2, num_step:num_step + N + 1] for i in range(V)])


def c4_scenarios(S=4096, N=15, seed=4096, step=12, T=50):
    """Config 4: S perturbed copies of the two-vehicle intersection, each started at MPC step
    `step` of its nominal run so the collision edge is active inside the horizon.
    Returns xt (S,2,3), spd (S,2), refwin (S,2,2,N+1)."""
    rng = np.random.default_rng(seed)
    xA0 = -10.0 + rng.uniform(-1, 1, S)
    yB0 = 20.0 + rng.uniform(-2, 2, S)
    spd = np.array([4.0, 8.0])[None, :] * (1.0 + rng.uniform(-0.1, 0.1, (S, 2)))
    xt = np.zeros((S, 2, 3))
    xt[:, 0, 0] = xA0 + 0.1 * step * spd[:, 0]
    xt[:, 1, 1] = yB0 - 0.1 * step * spd[:, 1]
    xt[:, 1, 2] = -np.pi / 2
    ref = _c4_reference(xA0, yB0, T)
    refwin = np.ascontiguousarray(ref[:, :, :, step:step + N + 1])
    return xt, spd, refwin


def _c4_reference(xA0, yB0, T):
    lam = np.linspace(0.0, 1.0, T)[None, :]
    ref = np.zeros((len(xA0), 2, 2, T))
    ref[:, 0, 0, :] = xA0[:, None] * (1 - 2 * lam)       # straight line from the start through the mirrored end point
    ref[:, 1, 1, :] = yB0[:, None] * (1 - 2 * lam)
    return ref


def c4_reference(S=4096, seed=4096, T=50):
    """Whole reference trajectories (S,2,2,T) of the scenarios c4_scenarios(S, seed=seed) draws: what
    PI_ADMM_CASADI.ref_traj is for config 1; c4_scenarios returns the window [step, step+N] of it."""
    rng = np.random.default_rng(seed)
    xA0 = -10.0 + rng.uniform(-1, 1, S)
    yB0 = 20.0 + rng.uniform(-2, 2, S)
    return _c4_reference(xA0, yB0, T)


def c3_intersection(V=64, N=30, seed=64, dt=0.1, r0=12.0, gap=6.0):
    """Config 3: 4-arm intersection, V/4 vehicles per arm heading to the origin."""
    rng = np.random.default_rng(seed)
    per = V // 4
    xt = np.zeros((1, V, 3)); spd = np.zeros((1, V)); refwin = np.zeros((1, V, 2, N + 1))
    speeds = (4.0, 6.0, 8.0)
    i = 0
    for arm in range(4):
        ang = arm * np.pi / 2                    # position angle of the arm
        for q in range(per):
            r = r0 + gap * q + rng.uniform(-0.5, 0.5)
            v = speeds[(arm + q) % 3]
            px, py = r * np.cos(ang), r * np.sin(ang)
            th = np.arctan2(-py, -px)            # heading to the origin
            xt[0, i] = (px, py, th)
            spd[0, i] = v
            k = np.arange(N + 1)
            refwin[0, i, 0] = px + k * dt * v * np.cos(th)
            refwin[0, i, 1] = py + k * dt * v * np.sin(th)
            i += 1
    return xt, spd, refwin


def c5_grid(lanes=64, per_lane=32, N=30, seed=4096, dt=0.1, lane_gap=20.0, clear=3.0):
    """Config 5: `lanes` east-bound lanes (y = gap*r) and `lanes` south-bound lanes (x = gap*c),
    `per_lane` vehicles each, along-lane spacing 20 + U(-3,3).

    clear > 0 (default): rejection rule -- a spacing draw is repeated while it would START the vehicle within `clear`
    metres (along its lane) of a crossing point.  Without it (clear=0, the round-1 generator) 15 east/south pairs start
    already inside dis_thres (minimum 0.84 m); the first two trajectory points do not depend on the steering sequence,
    so no plan can repair those pairs and their hinge terms stay active for ever.  With clear=3 every pair is at
    least 3*sqrt(2) = 4.2 m apart at t=0; the conflicts that remain are the ones steering can resolve."""
    rng = np.random.default_rng(seed)
    V = 2 * lanes * per_lane
    xt = np.zeros((1, V, 3)); spd = np.zeros((1, V)); refwin = np.zeros((1, V, 2, N + 1))
    speeds = (4.0, 6.0, 8.0)
    k = np.arange(N + 1)

    def advance(pos, sign):
        while True:
            nxt = pos + sign * (20.0 + rng.uniform(-3, 3))
            c = round(nxt / lane_gap)
            if clear <= 0 or not (0 <= c < lanes) or abs(nxt - lane_gap * c) >= clear:
                return nxt

    i = 0
    for r in range(lanes):                       # east-bound
        pos = -10.0
        for q in range(per_lane):
            pos = advance(pos, +1.0)
            v = speeds[(r + q) % 3]
            xt[0, i] = (pos, lane_gap * r, 0.0); spd[0, i] = v
            refwin[0, i, 0] = pos + k * dt * v; refwin[0, i, 1] = lane_gap * r
            i += 1
    for c in range(lanes):                       # south-bound
        pos = lane_gap * lanes + 10.0
        for q in range(per_lane):
            pos = advance(pos, -1.0)
            v = speeds[(c + q) % 3]
            xt[0, i] = (lane_gap * c, pos, -np.pi / 2); spd[0, i] = v
            refwin[0, i, 0] = lane_gap * c; refwin[0, i, 1] = pos - k * dt * v
            i += 1
    return xt, spd, refwin
