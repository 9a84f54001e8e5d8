"""Runs the reference's MATLAB script matlab_old_files/ADMM_CVX_two_veh_intesection_PI_antiwindup.m -- the source of the
`matlab_pi` semantics every bench configuration uses -- in the build container, from its own text:

    python tests/golden/run_reference_matlab.py        # writes tests/golden/reference_matlab_run.json

There is no MATLAB / Octave and no CVX here.  The script is executed by the MATLAB-subset interpreter of
tests/golden/_matlab_subset.py (statements, its four local functions, cell arrays, the whole receding-horizon loop :52-238);
the two `cvx_begin ... cvx_end` blocks (:86-99 vehicle problem, :131-145 edge problem) are replaced by hooks that solve the
problem the block states with the oracle's OSQP restatement at tight tolerance (CVX solves to optimality, so the result is
the minimiser, whatever the solver):

  * vehicle: minimise the script's own cost_function_primal (its (H, q): the oracle's closed form, checked against the
    MATLAB function's value at a random point in every call), rows of the block;
  * edge: minimise the script's own cost_function_edge (convex: hinge + effort + augmented Lagrangian behind the linearised
    roll-out), same check, rows of the block.

`steer` selects the steering rows: "two_sided" = |u| <= pi/6 (what this repository's presets carry),
"as_written" = the block's own `veh_u <= pi/6` with NO lower bound (:95, :141).  Both are run; the JSON says whether they
differ on this scenario.  Plot / timing / print statements are no-ops.  Everything else -- seed points, re-initialisation,
collision test, early exit, PI update with saturation and back-calculation, residual sums, stop test, plant step with the
non-linear roll-out, histories -- is the script's text executing."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
import _matlab_subset as ms  # noqa: E402
from oracle.problem import Params, vehicle_qp, edge_qp_blocks, constraint_matrix, constraint_lower, diff1, roll_matrix  # noqa: E402
from oracle.solver import InnerSettings, solve_box_qp, solve_edge_pair, solve_dense_qp  # noqa: E402

REF = "/root/reference/matlab_old_files/ADMM_CVX_two_veh_intesection_PI_antiwindup.m"
LOG = dict(veh=0, edge=0, max_cost_mismatch=0.0)


def params_of(P):
    return Params(dt=P["dt"], L=P["L"], num_ho=int(P["num_ho"]), dis_thres=P["dis_thres"], spd=np.asarray(P["spd"], dtype=float).ravel(),
                  beta=P["beta"], Pnorm=P["Pnorm"], Pcost=P["Pcost"], rho=P["rho"])


def make_hooks(functions, steer, inner):
    rng = np.random.default_rng(5)

    def vehicle(ns):
        P = ns["param"]; p = params_of(P); N = p.num_ho
        v = int(ns["i_veh"]) - 1; o = 1 - v
        s0 = int(ns["num_step"]) - 1
        xt = np.asarray(ns["xt"], dtype=float)
        ref = np.asarray(ns["ref_traj"], dtype=float)[2 * v:2 * v + 2, s0:s0 + N + 1]
        hat, dual = ns["hat_pos_old"][(v + 1, o + 1)], ns["dual_var_old"][(v + 1, o + 1)]
        H, q = vehicle_qp(p, xt[v], p.spd[v], ref[0], ref[1], [hat], [dual])
        # the quadratic IS the script's cost function: value check at a random point, through the MATLAB function itself
        z = np.zeros((1, N + 1))
        I = ms.Interp({}, functions)
        f = lambda u: float(I.call_user("cost_function_primal", [float(s0 + 1), u.reshape(1, N), xt[v:v + 1], ns["ref_traj"], P, z, z, z,
                                                                 ns["hat_pos_old"], ns["dual_var_old"], float(v + 1)], 1)[0])
        u = rng.normal(0, 0.3, N)
        LOG["max_cost_mismatch"] = max(LOG["max_cost_mismatch"], abs((f(u) - f(np.zeros(N))) - (0.5 * u @ H @ u + q @ u)) / max(1.0, abs(f(u))))
        if steer == "two_sided":
            x, info = solve_box_qp(H, q, inner)
        else:
            D1 = diff1(N)
            A = np.vstack((np.eye(N), D1))
            l = np.concatenate((np.full(N, -np.inf), np.full(N - 1, -np.pi / 9)))
            uu = np.concatenate((np.full(N, np.pi / 6), np.full(N - 1, np.pi / 9)))
            x, _, info = solve_dense_qp(H, q, A, l, uu, inner)
        assert info["status"] == 0
        ns["veh_u"] = x.reshape(1, N)
        LOG["veh"] += 1

    def edge(ns):
        P = ns["param"]; p = params_of(P); N = p.num_ho
        xt_e = np.asarray(ns["xt_edge"], dtype=float)
        v1, v2 = int(ns["veh1"]) - 1, int(ns["veh2"]) - 1
        spd_e = np.array([p.spd[0], p.spd[1]])            # dynamic_update_edge takes param.spd by LOCAL row (:385), as the Python class does
        blk = edge_qp_blocks(p, xt_e, spd_e, np.asarray(ns["pos_old_edge"], dtype=float), np.asarray(ns["dual_var_old_edge"], dtype=float),
                             np.asarray(ns["prev_pred_pos"], dtype=float))
        z2, z4 = np.zeros((2, N + 1)), np.zeros((4, N + 1))
        I = ms.Interp({}, functions)
        f = lambda hu: float(I.call_user("cost_function_edge", [hu, xt_e, P, z2, z2, z2, z4, ns["pos_old_edge"], ns["dual_var_old_edge"],
                                                                ns["prev_pred_pos"]], 1)[0])
        M = roll_matrix(N)[1:]
        model = lambda u: (sum(0.5 * u[r] @ blk["H"][r] @ u[r] + blk["q"][r] @ u[r] for r in range(2))
                           + p.beta * np.sum(np.maximum(0.0, blk["c"] - blk["a"][0] * (M @ u[0]) - blk["a"][1] * (M @ u[1]))))
        hu = rng.normal(0, 0.3, (2, N)); z = np.zeros((2, N))
        LOG["max_cost_mismatch"] = max(LOG["max_cost_mismatch"], abs((f(hu) - f(z)) - (model(hu) - model(z))) / max(1.0, abs(f(hu))))
        if steer == "two_sided":
            uh, info = solve_edge_pair(blk["H"], blk["q"], blk["a"], blk["c"], p.beta, inner)
        else:                                               # the block as written: slack form, u <= pi/6 only
            D1 = diff1(N); Z = np.zeros
            n = 3 * N
            Pm = np.zeros((n, n)); Pm[:N, :N] = blk["H"][0]; Pm[N:2 * N, N:2 * N] = blk["H"][1]
            qv = np.concatenate((blk["q"][0], blk["q"][1], np.full(N, p.beta)))
            rows, lo, up = [], [], []
            for r in range(2):
                E = Z((N, n)); E[:, r * N:(r + 1) * N] = np.eye(N); rows.append(E); lo.append(np.full(N, -np.inf)); up.append(np.full(N, np.pi / 6))
                E = Z((N - 1, n)); E[:, r * N:(r + 1) * N] = D1; rows.append(E); lo.append(np.full(N - 1, -np.pi / 9)); up.append(np.full(N - 1, np.pi / 9))
            E = Z((N, n)); E[:, 2 * N:] = np.eye(N); rows.append(E); lo.append(np.zeros(N)); up.append(np.full(N, np.inf))          # t >= 0
            E = Z((N, n)); E[:, :N] = blk["a"][0] * M; E[:, N:2 * N] = blk["a"][1] * M; E[:, 2 * N:] = np.eye(N)                     # t + a.s >= c
            rows.append(E); lo.append(blk["c"]); up.append(np.full(N, np.inf))
            x, _, info = solve_dense_qp(Pm, qv, np.vstack(rows), np.concatenate(lo), np.concatenate(up),
                                        InnerSettings(eps_abs=1e-9, eps_rel=1e-9, max_iter=200000))
            uh = x[:2 * N].reshape(2, N)
        assert info["status"] == 0
        ns["hat_u"] = uh
        LOG["edge"] += 1

    return {"cvx_vehicle_hook": vehicle, "cvx_edge_hook": edge}


def run(steer):
    text = open(REF).read()
    L = text.split("\n")
    assert L[85].strip().startswith("cvx_begin") and L[98].strip() == "cvx_end" and L[130].strip().startswith("cvx_begin") and L[144].strip() == "cvx_end"
    assert L[242].startswith("%% function") and "veh_u <= pi/6" in L[94] and "hat_u <= pi/6" in L[140]
    script = L[:85] + ["cvx_vehicle_hook"] + L[99:130] + ["cvx_edge_hook"] + L[145:242]
    functions = ms.parse_functions(text)
    LOG.update(veh=0, edge=0, max_cost_mismatch=0.0)
    it = ms.Interp({}, functions, make_hooks(functions, steer, InnerSettings.tight()))
    it.run("\n".join(script))
    ns = it.ns
    T = int(round(ns["Nt"] / ns["dt"] - ns["param"]["num_ho"]))
    xv = np.asarray(ns["x_vec"]).reshape(T, 2, 2)            # rows appended per step: [x_A x_B; y_A y_B]
    return dict(steer=steer, steps=T, iter_his=[int(v) for v in np.asarray(ns["iter_his"]).ravel()],
                sum_iter_num=int(ns["sum_iter_num"]), x_vec=xv.tolist(), theta_vec=np.asarray(ns["theta_vec"]).tolist(),
                u_vec=np.asarray(ns["u_vec"]).tolist(), vehicle_solves=LOG["veh"], edge_solves=LOG["edge"],
                max_cost_mismatch=LOG["max_cost_mismatch"], min_u=float(np.asarray(ns["u_vec"]).min()))


def main():
    out = {"generator": "tests/golden/run_reference_matlab.py: the .m script executed from its text (MATLAB-subset interpreter); CVX blocks "
                        "solved by the oracle's OSQP restatement at tight tolerance"}
    for steer in ("two_sided", "as_written"):
        out[steer] = run(steer)
        r = out[steer]
        print(steer, "steps", r["steps"], "iter_his", r["iter_his"], "vehicle / edge solves", r["vehicle_solves"], r["edge_solves"],
              "cost mismatch", r["max_cost_mismatch"])
    a, b = out["two_sided"], out["as_written"]
    out["as_written_differs"] = dict(iter_his=a["iter_his"] != b["iter_his"],
                                     max_abs_x=float(np.abs(np.array(a["x_vec"]) - np.array(b["x_vec"])).max()),
                                     max_abs_u=float(np.abs(np.array(a["u_vec"]) - np.array(b["u_vec"])).max()))
    print("one-sided steering bound vs two-sided:", out["as_written_differs"])
    json.dump(out, open(os.path.join(HERE, "reference_matlab_run.json"), "w"))


if __name__ == "__main__":
    main()
