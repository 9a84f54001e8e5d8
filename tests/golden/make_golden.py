"""Generates tests/golden/*.json from the numpy oracle (oracle/driver.py).

These files freeze the ORACLE's outputs; they guard the oracle, its C twin and the CUDA path against regressions.
c1_python_current_default.json is additionally equal (difference 0.0) to what the reference's own casadi/main.py computed
when run unmodified here with the oracle's solver standing in for OSQP (tests/golden/run_reference_main.py ->
reference_main_run.json); the PI variants have no runnable reference script (MATLAB / IPOPT).
   python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle.problem import Params, reference_trajectories  # noqa: E402
from oracle.driver import Modes, run_closed_loop, run_step  # noqa: E402
from oracle.solver import InnerSettings  # noqa: E402

CASES = {
    # name: (Params overrides, Modes factory, inner)
    "c1_python_current_default": ({}, Modes.python_current, "default"),
    "c1_python_current_tight": ({}, Modes.python_current, "tight"),
    "c1_matlab_pi_default": (dict(beta=1000.0, Pnorm=5.0, rho=3.5, eps_pri=0.1, eps_dual=0.1, num_ho=8),
                             Modes.matlab_pi, "default"),
    "c1_python_old_pi_default": (dict(rho=1.0, eps_pri=1.0, eps_dual=1.0, num_ho=5, beta=10000.0, Pnorm=5.0,
                                      dis_thres=1.5), Modes.python_old_pi, "default"),
}


def inner_of(name):
    return InnerSettings.tight() if name == "tight" else InnerSettings.osqp_default()


def closed_loop(name):
    pkw, mf, inner = CASES[name]
    p = Params(**pkw)
    ref = reference_trajectories(p)
    xt0 = np.array([[-10, 0, 0], [0, 20, -np.pi / 2]], dtype=np.float64)
    out = run_closed_loop(p, mf(), inner_of(inner), xt0, p.spd.astype(np.float64), ref)
    return dict(
        params=pkw, inner=inner,
        iter_his=[int(v) for v in out["iter_his"]], n_outer=[int(v) for v in out["n_outer"]],
        err_rk=[[float(x) for x in e] for e in out["err_rk"]],
        err_sk=[[float(x) for x in e] for e in out["err_sk"]],
        x_vec=[a.tolist() for a in out["x_vec"]], theta_vec=[a.tolist() for a in out["theta_vec"]],
        u_vec=[a.tolist() for a in out["u_vec"]],
        primal_u_step12=out["steps"][min(12, len(out["steps"]) - 1)]["primal_u"].tolist(),
        veh_inner=[int(sum(s["veh_iters"])) for s in out["steps"]],
        edge_inner=[int(sum(s["edge_iters"])) for s in out["steps"]])


def main():
    for name in CASES:
        with open(os.path.join(HERE, name + ".json"), "w") as f:
            json.dump(closed_loop(name), f)
        print("wrote", name)


if __name__ == "__main__":
    main()
