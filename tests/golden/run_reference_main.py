"""Runs the REFERENCE'S OWN SCRIPT /root/reference/casadi/main.py, unmodified, in the build container and records what it
computed:

    python tests/golden/run_reference_main.py        # writes tests/golden/reference_main_run.json

The script needs four things this container lacks; each gets a stand-in that is as thin as possible:

  casadi (containers)   tests/golden/_ref_import_stubs.py (numpy-backed, see there)
  casadi.SX.sym         a placeholder for the decision variable; the script only slices / reshapes it and passes it on
  casadi.qpsol("osqp")  the QP SOLVE.  OSQP is not installed, so the solve is done by the oracle's restatement of the OSQP
                        iteration (oracle/solver.py).  What is solved:
                          * vehicle QP: the reference's own cost_function_primal, called on numeric arrays with the
                            script's arguments, probed at 0, +-e_i, e_i+e_j -> its exact (H, q); rows of nonlcon_function;
                          * edge problem: the script hands the Python (non-convex) cost to a QP solver (casadi/main.py:133-151);
                            this repository solves the CONVEX statement instead -- the stand-in
                            does the same from the script's own numeric arguments (xt_edge, pos_old_edge, dual_var_old_edge,
                            prev_pred_pos).
  matplotlib, bunch     no-ops / attribute dict

To see the numeric arguments of the two cost functions, the class the script imports (`from PI_ADMM_class import
PI_ADMM_CASADI`) is a subclass of the reference's class that overrides cost_function_primal / cost_function_edge /
nonlcon_function to RECORD their arguments and delegate to the reference's bodies when the solver asks for values.
Everything else the script does -- seed points, re-initialisation, the outer loop, pos_old roll-outs, edge_mat with the
un-squared threshold (:113), the early exit (:115-118), np.where edge order, z-hat roll-out (:156-158), the dual update
(:161-162), the residual sums and the aliasing of last_iter_hat_pos (:165-180), the stop test, the plant step and the
histories (:185-201) -- is the reference's own statements executing.  The recorded histories therefore pin the OUTER LOOP
of oracle/driver.py (and, through it, of the C twin and the CUDA closed loop) to the reference script; the inner solver's
iterates are this repository's on both sides (OSQP itself stays unpinned)."""
import importlib.util
import json
import os
import runpy
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
import _ref_import_stubs as stubs  # noqa: E402
from oracle.problem import Params, edge_qp_blocks, constraint_matrix, constraint_lower  # noqa: E402
from oracle.solver import InnerSettings, solve_box_qp, solve_edge_pair  # noqa: E402

REF = "/root/reference"
LOG = dict(veh_solves=0, edge_solves=0, veh_inner=[], edge_inner=[], cieq_checked=0)


class SymTok:
    """Placeholder for ca.SX.sym: the script slices it, stacks it, reshapes it and hands it to the class methods."""

    def __init__(self, shape):
        self.shape = shape

    def __getitem__(self, key):
        return SymTok(self.shape)


class Token:
    def __init__(self, kind, **kw):
        self.kind = kind
        self.__dict__.update(kw)


def make_wrapped_class(ref_cls):
    class PI_ADMM_CASADI(ref_cls):
        def cost_function_primal(self, num_step, u, xt, hat_pos, dual_var, veh_index):
            hat = np.array([[np.array(h, dtype=float) for h in row] for row in hat_pos], dtype=object) if False else hat_pos
            f = lambda un: float(np.asarray(ref_cls.cost_function_primal(self, num_step, stubs.Mat(np.asarray(un).reshape(1, -1)),
                                                                          xt, hat, dual_var, veh_index)).reshape(-1)[0])
            return Token("veh", f=f, owner=self)

        def nonlcon_function(self, u):
            return [], Token("con", owner=self, rows=ref_cls.nonlcon_function)

        def cost_function_edge(self, hat_u, xt, pos_old, dual_var_old, prev_pred_pos):
            return Token("edge", owner=self, xt=np.array(xt, dtype=float), pos_old=np.array(pos_old, dtype=float),
                         dual=np.array(dual_var_old, dtype=float), prev=np.array(prev_pred_pos, dtype=float))
    return PI_ADMM_CASADI


def oracle_params(P):
    pr = P.param
    return Params(dt=pr.dt, Nt=pr.Nt, L=pr.L, num_ho=pr.num_ho, num_veh=pr.num_veh, dis_thres=pr.dis_thres,
                  spd=np.asarray(pr.spd, dtype=float), beta=pr.beta, Pnorm=pr.Pnorm, Pcost=pr.Pcost, iter_num=pr.iter_num,
                  rho=pr.rho, eps_pri=pr.eps_pri, eps_dual=pr.eps_dual)


class QPSolver:
    """ca.qpsol("solver", "osqp", prob, options) -> F; F(lbg=0) -> {'x': ...} (casadi/main.py:96-103,146-153)."""

    def __init__(self, name, plugin, prob, options=None):
        assert plugin == "osqp"
        self.f, self.g = prob["f"], prob["g"]
        self.inner = InnerSettings.osqp_default()

    def __call__(self, lbg=0, **kw):
        assert lbg == 0
        P = self.f.owner
        N = int(P.param.num_ho)
        if self.f.kind == "veh":
            f = self.f.f
            E = np.eye(N)
            f0 = f(np.zeros(N))
            fp = np.array([f(E[i]) for i in range(N)]); fm = np.array([f(-E[i]) for i in range(N)])
            q = (fp - fm) / 2.0
            H = np.zeros((N, N))
            for i in range(N):
                H[i, i] = fp[i] + fm[i] - 2.0 * f0
                for j in range(i + 1, N):
                    H[i, j] = H[j, i] = f(E[i] + E[j]) - fp[i] - fp[j] + f0
            if LOG["cieq_checked"] < 3:      # the rows the oracle's solver uses ARE the rows nonlcon_function returns
                u = np.linspace(-0.4, 0.5, N)
                cieq = np.asarray(self.g.rows(P, stubs.Mat(u.reshape(1, N)))[1]).ravel()
                assert np.abs(cieq - (constraint_matrix(N) @ u - constraint_lower(N))).max() < 1e-13
                LOG["cieq_checked"] += 1
            x, info = solve_box_qp(H, q, self.inner)
            LOG["veh_solves"] += 1; LOG["veh_inner"].append(int(info["iters"]))
            return {"x": x.reshape(1, N)}
        t = self.f
        p = oracle_params(P)
        blk = edge_qp_blocks(p, t.xt, np.asarray(P.param.spd, dtype=float)[:2], t.pos_old, t.dual, t.prev)
        uh, info = solve_edge_pair(blk["H"], blk["q"], blk["a"], blk["c"], p.beta, self.inner)
        LOG["edge_solves"] += 1; LOG["edge_inner"].append(int(info["iters"]))
        return {"x": uh.reshape(1, 2 * N)}


def install_script_environment():
    ca, _ = stubs.install()
    ca.SX.sym = staticmethod(lambda name, r=1, c=1: SymTok((r, c)))
    real_vcat, real_reshape = ca.vcat, ca.reshape
    ca.vcat = lambda xs: SymTok(None) if any(isinstance(x, SymTok) for x in xs) else real_vcat(xs)
    ca.reshape = lambda x, r, c: x if isinstance(x, (SymTok, Token)) else real_reshape(x, r, c)
    ca.qpsol = QPSolver
    # matplotlib.pyplot: the script plots at the end (casadi/main.py:205-222)
    class _Any:
        def __getattr__(self, k):
            return lambda *a, **kw: (_Any(), _Any()) if k == "subplots" else _Any()
    mpl = types.ModuleType("matplotlib"); plt = types.ModuleType("matplotlib.pyplot")
    anyo = _Any()
    for k in ("subplots", "show", "figure", "plot", "scatter"):
        setattr(plt, k, getattr(anyo, k))
    mpl.pyplot = plt
    sys.modules["matplotlib"] = mpl; sys.modules["matplotlib.pyplot"] = plt
    # the class module the script imports: the reference's class, with the three symbolic methods recording their arguments
    spec = importlib.util.spec_from_file_location("ref_PI_ADMM_class", os.path.join(REF, "casadi", "PI_ADMM_class.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    wrapped = types.ModuleType("PI_ADMM_class")
    wrapped.PI_ADMM_CASADI = make_wrapped_class(m.PI_ADMM_CASADI)
    sys.modules["PI_ADMM_class"] = wrapped


def main():
    install_script_environment()
    g = runpy.run_path(os.path.join(REF, "casadi", "main.py"), run_name="__reference_main__")
    out = dict(generator="tests/golden/run_reference_main.py: /root/reference/casadi/main.py executed unmodified; QP solves by the "
                         "oracle's OSQP restatement (vehicle QP = the reference's own cost function probed numerically; edge = the "
                         "convex statement)",
               iter_his=[int(v) for v in g["iter_his"]], total_iter=int(g["total_iter"]), sum_iter_num=int(g["sum_iter_num"]),
               x_vec=[np.asarray(a).tolist() for a in g["x_vec"]], theta_vec=[np.asarray(a).tolist() for a in g["theta_vec"]],
               u_vec=[np.asarray(a).tolist() for a in g["u_vec"]], final_xt=np.asarray(g["xt"]).tolist(),
               final_primal_u=np.asarray(g["primal_u"]).tolist(), final_pos_old=np.asarray(g["pos_old"]).tolist(),
               veh_solves=LOG["veh_solves"], edge_solves=LOG["edge_solves"], veh_inner=LOG["veh_inner"], edge_inner=LOG["edge_inner"])
    path = os.path.join(HERE, "reference_main_run.json")
    json.dump(out, open(path, "w"))
    print("wrote", path, "steps", len(out["iter_his"]), "vehicle QPs", out["veh_solves"], "edge problems", out["edge_solves"],
          "iter_his", out["iter_his"])


if __name__ == "__main__":
    main()
