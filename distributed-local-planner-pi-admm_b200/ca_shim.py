"""The handful of CasADi entry points the reference driver uses, over the CUDA solver.

casadi/main.py builds every QP symbolically and hands it to ``ca.qpsol("solver", "osqp", {...})`` (:83-103 for a
vehicle, :133-153 for an edge).  With

    from pi_admm_b200 import ca_shim as ca
    from pi_admm_b200.PI_ADMM_class import PI_ADMM_CASADI

in place of ``import casadi as ca`` / ``from PI_ADMM_class import PI_ADMM_CASADI`` those statements stay as they are:
``ca.SX.sym`` returns a shape-only placeholder, ``cost_function_primal`` / ``cost_function_edge`` / ``nonlcon_function``
return descriptors of the numeric data the problem depends on, and the callable ``qpsol`` returns solves the described
problem in the kernels of libpiadmm.so (the vehicle problem through piadmm_x_update, the edge problem through
piadmm_solve_edges) and hands back ``{'x': ...}`` like a CasADi function call.  Nothing here is a CPU solver: without
the CUDA library the first solve raises.

Only what the reference scripts call is provided (SX.sym, reshape, vcat, qpsol): this is an adapter for that one
call sequence, not a CasADi replacement.  The batched fast path (one persistent kernel per MPC step) is
``PI_ADMM_CASADI.step`` / ``PIADMMSolver``; this module exists so that a script in casadi/main.py's shape runs unchanged.
"""
from __future__ import annotations

import numpy as np


class _Sym:
    """Shape-only stand-in for an SX matrix (the decision variable of a QP)."""

    def __init__(self, name, rows, cols):
        self.name, self.shape = name, (int(rows), int(cols))

    def __getitem__(self, idx):
        n = self.shape[0] * self.shape[1]
        k = len(range(*idx.indices(n))) if isinstance(idx, slice) else 1
        return _Sym(self.name, 1, k)

    def __repr__(self):
        return f"SX({self.name}, {self.shape[0]}x{self.shape[1]})"


class SX:
    @staticmethod
    def sym(name, rows=1, cols=1):
        return _Sym(name, rows, cols)


def reshape(x, *shape):
    """ca.reshape of a descriptor / placeholder is the identity (casadi/main.py:91); arrays are reshaped."""
    if isinstance(x, np.ndarray):
        return x.reshape(*shape)
    return x


def vcat(xs):
    """Vertical concatenation of placeholders (casadi/main.py:135-136: the two halves of hat_u stacked to 2 x N)."""
    xs = list(xs)
    if all(isinstance(x, _Sym) for x in xs):
        return _Sym("vcat", len(xs), xs[0].shape[0] * xs[0].shape[1])
    return np.vstack([np.asarray(x) for x in xs])


class _QPFunction:
    """What ``ca.qpsol`` returns: call it (``F(lbg=0)``) to solve; the result is ``{'x': ndarray}``."""

    def __init__(self, prob, opts):
        self.prob, self.opts = prob, dict(opts or {})
        f = prob.get("f")
        if not hasattr(f, "solve"):
            raise TypeError("qpsol: prob['f'] must come from PI_ADMM_CASADI.cost_function_primal / cost_function_edge")
        g = prob.get("g")
        if g is not None and not getattr(g, "is_steering_box", False):
            raise TypeError("qpsol: prob['g'] must come from PI_ADMM_CASADI.nonlcon_function")

    def __call__(self, lbg=0, x0=None, **kw):
        if np.any(np.asarray(lbg) != 0):
            raise ValueError("the constraint rows are stated as g(u) >= 0 (casadi/main.py:101,151): lbg must be 0")
        # CasADi's error_on_fail defaults to True (vehicle problem, casadi/main.py:96) and is switched off for the
        # edge problem (:145)
        x = self.prob["f"].solve(error_on_fail=bool(self.opts.get("error_on_fail", True)))
        return {"x": x}


def qpsol(name, plugin, prob, opts=None):
    if plugin != "osqp":
        raise ValueError("only the 'osqp' role is provided (the reference calls qpsol with nothing else)")
    return _QPFunction(prob, opts)
