"""Host-side mirror of the reference's problem class (casadi/PI_ADMM_class.py:12-192).

Same class name, constructor, ``param`` attribute bag, reference trajectories and method names /
argument meaning as ``PI_ADMM_CASADI``.  What differs is what the symbolic methods return: the
reference builds CasADi SX expressions that ``ca.qpsol("osqp")`` turns into a QP at every call;
here ``cost_function_primal`` / ``cost_function_edge`` / ``nonlcon_function`` return light
descriptors (the numeric data the QP depends on) and the solve itself runs in the CUDA kernels of
libpiadmm.so (``solve_vehicles`` / ``solve_edges`` / ``step``).  The numeric (``if_SX=0``) roll-outs
are the reference's own few lines of numpy index arithmetic and stay on the host, they are interface
glue for scripts that call them directly; the batched solver never uses them.
"""
from __future__ import annotations

import numpy as np

from . import _abi
from .solver import PIADMMSolver


class Bunch(dict):
    """Minimal stand-in for ``bunch.Bunch`` (attribute access + mutation), casadi/PI_ADMM_class.py:8."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


class VehicleQP:
    """Descriptor returned by cost_function_primal: everything the vehicle QP depends on."""

    def __init__(self, num_step, xt, hat_pos, dual_var, veh_index):
        self.num_step, self.xt, self.hat_pos, self.dual_var, self.veh_index = num_step, xt, hat_pos, dual_var, veh_index


class EdgeQP:
    """Descriptor returned by cost_function_edge."""

    def __init__(self, xt, pos_old, dual_var_old, prev_pred_pos):
        self.xt, self.pos_old, self.dual_var_old, self.prev_pred_pos = xt, pos_old, dual_var_old, prev_pred_pos


class Constraints:
    """Descriptor returned by nonlcon_function: g(u) = [u+pi/6; -u+pi/6; -D1u+pi/9; D1u+pi/9] >= 0."""

    def __init__(self, rows, N):
        self.rows, self.N = rows, N
        self.u_max, self.du_max = np.pi / 6, np.pi / 9
        self.m = 4 * rows * N - 2 * rows


def check_status(status: int, error_on_fail: bool = True, where: str = ""):
    """The reference's failure behaviour for one scenario's PIADMM_ST_* bits: a vehicle QP that did not converge
    raises (CasADi's error_on_fail=True default, casadi/main.py:96-103), an edge problem that did not converge is
    tolerated (error_on_fail=False, casadi/main.py:145-153); running out of neighbour slots is a sizing error of
    this implementation and always raises."""
    if status & _abi.ST_SLOT_OVERFLOW:
        raise _abi.PIADMMError(f"{where}: a vehicle needed more than D neighbour slots (PIADMM_ST_SLOT_OVERFLOW)")
    if error_on_fail and status & _abi.ST_VEH_MAXITER:
        raise RuntimeError(f"{where}: vehicle QP did not converge within in_max_iter inner iterations "
                           "(PIADMM_ST_VEH_MAXITER; the reference's qpsol raises here, error_on_fail=True)")


class PI_ADMM_CASADI(object):
    def __init__(self, semantics: str = "python_current", inner: str = "osqp_default", **overrides):
        # ADMM parameter setting -- casadi/PI_ADMM_class.py:15-28
        self.param = Bunch(dt=0.1, Nt=5, L=1, num_ho=15, num_veh=2, dis_thres=2, spd=np.array([4, 8]),
                           beta=10, Pnorm=1, Pcost=1, iter_num=100, rho=2, eps_pri=20, eps_dual=1)
        Nt, dt = self.param.Nt, self.param.dt
        x_A = np.linspace(-10, 10, int(Nt / dt))
        y_B = np.linspace(20, -20, int(Nt / dt))
        self.ref_traj_A = np.vstack((x_A, np.zeros_like(x_A)))
        self.ref_traj_B = np.vstack((np.zeros_like(y_B), y_B))
        self.ref_traj = np.vstack((self.ref_traj_A, self.ref_traj_B))
        self.semantics, self.inner, self.overrides = semantics, inner, dict(overrides)
        # casadi/main.py:96 builds the vehicle solver with CasADi's default error_on_fail=True (a failed vehicle QP
        # raises RuntimeError) and the edge solver with error_on_fail=False (:145-146, a failed edge solve is used
        # as it is).  step() mirrors that: PIADMM_ST_VEH_MAXITER raises, PIADMM_ST_EDGE_MAXITER is only recorded.
        self.error_on_fail = True
        self._solver = None
        self._solver_key = None

    # ------------------------------------------------------------------ numeric roll-outs (if_SX=0)
    def dynamic_update_local(self, xt, u, veh_index, if_SX=0):
        """casadi/PI_ADMM_class.py:45-70, numeric branch: u is 2-D (1, N), indexed u[0][k]."""
        if if_SX:
            raise NotImplementedError("symbolic roll-outs are replaced by QP descriptors; use cost_function_primal")
        p = self.param
        v = p.spd[veh_index]
        N = p.num_ho
        x = np.zeros(N + 1); y = np.zeros(N + 1); th = np.zeros(N + 1)
        x[0], y[0], th[0] = xt[0], xt[1], xt[2]
        for k in range(N):
            x[k + 1] = x[k] + (-v * np.sin(xt[2]) * th[k] + (v * np.cos(xt[2]) + v * xt[2] * np.sin(xt[2]))) * p.dt
            y[k + 1] = y[k] + (v * np.cos(xt[2]) * th[k] + (v * np.sin(xt[2]) - v * xt[2] * np.cos(xt[2]))) * p.dt
            th[k + 1] = th[k] + v / p.L * u[0][k] * p.dt
        return x, y, th

    def dynamic_update_edge(self, xt, u, if_SX=0):
        """casadi/PI_ADMM_class.py:77-105, numeric branch (non-linear roll-out, local speed index)."""
        if if_SX:
            raise NotImplementedError("symbolic roll-outs are replaced by QP descriptors; use cost_function_edge")
        p = self.param
        R, N = p.num_veh, p.num_ho
        x = np.zeros((R, N + 1)); y = np.zeros((R, N + 1)); th = np.zeros((R, N + 1))
        x[:, 0], y[:, 0], th[:, 0] = xt[:, 0], xt[:, 1], xt[:, 2]
        for k in range(N):
            for i in range(R):
                v = p.spd[i]
                t = th[i, k]
                x[i, k + 1] = x[i, k] + (-v * np.sin(t) * t + (v * np.cos(t) + v * t * np.sin(t))) * p.dt
                y[i, k + 1] = y[i, k] + (v * np.cos(t) * t + (v * np.sin(t) - v * t * np.cos(t))) * p.dt
                th[i, k + 1] = t + v / p.L * u[i, k] * p.dt
        return x, y, th

    # ------------------------------------------------------------------ problem descriptors
    def cost_function_primal(self, num_step, u, xt, hat_pos, dual_var, veh_index):
        """casadi/PI_ADMM_class.py:114-135 -> VehicleQP descriptor (u is ignored: the decision variable)."""
        return VehicleQP(num_step, np.asarray(xt, dtype=np.float64), hat_pos, dual_var, veh_index)

    def cost_function_edge(self, hat_u, xt, pos_old, dual_var_old, prev_pred_pos):
        """casadi/PI_ADMM_class.py:145-169 -> EdgeQP descriptor (convex statement)."""
        return EdgeQP(np.asarray(xt, dtype=np.float64), pos_old, dual_var_old, prev_pred_pos)

    def nonlcon_function(self, u):
        """casadi/PI_ADMM_class.py:172-192: returns (ceq, cieq) with cieq a Constraints descriptor."""
        shape = getattr(u, "shape", (1, self.param.num_ho))
        rows = shape[0] if len(shape) == 2 else 1
        return [], Constraints(rows, self.param.num_ho)

    # ------------------------------------------------------------------ solver plumbing
    def abi_params(self) -> _abi.Params:
        """self.param (mutable, like the reference's Bunch) -> struct piadmm_params."""
        p = _abi.preset(self.semantics)
        pr = self.param
        p.update(dt=float(pr.dt), L=float(pr.L), dis_thres=float(pr.dis_thres), beta=float(pr.beta),
                 Pnorm=float(pr.Pnorm), Pcost=float(pr.Pcost), rho=float(pr.rho), eps_pri=float(pr.eps_pri),
                 eps_dual=float(pr.eps_dual), num_ho=int(pr.num_ho), iter_num=int(pr.iter_num))
        if self.inner == "tight":
            _abi.inner_tight(p)
        p.update(**self.overrides)
        return p

    def solver(self, S=1) -> PIADMMSolver:
        """The batched CUDA solver for S copies of this problem (created lazily, rebuilt when the
        horizon / iteration cap / batch size changes, otherwise parameters are re-sent)."""
        p = self.abi_params()
        key = (S, int(self.param.num_veh), p.num_ho, p.iter_num)
        if self._solver is None or self._solver_key != key:
            if self._solver is not None:
                self._solver.close()
            self._solver = PIADMMSolver(p, S=S, V=int(self.param.num_veh))
            self._solver_key = key
        else:
            self._solver.set_params(**p.as_dict())
        return self._solver

    def ref_window(self, num_step):
        N, V = int(self.param.num_ho), int(self.param.num_veh)
        return np.stack([self.ref_traj[2 * i:2 * i + 2, num_step:num_step + N + 1] for i in range(V)])

    def step(self, num_step, xt, rho=None):
        """One whole MPC step of casadi/main.py:46-192 on the GPU (fused persistent kernel).
        Returns dict(primal_u, pos_old, iter, n_outer, err_rk, err_sk, xt_next, rho)."""
        sol = self.solver(1)
        sol.setup(np.asarray(xt, dtype=np.float64)[None], np.asarray(self.param.spd, dtype=np.float64)[None],
                  self.ref_window(num_step)[None], rho=self.param.rho if rho is None else rho)
        sol.iterate()
        sol.plant_step()
        o = sol.to_numpy()
        n = int(o["n_outer"][0])
        check_status(int(o["status"][0]), self.error_on_fail, f"MPC step {num_step}")
        return dict(primal_u=o["u"][0], pos_old=o["pos"][0].reshape(-1, o["pos"].shape[-1]), iter=int(o["curr_iter"][0]),
                    n_outer=n, err_rk=o["err_rk"][0, :n], err_sk=o["err_sk"][0, :n], xt_next=o["xt_next"][0],
                    rho=float(o["rho_s"][0]), hat=o["hat"][0], dual=o["dual"][0], status=int(o["status"][0]))
