"""Host-side mirror of the reference's problem class (casadi/PI_ADMM_class.py:12-192).

Same class name, constructor, ``param`` attribute bag, reference trajectories and method names /
argument meaning as ``PI_ADMM_CASADI``.  What differs is what the symbolic methods return: the
reference builds CasADi SX expressions that ``ca.qpsol("osqp")`` turns into a QP at every call;
here ``cost_function_primal`` / ``cost_function_edge`` / ``nonlcon_function`` return
descriptors (the numeric data the QP depends on) whose ``solve()`` runs the problem in the CUDA kernels of
libpiadmm.so -- one at a time through ``solve`` / ``ca_shim.qpsol`` (a script in casadi/main.py's shape keeps its
statements, see main_qpsol_shape.py), or the whole MPC step at once through ``step``.  The numeric (``if_SX=0``) roll-outs
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
    """What cost_function_primal returns: the numeric data the vehicle QP of casadi/PI_ADMM_class.py:114-135 depends on.
    ``solve()`` runs it through piadmm_x_update (one owned vehicle of a V-vehicle handle, AL term over every j != i as the
    reference states it, :126-129) and returns the UNROUNDED steering sequence, shape (1, N) -- the script rounds
    (casadi/main.py:103)."""

    def __init__(self, owner, num_step, xt, hat_pos, dual_var, veh_index):
        self.owner, self.num_step, self.veh_index = owner, int(num_step), int(veh_index)
        self.xt, self.hat_pos, self.dual_var = np.asarray(xt, dtype=np.float64), hat_pos, dual_var

    def solve(self, error_on_fail=True):
        import torch
        P, i = self.owner, self.veh_index
        V, N = int(P.param.num_veh), int(P.param.num_ho)
        sol = P._descriptor_solver("vehicle", V, max(1, V - 1))
        xt = self.xt if self.xt.ndim == 2 else np.tile(self.xt, (V, 1))
        sol.setup(xt[None], np.asarray(P.param.spd, dtype=np.float64)[None], P.ref_window(self.num_step)[None],
                  rho=float(P.param.rho))
        sol.begin_step()                                   # zero state, slot d of vehicle i <-> the d-th j != i
        for d, j in enumerate([j for j in range(V) if j != i]):
            sol.buf["hat"][0, i, d].copy_(torch.from_numpy(np.ascontiguousarray(self.hat_pos[i, j], dtype=np.float64)))
            sol.buf["dual"][0, i, d].copy_(torch.from_numpy(np.ascontiguousarray(self.dual_var[i, j], dtype=np.float64)))
        sol.set_ownership(i, i + 1)
        sol.x_update()
        status = int(sol.buf["status"][0].item())
        if error_on_fail and status & _abi.ST_VEH_MAXITER:
            raise RuntimeError("vehicle QP did not converge within in_max_iter inner iterations (qpsol error_on_fail)")
        return sol.buf["u"][0, i].cpu().numpy().reshape(1, N).copy()


class EdgeQP:
    """What cost_function_edge returns (casadi/PI_ADMM_class.py:145-169; convex statement).  ``solve()`` runs
    the edge problem ALONE through piadmm_solve_edges and returns the unrounded controls of both vehicles, shape
    (1, 2N) = [veh1 | veh2]; the script reshapes / rounds (casadi/main.py:153) and does the roll-out of z-hat and the dual
    update itself.  Speeds are taken by LOCAL row index (spd[0], spd[1]) as dynamic_update_edge does (:93)."""

    def __init__(self, owner, xt, pos_old, dual_var_old, prev_pred_pos):
        self.owner = owner
        self.xt = np.asarray(xt, dtype=np.float64).reshape(2, 3)
        self.pos_old = np.asarray(pos_old, dtype=np.float64)
        self.dual_var_old = np.asarray(dual_var_old, dtype=np.float64)
        self.prev_pred_pos = np.asarray(prev_pred_pos, dtype=np.float64)

    def solve(self, error_on_fail=False):
        import ctypes as C
        import torch
        P = self.owner
        N = int(P.param.num_ho)
        sol = P._descriptor_solver("edge", 2, 1)
        spd = np.asarray(P.param.spd, dtype=np.float64)[:2]
        # the kernel forms the seed point itself (casadi/main.py:48-49); the script's value must be that very point
        seed = np.around(self.xt[:, :2] + float(P.param.dt) * spd[:, None] * np.stack((np.cos(self.xt[:, 2]), np.sin(self.xt[:, 2])), 1), 4)
        if not np.allclose(seed, self.prev_pred_pos.reshape(2, 2), atol=1e-9):
            raise ValueError("prev_pred_pos is not the one-step-ahead point of xt_edge (casadi/main.py:48-49)")
        put = lambda name, val: sol.buf[name].copy_(torch.from_numpy(np.ascontiguousarray(val)).reshape(sol.buf[name].shape))
        put("xt", self.xt); put("spd", spd)
        put("pos", self.pos_old.reshape(2, 2, N + 1))
        put("dual", self.dual_var_old.reshape(2, 1, 2, N + 1))
        sol.buf["rho_s"].fill_(float(P.param.rho))
        put("nbr", np.array([1, 0], dtype=np.int32)); put("edge_n", np.array([1], dtype=np.int32))
        put("edge_ij", np.array([0, 1], dtype=np.int32)); put("edge_slot", np.array([0, 0], dtype=np.int32))
        sol.buf["done"].zero_(); sol.buf["status"].zero_()
        sol._check(sol.lib.piadmm_solve_edges(sol._h, sol._stream()), "piadmm_solve_edges")
        out = np.zeros((1, sol.E, 2, N))
        sol._check(sol.lib.piadmm_get_edge_u(sol._h, out.ctypes.data_as(C.c_void_p), sol._stream()), "piadmm_get_edge_u")
        if error_on_fail and int(sol.buf["status"][0].item()) & _abi.ST_EDGE_MAXITER:
            raise RuntimeError("edge problem did not converge within in_max_iter inner iterations (qpsol error_on_fail)")
        return out[0, 0].reshape(1, 2 * N).copy()


class Constraints:
    """What nonlcon_function returns: g(u) = [u+pi/6; -u+pi/6; -D1u+pi/9; D1u+pi/9] >= 0 (casadi/PI_ADMM_class.py:172-192).
    The kernels carry exactly these rows (csrc/piadmm_dev.cuh); the descriptor lets qpsol check that the problem it is
    given has them."""
    is_steering_box = True

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
        self._desc_solvers = {}

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
        return VehicleQP(self, num_step, xt, hat_pos, dual_var, veh_index)

    def cost_function_edge(self, hat_u, xt, pos_old, dual_var_old, prev_pred_pos):
        """casadi/PI_ADMM_class.py:145-169 -> EdgeQP descriptor (convex statement)."""
        return EdgeQP(self, xt, pos_old, dual_var_old, prev_pred_pos)

    def nonlcon_function(self, u):
        """casadi/PI_ADMM_class.py:172-192: returns (ceq, cieq) with cieq a Constraints descriptor."""
        shape = getattr(u, "shape", (1, self.param.num_ho))
        rows = shape[0] if len(shape) == 2 else 1
        return [], Constraints(rows, self.param.num_ho)

    # ------------------------------------------------------------------ descriptor solves (what ca_shim.qpsol calls)
    def _descriptor_solver(self, kind, V, D) -> PIADMMSolver:
        """Handle for single-problem solves in casadi/main.py's own statement order: the script rounds (:103,153), rolls
        out and updates the duals itself, so rounding is off and the AL term runs over every j != i (:126-129)."""
        p = self.abi_params()
        p.update(round_u=-1, round_edge=-1, al_all=1)
        key = (kind, V, D, p.num_ho, p.iter_num)
        hit = self._desc_solvers.get(kind)
        if hit is None or hit[0] != key:
            if hit is not None:
                hit[1].close()
            hit = (key, PIADMMSolver(p, S=1, V=V, D=D))
            self._desc_solvers[kind] = hit
        else:
            hit[1].set_params(**p.as_dict())
        return hit[1]

    def solve(self, desc, cieq=None, error_on_fail=None):
        """Solve ONE described problem on the GPU (a VehicleQP or an EdgeQP); returns the unrounded solution."""
        if cieq is not None and not getattr(cieq, "is_steering_box", False):
            raise TypeError("cieq must come from nonlcon_function")
        if error_on_fail is None:
            error_on_fail = isinstance(desc, VehicleQP)
        return desc.solve(error_on_fail=error_on_fail)

    def solve_vehicles(self, descs):
        return [self.solve(d) for d in descs]

    def solve_edges(self, descs):
        return [self.solve(d) for d in descs]

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
