"""Host mirror of the consensus pieces of ``OBCAOptimizer`` (BASELINE config 2),
batched over S independent scenarios on one B200.

Same names and argument meaning as the reference
(Distributed_planner/decentralized/optimizer.py): ``create_bar_state`` / ``mid_state`` (:347-373),
``lambda_update(rho)`` (:330-335), ``check_converge(thres)`` (:225-235), ``iterate_next_state(bar_state)``
(:337-344), plus ``residuals`` and ``consensus_tail`` for the statements the driver script writes inline
(Distributed_planner/decentralized_overtaking_ADMM.py:79-97).  The local / RSU problems between them are
IPOPT NLPs and are NOT part of this package.

Every array is a ``torch.float64`` CUDA tensor with a leading scenario axis; the arithmetic runs in
csrc/obca_consensus.cu through the C ABI of include/obca_consensus.h.  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _abi

N_STATES, N_CONTROLS, N_LOC_LAMBDA, N_DUAL = 5, 2, 4, 4      # optimizer.py:15-18
FIELDS = ("Z_bar", "A", "b", "lamb_bar", "lamb_ij", "local_x")

# warm-start multipliers of optimizer.py:357-371 (3-decimal values copied by the reference's author from a
# centralised run; they are input DATA of config 2)
_LAMB_IJ0 = (
    ((1.49, 0.566, 0.566, 1.49), (1.438, 0.514, 0.514, 1.438), (1.387, 0.462, 0.462, 1.387),
     (1.336, 0.411, 0.411, 1.336), (1.287, 0.361, 0.361, 1.287), (1.238, 0.312, 0.312, 1.238),
     (1.191, 0.263, 0.263, 1.191)),
    ((1.436,) * 4, (1.325,) * 4, (1.213,) * 4, (1.1,) * 4, (0.986,) * 4, (0.871,) * 4, (0.755,) * 4))


class Dims(C.Structure):
    """struct obca_dims"""
    _fields_ = [(n, C.c_int32) for n in ("S", "V", "H", "ns", "nl", "C")]


class State(C.Structure):
    """struct obca_state"""
    _fields_ = [(n, C.c_void_p) for n in FIELDS]


class Vehicle(C.Structure):
    """struct obca_vehicle: Distributed_planner/decentralized/veh_config.py:9-27"""
    _fields_ = [(n, C.c_double) for n in ("length", "width", "avg_delay", "var_delay", "prob")]


class mid_state:
    """optimizer.py:351-373 with a leading scenario axis."""

    def __init__(self, outer: "OBCAConsensus"):
        S, V, H, dev = outer.S, outer.num_veh, outer.N_horz - 1, outer.device
        nz = outer.n_states + outer.n_loc_lambda
        kw = dict(dtype=torch.float64, device=dev)
        self.Z_bar = torch.zeros((S, V, H, nz), **kw)
        self.A = torch.zeros((S, V, H, outer.n_dual_variable, 2), **kw)
        self.b = torch.zeros((S, V, H, outer.n_dual_variable), **kw)
        self.lamb_bar = torch.full((S, V, H, nz), 1e-3, **kw)
        if (V, H, outer.n_dual_variable) == (2, 7, 4):
            self.lamb_ij = torch.tensor(_LAMB_IJ0, **kw).expand(S, V, H, 4).contiguous()
        else:
            self.lamb_ij = torch.full((S, V, H, outer.n_dual_variable), 1e-3, **kw)
        self.local_x = torch.zeros((S, V, H, outer.n_states), **kw)

    def clone(self):
        o = object.__new__(mid_state)
        for f in FIELDS:
            setattr(o, f, getattr(self, f).clone())
        return o

    def _c(self) -> State:
        st = State()
        for f in FIELDS:
            t = getattr(self, f)
            if not (t.is_cuda and t.dtype == torch.float64 and t.is_contiguous()):
                raise _abi.PIADMMError(f"bar_state.{f} must be a contiguous float64 CUDA tensor")
            setattr(st, f, t.data_ptr())
        return st


class OBCAConsensus:
    """The ``OBCAOptimizer`` members the ADMM loop touches between the NLP solves."""

    def __init__(self, S=1, num_veh=2, N_horz=8, device=0, primal_thres=0.01, dual_thres=0.01, max_iter=50):
        if not torch.cuda.is_available():
            raise _abi.PIADMMError("OBCAConsensus needs a CUDA device (there is no CPU fallback)")
        self.lib = _abi.load_library()
        self.S, self.num_veh, self.N_horz = int(S), int(num_veh), int(N_horz)
        self.n_states, self.n_controls = N_STATES, N_CONTROLS
        self.n_loc_lambda, self.n_dual_variable = N_LOC_LAMBDA, N_DUAL
        self.primal_thres, self.dual_thres, self.max_iter = primal_thres, dual_thres, max_iter
        self.device = torch.device("cuda", device)
        self.dims = Dims(self.S, self.num_veh, self.N_horz - 1, self.n_states, self.n_loc_lambda,
                         (self.N_horz - 1) * self.n_controls)
        self.bar_state = self.create_bar_state()
        kw = dict(device=self.device)
        self.primal_res = torch.zeros(self.S, dtype=torch.float64, **kw)
        self.dual_res = torch.zeros(self.S, dtype=torch.float64, **kw)
        self.n_iter = torch.zeros(self.S, dtype=torch.int32, **kw)
        self.stop = torch.zeros(self.S, dtype=torch.int32, **kw)
        self.launch_count = 0

    # -- plumbing
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _ck(self, rc, what):
        if rc != 0:
            raise _abi.PIADMMError(f"{what} failed with code {rc}")
        self.launch_count += 1

    # -- reference surface
    def create_bar_state(self) -> mid_state:
        return mid_state(self)

    def bar_state_update(self, bar_fullx_vehs, bar_lambda=None, prob=1, cfg: Vehicle | None = None):
        """optimizer.py:205-222: half-spaces (A, b) of every vehicle rectangle and local_x from the local solutions
        bar_fullx_vehs [S, V, N_horz-1, 9] (the reference passes a list over vehicles of (N_horz-1) x 9 arrays).
        prob=1: the delay-tightened half-spaces the driver runs (util.py:76-101); lamb_ij is not touched (:219-220)."""
        fx = bar_fullx_vehs if torch.is_tensor(bar_fullx_vehs) else torch.as_tensor(bar_fullx_vehs, dtype=torch.float64)
        fx = fx.to(self.device, torch.float64).reshape(self.S, self.num_veh, self.N_horz - 1, self.n_states + self.n_loc_lambda).contiguous()
        cfg = cfg or Vehicle(3.5, 2.0, 0.05, 0.025, 0.95)
        st = self.bar_state._c()
        self._ck(self.lib.obca_bar_state_update(C.byref(self.dims), C.byref(st), C.c_void_p(fx.data_ptr()), C.byref(cfg), int(prob),
                                                self._stream()), "obca_bar_state_update")
        self._fx_keep = fx

    def lambda_update(self, rho):
        st = self.bar_state._c()
        self._ck(self.lib.obca_lambda_update(C.byref(self.dims), C.byref(st), float(rho), self._stream()), "obca_lambda_update")

    def check_converge(self, thres, min_dis=0.1):
        """Per-scenario bool tensor (the reference returns one Python bool for its single scenario)."""
        ok = torch.empty(self.S, dtype=torch.int32, device=self.device)
        st = self.bar_state._c()
        self._ck(self.lib.obca_check_converge(C.byref(self.dims), C.byref(st), float(thres), float(min_dis),
                                              C.c_void_p(ok.data_ptr()), self._stream()), "obca_check_converge")
        return ok.bool()

    def iterate_next_state(self, bar_state: mid_state) -> mid_state:
        st = bar_state._c()
        self._ck(self.lib.obca_iterate_next_state(C.byref(self.dims), C.byref(st), self._stream()), "obca_iterate_next_state")
        return bar_state

    def residuals(self, bar_ctrl, bar_ctrl_prev, bar_state_prev: mid_state):
        """decentralized_overtaking_ADMM.py:82-85; bar_ctrl / bar_ctrl_prev: [S, V, (N_horz-1)*n_controls]."""
        self._check_ctrl(bar_ctrl, bar_ctrl_prev)
        self._ck(self.lib.obca_residuals(C.byref(self.dims), C.c_void_p(bar_ctrl.data_ptr()),
                                         C.c_void_p(bar_ctrl_prev.data_ptr()),
                                         C.c_void_p(self.bar_state.lamb_bar.data_ptr()),
                                         C.c_void_p(bar_state_prev.lamb_bar.data_ptr()),
                                         C.c_void_p(self.primal_res.data_ptr()), C.c_void_p(self.dual_res.data_ptr()),
                                         self._stream()), "obca_residuals")
        return self.primal_res, self.dual_res

    def consensus_tail(self, bar_state_prev: mid_state, bar_ctrl, bar_ctrl_prev, rho=1.0):
        """decentralized_overtaking_ADMM.py:79-97 for every scenario with stop == 0, one kernel:
        lambda update, residuals, stop test, then shift of ``bar_state_prev`` (stopped) or
        ``bar_state_prev <- bar_state``, ``bar_ctrl_prev <- bar_ctrl`` (continuing), all in place."""
        self._check_ctrl(bar_ctrl, bar_ctrl_prev)
        cur, prev = self.bar_state._c(), bar_state_prev._c()
        self._ck(self.lib.obca_consensus_tail(
            C.byref(self.dims), C.byref(cur), C.byref(prev), C.c_void_p(bar_ctrl.data_ptr()),
            C.c_void_p(bar_ctrl_prev.data_ptr()), float(rho), float(self.primal_thres), float(self.dual_thres),
            int(self.max_iter), C.c_void_p(self.primal_res.data_ptr()), C.c_void_p(self.dual_res.data_ptr()),
            C.c_void_p(self.n_iter.data_ptr()), C.c_void_p(self.stop.data_ptr()), self._stream()), "obca_consensus_tail")
        return self.stop

    def begin_time_step(self):
        """decentralized_overtaking_ADMM.py:37-39: fresh iteration counters for the next MPC step."""
        self.n_iter.zero_()
        self.stop.zero_()

    def _check_ctrl(self, *ts):
        shape = (self.S, self.num_veh, self.dims.C)
        for t in ts:
            if tuple(t.shape) != shape or t.dtype != torch.float64 or not t.is_cuda or not t.is_contiguous():
                raise _abi.PIADMMError(f"bar_ctrl tensors must be contiguous float64 CUDA tensors of shape {shape}")
