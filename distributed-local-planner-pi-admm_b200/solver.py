"""Batched PI-ADMM solver: PyTorch owns the fp64 device buffers, libpiadmm.so (hand-written
sm_100a CUDA behind the C ABI of include/piadmm.h) does all the arithmetic.

This is the batched form of the loop body of the reference's casadi/main.py:46-192:
``setup`` = the per-step inputs (xt, reference window), ``iterate`` = the outer ADMM iterations
(x-update QPs, collision test, edge problems, dual/PI update, residuals), ``residuals`` and
``trajectories`` = the outputs the script keeps (err_rk_vec, iter_his, primal_u, pos_old).

There is no CPU path: constructing a solver without a CUDA device raises.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _abi

_TORCH_DT = {C.c_double: torch.float64, C.c_int32: torch.int32, C.c_int64: torch.int64}


class PIADMMSolver:
    def __init__(self, params: _abi.Params | None = None, S: int = 1, V: int = 2, D: int | None = None,
                 device: int | torch.device | None = None):
        self.lib = _abi.load_library()
        if not torch.cuda.is_available():
            raise _abi.PIADMMError("PIADMMSolver needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.params = (params or _abi.default_params()).copy()
        if device is None:
            device = torch.cuda.current_device()
        self.device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        self.S, self.V = int(S), int(V)
        self.D = int(D) if D is not None else max(1, (self.V - 1) if self.params.al_all else 1)
        self.N, self.I = int(self.params.num_ho), int(self.params.iter_num)
        self.E = _abi.max_edges(self.V, self.D)
        self._h = C.c_void_p()
        rc = self.lib.piadmm_create(C.byref(self.params), self.device.index or 0, self.S, self.V, self.D,
                                    C.byref(self._h))
        if rc != 0:
            raise _abi.PIADMMError(f"piadmm_create failed ({rc}): {self.lib.piadmm_last_error(None).decode()}")
        self.buf = {}
        b = _abi.Buffers()
        for name, (ct, shape) in _abi.buffer_shapes(self.S, self.V, self.D, self.N, self.I).items():
            t = torch.zeros(shape, dtype=_TORCH_DT[ct], device=self.device)
            self.buf[name] = t
            setattr(b, name, t.data_ptr())
        self.buf["rho_s"].fill_(self.params.rho)
        self._buffers = b
        self._check(self.lib.piadmm_bind(self._h, C.byref(b)), "piadmm_bind")

    # ------------------------------------------------------------------ plumbing
    def _check(self, rc, what):
        if rc != 0:
            raise _abi.PIADMMError(f"{what} failed ({rc}): {self.lib.piadmm_last_error(self._h).decode()}")

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self.lib.piadmm_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_params(self, **kw):
        self.params.update(**kw)
        self._check(self.lib.piadmm_set_params(self._h, C.byref(self.params)), "piadmm_set_params")

    @property
    def launch_count(self) -> int:
        return int(self.lib.piadmm_launch_count(self._h))

    # ------------------------------------------------------------------ per-step API
    def setup(self, xt, spd, refwin, rho=None):
        """Inputs of one MPC step.  xt (S,V,3), spd (S,V) or (V,), refwin (S,V,2,N+1); torch or numpy."""
        def put(name, val):
            t = torch.as_tensor(val, dtype=torch.float64)
            self.buf[name].copy_(t.to(self.device, non_blocking=True).expand_as(self.buf[name]))
        put("xt", xt)
        put("spd", spd)
        put("refwin", refwin)
        if rho is not None:
            put("rho_s", rho)

    def begin_step(self):
        self._check(self.lib.piadmm_begin_step(self._h, self._stream()), "piadmm_begin_step")

    def x_update(self):
        self._check(self.lib.piadmm_x_update(self._h, self._stream()), "piadmm_x_update")

    def detect_edges(self):
        self._check(self.lib.piadmm_detect_edges(self._h, self._stream()), "piadmm_detect_edges")

    def z_update(self):
        self._check(self.lib.piadmm_z_update(self._h, self._stream()), "piadmm_z_update")

    def update_residuals(self):
        self._check(self.lib.piadmm_residuals(self._h, self._stream()), "piadmm_residuals")

    def solve_joint(self, pair_n, pair_ij, hinge_weight=None):
        """ONE joint RSU edge problem per scenario over all V vehicles (piadmm_solve_joint; cost_function_edge of
        matlab_old_files/ADMM_fmincon_multi_veh_intesection_static.m:250-291 in its convex statement).  Uses the bound xt,
        spd, pos_old, rho_s and neighbour slot 0 of dual as the vehicles' ONE multiplier.  pair_n (S,), pair_ij (S,P,2) with
        veh_lo < veh_hi, at most 8 pairs per vehicle; hinge_weight defaults to 2*beta (the reference's ordered-pair sum).
        Returns device tensors u (S,V,N), iters (S,), status (S,)."""
        dev = self.device
        pn = torch.as_tensor(pair_n, dtype=torch.int32).to(dev).contiguous().reshape(self.S)
        pij = torch.as_tensor(pair_ij, dtype=torch.int32).to(dev).contiguous()
        if pij.dim() != 3 or pij.shape[0] != self.S or pij.shape[2] != 2 or pij.shape[1] < 1:
            raise _abi.PIADMMError(f"pair_ij must have shape ({self.S},P,2) with P >= 1, got {tuple(pij.shape)}")
        if int(pn.max()) > pij.shape[1]:
            raise _abi.PIADMMError("pair_n exceeds the pair list")
        u = torch.zeros((self.S, self.V, self.N), dtype=torch.float64, device=dev)
        iters = torch.zeros(self.S, dtype=torch.int32, device=dev)
        status = torch.zeros(self.S, dtype=torch.int32, device=dev)
        w = 2.0 * float(self.params.beta) if hinge_weight is None else float(hinge_weight)
        self._check(self.lib.piadmm_solve_joint(self._h, C.c_void_p(pn.data_ptr()), C.c_void_p(pij.data_ptr()),
                                                int(pij.shape[1]), w, C.c_void_p(u.data_ptr()),
                                                C.c_void_p(iters.data_ptr()), C.c_void_p(status.data_ptr()),
                                                self._stream()), "piadmm_solve_joint")
        self._joint_keep = (pn, pij)      # the launch is asynchronous: keep the inputs alive
        return u, iters, status

    def set_ownership(self, lo: int, hi: int):
        """Vehicle sharding: this instance x-updates vehicles [lo, hi) only (see parallel.py)."""
        self._check(self.lib.piadmm_set_ownership(self._h, int(lo), int(hi)), "piadmm_set_ownership")
        self.own = (int(lo), int(hi))

    def peer_export(self) -> bytes:
        """CUDA IPC handle of this rank's exchange block (piadmm_peer_export)."""
        buf = C.create_string_buffer(_abi.PEER_HANDLE_BYTES)
        self._check(self.lib.piadmm_peer_export(self._h, buf), "piadmm_peer_export")
        return buf.raw

    def peer_connect(self, rank: int, world: int, handles: bytes):
        """Map the exchange blocks of all ranks (`handles` = world x PEER_HANDLE_BYTES, rank order)."""
        assert len(handles) == world * _abi.PEER_HANDLE_BYTES
        self._check(self.lib.piadmm_peer_connect(self._h, int(rank), int(world), C.c_char_p(handles)),
                    "piadmm_peer_connect")

    def peer_disconnect(self):
        self._check(self.lib.piadmm_peer_disconnect(self._h), "piadmm_peer_disconnect")

    def coop_trace(self):
        """Phase timeline (ns, relative to the kernel start) of the last cooperative launch; the first call enables it."""
        buf = np.zeros(1024, dtype=np.int64)
        self._check(self.lib.piadmm_debug_coop_trace(self._h, buf.ctypes.data_as(C.c_void_p), 1024), "piadmm_debug_coop_trace")
        n = int(buf[0])
        return (buf[1:1 + n] - buf[1]) if n else buf[:0]

    def residual_partials(self):
        self._check(self.lib.piadmm_residual_partials(self._h, self._stream()), "piadmm_residual_partials")

    def residual_decide(self):
        self._check(self.lib.piadmm_residual_decide(self._h, self._stream()), "piadmm_residual_decide")

    def num_active(self) -> int:
        out = C.c_int32(0)
        self._check(self.lib.piadmm_num_active(self._h, self._stream(), C.byref(out)), "piadmm_num_active")
        return int(out.value)

    def iterate(self, fused: bool | None = None):
        """Run the whole outer loop of one MPC step (begin_step included).

        fused=None picks the persistent single-kernel path when the scenario shape allows it
        (V == 2, D == 1) and the phase kernels otherwise; fused=False forces the phase kernels."""
        if fused is None or fused:
            if fused and not (self.V == 2 and self.D == 1):
                raise _abi.PIADMMError("the fused persistent kernel needs V == 2 and D == 1")
            self._check(self.lib.piadmm_step(self._h, self._stream()), "piadmm_step")
            return
        self.begin_step()
        for _ in range(self.I):
            self.x_update()
            self.detect_edges()
            self.z_update()
            self.update_residuals()
            if self.num_active() == 0:
                break

    def plant_step(self):
        self._check(self.lib.piadmm_plant_step(self._h, self._stream()), "piadmm_plant_step")

    def step_host(self, xt, spd, refwin):
        """Host-buffer entry point (numpy in, numpy out; copies happen inside the C call)."""
        xt = np.ascontiguousarray(xt, dtype=np.float64).reshape(self.S, self.V, 3)
        spd = np.ascontiguousarray(np.broadcast_to(spd, (self.S, self.V)), dtype=np.float64)
        refwin = np.ascontiguousarray(refwin, dtype=np.float64).reshape(self.S, self.V, 2, self.N + 1)
        out = dict(u=np.empty((self.S, self.V, self.N)), pos=np.empty((self.S, self.V, 2, self.N + 1)),
                   xt_next=np.empty((self.S, self.V, 3)), curr_iter=np.empty(self.S, dtype=np.int32),
                   n_outer=np.empty(self.S, dtype=np.int32))
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        self._check(self.lib.piadmm_step_host(self._h, p(xt), p(spd), p(refwin), p(out["u"]), p(out["pos"]),
                                              p(out["xt_next"]), p(out["curr_iter"]), p(out["n_outer"]),
                                              self._stream()), "piadmm_step_host")
        return out

    def alloc_host_io(self):
        """Pinned host tensors for the host-buffer entry point (inputs and outputs of one step)."""
        S, V, N = self.S, self.V, self.N
        mk = lambda shape, dt=torch.float64: torch.zeros(shape, dtype=dt).pin_memory()
        return dict(xt=mk((S, V, 3)), spd=mk((S, V)), refwin=mk((S, V, 2, N + 1)), u=mk((S, V, N)),
                    pos=mk((S, V, 2, N + 1)), xt_next=mk((S, V, 3)), curr_iter=mk((S,), torch.int32),
                    n_outer=mk((S,), torch.int32))

    def step_host_tensors(self, io):
        """piadmm_step_host on (pinned) host tensors from alloc_host_io(): H2D of the inputs, the whole
        MPC step, D2H of u / pos / xt_next / curr_iter / n_outer, all inside the one C call."""
        p = lambda t: C.c_void_p(t.data_ptr())
        self._check(self.lib.piadmm_step_host(self._h, p(io["xt"]), p(io["spd"]), p(io["refwin"]), p(io["u"]),
                                              p(io["pos"]), p(io["xt_next"]), p(io["curr_iter"]),
                                              p(io["n_outer"]), self._stream()), "piadmm_step_host")

    def set_reference(self, ref):
        """Resident reference trajectories (S,V,2,T) -- numpy or torch, host or device: uploaded once, like
        ``PI_ADMM_CASADI.ref_traj`` (casadi/PI_ADMM_class.py:31-37); see piadmm_set_reference."""
        if not torch.is_tensor(ref):
            ref = torch.from_numpy(np.ascontiguousarray(ref, dtype=np.float64))
        ref = ref.to(torch.float64).contiguous()
        if ref.dim() != 4 or tuple(ref.shape[:3]) != (self.S, self.V, 2):
            raise _abi.PIADMMError(f"reference must have shape ({self.S},{self.V},2,T), got {tuple(ref.shape)}")
        self._check(self.lib.piadmm_set_reference(self._h, C.c_void_p(ref.data_ptr()), int(ref.shape[3]),
                                                  self._stream()), "piadmm_set_reference")

    def step_host_at(self, num_step, io):
        """piadmm_step_host_at on the (pinned) host tensors of alloc_host_io(): per step only io["xt"] goes to
        the device (the reference window is gathered there from the resident reference, spd is resident);
        u / pos / xt_next / curr_iter / n_outer come back, all inside the one C call."""
        p = lambda t: C.c_void_p(t.data_ptr())
        self._check(self.lib.piadmm_step_host_at(self._h, int(num_step), p(io["xt"]), p(io["u"]), p(io["pos"]),
                                                 p(io["xt_next"]), p(io["curr_iter"]), p(io["n_outer"]),
                                                 self._stream()), "piadmm_step_host_at")

    def run_closed_loop(self, first_step, n_steps):
        """n_steps receding-horizon steps of casadi/main.py:43-201 on the device, no host round trip per step
        (needs set_reference; starts from the plant state given to setup()).  Returns the script's histories as
        device tensors: x_vec (n,S,V,2), theta_vec (n,S,V), u_vec (n,S,V), iter_his (n,S), n_outer_his (n,S),
        err_rk_his (n,S)."""
        n, S, V, dev = int(n_steps), self.S, self.V, self.device
        f64 = lambda *sh: torch.zeros(sh, dtype=torch.float64, device=dev)
        i32 = lambda *sh: torch.zeros(sh, dtype=torch.int32, device=dev)
        out = dict(x_vec=f64(n, S, V, 2), theta_vec=f64(n, S, V), u_vec=f64(n, S, V), err_rk_his=f64(n, S),
                   iter_his=i32(n, S), n_outer_his=i32(n, S), status_his=i32(n, S))
        hs = _abi.Histories(**{k: v.data_ptr() for k, v in out.items()})
        self._check(self.lib.piadmm_run_closed_loop(self._h, int(first_step), n, C.byref(hs), self._stream()),
                    "piadmm_run_closed_loop")
        return out

    @staticmethod
    def host_io_bytes(io, resident_reference=False):
        nb = lambda k: io[k].numel() * io[k].element_size()
        return (sum(nb(k) for k in (("xt",) if resident_reference else ("xt", "spd", "refwin"))) + (4 if resident_reference else 0),
                sum(nb(k) for k in ("u", "pos", "xt_next", "curr_iter", "n_outer")))

    # ------------------------------------------------------------------ outputs
    def residuals(self):
        """(err_rk, err_sk) histories (S, iter_num) and the per-scenario iteration counters."""
        return dict(err_rk=self.buf["err_rk"], err_sk=self.buf["err_sk"], curr_iter=self.buf["curr_iter"],
                    n_outer=self.buf["n_outer"], status=self.buf["status"])

    def trajectories(self):
        """primal_u (S,V,N), pos_old (S,V,2,N+1), hat/dual per neighbour slot, next plant state."""
        return dict(u=self.buf["u"], pos=self.buf["pos"], hat=self.buf["hat"], dual=self.buf["dual"],
                    nbr=self.buf["nbr"], xt_next=self.buf["xt_next"])

    def to_numpy(self):
        torch.cuda.synchronize(self.device)
        return {k: v.detach().cpu().numpy() for k, v in self.buf.items()}
