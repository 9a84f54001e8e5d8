"""ctypes mirror of include/piadmm.h (struct layouts, enums, prototypes) and the loader of
libpiadmm.so.  There is no CPU fallback: ``load_library()`` raises if the CUDA extension has not
been built, and every compute call raises ``PIADMMError`` on a non-zero return code."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PIADMM_LIB") or os.path.join(_HERE, "csrc", "libpiadmm.so")   # PIADMM_LIB: A/B builds

ABI_VERSION = 3
DUAL_PLAIN, DUAL_PI, DUAL_PI_OLD = 0, 1, 2
PEER_HANDLE_BYTES = 64
ST_VEH_MAXITER, ST_EDGE_MAXITER, ST_SLOT_OVERFLOW, ST_OUTER_MAXITER, ST_PEER_TIMEOUT = 1, 2, 4, 8, 16

_D = C.c_double
_I = C.c_int32


class Params(C.Structure):
    """struct piadmm_params (include/piadmm.h)."""
    _fields_ = [(n, _D) for n in (
        "dt", "L", "dis_thres", "beta", "Pnorm", "Pcost", "rho", "eps_pri", "eps_dual",
        "theta1", "theta2", "kI", "windup_sat", "backcalc",
        "K_I", "KP_num", "KP_max", "rho_num", "rho_min", "rho_max",
        "in_rho0", "in_sigma", "in_alpha", "in_eps_abs", "in_eps_rel", "in_adapt_tol")] + [(n, _I) for n in (
        "num_ho", "iter_num", "dual_mode", "thr_squared", "dual_res_alias", "stop_dist_test",
        "pos_nonlinear", "curr_iter_latest", "round_u", "round_edge", "al_all", "always_edge",
        "residual_global", "PI", "windup", "trad", "in_max_iter", "in_check", "in_adaptive", "x_memo")]

    def copy(self):
        o = Params()
        C.memmove(C.byref(o), C.byref(self), C.sizeof(Params))
        return o

    def update(self, **kw):
        for k, v in kw.items():
            if k not in PARAM_FIELDS:
                raise AttributeError(f"piadmm_params has no field {k!r}")
            setattr(self, k, v)
        return self

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


PARAM_FIELDS = tuple(n for n, _ in Params._fields_)

# name -> (ctype, shape expression in S,V,D,N,I,E)
BUFFER_SPECS = (
    ("xt", _D, "S,V,3"), ("spd", _D, "S,V"), ("refwin", _D, "S,V,2,N+1"),
    ("u", _D, "S,V,N"), ("pos", _D, "S,V,2,N+1"),
    ("nbr", _I, "S,V,D"),
    ("hat", _D, "S,V,D,2,N+1"), ("dual", _D, "S,V,D,2,N+1"), ("integ", _D, "S,V,D,2,N+1"),
    ("diff", _D, "S,V,D,2,N+1"), ("last_hat", _D, "S,V,D,2,N+1"),
    ("rho_s", _D, "S"), ("flag", _I, "S"), ("done", _I, "S"), ("aliased", _I, "S"),
    ("curr_iter", _I, "S"), ("n_outer", _I, "S"), ("status", _I, "S"),
    ("err_rk", _D, "S,I"), ("err_sk", _D, "S,I"),
    ("veh_inner", C.c_int64, "S"), ("edge_inner", C.c_int64, "S"),
    ("n_veh_solves", _I, "S"), ("n_edge_solves", _I, "S"),
    ("edge_n", _I, "S"), ("edge_ij", _I, "S,E,2"), ("edge_slot", _I, "S,E,2"),
    ("res_part", _D, "S,2"),
    ("xt_next", _D, "S,V,3"),
)


class Buffers(C.Structure):
    """struct piadmm_buffers (include/piadmm.h)."""
    _fields_ = [(n, C.c_void_p) for n, _, _ in BUFFER_SPECS]


class Histories(C.Structure):
    """struct piadmm_histories (device pointers, any may be NULL)."""
    _fields_ = [(n, C.c_void_p) for n in ("x_vec", "theta_vec", "u_vec", "err_rk_his", "iter_his", "n_outer_his",
                                            "status_his")]


def max_edges(V: int, D: int) -> int:
    return max(1, V * D // 2)


def buffer_shapes(S, V, D, N, I):
    E = max_edges(V, D)
    env = dict(S=S, V=V, D=D, N=N, I=I, E=E)
    out = {}
    for name, ct, expr in BUFFER_SPECS:
        shape = tuple(int(eval(tok, {}, env)) for tok in expr.split(","))
        out[name] = (ct, shape)
    return out


def default_params() -> Params:
    """casadi/PI_ADMM_class.py:15-28 + the semantics of casadi/main.py as written
    (same values as piadmm_default_params in csrc/piadmm.cu)."""
    p = Params()
    p.update(dt=0.1, L=1.0, dis_thres=2.0, beta=10.0, Pnorm=1.0, Pcost=1.0, rho=2.0, eps_pri=20.0,
             eps_dual=1.0, theta1=5.0, theta2=3.0, kI=-1.0, windup_sat=30.0, backcalc=1.0,
             K_I=3.0, KP_num=5.0, KP_max=2.5, rho_num=4.0, rho_min=1.0, rho_max=5.0,
             in_rho0=0.1, in_sigma=1e-6, in_alpha=1.6, in_eps_abs=1e-3, in_eps_rel=1e-3, in_adapt_tol=5.0,
             num_ho=15, iter_num=100, dual_mode=DUAL_PLAIN, thr_squared=0, dual_res_alias=1,
             stop_dist_test=0, pos_nonlinear=0, curr_iter_latest=0, round_u=4, round_edge=4, al_all=1,
             always_edge=0, residual_global=0, PI=1, windup=1, trad=0, in_max_iter=4000, in_check=25,
             in_adaptive=1, x_memo=1)
    return p


# ---- semantic presets ----
def preset(name: str, **kw) -> Params:
    p = default_params()
    if name in ("python_current", "plain"):
        pass
    elif name in ("matlab_pi", "pi"):
        p.update(dual_mode=DUAL_PI, thr_squared=1, dual_res_alias=0, stop_dist_test=1, pos_nonlinear=1,
                 curr_iter_latest=1, round_u=-1, round_edge=-1, windup_sat=30.0, backcalc=1.0)
    elif name in ("python_old_pi", "pi_old"):
        p.update(dual_mode=DUAL_PI_OLD, always_edge=1, stop_dist_test=1, residual_global=1,
                 windup_sat=20.0, backcalc=2.0)
    else:
        raise ValueError(name)
    p.update(**kw)
    return p


def inner_tight(p: Params) -> Params:
    return p.update(in_eps_abs=1e-9, in_eps_rel=1e-9, in_max_iter=20000)


class PIADMMError(RuntimeError):
    pass


_PROTOS = {
    "piadmm_abi_version": (C.c_int, []),
    "piadmm_default_params": (None, [C.POINTER(Params)]),
    "piadmm_max_edges": (C.c_int, [C.c_int, C.c_int]),
    "piadmm_create": (C.c_int, [C.POINTER(Params), C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "piadmm_set_params": (C.c_int, [C.c_void_p, C.POINTER(Params)]),
    "piadmm_bind": (C.c_int, [C.c_void_p, C.POINTER(Buffers)]),
    "piadmm_destroy": (None, [C.c_void_p]),
    "piadmm_last_error": (C.c_char_p, [C.c_void_p]),
    "piadmm_begin_step": (C.c_int, [C.c_void_p, C.c_void_p]),
    "piadmm_x_update": (C.c_int, [C.c_void_p, C.c_void_p]),
    "piadmm_detect_edges": (C.c_int, [C.c_void_p, C.c_void_p]),
    "piadmm_z_update": (C.c_int, [C.c_void_p, C.c_void_p]),
    "piadmm_residuals": (C.c_int, [C.c_void_p, C.c_void_p]),
    "piadmm_solve_edges": (C.c_int, [C.c_void_p, C.c_void_p]),
    "piadmm_get_edge_u": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "piadmm_solve_joint": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p]),
    "piadmm_set_ownership": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "piadmm_residual_partials": (C.c_int, [C.c_void_p, C.c_void_p]),
    "piadmm_residual_decide": (C.c_int, [C.c_void_p, C.c_void_p]),
    "piadmm_num_active": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32)]),
    "piadmm_step": (C.c_int, [C.c_void_p, C.c_void_p]),
    "piadmm_peer_export": (C.c_int, [C.c_void_p, C.c_void_p]),
    "piadmm_peer_connect": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "piadmm_peer_disconnect": (C.c_int, [C.c_void_p]),
    "piadmm_plant_step": (C.c_int, [C.c_void_p, C.c_void_p]),
    "piadmm_step_host": (C.c_int, [C.c_void_p] + [C.c_void_p] * 8 + [C.c_void_p]),
    "piadmm_set_reference": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "piadmm_step_host_at": (C.c_int, [C.c_void_p, C.c_int] + [C.c_void_p] * 6 + [C.c_void_p]),
    "piadmm_run_closed_loop": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "piadmm_launch_count": (C.c_int64, [C.c_void_p]),
    "piadmm_debug_coop_trace": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "piadmm_fp64_peak": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    # allocation-free collision-edge test (csrc/piadmm_static.cu; no reference counterpart, operators' request)
    "piadmm_edge_list_static": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int,
                                          C.c_void_p, C.c_void_p, C.c_void_p]),
    "piadmm_static_last_error": (C.c_char_p, []),
    # include/obca_consensus.h (struct pointers passed as void*: obca_consensus.py owns the ctypes mirrors)
    "obca_bar_state_update": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "obca_lambda_update": (C.c_int, [C.c_void_p, C.c_void_p, C.c_double, C.c_void_p]),
    "obca_residuals": (C.c_int, [C.c_void_p] * 7 + [C.c_void_p]),
    "obca_check_converge": (C.c_int, [C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_void_p, C.c_void_p]),
    "obca_iterate_next_state": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "obca_consensus_tail": (C.c_int, [C.c_void_p] * 5 + [C.c_double] * 3 + [C.c_int32] + [C.c_void_p] * 5),
}
EXPORTED_SYMBOLS = tuple(_PROTOS)

_lib = None


def load_library(path: str | None = None):
    """dlopen libpiadmm.so and attach prototypes.  Raises (never falls back) when missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise PIADMMError(
            f"{path} not found: build the CUDA extension first (python -c 'import __graft_entry__ as g; "
            "g.build()').  There is no CPU fallback.")
    lib = C.CDLL(path)
    for name, (res, args) in _PROTOS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.piadmm_abi_version() != ABI_VERSION:
        raise PIADMMError("libpiadmm.so ABI version mismatch")
    _lib = lib
    return lib
