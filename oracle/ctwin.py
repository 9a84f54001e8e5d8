"""CPU ORACLE (test infrastructure) -- ctypes front-end of the C twin (oracle/piadmm_ref.c).

Allocates the piadmm_buffers set as numpy arrays and runs whole MPC steps on the host cores.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm import this."""
from __future__ import annotations

import ctypes as C
import importlib
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libpiadmm_ref.so")
_abi = importlib.import_module("distributed-local-planner-pi-admm_b200._abi")

_NP = {C.c_double: np.float64, C.c_int32: np.int32, C.c_int64: np.int64}
_lib = None


def build(force: bool = False):
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(os.path.join(_HERE, "piadmm_ref.c")):
        subprocess.check_call(["make", "-C", _HERE, "-s"], stdout=sys.stderr)
    return _LIB


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        _lib = C.CDLL(_LIB)
        _lib.piadmm_ref_step.restype = C.c_int
        _lib.piadmm_ref_step.argtypes = [C.POINTER(_abi.Params), C.POINTER(_abi.Buffers), C.c_int, C.c_int,
                                         C.c_int, C.c_int]
        _lib.piadmm_ref_default_params.argtypes = [C.POINTER(_abi.Params)]
        _lib.piadmm_ref_max_threads.restype = C.c_int
    return _lib


def max_threads() -> int:
    return int(lib().piadmm_ref_max_threads())


def c_default_params():
    p = _abi.Params()
    lib().piadmm_ref_default_params(C.byref(p))
    return p


def alloc(S, V, D, N, I):
    arrs = {}
    for name, (ct, shape) in _abi.buffer_shapes(S, V, D, N, I).items():
        arrs[name] = np.zeros(shape, dtype=_NP[ct])
    return arrs


def step(params, xt, spd, refwin, D=None, rho_s=None, nthreads=0, arrs=None):
    """One MPC step for S scenarios on the CPU.  xt (S,V,3), spd (S,V), refwin (S,V,2,N+1)."""
    xt = np.ascontiguousarray(xt, dtype=np.float64)
    S, V, _ = xt.shape
    N = params.num_ho
    D = (V - 1 if params.al_all else 1) if D is None else D
    if arrs is None:
        arrs = alloc(S, V, D, N, params.iter_num)
    arrs["xt"][...] = xt
    arrs["spd"][...] = spd
    arrs["refwin"][...] = refwin
    arrs["rho_s"][...] = params.rho if rho_s is None else rho_s
    b = _abi.Buffers()
    for name in arrs:
        setattr(b, name, arrs[name].ctypes.data)
    rc = lib().piadmm_ref_step(C.byref(params), C.byref(b), S, V, D, int(nthreads))
    if rc != 0:
        raise RuntimeError(f"piadmm_ref_step failed: {rc}")
    return arrs
