"""The C-ABI library loads and exports every symbol include/piadmm.h declares; struct layouts of the
ctypes mirror match the C compiler's; without a GPU the library refuses to work (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

import pi_admm_b200 as pa
from pi_admm_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "piadmm.h")


HEADER_OBCA = os.path.join(ROOT, "include", "obca_consensus.h")


def declared_functions():
    out = set()
    for path, prefix in ((HEADER, "piadmm_"), (HEADER_OBCA, "obca_")):
        src = open(path).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        out |= set(re.findall(r"\b(%s[a-z0-9_]+)\s*\(" % prefix, src))
    return sorted(out)


def test_header_symbols_are_exported_and_bound():
    lib = pa.load_library()
    names = declared_functions()
    assert len(names) >= 23 and "obca_consensus_tail" in names
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/piadmm.h but not exported"
    assert set(names) == set(_abi.EXPORTED_SYMBOLS)
    assert lib.piadmm_abi_version() == _abi.ABI_VERSION


def test_struct_layouts_match_the_c_compiler(tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "piadmm.h"\nint main(void){printf("%zu %zu %zu %zu\\n",'
                   "sizeof(piadmm_params), sizeof(piadmm_buffers), offsetof(piadmm_params, num_ho),"
                   "offsetof(piadmm_buffers, xt_next));return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.check_call(["/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc", "-I", os.path.join(ROOT, "include"),
                           str(src), "-o", str(exe)])
    a, b, c, d = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert a == C.sizeof(_abi.Params) and b == C.sizeof(_abi.Buffers)
    assert c == _abi.Params.num_ho.offset and d == _abi.Buffers.xt_next.offset


def test_obca_struct_layouts_match_the_c_compiler(tmp_path):
    from pi_admm_b200 import obca_consensus as oc
    src = tmp_path / "szo.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "obca_consensus.h"\nint main(void){printf("%zu %zu %zu %zu\\n",'
                   "sizeof(obca_dims), sizeof(obca_state), offsetof(obca_dims, C), offsetof(obca_state, local_x));return 0;}\n")
    exe = tmp_path / "szo"
    subprocess.check_call(["/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc", "-I", os.path.join(ROOT, "include"),
                           str(src), "-o", str(exe)])
    a, b, c, d = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert a == C.sizeof(oc.Dims) and b == C.sizeof(oc.State)
    assert c == oc.Dims.C.offset and d == oc.State.local_x.offset


def test_obca_no_cpu_fallback_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from pi_admm_b200.obca_consensus import OBCAConsensus, Dims, State
    with pytest.raises(pa.PIADMMError):
        OBCAConsensus(S=2)
    lib = pa.load_library()
    d, st = Dims(1, 2, 7, 5, 4, 14), State(*([8] * 6))    # non-NULL dummies: refused before any dereference
    assert lib.obca_lambda_update(C.byref(d), C.byref(st), 1.0, None) == -2     # PIADMM_E_CUDA
    assert lib.obca_lambda_update(None, C.byref(st), 1.0, None) == -1           # PIADMM_E_ARG


def test_default_params_agree_in_c_and_python():
    lib = pa.load_library()
    p = _abi.Params()
    lib.piadmm_default_params(C.byref(p))
    assert p.as_dict() == _abi.default_params().as_dict()
    from oracle import ctwin
    assert ctwin.c_default_params().as_dict() == p.as_dict()
    assert lib.piadmm_max_edges(64, 8) == 256 and lib.piadmm_max_edges(2, 1) == 1


def test_no_cpu_fallback_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = pa.load_library()
    h = C.c_void_p()
    p = _abi.default_params()
    rc = lib.piadmm_create(C.byref(p), 0, 4, 2, 1, C.byref(h))
    assert rc == -2 and not h.value                     # PIADMM_E_CUDA
    assert b"no CPU fallback" in lib.piadmm_last_error(None)
    from pi_admm_b200.solver import PIADMMSolver
    with pytest.raises(pa.PIADMMError):
        PIADMMSolver(p, S=4, V=2)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "distributed-local-planner-pi-admm_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "oracle" not in txt.replace("oracle/solver.py", "").replace("oracle/", "") or f.endswith((".cu", ".cuh")), f
                assert "import oracle" not in txt and "from oracle" not in txt, f


def test_class_mirror_surface():
    """Attribute / method surface of casadi/PI_ADMM_class.py (no GPU needed for the host-side glue)."""
    import numpy as np
    from pi_admm_b200.PI_ADMM_class import PI_ADMM_CASADI
    P = PI_ADMM_CASADI()
    assert P.param.num_ho == 15 and P.param.rho == 2 and P.ref_traj.shape == (4, 50)
    P.param.rho = 3.5                                    # the old script mutates param.rho
    assert P.abi_params().rho == 3.5
    for m in ("dynamic_update_local", "dynamic_update_edge", "cost_function_primal", "cost_function_edge", "nonlcon_function"):
        assert callable(getattr(P, m))
    xt = np.array([[-10, 0, 0], [0, 20, -np.pi / 2]], dtype=float)
    x, y, th = P.dynamic_update_local(xt[1], np.zeros((1, 15)), 1, if_SX=0)
    assert x.shape == (16,) and abs(y[1] - (20 - 0.8)) < 1e-12
    xe, ye, _ = P.dynamic_update_edge(xt, np.zeros((2, 15)), if_SX=0)
    assert xe.shape == (2, 16) and abs(xe[0, 1] - (-9.6)) < 1e-12
    _, c = P.nonlcon_function(np.zeros((1, 15)))
    assert c.m == 58
