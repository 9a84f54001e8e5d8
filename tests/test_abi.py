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


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(piadmm_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported_and_bound():
    lib = pa.load_library()
    names = declared_functions()
    assert len(names) >= 18
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
