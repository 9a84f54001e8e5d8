"""Stand-ins for the two third-party imports the reference's classes need and this container lacks (`casadi`, `bunch`),
so that the reference's OWN code can be imported from /root/reference and evaluated NUMERICALLY:

* `bunch.Bunch`  -- an attribute dict (what casadi/PI_ADMM_class.py:15-28 uses it for);
* `casadi`       -- numpy-backed versions of the dozen functions the reference's class bodies call (SX.zeros / SX.ones /
                    SX(...), sum1, sum2, sumsqr, transpose, horzcat, vertcat, vcat, fmax, reshape (column major, as CasADi),
                    DM, norm_2, cos / sin / tan / atan).  No symbolic engine, no solver: `qpsol` / `nlpsol` raise.

With them, `cost_function_primal`, `nonlcon_function`, `dynamic_update_local / _edge`, `cost_function_edge` of the reference
evaluate on numeric arrays exactly the arithmetic statements they contain -- the statements are the reference's, only the
array container is numpy.  Used ONLY by tests/golden/make_reference_vectors.py (run in the build container, where
/root/reference exists); the committed fixtures are what the tests read.  Test infrastructure, never a product path."""
import sys
import types

import numpy as np


class Bunch(dict):
    def __init__(self, **kw):
        super().__init__(**kw)
        self.__dict__ = self


class Mat(np.ndarray):
    """CasADi's matrix container on top of numpy: ALWAYS two-dimensional (vectors are columns, scalars 1 x 1), a single
    index is a LINEAR (column-major) index, `A[i, :]` stays a 1 x n row and `A[:, j]` an n x 1 column."""

    def __new__(cls, a):
        a = np.array(a, dtype=np.float64)
        if a.ndim == 0:
            a = a.reshape(1, 1)
        elif a.ndim == 1:
            a = a.reshape(-1, 1)
        return a.view(cls)

    def __getitem__(self, key):
        base = np.asarray(self)
        if not isinstance(key, tuple):
            r = base.reshape(-1, order="F")[key]
            return r if np.ndim(r) == 0 else Mat(r)
        r = base[key]
        if np.ndim(r) == 0:
            return r
        if np.ndim(r) == 1:
            return Mat(r.reshape(1, -1) if isinstance(key[0], (int, np.integer)) else r.reshape(-1, 1))
        return Mat(r)

    def __setitem__(self, key, val):
        base = np.asarray(self)
        if not isinstance(key, tuple) and base.ndim == 2 and base.shape[1] == 1:
            base[key, 0] = np.asarray(val, dtype=np.float64).reshape(-1) if np.ndim(val) else val
            return
        v = np.asarray(val, dtype=np.float64)
        base[key] = v.reshape(base[key].shape) if v.ndim and v.size == np.size(base[key]) else v

    def __array_finalize__(self, obj):
        pass


def _m(x):
    return x if isinstance(x, Mat) else Mat(x)


class SX:
    def __new__(cls, x=0.0):
        return Mat(x)

    @staticmethod
    def _shape(shape):
        if len(shape) == 1 and isinstance(shape[0], tuple):
            shape = shape[0]
        return (shape[0], 1) if len(shape) == 1 else tuple(shape)

    @staticmethod
    def zeros(*shape):
        return Mat(np.zeros(SX._shape(shape)))

    @staticmethod
    def ones(*shape):
        return Mat(np.ones(SX._shape(shape)))

    @staticmethod
    def sym(*a, **k):
        raise RuntimeError("symbolic variables are not available in the numeric stand-in")


def DM(x):
    return Mat(x)


def sum1(x):                       # sum over rows -> 1 x n
    return Mat(np.asarray(_m(x)).sum(axis=0, keepdims=True))


def sum2(x):                       # sum over columns -> n x 1
    return Mat(np.asarray(_m(x)).sum(axis=1, keepdims=True))


def sumsqr(x):
    return float(np.sum(np.asarray(x, dtype=np.float64) ** 2))


def transpose(x):
    return Mat(np.asarray(_m(x)).T)


def horzcat(*xs):
    return Mat(np.hstack([np.asarray(_m(x)) for x in xs]))


def vertcat(*xs):
    return Mat(np.vstack([np.asarray(_m(x)) for x in xs]))


def vcat(xs):
    return vertcat(*xs)


def fmax(a, b):
    return Mat(np.maximum(np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)))


def reshape(x, r, c):
    return Mat(np.reshape(np.asarray(_m(x)), (r, c), order="F"))


def norm_2(x):
    return float(np.sqrt(np.sum(np.asarray(x, dtype=np.float64) ** 2)))


def _no_solver(*a, **k):
    raise RuntimeError("no solver in the numeric stand-in (CasADi / OSQP / IPOPT are not installed here)")


def install():
    ca = types.ModuleType("casadi")
    for name, obj in dict(SX=SX, DM=DM, sum1=sum1, sum2=sum2, sumsqr=sumsqr, transpose=transpose, horzcat=horzcat, vertcat=vertcat,
                          vcat=vcat, fmax=fmax, reshape=reshape, norm_2=norm_2, cos=np.cos, sin=np.sin, tan=np.tan, atan=np.arctan,
                          qpsol=_no_solver, nlpsol=_no_solver, Function=_no_solver).items():
        setattr(ca, name, obj)
    b = types.ModuleType("bunch")
    b.Bunch = Bunch
    sys.modules["casadi"] = ca
    sys.modules["bunch"] = b
    return ca, b
