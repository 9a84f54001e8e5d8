"""Shared helpers of the parity tests."""
import numpy as np

import pi_admm_b200 as pa

REL_TOL = 1e-6   # north_star: "within 1e-6 relative in fp64"


def make_params(preset="python_current", inner="default", **kw):
    p = pa.preset(preset)
    if preset == "matlab_pi":          # constants of the MATLAB script :6-25
        p.update(beta=1000.0, Pnorm=5.0, rho=3.5, eps_pri=0.1, eps_dual=0.1, num_ho=8)
    elif preset == "python_old_pi":    # constants of casadi_old_PI_ADMM/PI_ADMM_class.py:15-28
        p.update(rho=1.0, eps_pri=1.0, eps_dual=1.0, num_ho=5, beta=10000.0, Pnorm=5.0, dis_thres=1.5)
    if inner == "tight":
        pa.inner_tight(p)
    p.update(**kw)
    return p


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(1.0, np.abs(b).max())) if a.size else 0.0


INT_KEYS = ("n_outer", "curr_iter", "flag", "status", "n_veh_solves", "n_edge_solves", "edge_n", "nbr")
INNER_KEYS = ("veh_inner", "edge_inner")
FLT_KEYS = ("u", "pos", "hat", "dual", "integ", "diff", "err_rk", "err_sk", "rho_s", "xt_next")


def assert_same(got, ref, tol=REL_TOL, what="", inner_exact=True):
    """Integer outputs identical, floats within tol.  The inner-ADMM iteration totals are diagnostics:
    with the 1e-9 ("tight") inner tolerance the termination test sits at the fp64 noise floor of the
    residual, so a solve may stop one check (25 iterations) earlier or later on the GPU than on the
    CPU without changing the result; then only a small fraction of scenarios may differ."""
    for k in INNER_KEYS:
        ne = np.asarray(got[k]) != np.asarray(ref[k])
        if inner_exact:
            assert not ne.any(), f"{what}: {k} differs in scenarios {np.flatnonzero(ne)[:8]}"
        else:
            assert ne.mean() <= 0.1, f"{what}: {k} differs in {ne.mean():.1%} of the scenarios"
    for k in INT_KEYS:
        assert np.array_equal(got[k], ref[k]), f"{what}: integer field {k} differs: " \
            f"{np.flatnonzero((np.asarray(got[k]) != np.asarray(ref[k])).reshape(len(got[k]), -1).any(1))[:8]}"
    for k in FLT_KEYS:
        e = rel_err(got[k], ref[k])
        assert e <= tol, f"{what}: {k} differs, rel err {e:.3e}"
