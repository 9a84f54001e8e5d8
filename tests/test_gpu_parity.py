"""GPU parity: the CUDA path (through the C ABI) against the oracle's C twin on identical seeded
inputs.  Integer outputs (iteration counts, edge lists, flags) must be identical, floating point
within 1e-6 relative."""
import numpy as np
import pytest
import torch

import pi_admm_b200 as pa
from pi_admm_b200.scenarios import c4_scenarios, c1_initial_state, reference_trajectories, ref_window
from pi_admm_b200.solver import PIADMMSolver
from oracle import ctwin
from util import make_params, assert_same, rel_err

pytestmark = pytest.mark.gpu

CASES = [
    ("python_current", "default", {}),
    ("python_current", "tight", {}),
    ("python_current", "default", dict(num_ho=30)),
    ("python_current", "tight", dict(num_ho=30)),
    ("python_current", "default", dict(num_ho=20, eps_pri=1.0, dual_res_alias=0)),
    ("matlab_pi", "default", {}),
    ("matlab_pi", "tight", {}),
    ("matlab_pi", "default", dict(num_ho=15, eps_pri=1.0, eps_dual=1.0)),
    ("matlab_pi", "default", dict(num_ho=30, eps_pri=1.0, eps_dual=1.0, iter_num=30)),
    ("matlab_pi", "default", dict(PI=0)),
    ("matlab_pi", "default", dict(windup=0)),
    ("python_old_pi", "default", dict(iter_num=20)),
    ("python_old_pi", "default", dict(iter_num=20, trad=1, round_edge=4, num_ho=15)),
    ("python_old_pi", "tight", dict(iter_num=10, num_ho=24)),
]


def _run_gpu(p, xt, spd, refwin, fused):
    S = xt.shape[0]
    sol = PIADMMSolver(p, S=S, V=2)
    sol.setup(xt, spd, refwin)
    sol.iterate(fused=fused)
    sol.plant_step()
    out = sol.to_numpy()
    out["_launches"] = sol.launch_count
    sol.close()
    return out


@pytest.mark.parametrize("preset,inner,kw", CASES)
@pytest.mark.parametrize("fused", [True, False])
def test_step_matches_oracle(preset, inner, kw, fused):
    p = make_params(preset, inner, **kw)
    xt, spd, refwin = c4_scenarios(S=96, N=p.num_ho, seed=7)
    ref = ctwin.step(p, xt, spd, refwin)
    got = _run_gpu(p, xt, spd, refwin, fused)
    assert got["_launches"] > 0
    assert_same(got, ref, what=f"{preset}/{inner}/{kw}/fused={fused}", inner_exact=(inner != "tight"))
    assert ref["n_edge_solves"].sum() > 0      # the case exercises the edge step


def test_step_host_entry_point():
    p = make_params("python_current", "default")
    xt, spd, refwin = c4_scenarios(S=33, N=p.num_ho, seed=3)
    ref = ctwin.step(p, xt, spd, refwin)
    sol = PIADMMSolver(p, S=33, V=2)
    out = sol.step_host(xt, spd, refwin)
    assert np.array_equal(out["n_outer"], ref["n_outer"])
    assert np.array_equal(out["curr_iter"], ref["curr_iter"])
    for k in ("u", "pos", "xt_next"):
        assert rel_err(out[k], ref[k]) <= 1e-6


@pytest.mark.parametrize("preset,inner", [("python_current", "default"), ("python_current", "tight"),
                                          ("matlab_pi", "default"), ("python_old_pi", "default")])
def test_closed_loop_config1(preset, inner):
    """casadi/main.py:43-201 for all 35 (50 - num_ho) receding-horizon steps, GPU vs C twin."""
    p = make_params(preset, inner)
    N = p.num_ho
    ref_traj = reference_trajectories()
    spd = np.array([[4.0, 8.0]])
    xt_g = c1_initial_state()[None].copy(); xt_c = xt_g.copy()
    rho_g = rho_c = p.rho
    sol = PIADMMSolver(p, S=1, V=2)
    for step in range(ref_traj.shape[1] - N):
        rw = ref_window(ref_traj, step, N)[None]
        ref = ctwin.step(p, xt_c, spd, rw, rho_s=rho_c)
        sol.setup(xt_g, spd, rw, rho=rho_g)
        sol.iterate()
        sol.plant_step()
        got = sol.to_numpy()
        assert_same(got, ref, what=f"{preset}/{inner} step {step}", inner_exact=(inner != "tight"))
        xt_g = got["xt_next"].copy(); xt_c = ref["xt_next"].copy()
        if p.dual_mode == pa.DUAL_PI_OLD:
            rho_g = float(got["rho_s"][0]); rho_c = float(ref["rho_s"][0])


def test_no_collision_breaks_after_one_iteration():
    p = make_params("python_current", "default")
    xt, spd, refwin = c4_scenarios(S=8, N=p.num_ho, seed=1, step=0)   # far apart
    got = _run_gpu(p, xt, spd, refwin, True)
    assert (got["n_outer"] == 1).all() and (got["curr_iter"] == 1).all() and (got["flag"] == 0).all()


def test_create_rejects_bad_arguments():
    lib = pa.load_library()
    p = make_params()
    with pytest.raises(pa.PIADMMError):
        PIADMMSolver(p.copy().update(num_ho=40), S=1, V=2)
    with pytest.raises(pa.PIADMMError):
        PIADMMSolver(p, S=1, V=4, D=1)          # al_all needs D >= V-1
    assert lib.piadmm_max_edges(2, 1) == 1
