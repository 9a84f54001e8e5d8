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


# ------------------------------------------------------------------ generic V (configs 3 / 5 shapes)
def _generic_params(N, inner="default", **kw):
    p = make_params("matlab_pi", inner, num_ho=N, beta=10.0, Pnorm=1.0, rho=2.0, eps_pri=0.5, eps_dual=0.5,
                    al_all=0, iter_num=12)
    p.update(**kw)
    return p


@pytest.mark.parametrize("V,N,D,r0,gap", [(16, 15, 8, 7.0, 5.0), (64, 30, 8, 12.0, 6.0), (12, 20, 4, 12.0, 6.0),
                                          (32, 15, 8, 8.0, 4.0)])
def test_coupled_intersection_matches_oracle(V, N, D, r0, gap):
    """Config-3-shaped scene (4-arm intersection, active-edge AL term, dual=pi) through the phase kernels."""
    from pi_admm_b200.scenarios import c3_intersection
    p = _generic_params(N)
    xt, spd, refwin = c3_intersection(V=V, N=N, r0=r0, gap=gap)
    ref = ctwin.step(p, xt, spd, refwin, D=D)
    sol = PIADMMSolver(p, S=1, V=V, D=D)
    sol.setup(xt, spd, refwin)
    sol.iterate()
    sol.plant_step()
    got = sol.to_numpy()
    assert ref["n_edge_solves"].sum() > 0 and (ref["status"] & 4).sum() == 0
    for k in ("edge_ij", "edge_slot"):
        n = int(ref["edge_n"][0])
        assert np.array_equal(got[k][0, :n], ref[k][0, :n])
    assert_same(got, ref, what=f"c3 V={V} N={N}")


def test_vehicle_sharded_lockstep_equals_single(monkeypatch):
    """Config-5 algorithm (vehicle sharding, all-gather of pos, redundant cross-rank edges, all-reduce of
    the residual sums) emulated with 2 and 4 'ranks' on one GPU == the unsharded run == the oracle."""
    from pi_admm_b200.scenarios import c3_intersection
    from pi_admm_b200.parallel import iterate_lockstep, shard_range
    V, N, D = 32, 15, 8
    p = _generic_params(N)
    xt, spd, refwin = c3_intersection(V=V, N=N, r0=8.0, gap=4.0)
    ref = ctwin.step(p, xt, spd, refwin, D=D)
    assert ref["n_edge_solves"].sum() > 100 and (ref["status"][0] & ~8) == 0      # (8 = outer cap of 12 reached)
    for world in (2, 4):
        sols = []
        for r in range(world):
            s = PIADMMSolver(p, S=1, V=V, D=D)
            s.setup(xt, spd, refwin)
            s.set_ownership(*shard_range(V, r, world))
            sols.append(s)
        iterate_lockstep(sols)
        for r, s in enumerate(sols):
            s.plant_step()
            got = s.to_numpy()
            lo, hi = s.own
            assert np.array_equal(got["n_outer"], ref["n_outer"]) and np.array_equal(got["curr_iter"], ref["curr_iter"])
            assert np.array_equal(got["nbr"], ref["nbr"]) and np.array_equal(got["edge_n"], ref["edge_n"])
            for k in ("u", "pos", "xt_next", "err_rk", "err_sk"):
                assert rel_err(got[k], ref[k]) <= 1e-6, (world, r, k)
            for k in ("hat", "dual"):     # slots of the owned vehicles are authoritative on this rank
                assert rel_err(got[k][:, lo:hi], ref[k][:, lo:hi]) <= 1e-6, (world, r, k)
            s.close()


# ------------------------------------------------------------------ drop-in script vs golden fixtures
@pytest.mark.parametrize("fast", [False, True])
def test_dropin_main_reproduces_golden(fast):
    """pi_admm_b200/main_intersection.py (the casadi/main.py shaped driver over the class mirror)
    against tests/golden/c1_python_current_default.json (frozen oracle output)."""
    import json, os
    from pi_admm_b200.main_intersection import run
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "c1_python_current_default.json")))
    out = run(semantics="python_current", inner="osqp_default", fast=fast)
    assert [int(v) for v in out["iter_his"]] == g["iter_his"]
    assert out["n_outer"] == g["n_outer"]
    assert np.allclose(np.array(out["x_vec"]), np.array(g["x_vec"]), rtol=1e-6, atol=1e-9)
    assert np.allclose(np.array(out["u_vec"]), np.array(g["u_vec"]), rtol=1e-6, atol=1e-9)
    for a, b in zip(out["err_rk"], g["err_rk"]):
        assert np.allclose(a, np.array(b), rtol=1e-6, atol=1e-6)
    assert out["launches"] > 0


def test_class_mirror_step_and_mutable_rho():
    from pi_admm_b200.PI_ADMM_class import PI_ADMM_CASADI
    P = PI_ADMM_CASADI(semantics="matlab_pi", eps_pri=1.0, eps_dual=1.0)
    xt = np.array([[-5.2, 0, 0], [0, 10.4, -np.pi / 2]])
    r1 = P.step(12, xt)
    P.param.rho = 3.5                      # scripts mutate param.rho (casadi_old_PI_ADMM/main.py:137)
    r2 = P.step(12, xt)
    p = P.abi_params()
    ref = ctwin.step(p, xt[None], np.array([[4.0, 8.0]]), P.ref_window(12)[None])
    assert r2["n_outer"] == int(ref["n_outer"][0]) and rel_err(r2["primal_u"], ref["u"][0]) <= 1e-6
    assert rel_err(r1["primal_u"], r2["primal_u"]) > 1e-6


# ------------------------------------------------------------------ edge cases
EDGE_CASES = [
    ("python_current", dict(num_ho=3), 5),                       # smallest horizon
    ("python_current", dict(num_ho=31, iter_num=6), 7),          # largest horizon (32-lane groups, all lanes used)
    ("python_current", dict(num_ho=16), 9),                      # first horizon that needs 32-lane groups
    ("matlab_pi", dict(num_ho=15, iter_num=1), 6),               # iteration cap reached at once
    ("matlab_pi", dict(num_ho=15, eps_pri=1e-6, eps_dual=1e-6, iter_num=5), 4),   # never converges: cap
    ("python_current", dict(in_max_iter=30), 8),                 # inner solver hits max_iter -> status bits
    ("python_current", dict(in_adaptive=0, in_max_iter=500), 8),
    ("python_current", dict(round_u=2, round_edge=1), 5),
    ("matlab_pi", dict(num_ho=12, in_check=10, eps_pri=0.5, eps_dual=0.5), 5),
    ("python_old_pi", dict(num_ho=9, iter_num=7, round_edge=0), 3),   # old main.py:119 rounds the edge controls to 0 decimals
]


@pytest.mark.parametrize("preset,kw,S", EDGE_CASES)
@pytest.mark.parametrize("fused", [True, False])
def test_edge_cases_match_oracle(preset, kw, S, fused):
    p = make_params(preset, "default", **kw)
    xt, spd, refwin = c4_scenarios(S=S, N=p.num_ho, seed=21)
    ref = ctwin.step(p, xt, spd, refwin)
    got = _run_gpu(p, xt, spd, refwin, fused)
    assert_same(got, ref, what=f"{preset}/{kw}/fused={fused}")
    if "in_max_iter" in kw and kw["in_max_iter"] == 30:
        assert (ref["status"] & 1).any()                          # PIADMM_ST_VEH_MAXITER reported


def test_slot_overflow_is_reported():
    from pi_admm_b200.scenarios import c3_intersection
    p = _generic_params(15)
    xt, spd, refwin = c3_intersection(V=16, N=15, r0=5.0, gap=3.0)
    ref = ctwin.step(p, xt, spd, refwin, D=2)
    sol = PIADMMSolver(p, S=1, V=16, D=2)
    sol.setup(xt, spd, refwin)
    sol.iterate()
    got = sol.to_numpy()
    assert ref["status"][0] & 4 and got["status"][0] & 4          # PIADMM_ST_SLOT_OVERFLOW on both


def test_large_batch_properties():
    """BASELINE config-4 size (4096 scenarios): size-independent properties instead of the oracle --
    steering / rate limits hold, roll-out is consistent with u, a repeated run is bit-identical,
    and a random subset equals the oracle."""
    p = make_params("matlab_pi", "default", num_ho=15, eps_pri=0.1, eps_dual=0.1, round_u=4, round_edge=4)
    xt, spd, refwin = c4_scenarios(S=4096, N=15)
    a = _run_gpu(p, xt, spd, refwin, True)
    b = _run_gpu(p, xt, spd, refwin, True)
    for k in ("u", "pos", "hat", "dual", "err_rk", "n_outer", "veh_inner"):
        assert np.array_equal(a[k], b[k]), k                       # deterministic
    u = a["u"]
    assert np.abs(u).max() <= np.pi / 6 + 2e-3 and np.abs(np.diff(u, axis=2)).max() <= np.pi / 9 + 2e-3  # eps 1e-3 solver
    assert (a["status"] == 0).all() and (a["n_outer"] >= 1).all()
    idx = np.random.default_rng(0).choice(4096, 64, replace=False)
    ref = ctwin.step(p, xt[idx], spd[idx], refwin[idx])
    for k in ("u", "pos", "err_rk", "xt_next"):
        assert rel_err(a[k][idx], ref[k]) <= 1e-6, k
    assert np.array_equal(a["n_outer"][idx], ref["n_outer"])


@pytest.mark.parametrize("N", [15, 30])
def test_bench_workload_every_scenario_equals_the_oracle(N):
    """The exact bench workload (BASELINE config 4, 4096 scenarios; N=15 and the N=30 variant, bench parameters):
    EVERY scenario against the C twin (all host threads): identical outer / inner iteration counts and edge flags,
    floating point within 1e-6."""
    import os
    import bench
    p = bench.bench_params(N)
    xt, spd, refwin = c4_scenarios(S=4096, N=N)
    a = _run_gpu(p, xt, spd, refwin, True)
    try:
        nt = len(os.sched_getaffinity(0))
    except AttributeError:
        nt = os.cpu_count() or 1
    ref = ctwin.step(p, xt, spd, refwin, nthreads=nt)
    for k in ("n_outer", "curr_iter", "veh_inner", "edge_inner", "n_veh_solves", "n_edge_solves", "flag", "status"):
        assert np.array_equal(a[k], ref[k]), k
    for k in ("u", "pos", "hat", "dual", "err_rk", "err_sk", "xt_next"):
        assert rel_err(a[k], ref[k]) <= 1e-6, k


def test_step_host_chunked_pipeline_equals_device_path():
    """S >= 256 makes piadmm_step_host split the batch over internal streams (copy/compute overlap)."""
    p = make_params("matlab_pi", "default", num_ho=15, eps_pri=0.1, eps_dual=0.1)
    xt, spd, refwin = c4_scenarios(S=301, N=15, seed=5)
    ref = ctwin.step(p, xt, spd, refwin)
    sol = PIADMMSolver(p, S=301, V=2)
    for _ in range(2):
        out = sol.step_host(xt, spd, refwin)
        assert np.array_equal(out["n_outer"], ref["n_outer"]) and np.array_equal(out["curr_iter"], ref["curr_iter"])
        for k in ("u", "pos", "xt_next"):
            assert rel_err(out[k], ref[k]) <= 1e-6


def test_step_host_at_resident_reference_closed_loop():
    """piadmm_set_reference + piadmm_step_host_at: per step only xt goes to the device.  Pinned buffers, so the
    cached CUDA graph is captured and then replayed with a changing step index and a changing scheduling order;
    every step is compared with the C twin fed the host-sliced window."""
    import torch
    from pi_admm_b200.scenarios import c4_reference
    S, N = 301, 15
    p = make_params("matlab_pi", "default", num_ho=N, eps_pri=0.1, eps_dual=0.1)
    xt, spd, _ = c4_scenarios(S=S, N=N, seed=5, step=10)
    ref_traj = c4_reference(S=S, seed=5)
    sol = PIADMMSolver(p, S=S, V=2)
    sol.setup(xt, spd, np.zeros((S, 2, 2, N + 1)))          # spd becomes resident; the window buffer is overwritten
    sol.set_reference(ref_traj)
    io = sol.alloc_host_io()
    io["xt"].copy_(torch.from_numpy(xt))
    for step in range(10, 16):
        sol.step_host_at(step, io)
        ref = ctwin.step(p, io["xt"].numpy(), spd, np.ascontiguousarray(ref_traj[:, :, :, step:step + N + 1]))
        assert np.array_equal(io["n_outer"].numpy(), ref["n_outer"]), step
        assert np.array_equal(io["curr_iter"].numpy(), ref["curr_iter"]), step
        for k in ("u", "pos", "xt_next"):
            assert rel_err(io[k].numpy(), ref[k]) <= 1e-6, (step, k)
        io["xt"].copy_(io["xt_next"])                        # plant step closes the loop (casadi/main.py:185-192)
    assert ref["n_edge_solves"].sum() > 0


def test_step_host_at_argument_errors():
    import torch
    from pi_admm_b200.scenarios import c4_reference
    p = make_params("python_current", "default")
    xt, spd, refwin = c4_scenarios(S=8, N=p.num_ho, seed=1)
    sol = PIADMMSolver(p, S=8, V=2)
    sol.setup(xt, spd, refwin)
    io = sol.alloc_host_io()
    with pytest.raises(pa.PIADMMError):                      # no resident reference yet
        sol.step_host_at(0, io)
    with pytest.raises(pa.PIADMMError):                      # wrong shape
        sol.set_reference(np.zeros((8, 2, 50)))
    sol.set_reference(c4_reference(S=8, seed=1))
    io["xt"].copy_(torch.from_numpy(xt))
    sol.step_host_at(34, io)                                 # last admissible step: 34 + 15 + 1 = 50 columns
    with pytest.raises(pa.PIADMMError):                      # the reference script stops at Nt/dt - num_ho
        sol.step_host_at(35, io)
    with pytest.raises(pa.PIADMMError):
        sol.step_host_at(-1, io)
    # small batch (not chunked, no graph) agrees with the plain host entry point
    sol.step_host_at(12, io)
    out = sol.step_host(xt, spd, refwin)
    for k in ("u", "pos", "xt_next", "n_outer"):
        assert np.array_equal(io[k].numpy(), out[k]), k


# ------------------------------------------------------------------ whole receding-horizon run on the device
GOLDEN_PRESET = {"c1_python_current_default": ("python_current", "default"), "c1_matlab_pi_default": ("matlab_pi", "default"),
                 "c1_python_old_pi_default": ("python_old_pi", "default"), "c1_python_current_tight": ("python_current", "tight")}


@pytest.mark.parametrize("name", sorted(GOLDEN_PRESET))
def test_run_closed_loop_reproduces_golden_config1(name):
    """piadmm_run_closed_loop: all 50 - num_ho steps of casadi/main.py:43-201 in one asynchronous call, against the
    frozen oracle closed loops (tests/golden)."""
    import json, os
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", name + ".json")))
    preset, inner = GOLDEN_PRESET[name]
    p = make_params(preset, inner)
    N = p.num_ho
    ref_traj = reference_trajectories()                                  # (2V, 50) rows [x_A; 0; 0; y_B]
    n_steps = ref_traj.shape[1] - N
    assert n_steps == len(g["iter_his"])
    sol = PIADMMSolver(p, S=1, V=2)
    sol.setup(c1_initial_state()[None], np.array([[4.0, 8.0]]), np.zeros((1, 2, 2, N + 1)))
    sol.set_reference(ref_traj.reshape(1, 2, 2, -1))
    h = {k: v.cpu().numpy() for k, v in sol.run_closed_loop(0, n_steps).items()}
    assert h["iter_his"][:, 0].tolist() == g["iter_his"] and h["n_outer_his"][:, 0].tolist() == g["n_outer"]
    assert np.allclose(h["x_vec"][:, 0].transpose(0, 2, 1), np.array(g["x_vec"]), rtol=1e-6, atol=1e-9)
    assert np.allclose(h["theta_vec"][:, 0], np.array(g["theta_vec"]), rtol=1e-6, atol=1e-9)
    assert np.allclose(h["u_vec"][:, 0], np.array(g["u_vec"]), rtol=1e-6, atol=1e-9)
    last_rk = [r[-1] if len(r) else 0.0 for r in g["err_rk"]]
    assert np.allclose(h["err_rk_his"][:, 0], last_rk, rtol=1e-6, atol=1e-6)
    with pytest.raises(pa.PIADMMError):                                  # one step too many leaves the reference
        sol.run_closed_loop(0, n_steps + 1)


def test_run_closed_loop_batch_equals_stepping():
    """A batch of perturbed scenarios: the on-device run equals stepping the host entry point (and the C twin)."""
    from pi_admm_b200.scenarios import c4_reference
    S, N, first, n = 96, 15, 9, 7
    p = make_params("matlab_pi", "default", num_ho=N, eps_pri=0.1, eps_dual=0.1)
    xt0, spd, _ = c4_scenarios(S=S, N=N, seed=8, step=first)
    ref_traj = c4_reference(S=S, seed=8)
    sol = PIADMMSolver(p, S=S, V=2)
    sol.setup(xt0, spd, np.zeros((S, 2, 2, N + 1)))
    sol.set_reference(ref_traj)
    h = {k: v.cpu().numpy() for k, v in sol.run_closed_loop(first, n).items()}
    xt = xt0.copy()
    for t in range(n):
        ref = ctwin.step(p, xt, spd, np.ascontiguousarray(ref_traj[:, :, :, first + t:first + t + N + 1]))
        assert np.array_equal(h["n_outer_his"][t], ref["n_outer"]) and np.array_equal(h["iter_his"][t], ref["curr_iter"]), t
        assert rel_err(h["x_vec"][t], ref["xt_next"][:, :, :2]) <= 1e-6 and rel_err(h["theta_vec"][t], ref["xt_next"][:, :, 2]) <= 1e-6
        assert rel_err(h["u_vec"][t], ref["u"][:, :, 0]) <= 1e-6
        xt = ref["xt_next"].copy()
    assert rel_err(sol.buf["xt"].cpu().numpy(), xt) <= 1e-6             # the bound state buffer holds the final state
    assert h["n_outer_his"].max() >= 2


def test_run_closed_loop_generic_v():
    """V > 2 goes through the phase kernels; compared with stepping the same solver by hand."""
    from pi_admm_b200.scenarios import c3_intersection
    V, N, D, n = 16, 15, 8, 3
    p = _generic_params(N)
    xt, spd, rw0 = c3_intersection(V=V, N=N, r0=7.0, gap=5.0)
    T = N + 1 + n
    k = np.arange(T)
    ref_traj = np.zeros((1, V, 2, T))                                    # constant-speed straight lines
    ref_traj[0, :, 0] = xt[0, :, 0:1] + k * 0.1 * spd[0, :, None] * np.cos(xt[0, :, 2:3])
    ref_traj[0, :, 1] = xt[0, :, 1:2] + k * 0.1 * spd[0, :, None] * np.sin(xt[0, :, 2:3])
    assert np.allclose(ref_traj[:, :, :, :N + 1], rw0)
    a = PIADMMSolver(p, S=1, V=V, D=D)
    a.setup(xt, spd, rw0); a.set_reference(ref_traj)
    h = {kk: v.cpu().numpy() for kk, v in a.run_closed_loop(0, n).items()}
    b = PIADMMSolver(p, S=1, V=V, D=D)
    x = xt.copy()
    for t in range(n):
        b.setup(x, spd, np.ascontiguousarray(ref_traj[:, :, :, t:t + N + 1]))
        b.iterate(); b.plant_step()
        o = b.to_numpy()
        assert np.array_equal(h["n_outer_his"][t], o["n_outer"])
        assert np.array_equal(h["x_vec"][t], o["xt_next"][:, :, :2]) and np.array_equal(h["u_vec"][t], o["u"][:, :, 0])
        x = o["xt_next"].copy()


@pytest.mark.parametrize("N,V", [(15, 2), (30, 2), (15, 8)])
def test_kinv_cache_is_invalidated_by_changed_coefficients(N, V):
    """The K^-1(rho0) cache is keyed by the coefficients that define K (speed, outer penalty, inner settings).  A handle
    that has cached one problem set and is then given other speeds / rho / inner sigma must give, bit for bit, what a
    fresh handle gives (a stale entry would change iterates and iteration counts)."""
    from pi_admm_b200.scenarios import c3_intersection
    keys = ("u", "pos", "err_rk", "n_outer", "veh_inner", "edge_inner")

    def inputs(seed, vscale):
        if V == 2:
            xt, spd, rw = c4_scenarios(S=48, N=N, seed=seed)
        else:
            xt, spd, rw = c3_intersection(V=V, N=N, seed=seed)
        return xt, np.asarray(spd) * vscale, rw

    def fresh(p, inp, rho=None):
        s = PIADMMSolver(p, S=inp[0].shape[0], V=V, D=None if V == 2 else 7)
        s.setup(*inp, rho=rho)
        s.iterate()
        out = s.to_numpy()
        s.close()
        return out

    kw = dict(num_ho=N, eps_pri=0.1, eps_dual=0.1)
    if V != 2:
        kw.update(al_all=0, iter_num=4, eps_pri=0.5, eps_dual=0.5, beta=10.0, Pnorm=1.0, rho=2.0)
    p = make_params("matlab_pi", "default", **kw)
    a, b = inputs(3, 1.0), inputs(3, 1.07)
    sol = PIADMMSolver(p, S=a[0].shape[0], V=V, D=None if V == 2 else 7)
    sol.setup(*a)
    sol.iterate(); sol.iterate()                         # second pass runs on cache hits
    want = fresh(p, a)
    got = sol.to_numpy()
    for k in keys:
        assert np.array_equal(got[k], want[k]), ("hit", k)
    sol.setup(*b)                                        # other speeds: every vehicle key changes
    sol.iterate()
    want, got = fresh(p, b), sol.to_numpy()
    for k in keys:
        assert np.array_equal(got[k], want[k]), ("speed", k)
    rho = np.full(a[0].shape[0], 3.1)
    sol.setup(*b, rho=rho)                               # other outer penalty
    sol.iterate()
    want, got = fresh(p, b, rho=rho), sol.to_numpy()
    for k in keys:
        assert np.array_equal(got[k], want[k]), ("rho", k)
    sol.set_params(in_sigma=2e-6)                        # other inner setting
    sol.setup(*b, rho=rho)
    sol.iterate()
    p2 = make_params("matlab_pi", "default", **kw)
    p2.update(in_sigma=2e-6)
    want, got = fresh(p2, b, rho=rho), sol.to_numpy()
    for k in keys:
        assert np.array_equal(got[k], want[k]), ("sigma", k)
    sol.close()


def _ragged_scene(V, N, seeds, far):
    """S = len(seeds) scenes of V vehicles cut out of c3 intersections; scene `far` is spread out (no collision)."""
    from pi_admm_b200.scenarios import c3_intersection
    V4 = 4 * ((V + 3) // 4)
    parts = []
    for n, seed in enumerate(seeds):
        xt, spd, rw = c3_intersection(V=V4, N=N, seed=seed, r0=60.0 if n == far else 7.0, gap=25.0 if n == far else 5.0)
        parts.append((xt[:, :V], spd[:, :V], rw[:, :V]))
    return tuple(np.ascontiguousarray(np.concatenate([p[i] for p in parts], axis=0)) for i in range(3))


@pytest.mark.parametrize("V,N", [(5, 15), (7, 12), (3, 20)])
def test_ragged_generic_batch_matches_oracle(V, N):
    """Odd V and a batch in which one scenario stops after its first outer iteration (no collision): at N <= 15 a warp
    then holds groups of two different scenarios, one finished and one active (groups masked inside a running warp)."""
    p = _generic_params(N)
    xt, spd, refwin = _ragged_scene(V, N, seeds=(3, 4, 5), far=1)
    ref = ctwin.step(p, xt, spd, refwin, D=V - 1)
    assert ref["n_outer"][1] == 1 and ref["n_edge_solves"][0] > 0 and ref["n_edge_solves"][2] > 0
    sol = PIADMMSolver(p, S=3, V=V, D=V - 1)
    sol.setup(xt, spd, refwin)
    sol.iterate()
    sol.plant_step()
    assert_same(sol.to_numpy(), ref, what=f"ragged V={V} N={N}")
    sol.close()


def test_vehicle_sharding_with_odd_ownership_ranges():
    """Ownership ranges that split a warp's two groups between ranks (V=6 over ranks of 3 + 3, V=5 over 3 + 2)."""
    from pi_admm_b200.parallel import iterate_lockstep
    for V, cuts in ((6, (0, 3, 6)), (5, (0, 3, 5))):
        N = 15
        p = _generic_params(N)
        xt, spd, refwin = _ragged_scene(V, N, seeds=(3,), far=-1)
        ref = ctwin.step(p, xt, spd, refwin, D=V - 1)
        assert ref["n_edge_solves"].sum() > 0
        sols = []
        for r in range(len(cuts) - 1):
            s = PIADMMSolver(p, S=1, V=V, D=V - 1)
            s.setup(xt, spd, refwin)
            s.set_ownership(cuts[r], cuts[r + 1])
            sols.append(s)
        iterate_lockstep(sols)
        for r, s in enumerate(sols):
            got = s.to_numpy()
            lo, hi = s.own
            assert np.array_equal(got["n_outer"], ref["n_outer"]) and np.array_equal(got["nbr"], ref["nbr"])
            for k in ("u", "pos", "err_rk", "err_sk"):
                assert rel_err(got[k], ref[k]) <= 1e-6, (V, r, k)
            for k in ("hat", "dual"):
                assert rel_err(got[k][:, lo:hi], ref[k][:, lo:hi]) <= 1e-6, (V, r, k)
            s.close()


GENERIC_PRESETS = [
    ("python_current", dict(al_all=1, eps_pri=1.0, eps_dual=0.5, dual_res_alias=0), 3, 15),   # AL over every j != i
    ("python_current", dict(al_all=1), 4, 12),                                    # C1 semantics incl. alias quirk
    ("python_current", dict(al_all=0, round_u=3, round_edge=2, eps_pri=0.5), 5, 15),
    ("matlab_pi", dict(al_all=1, beta=10.0, Pnorm=1.0, rho=2.0, eps_pri=0.5, eps_dual=0.5), 3, 20),
    ("matlab_pi", dict(al_all=0, beta=10.0, Pnorm=1.0, rho=2.0, eps_pri=0.5, eps_dual=0.5, PI=0), 6, 15),
    ("python_old_pi", dict(al_all=0, num_ho=10, iter_num=6), 4, 10),              # scene-wide rho set by the last edge
    ("python_old_pi", dict(al_all=1, num_ho=15, iter_num=5, trad=1), 3, 15),
    ("matlab_pi", dict(al_all=0, beta=10.0, Pnorm=1.0, rho=2.0, eps_pri=0.5, eps_dual=0.5), 1, 15),   # V = 1: no pair at all
]


@pytest.mark.parametrize("preset,kw,V,N", GENERIC_PRESETS)
def test_generic_v_semantic_presets_match_oracle(preset, kw, V, N):
    """The three dual-update laws and the AL neighbourhoods through the phase kernels (generic V), on a ragged batch
    of three scenes (the middle one collision free)."""
    p = make_params(preset, "default", **dict(dict(num_ho=N, iter_num=8), **kw))
    xt, spd, refwin = _ragged_scene(V, p.num_ho, seeds=(11, 12, 13), far=1)
    D = max(1, V - 1)
    ref = ctwin.step(p, xt, spd, refwin, D=D)
    if V > 1 and not p.always_edge:
        assert ref["n_edge_solves"][0] + ref["n_edge_solves"][2] > 0
    sol = PIADMMSolver(p, S=3, V=V, D=D)
    sol.setup(xt, spd, refwin)
    sol.iterate()
    sol.plant_step()
    assert_same(sol.to_numpy(), ref, what=f"generic {preset} {kw} V={V}")
    sol.close()


def test_run_closed_loop_generic_v_ragged_batch_equals_the_twin():
    """On-device closed loop through the phase kernels for a ragged batch (odd V, three scenes, one collision free)
    against stepping the C twin on the host."""
    V, N, n = 5, 15, 3
    p = _generic_params(N, iter_num=6)
    xt, spd, rw0 = _ragged_scene(V, N, seeds=(3, 4, 5), far=1)
    T = N + 1 + n
    k = np.arange(T)
    ref_traj = np.zeros((3, V, 2, T))
    ref_traj[:, :, 0] = xt[:, :, 0:1] + k * 0.1 * spd[:, :, None] * np.cos(xt[:, :, 2:3])
    ref_traj[:, :, 1] = xt[:, :, 1:2] + k * 0.1 * spd[:, :, None] * np.sin(xt[:, :, 2:3])
    assert np.allclose(ref_traj[:, :, :, :N + 1], rw0)
    sol = PIADMMSolver(p, S=3, V=V, D=V - 1)
    sol.setup(xt, spd, rw0)
    sol.set_reference(ref_traj)
    h = {kk: v.cpu().numpy() for kk, v in sol.run_closed_loop(0, n).items()}
    sol.close()
    x = xt.copy()
    for t in range(n):
        ref = ctwin.step(p, x, spd, np.ascontiguousarray(ref_traj[:, :, :, t:t + N + 1]), D=V - 1)
        assert np.array_equal(h["n_outer_his"][t], ref["n_outer"]) and np.array_equal(h["iter_his"][t], ref["curr_iter"]), t
        assert rel_err(h["x_vec"][t], ref["xt_next"][:, :, :2]) <= 1e-6 and rel_err(h["u_vec"][t], ref["u"][:, :, 0]) <= 1e-6
        x = ref["xt_next"].copy()
    assert h["n_outer_his"].max() >= 2


def test_reference_error_behaviour_vehicle_failure_raises_edge_failure_is_recorded():
    """casadi/main.py:96-103 builds the vehicle solver with CasADi's error_on_fail default (a failed QP raises
    RuntimeError) and the edge solver with error_on_fail=False (:145-153).  The class mirror and the drop-in driver
    do the same with the solver's status bits; the on-device closed loop reports them per step (status_his)."""
    from pi_admm_b200 import main_intersection
    from pi_admm_b200.PI_ADMM_class import PI_ADMM_CASADI
    xt = np.array([[-5.2, 0, 0], [0, 10.4, -np.pi / 2]])
    P = PI_ADMM_CASADI(in_max_iter=30)                     # no vehicle QP converges within 30 inner iterations
    with pytest.raises(RuntimeError, match="vehicle QP"):
        P.step(12, xt)
    P.error_on_fail = False                                # error_on_fail=False: the step goes through, bits recorded
    r = P.step(12, xt)
    assert r["status"] & pa._abi.ST_VEH_MAXITER
    for kw in (dict(fast=True), dict(fast=False), dict(device_loop=True)):
        with pytest.raises(RuntimeError, match="vehicle QP"):
            main_intersection.run(n_steps=14, in_max_iter=30, **kw)
    out = main_intersection.run(n_steps=3, device_loop=True)             # healthy run: nothing raised
    assert len(out["u_vec"]) == 3
    # status_his of the device loop equals the status of stepping
    S, N = 8, 15
    p = make_params("python_current", "default", in_max_iter=50)
    from pi_admm_b200.scenarios import c4_reference
    xt0, spd, _ = c4_scenarios(S=S, N=N, seed=8, step=10)
    ref_traj = c4_reference(S=S, seed=8)
    sol = PIADMMSolver(p, S=S, V=2)
    sol.setup(xt0, spd, np.zeros((S, 2, 2, N + 1)))
    sol.set_reference(ref_traj)
    h = {k: v.cpu().numpy() for k, v in sol.run_closed_loop(10, 3).items()}
    sol.close()
    x = xt0.copy()
    for t in range(3):
        ref = ctwin.step(p, x, spd, np.ascontiguousarray(ref_traj[:, :, :, 10 + t:10 + t + N + 1]))
        assert np.array_equal(h["status_his"][t], ref["status"]), t
        x = ref["xt_next"].copy()
    assert h["status_his"].any()


def test_qpsol_shaped_script_equals_fused_step():
    """VERDICT r1 #8 / SURVEY 8(b): a script in casadi/main.py's own shape -- ca.SX.sym, cost_function_primal / _edge,
    nonlcon_function, ca.qpsol("solver", "osqp", ...), F(lbg=0) per vehicle and per edge, numpy for everything around --
    runs over the descriptors (pi_admm_b200.ca_shim) and gives the same closed loop as the fused persistent kernel."""
    from pi_admm_b200 import main_qpsol_shape, main_intersection
    n = 22                                              # through the steps with an active edge
    a = main_qpsol_shape.run(n_steps=n)
    b = main_intersection.run(semantics="python_current", fast=True, n_steps=n)
    assert [int(v) for v in a["iter_his"][:n]] == [int(v) for v in b["iter_his"][:n]]
    assert a["n_outer"] == b["n_outer"] and max(a["n_outer"]) >= 2
    assert np.allclose(np.array(a["x_vec"]), np.array(b["x_vec"]), rtol=1e-6, atol=1e-9)
    assert np.allclose(np.array(a["u_vec"]), np.array(b["u_vec"]), rtol=1e-6, atol=1e-9)
    for ra, rb in zip(a["err_rk"], b["err_rk"]):
        if len(ra):
            assert np.allclose(ra, rb[:len(ra)], rtol=1e-6, atol=1e-6)


def test_descriptor_solves_match_the_oracle_problems():
    """One vehicle QP and one edge problem through PI_ADMM_CASADI.solve(descriptor) against the numpy oracle's solver on
    the same data (tight inner tolerance: the minimiser is solver independent)."""
    from pi_admm_b200.PI_ADMM_class import PI_ADMM_CASADI
    from oracle import problem as op, solver as osol
    P = PI_ADMM_CASADI(semantics="python_current", inner="tight")
    N, V = 15, 2
    rng = np.random.default_rng(5)
    xt = np.array([[-5.2, 0.1, 0.05], [0.2, 10.4, -np.pi / 2 + 0.03]])
    hat = np.empty((V, V), dtype=object); dual = np.empty((V, V), dtype=object)
    for i in range(V):
        for j in range(V):
            hat[i, j] = rng.normal(size=(2, N + 1)); dual[i, j] = 0.1 * rng.normal(size=(2, N + 1))
    st = osol.InnerSettings.tight()
    par = op.Params()
    rw = P.ref_window(12)
    for i in range(V):
        desc = P.cost_function_primal(12, None, xt, hat, dual, i)
        u = P.solve(desc, P.nonlcon_function(np.zeros((1, N)))[1])
        H, q = op.vehicle_qp(par, xt[i], float(P.param.spd[i]), rw[i, 0], rw[i, 1], [hat[i, 1 - i]], [dual[i, 1 - i]], float(P.param.rho))
        x_ref, info = osol.solve_box_qp(H, q, st)
        assert info["status"] == 0 and np.abs(u[0] - x_ref).max() <= 1e-6
    # edge problem of the pair, data as casadi/main.py:127-131 builds it
    spd = np.asarray(P.param.spd, dtype=np.float64)
    seed = np.around(xt[:, :2] + 0.1 * spd[:, None] * np.stack((np.cos(xt[:, 2]), np.sin(xt[:, 2])), 1), 4)
    pos_edge = rng.normal(size=(4, N + 1)) + np.repeat(xt[:, :2].reshape(4, 1), N + 1, axis=1)
    dual_edge = 0.1 * rng.normal(size=(4, N + 1))
    ue = P.solve(P.cost_function_edge(None, xt, pos_edge, dual_edge, seed)).reshape(2, N)
    blk = op.edge_qp_blocks(par, xt, spd, pos_edge, dual_edge, seed, float(P.param.rho))
    xs, info = osol.solve_edge_pair(blk["H"], blk["q"], blk["a"], blk["c"], par.beta, st)
    assert info["status"] == 0 and np.abs(ue - np.asarray(xs)).max() <= 1e-6
