"""Parity against vectors produced by the REFERENCE'S OWN CODE (tests/golden/reference_vectors.json, generated in the build
container by tests/golden/make_reference_vectors.py: casadi/PI_ADMM_class.py and Distributed_planner/decentralized/* imported
unmodified, numpy stand-ins for the casadi / bunch containers, no solver).  They pin the problem construction -- constants,
reference trajectories, (H, q) of the vehicle QP, the constraint rows, both roll-outs, the Python edge cost, the config-2
consensus arithmetic and half-spaces -- of the oracle, of the product's host mirror and of the CUDA path to the reference's
statements.  Solver iterates (OSQP / IPOPT) are not in the vectors: see DESIGN.md section 3."""
import json
import os

import numpy as np
import pytest

from oracle.problem import (Params, reference_trajectories, vehicle_qp, constraint_matrix, constraint_lower, affine_rollout,
                            literal_dynamic_update_local, literal_dynamic_update_edge)
from oracle.solver import InnerSettings, solve_box_qp
from oracle import obca_consensus as oc

HERE = os.path.dirname(os.path.abspath(__file__))
VEC = json.load(open(os.path.join(HERE, "golden", "reference_vectors.json")))
A_ = lambda x: np.asarray(x, dtype=np.float64)


def _params(case):
    return Params(num_ho=case["N"], spd=A_(case["spd"]))


def test_constants_and_reference_trajectories():
    rp = VEC["param"]
    p = Params()
    for k in ("dt", "Nt", "L", "num_ho", "num_veh", "dis_thres", "beta", "Pnorm", "Pcost", "iter_num", "rho", "eps_pri", "eps_dual"):
        assert float(getattr(p, k)) == float(rp[k]), k
    assert np.array_equal(A_(p.spd), A_(rp["spd"]))
    assert np.array_equal(reference_trajectories(p), A_(VEC["ref_traj"]))
    import pi_admm_b200 as pa
    from pi_admm_b200.PI_ADMM_class import PI_ADMM_CASADI
    P = PI_ADMM_CASADI()
    assert np.array_equal(P.ref_traj, A_(VEC["ref_traj"]))
    for k, v in rp.items():
        assert np.array_equal(np.asarray(P.param[k], dtype=np.float64), np.asarray(v, dtype=np.float64)), k
    d = pa.default_params()
    for k in ("dt", "L", "dis_thres", "beta", "Pnorm", "Pcost", "rho", "eps_pri", "eps_dual"):
        assert float(getattr(d, k)) == float(rp[k]), k
    assert d.num_ho == rp["num_ho"] and d.iter_num == rp["iter_num"]


@pytest.mark.parametrize("ci", range(len(VEC["vehicle"])))
def test_vehicle_qp_equals_the_reference_cost_function(ci):
    """(H, q) extracted EXACTLY from cost_function_primal (casadi/PI_ADMM_class.py:114-135; it is a quadratic: values at 0,
    +-e_i, e_i+e_j) equal the oracle's closed form; the cost at random u and the constraint rows of nonlcon_function
    (:172-192) agree too."""
    c = VEC["vehicle"][ci]
    p, N, v = _params(c), c["N"], c["veh"]
    xt = A_(c["xt"])
    ref = A_(VEC["ref_traj"])[2 * v:2 * v + 2, c["num_step"]:c["num_step"] + N + 1]
    H, q = vehicle_qp(p, xt[v], p.spd[v], ref[0], ref[1], [A_(c["hat"])], [A_(c["dual"])])
    Hr, qr = A_(c["H"]), A_(c["q"])
    assert np.abs(H - Hr).max() <= 1e-8 * np.abs(Hr).max() and np.abs(q - qr).max() <= 1e-8 * max(1.0, np.abs(qr).max())
    A, l = constraint_matrix(N), constraint_lower(N)
    for u, f, g in zip(A_(c["u"]), c["cost"], c["cieq"]):
        quad = 0.5 * u @ H @ u + q @ u
        assert abs((f - c["f0"]) - quad) <= 1e-9 * max(1.0, abs(f))
        assert np.abs(A_(g) - (A @ u - l)).max() <= 1e-14
    for u, (x, y, th) in zip(A_(c["u"]), c["rollout_local_of_rounded_u"]):
        ur = np.around(u, 4)
        xa, ya = affine_rollout(p, xt[v], p.spd[v], ur)
        xl, yl, tl = literal_dynamic_update_local(p, xt[v], ur, p.spd[v])
        for got, want in ((xa, x), (ya, y), (xl, x), (yl, y), (tl, th)):
            assert np.abs(got - A_(want)).max() <= 1e-11 * max(1.0, np.abs(A_(want)).max())
        from pi_admm_b200.PI_ADMM_class import PI_ADMM_CASADI
        P = PI_ADMM_CASADI(); P.param.num_ho = N; P.param.spd = p.spd
        xh, yh, thh = P.dynamic_update_local(xt[v], ur.reshape(1, N), v, if_SX=0)
        assert np.array_equal(xh, A_(x)) and np.array_equal(yh, A_(y)) and np.array_equal(thh, A_(th))


@pytest.mark.parametrize("ci", range(len(VEC["edge"])))
def test_edge_rollout_and_python_edge_cost(ci):
    """dynamic_update_edge numeric branch (casadi/PI_ADMM_class.py:77-105) and the value of the Python cost_function_edge
    (:145-169, the non-convex statement: same hinge / effort / AL terms as the convex one, non-linear roll-out inside)."""
    c = VEC["edge"][ci]
    p, N = _params(c), c["N"]
    xt, pos_old, dual, seed = A_(c["xt"]), A_(c["pos_old"]), A_(c["dual"]), A_(c["seed"])
    from pi_admm_b200.PI_ADMM_class import PI_ADMM_CASADI
    from pi_admm_b200.joint import nonlinear_rollout
    import torch
    P = PI_ADMM_CASADI(); P.param.num_ho = N; P.param.spd = p.spd
    for u, (x, y, th) in zip(A_(c["hat_u"]), c["rollout_edge_of_rounded_u"]):
        ur = np.around(u, 4)
        xo, yo, to = literal_dynamic_update_edge(p, xt, ur, p.spd)
        assert np.abs(xo - A_(x)).max() <= 1e-11 and np.abs(yo - A_(y)).max() <= 1e-11 and np.abs(to - A_(th)).max() <= 1e-12
        xh, yh, thh = P.dynamic_update_edge(xt, ur, if_SX=0)
        assert np.array_equal(xh, A_(x)) and np.array_equal(yh, A_(y))
        r = nonlinear_rollout(torch.from_numpy(xt)[None], torch.from_numpy(p.spd)[None], torch.from_numpy(ur)[None], p.dt, p.L)[0].numpy()
        assert np.abs(r[:, 0] - A_(x)).max() <= 1e-11 and np.abs(r[:, 1] - A_(y)).max() <= 1e-11
    ell = seed[1] - seed[0]
    for u, f in zip(A_(c["hat_u"]), c["cost_python_nonconvex"]):
        x, y, _ = literal_dynamic_update_edge(p, xt, u, p.spd)
        dis = 2.0 * (ell[0] * (x[1, 1:] - x[0, 1:]) + ell[1] * (y[1, 1:] - y[0, 1:])) - np.sum(ell ** 2)
        edge_pos = np.vstack((x[0], y[0], x[1], y[1]))
        mine = p.beta * np.sum(np.maximum(0.0, p.dis_thres ** 2 - dis)) + p.Pcost * np.sum(u ** 2) + p.rho / 2 * np.sum((pos_old - edge_pos + dual) ** 2)
        assert abs(mine - f) <= 1e-10 * max(1.0, abs(f))


@pytest.mark.parametrize("ci", range(len(VEC["obca"])))
def test_obca_consensus_oracle_equals_the_reference_methods(ci):
    """mid_state, bar_state_update (+ util half-spaces), check_converge, lambda_update, iterate_next_state of
    Distributed_planner/decentralized/optimizer.py, run by the reference itself."""
    c = VEC["obca"][ci]
    st = oc.MidState()
    assert np.array_equal(st.lamb_ij, A_(c["lamb_ij0"])) and np.array_equal(st.lamb_bar, A_(c["lamb_bar0"]))
    oc.bar_state_update(st, [A_(f) for f in c["fullx"]], prob=1)
    assert np.abs(st.A - A_(c["A"])).max() <= 1e-15 and np.abs(st.b - A_(c["b"])).max() <= 1e-12 * np.abs(A_(c["b"])).max()
    assert np.array_equal(st.local_x, A_(c["local_x"]))
    st.A, st.b, st.lamb_ij = A_(c["check_A"]), A_(c["check_b"]), A_(c["check_lamb_ij"])
    assert oc.check_converge(st, c["thres"], c["min_dis"]) == c["check_converge"]
    st.Z_bar, st.lamb_bar = A_(c["Z_bar"]), A_(c["lamb_bar_before"])
    oc.lambda_update(st, c["rho"])
    assert np.array_equal(st.lamb_bar, A_(c["lamb_bar_after"]))
    for f in oc.MidState.FIELDS:
        setattr(st, f, A_(c["shift_before"][f]))
    oc.iterate_next_state(st)
    for f in oc.MidState.FIELDS:
        assert np.array_equal(getattr(st, f), A_(c["shift_after"][f])), f
    assert any(v["check_converge"] for v in VEC["obca"]) and not all(v["check_converge"] for v in VEC["obca"])


# ------------------------------------------------------------------ GPU: the CUDA path against the reference's vectors
@pytest.mark.gpu
@pytest.mark.parametrize("ci", range(len(VEC["vehicle"])))
def test_gpu_vehicle_qp_minimises_the_reference_cost(ci):
    """cost_function_primal -> VehicleQP descriptor -> piadmm_x_update (tight inner tolerance): the returned steering
    sequence is the minimiser of the reference's OWN quadratic (H, q from its cost function) under its own constraint rows
    (oracle ADMM on that (H, q), and the KKT conditions checked directly)."""
    from pi_admm_b200.PI_ADMM_class import PI_ADMM_CASADI
    c = VEC["vehicle"][ci]
    N, v = c["N"], c["veh"]
    P = PI_ADMM_CASADI(inner="tight")
    P.param.num_ho = N; P.param.spd = A_(c["spd"])
    xt = A_(c["xt"])
    hat = np.empty((2, 2), dtype=object); dual = np.empty((2, 2), dtype=object)
    for i in range(2):
        for j in range(2):
            hat[i, j] = A_(c["hat"]); dual[i, j] = A_(c["dual"])
    desc = P.cost_function_primal(c["num_step"], None, xt, hat, dual, v)
    u = P.solve(desc)[0]
    Hr, qr = A_(c["H"]), A_(c["q"])
    ur, info = solve_box_qp(Hr, qr, InnerSettings.tight())
    assert info["status"] == 0
    assert np.abs(u - ur).max() <= 1e-6 * max(1.0, np.abs(ur).max())
    A, l = constraint_matrix(N), constraint_lower(N)
    assert (A @ u - l).min() >= -1e-7                                    # feasible for nonlcon_function's rows
    f = lambda w: 0.5 * w @ Hr @ w + qr @ w
    assert f(u) <= f(ur) + 1e-7 * max(1.0, abs(f(ur)))                   # and as good as the oracle's minimiser of the reference cost


@pytest.mark.gpu
@pytest.mark.parametrize("ci", range(len(VEC["edge"])))
def test_gpu_plant_step_equals_the_reference_edge_rollout(ci):
    """k_plant (casadi/main.py:185-192: xt <- column 1 of dynamic_update_edge(xt, primal_u)) against the reference's own
    dynamic_update_edge output."""
    import torch
    from pi_admm_b200.solver import PIADMMSolver
    from util import make_params
    c = VEC["edge"][ci]
    N = c["N"]
    xt, spd = A_(c["xt"]), A_(c["spd"])
    sol = PIADMMSolver(make_params("python_current", num_ho=N), S=1, V=2)
    for u, (x, y, th) in zip(A_(c["hat_u"]), c["rollout_edge_of_rounded_u"]):
        sol.setup(xt[None], spd[None], np.zeros((1, 2, 2, N + 1)))
        sol.buf["u"].copy_(torch.from_numpy(np.around(u, 4))[None])
        sol.plant_step()
        nxt = sol.buf["xt_next"].cpu().numpy()[0]
        want = np.stack((A_(x)[:, 1], A_(y)[:, 1], A_(th)[:, 1]), 1)
        assert np.abs(nxt - want).max() <= 1e-12 * max(1.0, np.abs(want).max())
    sol.close()


@pytest.mark.gpu
def test_gpu_obca_kernels_equal_the_reference_methods():
    """obca_bar_state_update / obca_check_converge / obca_lambda_update / obca_iterate_next_state, all four recorded cases as
    one batch of scenarios, against what the reference's own methods returned."""
    import torch
    from pi_admm_b200.obca_consensus import OBCAConsensus, FIELDS
    cs = VEC["obca"]
    S = len(cs)
    o = OBCAConsensus(S=S)
    dev = o.device
    T = lambda key, sub=None: torch.from_numpy(np.stack([A_(c[key] if sub is None else c[key][sub]) for c in cs])).to(dev).contiguous()
    assert np.array_equal(o.bar_state.lamb_ij[0].cpu().numpy(), A_(cs[0]["lamb_ij0"]))
    o.bar_state_update(T("fullx"), prob=1)
    torch.cuda.synchronize()
    assert np.abs(o.bar_state.A.cpu().numpy() - T("A").cpu().numpy()).max() <= 1e-15
    bw = T("b").cpu().numpy()
    assert np.abs(o.bar_state.b.cpu().numpy() - bw).max() <= 1e-12 * np.abs(bw).max()
    assert np.array_equal(o.bar_state.local_x.cpu().numpy(), T("local_x").cpu().numpy())
    o.bar_state.A.copy_(T("check_A")); o.bar_state.b.copy_(T("check_b")); o.bar_state.lamb_ij.copy_(T("check_lamb_ij"))
    ok = o.check_converge(cs[0]["thres"], cs[0]["min_dis"]).cpu().numpy()
    assert ok.tolist() == [c["check_converge"] for c in cs]
    for s, c in enumerate(cs):                                           # lambda_update takes one rho: scenario by scenario
        o1 = OBCAConsensus(S=1)
        for f, key in (("Z_bar", "Z_bar"), ("lamb_bar", "lamb_bar_before"), ("local_x", "local_x"), ("lamb_ij", "check_lamb_ij")):
            getattr(o1.bar_state, f).copy_(torch.from_numpy(A_(c[key]))[None].to(dev))
        o1.lambda_update(c["rho"])
        assert np.array_equal(o1.bar_state.lamb_bar[0].cpu().numpy(), A_(c["lamb_bar_after"])), s      # bit exact
    for f in FIELDS:
        getattr(o.bar_state, f).copy_(T("shift_before", f))
    o.iterate_next_state(o.bar_state)
    for f in FIELDS:
        assert np.array_equal(getattr(o.bar_state, f).cpu().numpy(), T("shift_after", f).cpu().numpy()), f


@pytest.mark.parametrize("ci", range(len(VEC["edge"])))
def test_convex_edge_blocks_equal_the_reference_statements_composed(ci):
    """The convex edge statement (MATLAB ...PI_antiwindup.m:276-304,383-396) evaluated by the reference's OWN Python
    functions -- cost_function_edge (casadi/PI_ADMM_class.py:145-169) with its roll-out replaced by the reference's
    linearised dynamic_update_local (:45-70) -- against the oracle's literal transcription AND the split blocks
    (H, q, a, c) the kernels iterate on; the hinge is active in part of the recorded points."""
    from oracle.problem import literal_cost_edge_convex, edge_qp_blocks, roll_matrix
    c = VEC["edge"][ci]
    p, N = _params(c), c["N"]
    xt, dual, seed = A_(c["xt"]), A_(c["dual"]), A_(c["seed"])
    pos_old, us, cost = A_(c["convex"]["pos_old"]), A_(c["convex"]["hat_u"]), c["convex"]["cost"]
    B = edge_qp_blocks(p, xt, p.spd, pos_old, dual, seed)
    M = roll_matrix(N)[1:]
    model = lambda u: (sum(0.5 * u[r] @ B["H"][r] @ u[r] + B["q"][r] @ u[r] for r in range(2))
                       + p.beta * np.sum(np.maximum(0.0, B["c"] - B["a"][0] * (M @ u[0]) - B["a"][1] * (M @ u[1]))))
    assert np.abs(us[0]).max() == 0.0
    active = 0
    for u, f in zip(us, cost):
        lit = literal_cost_edge_convex(p, u, xt, p.spd, pos_old, dual, seed)
        assert abs(lit - f) <= 1e-10 * max(1.0, abs(f))
        assert abs((f - cost[0]) - (model(u) - model(us[0]))) <= 1e-9 * max(1.0, abs(f))
        active += int(np.sum(B["c"] - B["a"][0] * (M @ u[0]) - B["a"][1] * (M @ u[1]) > 0))
    assert active > 0


# ------------------------------------------------------------------ the reference's own SCRIPT, run in the build container
RUN = json.load(open(os.path.join(HERE, "golden", "reference_main_run.json")))


def test_reference_script_run_equals_the_oracle_closed_loop():
    """tests/golden/reference_main_run.json = /root/reference/casadi/main.py executed unmodified (35 MPC steps; QP solves by the
    oracle's OSQP restatement, see tests/golden/run_reference_main.py).  The oracle's driver (python_current semantics) must
    reproduce the script's histories: this pins the OUTER LOOP -- edge_mat rule, early exit, dual update, residual sums,
    aliasing quirk, stop test, plant step, histories -- to the reference's own statements."""
    from oracle.driver import Modes, run_closed_loop
    p = Params()
    out = run_closed_loop(p, Modes.python_current(), InnerSettings.osqp_default(),
                          np.array([[-10, 0, 0], [0, 20, -np.pi / 2]], dtype=float), p.spd.astype(float), reference_trajectories(p))
    assert [int(v) for v in out["iter_his"]] == RUN["iter_his"] and len(RUN["iter_his"]) == 35
    assert int(np.sum(out["n_outer"])) == RUN["total_iter"]
    for k in ("x_vec", "theta_vec", "u_vec"):
        assert np.abs(np.array(out[k]) - A_(RUN[k])).max() <= 1e-9, k
    assert sum(sum(s["veh_iters"]) for s in out["steps"]) == sum(RUN["veh_inner"])
    assert sum(sum(s["edge_iters"]) for s in out["steps"]) == sum(RUN["edge_inner"])
    assert sum(len(s["veh_iters"]) for s in out["steps"]) == RUN["veh_solves"] and sum(len(s["edge_iters"]) for s in out["steps"]) == RUN["edge_solves"]
    assert max(RUN["iter_his"]) == 2 and RUN["edge_solves"] > 0       # the collision edge is active in part of the run
    g = json.load(open(os.path.join(HERE, "golden", "c1_python_current_default.json")))   # the frozen oracle loop the GPU tests use
    assert g["iter_his"] == RUN["iter_his"] and np.abs(A_(g["x_vec"]) - A_(RUN["x_vec"])).max() <= 1e-12


def test_c_twin_closed_loop_equals_the_reference_script_run():
    from oracle import ctwin
    import pi_admm_b200 as pa
    from pi_admm_b200.scenarios import ref_window
    p = pa.preset("python_current")
    ref = reference_trajectories(Params())
    xt = np.array([[[-10, 0, 0], [0, 20, -np.pi / 2]]], dtype=np.float64)
    for step in range(35):
        r = ctwin.step(p, xt, np.array([[4.0, 8.0]]), ref_window(ref, step, p.num_ho)[None])
        xt = r["xt_next"].copy()
        assert int(r["curr_iter"][0]) == RUN["iter_his"][step]
        assert np.abs(xt[0, :, :2].T - A_(RUN["x_vec"][step])).max() <= 1e-9 and np.abs(r["u"][0, :, 0] - A_(RUN["u_vec"][step])).max() <= 1e-9


@pytest.mark.gpu
@pytest.mark.parametrize("how", ["device_loop", "host_steps"])
def test_gpu_closed_loop_equals_the_reference_script_run(how):
    """The CUDA closed loop of config 1 (piadmm_run_closed_loop, and the drop-in driver stepping the class mirror) against the
    histories the reference's own casadi/main.py produced in the build container."""
    from pi_admm_b200.main_intersection import run
    out = run(semantics="python_current", inner="osqp_default", fast=True, device_loop=(how == "device_loop"))
    assert [int(v) for v in out["iter_his"]] == RUN["iter_his"]
    assert np.abs(np.array(out["x_vec"]) - A_(RUN["x_vec"])).max() <= 1e-6
    assert np.abs(np.array(out["u_vec"]) - A_(RUN["u_vec"])).max() <= 1e-6
    assert np.abs(np.array(out["theta_vec"]) - A_(RUN["theta_vec"])).max() <= 1e-6


@pytest.mark.parametrize("ci", range(len(VEC["old_pi_block"])))
def test_pi_old_law_equals_the_old_scripts_own_lines(ci):
    """casadi_old_PI_ADMM/main.py:127-164, exec'd verbatim for six outer iterations in a row (gains from the minimum distance of
    the non-linear roll-outs, output-then-integrate, saturation +-20, back-calculation 2*diff, global residuals, the
    last_iter_hat_pos alias, the stop test with the distance condition), against the functions oracle/driver.py's pi_old
    branch is made of -- the law the kernels implement as PIADMM_DUAL_PI_OLD."""
    from oracle.driver import Modes, pi_old_gains, pi_old_output_and_integrate, clamp_with_backcalc
    c = VEC["old_pi_block"][ci]
    md = Modes.python_old_pi(trad=c["trad"], windup_sat=float(c["windup_sat"]))
    p = Params(num_ho=c["N"], spd=A_(c["spd"]), dt=c["dt"], L=c["L"])
    xt = A_(c["xt"])
    shape = A_(c["iterations"][0]["pos_old"]).shape
    dual, integ, diffv = np.zeros(shape), np.zeros(shape), np.zeros(shape)
    last_hat = np.zeros(shape)
    aliased = False
    clipped = stopped = 0
    for it in c["iterations"]:
        pos, hat = A_(it["pos_old"]), A_(it["hat_pos_old"])
        px, py, _ = literal_dynamic_update_edge(p, xt, A_(it["primal_u"]), p.spd)
        dis = np.sqrt((px[0] - px[1]) ** 2 + (py[0] - py[1]) ** 2)
        assert np.abs(dis - A_(it["dis_vec"])).max() <= 1e-12
        dmin = dis.min()
        rho = pi_old_gains(md, dmin)[1]
        assert abs(rho - it["rho"]) <= 1e-13 * rho          # dis_vec comes out of a matrix product in the script: last-bit differences
        dual, integ = pi_old_output_and_integrate(md, dmin, pos - hat, dual, integ, diffv)
        dual, diffv = clamp_with_backcalc(md, dual)
        assert np.abs(dual - A_(it["dual_var_old"])).max() <= 1e-12 and np.abs(diffv - A_(it["diff_val"])).max() <= 1e-12
        if not c["trad"]:
            assert np.abs(integ - A_(it["sum_err"])).max() <= 1e-12
        lh = hat if aliased else last_hat                           # the script's last_iter_hat_pos IS hat_pos_old from iteration 2 on
        sk = np.sqrt(np.sum((rho * (lh - hat)) ** 2)); rk = np.sqrt(np.sum((pos - hat) ** 2))
        assert abs(sk - it["error_sk"]) <= 1e-10 * max(1.0, sk) and abs(rk - it["error_rk"]) <= 1e-10 * max(1.0, rk)
        stop = bool(rk <= c["eps_pri"] and sk <= c["eps_dual"] and dis[1] > c["dis_thres"])
        assert stop == it["stopped"]
        aliased = True
        clipped += int(np.any(np.abs(dual) == c["windup_sat"])); stopped += int(stop)
    if ci == 0:
        assert stopped > 0
    if ci == 1:
        assert clipped > 0


@pytest.mark.parametrize("ci", range(len(VEC["matlab_pi_block"])))
def test_matlab_pi_law_equals_the_m_file_statements(ci):
    """The default dual law (dual="pi": every bench configuration) is MATLAB source.  Its statements -- parameter block,
    collision test with the squared threshold and the early exit, per-edge PI update (sigmoid kP from the minimum distance,
    kI = rho, integrate THEN output), saturation +-30 with back-calculation, residual sums, stop test with dis_vec(2), the
    COPY of last_iter_hat_pos, curr_iter every iteration -- were executed from the .m text
    (matlab_old_files/ADMM_CVX_two_veh_intesection_PI_antiwindup.m:6-25,43,105-124,152-209) by tests/golden/_matlab_subset.py
    for six outer iterations per case (PI on / off, back-calculation on / off, clipping, no collision at all, an edge that
    drops out).  The functions oracle/driver.py's "pi" branch is made of must reproduce every iteration."""
    from oracle.driver import Modes, pi_gain, pi_integrate_and_output, clamp_with_backcalc
    c = VEC["matlab_pi_block"][ci]
    P = c["param"]
    md = Modes.matlab_pi(PI=int(P["PI"]), windup=int(P["windup"]), windup_sat=float(c["windup_sat"]), theta1=P["theta1"], theta2=P["theta2"])
    assert P["kI"] == P["rho"] and md.kI is None and md.thr_squared == 1 and md.dual_res_alias == 0 and md.stop_dist_test == 1
    rho, thr = float(P["rho"]), float(P["dis_thres"]) ** 2
    shape = A_(c["iterations"][0]["pos_old"])[:2].shape
    z = lambda: np.zeros(shape)
    dual, integ, diffv, hat, last_hat = [z(), z()], [z(), z()], [z(), z()], [z(), z()], [z(), z()]
    flag, curr_iter, dis2 = 0, 1, None
    for k, it in enumerate(c["iterations"]):
        pos = A_(it["pos_old"]); p = [pos[:2], pos[2:]]
        d2 = (p[0][0] - p[1][0]) ** 2 + (p[0][1] - p[1][1]) ** 2
        edge = bool(np.any(d2 < thr))
        assert int(edge) == it["edges"]
        if not edge and flag == 0:                                   # :113-114 no collision so far: leave the loop
            assert it["flag"] == 0 and not it["went_on"]
            continue
        flag = 1
        assert it["flag"] == 1
        rk = sk = 0.0
        if edge:
            hat = [A_(it["hat12"]), A_(it["hat21"])]
            dis = np.sqrt(((p[0] - p[1]) ** 2).sum(axis=0)); dis2 = dis[1]
            e = [p[0] - hat[0], p[1] - hat[1]]
            for r in range(2):
                if md.PI:
                    dual[r], integ[r] = pi_integrate_and_output(md, rho, dis.min(), e[r], integ[r], diffv[r])
                else:
                    dual[r] = dual[r] + rho * e[r]
                if md.windup:
                    dual[r], diffv[r] = clamp_with_backcalc(md, dual[r])
            if md.PI:
                assert abs(pi_gain(md, dis.min()) - it["kP"]) <= 1e-13
            sk = 2 * np.sqrt(np.sum((rho * (last_hat[0] - hat[0])) ** 2)); rk = 2 * np.sqrt(np.sum((p[0] - hat[0]) ** 2))
        for got, key in ((dual[0], "dual12"), (dual[1], "dual21"), (diffv[0], "diff12"), (diffv[1], "diff21")):
            assert np.abs(got - A_(it[key])).max() <= 1e-12, (k, key)
        if md.PI:
            assert np.abs(integ[0] - A_(it["sum_err1"])).max() <= 1e-12 and np.abs(integ[1] - A_(it["sum_err2"])).max() <= 1e-12
        assert abs(rk - it["error_rk"]) <= 1e-11 * max(1.0, rk) and abs(sk - it["error_sk"]) <= 1e-11 * max(1.0, sk)
        stop = bool(rk <= P["eps_pri"] and sk <= P["eps_dual"] and dis2 > P["dis_thres"])
        assert stop == (not it["went_on"]), k
        curr_iter = k + 1                                            # :203 and :209: either way
        assert it["curr_iter"] == curr_iter
        last_hat = [hat[0].copy(), hat[1].copy()]                    # :207 is a COPY in MATLAB (no alias quirk here)
    assert (P["PI"], P["windup"]) in ((1.0, 1.0), (0.0, 1.0), (1.0, 0.0))


@pytest.mark.parametrize("ci", range(len(VEC["matlab_functions"])))
def test_matlab_function_bodies_equal_the_oracle(ci):
    """dynamic_update_local / dynamic_update_edge (both branches), cost_function_primal and cost_function_edge of
    matlab_old_files/ADMM_CVX_two_veh_intesection_PI_antiwindup.m (:254-273, :287-302, :312-397), executed from the .m text:
    the roll-outs, the exact (H, q) of the vehicle cost and the CONVEX edge statement -- literal transcription and the split
    blocks (H, q, a, c) the kernels iterate on -- as the MATLAB statements evaluate them (hinge active in the close cases)."""
    from oracle.problem import literal_cost_edge_convex, literal_edge_linear_rollout, edge_qp_blocks, roll_matrix
    c = VEC["matlab_functions"][ci]
    P, N = c["param"], c["N"]
    spd = A_(c["spd"])
    p = Params(num_ho=N, spd=spd, dt=P["dt"], L=P["L"], dis_thres=P["dis_thres"], beta=P["beta"], Pnorm=P["Pnorm"], Pcost=P["Pcost"], rho=P["rho"])
    xt = A_(c["xt"])
    for r in c["local_rollouts"]:
        v, u = r["veh"] - 1, A_(r["u"])[0]
        xl, yl, tl = literal_dynamic_update_local(p, xt[v], u, spd[v])
        xa, ya = affine_rollout(p, xt[v], spd[v], u)
        want = [A_(a)[0] for a in r["linear"]]
        for got, w in ((xl, want[0]), (yl, want[1]), (tl, want[2]), (xa, want[0]), (ya, want[1])):
            assert np.abs(got - w).max() <= 1e-11 * max(1.0, np.abs(w).max())
        xn, yn, tn = literal_dynamic_update_edge(p, xt[v:v + 1], u[None], spd[v:v + 1])
        wantn = [A_(a)[0] for a in r["nonlinear"]]
        assert np.abs(xn[0] - wantn[0]).max() <= 1e-11 and np.abs(yn[0] - wantn[1]).max() <= 1e-11 and np.abs(tn[0] - wantn[2]).max() <= 1e-12
    ue = A_(c["edge_rollouts"]["u"])
    xn, yn, tn = literal_dynamic_update_edge(p, xt, ue, spd)
    assert np.abs(xn - A_(c["edge_rollouts"]["nonlinear"][0])).max() <= 1e-11 and np.abs(yn - A_(c["edge_rollouts"]["nonlinear"][1])).max() <= 1e-11
    xl, yl = literal_edge_linear_rollout(p, xt, ue, spd)
    assert np.abs(xl - A_(c["edge_rollouts"]["linear"][0])).max() <= 1e-11 and np.abs(yl - A_(c["edge_rollouts"]["linear"][1])).max() <= 1e-11
    # vehicle QP
    q_ = c["primal"]
    v, s0 = q_["veh"] - 1, q_["num_step"] - 1                         # MATLAB indices are 1-based
    ref = A_(c["ref_traj"])[2 * v:2 * v + 2, s0:s0 + N + 1]
    H, q = vehicle_qp(p, xt[v], spd[v], ref[0], ref[1], [A_(q_["hat"])], [A_(q_["dual"])])
    assert np.abs(H - A_(q_["H"])).max() <= 1e-8 * np.abs(H).max() and np.abs(q - A_(q_["q"])).max() <= 1e-8 * max(1.0, np.abs(q).max())
    for u, f in zip(A_(q_["u"]), q_["cost"]):
        assert abs((f - q_["f0"]) - (0.5 * u @ H @ u + q @ u)) <= 1e-9 * max(1.0, abs(f))
    # convex edge statement
    e = c["edge_cost"]
    pos_old, dual, prev, hus = A_(e["pos_old"]), A_(e["dual"]), A_(e["prev_pred_pos"]), A_(e["hat_u"])
    B = edge_qp_blocks(p, xt, spd, pos_old, dual, prev)
    M = roll_matrix(N)[1:]
    model = lambda u: (sum(0.5 * u[r] @ B["H"][r] @ u[r] + B["q"][r] @ u[r] for r in range(2))
                       + p.beta * np.sum(np.maximum(0.0, B["c"] - B["a"][0] * (M @ u[0]) - B["a"][1] * (M @ u[1]))))
    active = 0
    for u, f in zip(hus, e["cost"]):
        lit = literal_cost_edge_convex(p, u, xt, spd, pos_old, dual, prev)
        assert abs(lit - f) <= 1e-10 * max(1.0, abs(f))
        assert abs((f - e["cost"][0]) - (model(u) - model(hus[0]))) <= 1e-9 * max(1.0, abs(f))
        active += int(np.sum(B["c"] - B["a"][0] * (M @ u[0]) - B["a"][1] * (M @ u[1]) > 0))
    assert (active > 0) == (ci < 2)                                      # the third case keeps the pair 40 m apart


@pytest.mark.parametrize("ci", range(len(VEC["matlab_joint_cost"])))
def test_matlab_joint_edge_cost_equals_the_joint_oracle(ci):
    """cost_function_edge of matlab_old_files/ADMM_fmincon_multi_veh_intesection_static.m (:262-289),
    executed from its text on the linearised roll-out of ...PI_antiwindup.m:357-397 -- the convex joint statement.
    V = 2: the oracle's literal cost AND the split blocks (H_i, q_i, a, c; hinge 2 beta per pair) k_joint iterates on equal
    the MATLAB values.  V > 2: the MATLAB value equals the oracle's `as_written` form (linear indices `last_dis(1)`,
    `last_dis(2)`, :270), and differs from the per-pair form the product implements whenever the hinge is active."""
    from oracle import joint as oj
    from oracle.problem import roll_matrix
    c = VEC["matlab_joint_cost"][ci]
    V, N = c["V"], c["N"]
    spd, xt, prev = A_(c["spd"]), A_(c["xt"]), A_(c["prev_pred_pos"])
    p = Params(num_ho=N, beta=1000.0, rho=2.0, Pcost=1.0, dis_thres=2.0, dt=0.1, L=1.0)
    pos, dual = A_(c["pos_old"]).reshape(V, 2, N + 1), A_(c["dual"]).reshape(V, 2, N + 1)
    pairs = oj.all_pairs(V)
    hus, want = A_(c["hat_u"]), c["cost"]
    per_pair = []
    for u, f in zip(hus, want):
        assert abs(oj.literal_cost_joint(p, u, xt, spd, pos, dual, pairs, prev_pred_pos=prev, as_written=True) - f) <= 1e-10 * max(1.0, abs(f))
        per_pair.append(oj.literal_cost_joint(p, u, xt, spd, pos, dual, pairs, prev_pred_pos=prev))
    if V > 2:
        assert max(abs(a - b) for a, b in zip(per_pair, want)) > 1.0          # the linear-index defect is visible
        return
    assert max(abs(a - b) / max(1.0, abs(b)) for a, b in zip(per_pair, want)) <= 1e-10
    # the blocks the kernel iterates on, on the MATLAB seed points: the same quadratic + hinge
    B = oj.joint_blocks(p, xt, spd, pos, dual, pairs, prev_pred_pos=prev)
    M = roll_matrix(N)[1:]
    (i, j, ai, aj, cc), = B["pairs"]
    model = lambda u: (sum(0.5 * u[r] @ B["H"][r] @ u[r] + B["q"][r] @ u[r] for r in range(2))
                       + 2.0 * p.beta * np.sum(np.maximum(0.0, cc - ai * (M @ u[i]) - aj * (M @ u[j]))))
    for u, f in zip(hus, want):
        assert abs((f - want[0]) - (model(u) - model(hus[0]))) <= 1e-9 * max(1.0, abs(f))


# ------------------------------------------------------------------ the reference's MATLAB script, run from its own text
MRUN = json.load(open(os.path.join(HERE, "golden", "reference_matlab_run.json")))
MATLAB_PARAMS = dict(beta=1000.0, Pnorm=5.0, rho=3.5, eps_pri=0.1, eps_dual=0.1, num_ho=8)      # ...PI_antiwindup.m:6-25
MTOL = 2e-4      # the matlab_pi preset keeps the Python script's 4-decimal rounding of the seed points (casadi/main.py:48-49); MATLAB (:56-57) has none


def test_matlab_script_run_equals_the_oracle_matlab_pi_closed_loop():
    """tests/golden/reference_matlab_run.json = matlab_old_files/ADMM_CVX_two_veh_intesection_PI_antiwindup.m executed from its
    own text for all 42 MPC steps (tests/golden/run_reference_matlab.py; the two CVX blocks solved to optimality by the
    oracle's solver, two-sided steering rows).  The oracle's driver in matlab_pi semantics with the tight inner tolerance
    (solver independent) must reproduce it: identical iter_his (incl. the 4-iteration steps), the same number of vehicle and
    edge solves, trajectories within the seed-rounding tolerance.  This pins the whole OUTER LOOP of the semantics every bench
    configuration uses to the MATLAB source."""
    from oracle.driver import Modes, run_closed_loop
    r = MRUN["two_sided"]
    p = Params(**MATLAB_PARAMS)
    out = run_closed_loop(p, Modes.matlab_pi(), InnerSettings.tight(), np.array([[-10, 0, 0], [0, 20, -np.pi / 2]], dtype=float),
                          p.spd.astype(float), reference_trajectories(p))
    assert [int(v) for v in out["iter_his"]] == r["iter_his"] and len(r["iter_his"]) == 42 and max(r["iter_his"]) == 4
    assert sum(len(s["veh_iters"]) for s in out["steps"]) == r["vehicle_solves"] and sum(len(s["edge_iters"]) for s in out["steps"]) == r["edge_solves"]
    for k in ("x_vec", "theta_vec", "u_vec"):
        assert np.abs(np.array(out[k]) - A_(r[k])).max() <= MTOL, k
    assert r["max_cost_mismatch"] < 1e-10                    # the quadratic / hinge model IS the MATLAB cost functions (checked in every solve)
    g = json.load(open(os.path.join(HERE, "golden", "c1_matlab_pi_default.json")))        # the osqp_default golden the GPU tests use
    assert g["iter_his"] == r["iter_his"]
    # the script's own steering rows are one-sided (veh_u <= pi/6, :95): on this scenario that changes the run -- recorded
    assert MRUN["as_written_differs"]["max_abs_u"] > 0.1 and MRUN["as_written"]["min_u"] < -np.pi / 6 - 1e-3


def test_c_twin_equals_the_matlab_script_run():
    from oracle import ctwin
    import pi_admm_b200 as pa
    from pi_admm_b200.scenarios import ref_window
    r = MRUN["two_sided"]
    p = pa.inner_tight(pa.preset("matlab_pi").update(**{k: (int(v) if k == "num_ho" else v) for k, v in MATLAB_PARAMS.items()}))
    ref = reference_trajectories(Params())
    xt = np.array([[[-10, 0, 0], [0, 20, -np.pi / 2]]], dtype=np.float64)
    for step in range(42):
        o = ctwin.step(p, xt, np.array([[4.0, 8.0]]), ref_window(ref, step, p.num_ho)[None])
        xt = o["xt_next"].copy()
        assert int(o["curr_iter"][0]) == r["iter_his"][step], step
        assert np.abs(xt[0, :, :2].T - A_(r["x_vec"][step])).max() <= MTOL and np.abs(o["u"][0, :, 0] - A_(r["u_vec"][step])).max() <= MTOL


@pytest.mark.gpu
def test_gpu_closed_loop_equals_the_matlab_script_run():
    """The CUDA closed loop in matlab_pi semantics (tight inner tolerance): piadmm_run_closed_loop, all 42 MPC steps on the
    device, against the histories of the MATLAB script itself."""
    from pi_admm_b200.solver import PIADMMSolver
    from pi_admm_b200.scenarios import c1_initial_state
    from util import make_params
    r = MRUN["two_sided"]
    p = make_params("matlab_pi", "tight")
    assert p.num_ho == 8 and p.beta == 1000.0 and p.rho == 3.5 and p.Pnorm == 5.0 and p.eps_pri == 0.1
    ref_traj = reference_trajectories(Params())
    sol = PIADMMSolver(p, S=1, V=2)
    sol.setup(c1_initial_state()[None], np.array([[4.0, 8.0]]), np.zeros((1, 2, 2, p.num_ho + 1)))
    sol.set_reference(ref_traj.reshape(1, 2, 2, -1))
    h = {k: v.cpu().numpy() for k, v in sol.run_closed_loop(0, 42).items()}
    sol.close()
    assert h["iter_his"][:, 0].tolist() == r["iter_his"]
    assert np.abs(h["x_vec"][:, 0].transpose(0, 2, 1) - A_(r["x_vec"])).max() <= MTOL
    assert np.abs(h["theta_vec"][:, 0] - A_(r["theta_vec"])).max() <= MTOL
    assert np.abs(h["u_vec"][:, 0] - A_(r["u_vec"])).max() <= MTOL
