"""GPU parity of the COUPLED scenes (BASELINE configs 3 and 5): the persistent cooperative kernel, the candidate-row
collision detection, the round-1 phase path and full pair-test kernels (A/B switches), slot overflow, convergence, and
the vehicle-sharded multi-GPU step -- all against the oracle's C twin on identical inputs.  Integer outputs (iteration
counts, edge lists, slot tables, status bits) must be identical, floating point within 1e-6 relative
(north_star: "within 1e-6 relative in fp64")."""
import os

import numpy as np
import pytest
import torch

import bench
from pi_admm_b200.scenarios import c3_intersection, c5_grid
from pi_admm_b200.solver import PIADMMSolver
from pi_admm_b200.parallel import VehicleShardedSolver, spatial_order, iterate_lockstep, shard_range
from oracle import ctwin
from util import make_params, assert_same, rel_err

pytestmark = pytest.mark.gpu


def _threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def _run(p, xt, spd, refwin, D, fused=None):
    sol = PIADMMSolver(p, S=xt.shape[0], V=xt.shape[1], D=D)
    sol.setup(xt, spd, refwin)
    sol.iterate(fused=fused)
    launches = sol.launch_count
    sol.plant_step()
    out = sol.to_numpy()
    out["_launches"] = launches
    sol.close()
    return out


def _assert_edges(got, ref):
    for s in range(len(ref["edge_n"])):
        n = int(ref["edge_n"][s])
        assert np.array_equal(got["edge_ij"][s, :n], ref["edge_ij"][s, :n])
        assert np.array_equal(got["edge_slot"][s, :n], ref["edge_slot"][s, :n])


def converged(r, p, s=0):
    """The reference's two ways out of the outer loop (casadi/main.py:115-118,165-178) -- not the iteration cap."""
    n = int(r["n_outer"][s])
    if int(r["status"][s]) & 8:
        return False
    if int(r["edge_n"][s]) == 0:
        return True                                      # empty edge list: residuals are exactly 0
    return r["err_rk"][s, n - 1] <= p.eps_pri and r["err_sk"][s, n - 1] <= p.eps_dual


def test_c5_grid_4096_vehicles_matches_twin():
    """BASELINE config 5 at full size on ONE GPU: 4096 coupled vehicles, N=30, D=8, bench parameters (outer cap 100),
    the cooperative kernel, against the C twin on all host threads."""
    (xt, spd, refwin), V, N, D = bench.coupled_scene("c5")
    p = bench.coupled_params(N)
    ref = ctwin.step(p, xt, spd, refwin, D=D, nthreads=_threads())
    got = _run(p, xt, spd, refwin, D)
    assert ref["n_edge_solves"][0] > 300 and ref["status"][0] == 0
    _assert_edges(got, ref)
    assert_same(got, ref, what="c5")
    assert converged(ref, p) and converged(got, p)
    assert got["_launches"] == 3                         # candidate build (k_disc, k_cand_direct) + ONE cooperative launch for the whole outer loop


@pytest.mark.parametrize("name,iters,exit_edges", [("c5x4", 4, 0), ("c5x8", 14, None)])
def test_capacity_scenes_match_twin(name, iters, exit_edges):
    """The largest scenes the bench measures: the config-5 generator at 16384 and 32768 coupled vehicles (N=30, D=8) on ONE
    GPU against the C twin on all host threads -- outer iterations, edge lists and slots, solve and inner-iteration counts
    identical, floats to 1e-6.  (The twin needs ~1 s / ~15 s of 16 cores per step; the GPU ~1 ms / ~4 ms.)"""
    from pi_admm_b200.parallel import spatial_order
    (xt, spd, refwin), V, N, D = bench.coupled_scene(name)
    # the bench's vehicle numbering (spatial strips): the residual sums take the veh1 side of every edge
    # (casadi/main.py:165-173), so the stop iteration depends on the numbering -- 32768 vehicles pass the stop test after 14
    # iterations in this order and run into the cap of 100 in the generator's lane-major order
    pm = spatial_order(xt)
    xt, spd, refwin = (np.ascontiguousarray(np.broadcast_to(a, (1, V) + np.shape(a)[2:])[:, pm]) for a in (xt, spd, refwin))
    p = bench.coupled_params(N)
    ref = ctwin.step(p, xt, spd, refwin, D=D, nthreads=_threads())
    got = _run(p, xt, spd, refwin, D)
    assert int(ref["n_outer"][0]) == iters and ref["status"][0] == 0
    if exit_edges is not None:
        assert int(ref["edge_n"][0]) == exit_edges
    _assert_edges(got, ref)
    assert_same(got, ref, what=name)
    assert converged(ref, p) and converged(got, p)
    assert got["_launches"] == 3


@pytest.mark.parametrize("detect", ["cand", "matrix_warp", "matrix_thread"])
@pytest.mark.parametrize("coop", ["1", "0"])
def test_midsize_grid_every_detection_path(monkeypatch, detect, coop):
    """640 vehicles (V * ceil(V/32) = 12800 words: past the 9472-word switch of the round-1 pair test, 256-thread
    k_edges launch): candidate rows and both full pair-test kernels, cooperative kernel and phase kernels."""
    monkeypatch.setenv("PIADMM_DETECT", detect)
    monkeypatch.setenv("PIADMM_COOP", coop)
    xt, spd, refwin = c5_grid(lanes=10, per_lane=32, N=30, seed=11, clear=0.0)     # incl. pairs that start inside dis_thres
    p = bench.coupled_params(30).update(iter_num=6)
    ref = ctwin.step(p, xt, spd, refwin, D=8, nthreads=_threads())
    assert ref["n_edge_solves"][0] > 20
    got = _run(p, xt, spd, refwin, 8)
    _assert_edges(got, ref)
    assert_same(got, ref, what=f"grid640 {detect} coop={coop}")
    if detect == "cand" and coop == "1":
        assert got["_launches"] == 3


@pytest.mark.parametrize("V,N,D,r0,gap", [(64, 30, 8, 12.0, 6.0), (32, 15, 8, 8.0, 4.0), (12, 20, 4, 12.0, 6.0)])
def test_cooperative_kernel_equals_phase_kernels_bit_for_bit(monkeypatch, V, N, D, r0, gap):
    xt, spd, refwin = c3_intersection(V=V, N=N, r0=r0, gap=gap)
    p = bench.coupled_params(N)
    a = _run(p, xt, spd, refwin, D)
    monkeypatch.setenv("PIADMM_COOP", "0")
    b = _run(p, xt, spd, refwin, D)
    assert a["_launches"] == 3 and b["_launches"] > 5
    for k in ("u", "pos", "hat", "dual", "integ", "diff", "err_rk", "err_sk", "n_outer", "curr_iter", "nbr", "edge_n",
              "veh_inner", "edge_inner", "status", "xt_next", "edge_ij", "edge_slot"):
        assert np.array_equal(a[k], b[k]), k
    ref = ctwin.step(p, xt, spd, refwin, D=D)
    assert_same(a, ref, what=f"coop c3 V={V}")


@pytest.mark.parametrize("coop", ["1", "0"])
def test_x_memo_is_exact_on_the_gpu(monkeypatch, coop):
    """x_memo=1 (skip vehicle QPs whose inputs did not change) gives bit for bit what x_memo=0 gives, with fewer solves."""
    monkeypatch.setenv("PIADMM_COOP", coop)
    xt, spd, refwin = c5_grid(lanes=8, per_lane=16, N=30, seed=2, clear=0.0)
    p1 = bench.coupled_params(30).update(iter_num=6)
    p0 = bench.coupled_params(30).update(iter_num=6, x_memo=0)
    a, b = _run(p1, xt, spd, refwin, 8), _run(p0, xt, spd, refwin, 8)
    for k in ("u", "pos", "hat", "dual", "err_rk", "err_sk", "n_outer", "edge_n", "nbr", "edge_inner", "xt_next"):
        assert np.array_equal(a[k], b[k]), k
    assert a["n_veh_solves"][0] < b["n_veh_solves"][0]
    assert_same(a, ctwin.step(p1, xt, spd, refwin, D=8), what="memo on")
    assert_same(b, ctwin.step(p0, xt, spd, refwin, D=8), what="memo off")


@pytest.mark.parametrize("coop", ["1", "0"])
def test_dense_scene_candidate_list_overflow(monkeypatch, coop):
    """96 vehicles packed around the intersection: most vehicles have more than PIADMM_CAND_CAP = 64 candidates (their
    slot phase walks the candidate bit row), the outer ones fewer (candidate list) -- both in one scene; neighbour
    tables overflow too (D = 8), as in the oracle."""
    monkeypatch.setenv("PIADMM_COOP", coop)
    xt, spd, refwin = c3_intersection(V=96, N=30, r0=10.0, gap=1.5)
    p = bench.coupled_params(30).update(iter_num=3)
    ref = ctwin.step(p, xt, spd, refwin, D=8, nthreads=_threads())
    assert ref["n_edge_solves"][0] > 50
    got = _run(p, xt, spd, refwin, 8)
    _assert_edges(got, ref)
    assert_same(got, ref, what=f"dense96 coop={coop}")


def test_cooperative_kernel_ragged_batch(monkeypatch):
    """S = 3 scenes of 5 vehicles (odd V, N=15: a warp holds groups of two scenarios), the middle one collision free."""
    from test_gpu_parity import _ragged_scene, _generic_params
    p = _generic_params(15)
    xt, spd, refwin = _ragged_scene(5, 15, seeds=(3, 4, 5), far=1)
    ref = ctwin.step(p, xt, spd, refwin, D=4)
    got = _run(p, xt, spd, refwin, 4)
    assert got["_launches"] == 3
    _assert_edges(got, ref)
    assert_same(got, ref, what="coop ragged")


ALL_KEYS = ("u", "pos", "hat", "dual", "integ", "diff", "err_rk", "err_sk", "n_outer", "curr_iter", "nbr", "edge_n", "veh_inner",
            "edge_inner", "status", "xt_next", "edge_ij", "edge_slot", "n_veh_solves", "n_edge_solves")


@pytest.mark.parametrize("V,N", [(640, 30), (40, 15)])
def test_cooperative_kernel_register_and_occupancy_builds_agree(monkeypatch, V, N):
    """k_coop exists in two builds (4 CTAs per SM / ~250 registers for small scenes, 6 CTAs per SM / 168 registers for scenes
    of several waves; launch_coop picks by size): the same arithmetic, bit-identical results, both horizons' kernels."""
    if V == 640:
        xt, spd, refwin = c5_grid(lanes=10, per_lane=32, N=N, seed=11, clear=0.0)
    else:
        xt, spd, refwin = c3_intersection(V=V, N=N, r0=9.0, gap=5.0)
    p = bench.coupled_params(N).update(iter_num=8)
    monkeypatch.setenv("PIADMM_COOP_OCC", "lo")
    a = _run(p, xt, spd, refwin, 8)
    monkeypatch.setenv("PIADMM_COOP_OCC", "hi")
    b = _run(p, xt, spd, refwin, 8)
    assert int(a["n_outer"].max()) >= 2 and int(a["n_edge_solves"][0]) > 0
    for k in ALL_KEYS:
        assert np.array_equal(a[k], b[k]), k


@pytest.mark.parametrize("scene", ["grid640", "dense96", "ragged", "c3"])
def test_slot_gate_and_direct_candidate_lists_are_exact(monkeypatch, scene):
    """The cooperative kernel's slot phase only serves the vehicles marked by the x phase (veh_need), its edge phase adds
    chunk totals, rejects candidates on the two trajectory boxes before the exact point test, and the candidate lists
    come straight from the reach discs (k_cand_direct): bit for bit what the un-gated slot phase (PIADMM_SLOT_GATE=0)
    without the boxes (PIADMM_TBOX=0) on lists read off the candidate bit matrix (PIADMM_CAND=matrix) gives -- on a
    scene with several outer iterations, one with list overflow (gate off by itself), a ragged batch with a scenario that
    finishes early, and BASELINE config 3 (stop test with active edges, MATLAB :202 distance flag in the decision)."""
    if scene == "grid640":
        xt, spd, refwin = c5_grid(lanes=10, per_lane=32, N=30, seed=11, clear=0.0)
        p, D = bench.coupled_params(30).update(iter_num=8), 8
    elif scene == "dense96":
        xt, spd, refwin = c3_intersection(V=96, N=30, r0=10.0, gap=1.5)
        p, D = bench.coupled_params(30).update(iter_num=3), 8
    elif scene == "ragged":
        from test_gpu_parity import _ragged_scene, _generic_params
        p, D = _generic_params(15), 4
        xt, spd, refwin = _ragged_scene(5, 15, seeds=(3, 4, 5), far=1)
    else:
        xt, spd, refwin = c3_intersection(V=64, N=30, r0=12.0, gap=6.0)          # BASELINE config 3: 7 outer iterations
        p, D = bench.coupled_params(30), 8
    a = _run(p, xt, spd, refwin, D)
    monkeypatch.setenv("PIADMM_SLOT_GATE", "0")
    b = _run(p, xt, spd, refwin, D)
    monkeypatch.setenv("PIADMM_CAND", "matrix")
    monkeypatch.setenv("PIADMM_TBOX", "0")
    c = _run(p, xt, spd, refwin, D)
    assert a["_launches"] == 3 and b["_launches"] == 3 and c["_launches"] == 4
    assert int(a["n_outer"].max()) >= 2
    for k in ALL_KEYS:
        assert np.array_equal(a[k], b[k]), ("gate", k)
        assert np.array_equal(a[k], c[k]), ("direct lists / trajectory boxes", k)
    assert_same(a, ctwin.step(p, xt, spd, refwin, D=D, nthreads=_threads()), what=f"gate {scene}")


@pytest.mark.parametrize("coop", ["1", "0"])
def test_slot_overflow_star_equals_twin(monkeypatch, coop):
    """D = 1, V = 4, vehicle 3 meets vehicles 0, 1 and 2: only the first pair gets slots on both sides, the
    others are dropped with PIADMM_ST_SLOT_OVERFLOW -- edge list, status and histories as the oracle's."""
    monkeypatch.setenv("PIADMM_COOP", coop)
    N = 30
    p = make_params("matlab_pi", "default", num_ho=N, beta=10.0, Pnorm=1.0, rho=2.0, eps_pri=0.5, eps_dual=0.5,
                    al_all=0, iter_num=5)
    xt = np.zeros((1, 4, 3)); spd = np.array([[4.0, 4.0, 4.0, 8.0]])
    for j, x in enumerate((6.0, 12.0, 18.0)):
        xt[0, j] = (x, 4.0 * x / 8.0, -np.pi / 2)            # reaches y = 0 when vehicle 3 reaches x
    k = np.arange(N + 1)
    refwin = np.zeros((1, 4, 2, N + 1))
    refwin[0, :, 0] = xt[0, :, 0:1] + k * 0.1 * spd[0, :, None] * np.cos(xt[0, :, 2:3])
    refwin[0, :, 1] = xt[0, :, 1:2] + k * 0.1 * spd[0, :, None] * np.sin(xt[0, :, 2:3])
    ref = ctwin.step(p, xt, spd, refwin, D=1)
    assert ref["status"][0] & 4 and ref["n_edge_solves"][0] >= 1
    got = _run(p, xt, spd, refwin, 1)
    assert got["edge_n"][0] <= 2                              # E = V*D/2: never past the edge arrays
    _assert_edges(got, ref)
    assert_same(got, ref, what=f"star overflow coop={coop}")


def test_coupled_configs_converge():
    """config 3 and a config-5-shaped scene leave the outer loop through the stop test or the
    empty-edge rule within the reference's cap of 100 (casadi/PI_ADMM_class.py:25), oracle and GPU alike."""
    for name in ("c3", "c5"):
        (xt, spd, refwin), V, N, D = bench.coupled_scene(name)
        if name == "c5":
            xt, spd, refwin = c5_grid(lanes=16, per_lane=32, N=N)      # 1024 vehicles
            V = xt.shape[1]
        p = bench.coupled_params(N)
        assert p.iter_num == 100
        ref = ctwin.step(p, xt, spd, refwin, D=D, nthreads=_threads())
        got = _run(p, xt, spd, refwin, D)
        assert converged(ref, p), (name, ref["err_rk"][0, :int(ref["n_outer"][0])])
        assert converged(got, p)
        assert np.array_equal(got["n_outer"], ref["n_outer"]) and ref["n_edge_solves"][0] > 0
        assert_same(got, ref, what=f"convergence {name}")


@pytest.mark.parametrize("scene", ["grid", "intersection"])
def test_spatial_order_sharded_lockstep_equals_unsharded(scene):
    """The sharded algorithm in the spatial numbering (contiguous strips of spatial_order) with 2 and 4 'ranks' emulated on
    one GPU == the unsharded cooperative run in the same numbering, bit for bit in the stop decisions, 1e-6 elsewhere.
    The intersection scene has most of its edges ACROSS the cuts (every vehicle heads for the origin)."""
    if scene == "grid":
        xt, spd, refwin = c5_grid(lanes=6, per_lane=16, N=30, seed=5, clear=0.0)
    else:
        xt, spd, refwin = c3_intersection(V=32, N=30, r0=8.0, gap=4.0)
    V = xt.shape[1]
    perm = spatial_order(xt)
    xt, spd, refwin = xt[:, perm], spd[:, perm], refwin[:, perm]
    p = bench.coupled_params(30).update(iter_num=8)
    ref = ctwin.step(p, xt, spd, refwin, D=8)
    one = _run(p, xt, spd, refwin, 8)
    assert_same(one, ref, what="spatial single")
    for world in (2, 4):
        sols = []
        for r in range(world):
            s = PIADMMSolver(p, S=1, V=V, D=8)
            s.setup(xt, spd, refwin)
            s.set_ownership(*shard_range(V, r, world))
            sols.append(s)
        iterate_lockstep(sols)
        for s in sols:
            got = s.to_numpy()
            lo, hi = s.own
            assert np.array_equal(got["n_outer"], ref["n_outer"]) and np.array_equal(got["nbr"], ref["nbr"])
            for k in ("u", "pos", "err_rk", "err_sk"):
                assert rel_err(got[k], ref[k]) <= 1e-6, (world, k)
            for k in ("hat", "dual"):
                assert rel_err(got[k][:, lo:hi], ref[k][:, lo:hi]) <= 1e-6, (world, k)
            s.close()


def test_sharded_handle_without_exchange_is_rejected():
    """piadmm_step on a handle with a partial ownership range and no exchange must fail, not decide on
    rank-local residual sums."""
    import pi_admm_b200 as pa
    xt, spd, refwin = c3_intersection(V=16, N=15, r0=7.0, gap=5.0)
    p = bench.coupled_params(15)
    s = PIADMMSolver(p, S=1, V=16, D=8)
    s.setup(xt, spd, refwin)
    s.set_ownership(0, 8)
    with pytest.raises(pa.PIADMMError):
        s.iterate()
    s.close()


# ------------------------------------------------------------------ real multi-process run (needs >= 2 GPUs)
def _mp_worker(rank, world, port, exchange, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        xt, spd, refwin = c5_grid(lanes=10, per_lane=32, N=30, seed=11)
        V = xt.shape[1]
        p = bench.coupled_params(30)
        vs = VehicleShardedSolver(p, V, 8, device=rank, exchange=exchange)
        vs.setup(xt, spd, refwin)
        for _ in range(2):                          # twice: epochs / flags carry over between steps
            vs.iterate()
        vs.plant_step()
        out = vs.results()
        out["exchange"] = vs.exchange
        q.put((rank, out))
        dist.barrier()
        vs.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("exchange", ["peer", "nccl"])
def test_vehicle_sharded_two_ranks_equal_twin(exchange):
    """The real thing on 2 GPUs: torch.distributed rendezvous, peer-mapped exchange blocks (or the NCCL baseline), the
    persistent kernel on both ranks; every rank's results equal the C twin and each other bit for bit."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (the driver's multi-GPU bench line carries the same check: coupled.parity)")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_mp_worker, args=(r, 2, port, exchange, q)) for r in range(2)]
    for pr in procs:
        pr.start()
    res = dict(q.get(timeout=600) for _ in range(2))
    for pr in procs:
        pr.join(timeout=120)
        assert pr.exitcode == 0
    xt, spd, refwin = c5_grid(lanes=10, per_lane=32, N=30, seed=11)
    p = bench.coupled_params(30)
    ref = ctwin.step(p, xt, spd, refwin, D=8, nthreads=_threads())
    for r in (0, 1):
        assert res[r]["exchange"] == exchange
        assert np.array_equal(res[r]["n_outer"], ref["n_outer"]) and np.array_equal(res[r]["edge_n"], ref["edge_n"])
        for k in ("u", "pos", "xt_next", "err_rk", "err_sk"):
            assert rel_err(res[r][k], ref[k]) <= 1e-6, (r, k)
    for k in ("u", "pos", "err_rk", "err_sk", "n_outer"):
        assert np.array_equal(res[0][k], res[1][k]), k
