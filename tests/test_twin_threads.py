"""The C twin (oracle/piadmm_ref.c) must not depend on the number of OpenMP threads: batches run scenarios in parallel,
coupled scenes run the vehicles / pair-test rows / edges of one scenario in parallel, scratch is per thread.  (A
serialised nested region reports thread 0 -- a shared scratch pair there corrupted results on boxes with many cores.)"""
import numpy as np
import pytest

import bench
from oracle import ctwin
from pi_admm_b200.scenarios import c4_scenarios, c3_intersection, c5_grid

KEYS = ("u", "pos", "hat", "dual", "err_rk", "err_sk", "n_outer", "curr_iter", "veh_inner", "edge_inner", "n_veh_solves",
        "n_edge_solves", "status", "edge_n", "nbr")


def _same(a, b):
    for k in KEYS:
        assert np.array_equal(a[k], b[k]), k


@pytest.mark.parametrize("N", [15, 30])
def test_batch_is_thread_count_invariant(N):
    p = bench.bench_params(N)
    xt, spd, refwin = c4_scenarios(S=192, N=N, seed=3)
    a = {k: v.copy() for k, v in ctwin.step(p, xt, spd, refwin, nthreads=1).items()}
    for nt in (3, 8, 16):
        _same(ctwin.step(p, xt, spd, refwin, nthreads=nt), a)
    assert a["n_edge_solves"].sum() > 0


def test_coupled_scene_is_thread_count_invariant():
    xt, spd, refwin = c3_intersection(V=64, N=30)
    p = bench.coupled_params(30)
    a = {k: v.copy() for k, v in ctwin.step(p, xt, spd, refwin, D=8, nthreads=1).items()}
    for nt in (5, 16):
        _same(ctwin.step(p, xt, spd, refwin, D=8, nthreads=nt), a)


def test_coupled_configs_converge_in_the_oracle():
    """config 3 and a config-5-shaped scene leave the outer loop through the stop test (with active
    edges) or the empty-edge rule within the reference's cap of 100 -- never through the cap."""
    for name, scene in (("c3", c3_intersection(V=64, N=30)), ("c5", c5_grid(lanes=16, per_lane=32, N=30))):
        p = bench.coupled_params(30)
        r = ctwin.step(p, *scene, D=8, nthreads=8)
        n = int(r["n_outer"][0])
        assert p.iter_num == 100 and n < 100 and not (int(r["status"][0]) & 8), name
        if int(r["edge_n"][0]) > 0:
            assert r["err_rk"][0, n - 1] <= p.eps_pri and r["err_sk"][0, n - 1] <= p.eps_dual, name
        assert r["n_edge_solves"][0] > 0


def test_capacity_scene_generator_and_convergence():
    """bench.py's capacity scenes: shapes, no pair inside dis_thres at t = 0 (the generator's rejection rule), and the
    16384-vehicle scene leaves the outer loop through the empty-edge rule after 4 iterations in the oracle."""
    for name, V in (("c5x4", 16384), ("c5x8", 32768)):
        (xt, spd, refwin), v, N, D = bench.coupled_scene(name)
        assert v == V and xt.shape == (1, V, 3) and refwin.shape == (1, V, 2, N + 1) and (N, D) == (30, 8)
        east, south = xt[0, :V // 2], xt[0, V // 2:]
        # east-bound vehicles sit on y = 20 r, south-bound ones on x = 20 c: the closest pair of a crossing is found per lane pair
        dmin = np.inf
        for r in range(0, V // 2, V // 2 // 128 * 16):                    # a sample of the lanes (every 16th)
            e = east[r:r + V // 2 // 128]
            d = np.hypot(e[:, None, 0] - south[None, :, 0], e[:, None, 1] - south[None, :, 1])
            dmin = min(dmin, d.min())
        assert dmin > 2.0
    (xt, spd, refwin), V, N, D = bench.coupled_scene("c5x4")
    p = bench.coupled_params(N)
    r = ctwin.step(p, xt, spd, refwin, D=D, nthreads=8)
    assert int(r["n_outer"][0]) == 4 and int(r["edge_n"][0]) == 0 and int(r["status"][0]) == 0 and r["n_edge_solves"][0] > 1000


def test_x_memo_is_exact():
    """x_memo skips vehicle QPs whose inputs did not change: identical results, fewer executed solves."""
    xt, spd, refwin = c5_grid(lanes=8, per_lane=16, N=30, seed=2, clear=0.0)
    p1 = bench.coupled_params(30).update(iter_num=6)
    p0 = bench.coupled_params(30).update(iter_num=6, x_memo=0)
    a = {k: v.copy() for k, v in ctwin.step(p1, xt, spd, refwin, D=8, nthreads=4).items()}
    b = ctwin.step(p0, xt, spd, refwin, D=8, nthreads=4)
    for k in ("u", "pos", "hat", "dual", "err_rk", "err_sk", "n_outer", "edge_n", "nbr", "edge_inner"):
        assert np.array_equal(a[k], b[k]), k
    assert a["n_outer"][0] >= 2 and a["n_veh_solves"][0] < b["n_veh_solves"][0] == xt.shape[1] * a["n_outer"][0]
