"""Config 2 consensus pieces: oracle vs golden fixture on CPU, CUDA path
(csrc/obca_consensus.cu through include/obca_consensus.h) vs oracle on the GPU."""
import json
import os

import numpy as np
import pytest

import obca_synth
from oracle import obca_consensus as oc

HERE = os.path.dirname(os.path.abspath(__file__))
FIELDS = oc.MidState.FIELDS


def golden():
    with open(os.path.join(HERE, "golden", "c2_consensus.json")) as f:
        return json.load(f)


# ------------------------------------------------------------------ CPU: the oracle itself
def test_mid_state_matches_reference_shapes_and_warm_start():
    st = oc.MidState()
    assert st.Z_bar.shape == (2, 7, 9) and st.A.shape == (2, 7, 4, 2) and st.b.shape == (2, 7, 4)
    assert st.lamb_bar.shape == (2, 7, 9) and np.all(st.lamb_bar == 1e-3)
    assert st.lamb_ij.shape == (2, 7, 4) and st.lamb_ij[0, 0, 0] == 1.49 and st.lamb_ij[1, 6, 3] == 0.755
    assert st.local_x.shape == (2, 7, 5)


def test_oracle_loops_equal_vectorised_forms():
    seq = obca_synth.sequence(5, 3, seed=11)
    for step in seq:
        lb = np.random.default_rng(0).normal(size=(5, 2, 7, 9))
        want = oc.lambda_update_batched(step["local_x"], step["lamb_ij"], step["Z_bar"], lb, 0.7)
        for s in range(5):
            st = oc.MidState()
            st.local_x, st.lamb_ij, st.Z_bar, st.lamb_bar = step["local_x"][s], step["lamb_ij"][s], step["Z_bar"][s], lb[s].copy()
            oc.lambda_update(st, 0.7)
            assert np.array_equal(st.lamb_bar, want[s])
            st.A, st.b = step["A"][s], step["b"][s]
            sh = oc.iterate_next_state(st)
            assert np.array_equal(sh.A, oc.shift_batched(step["A"])[s])
            assert np.array_equal(sh.A[:, -1], sh.A[:, -2])


def test_oracle_reproduces_golden():
    g = golden()
    seq = obca_synth.sequence(g["S"], g["n_iter_max"], seed=g["seed"])
    r = obca_synth.run_oracle(seq, g["S"])
    assert r["n_iter"].tolist() == g["n_iter"] and r["stop"].tolist() == g["stop"]
    assert 51 in g["n_iter"] and min(g["n_iter"]) < 20          # both exits of :86 are covered
    assert r["check"] == g["check"] and any(any(c) for c in g["check"])
    for s in range(g["S"]):
        assert np.allclose(r["primal"][s], g["primal"][s], rtol=1e-12, atol=0)
        assert np.allclose(r["dual"][s], g["dual"][s], rtol=1e-12, atol=0)
    assert np.allclose(r["prev"]["lamb_bar"], g["prev_lamb_bar"], rtol=1e-13, atol=0)
    assert np.allclose(r["cur_lamb_bar"], g["cur_lamb_bar"], rtol=1e-13, atol=0)


# ------------------------------------------------------------------ GPU: CUDA path vs oracle
def _solver(S):
    from pi_admm_b200.obca_consensus import OBCAConsensus
    return OBCAConsensus(S=S)


def _t(a):
    import torch
    return torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float64).cuda()


def _load(state, step):
    for f in ("local_x", "lamb_ij", "A", "b", "Z_bar"):
        getattr(state, f).copy_(_t(step[f]))


@pytest.mark.gpu
def test_gpu_mid_state_warm_start():
    o = _solver(3)
    ref = oc.MidState()
    for f in FIELDS:
        got = getattr(o.bar_state, f).cpu().numpy()
        assert got.shape == (3,) + getattr(ref, f).shape
        assert all(np.array_equal(got[s], getattr(ref, f)) for s in range(3)), f


@pytest.mark.gpu
def test_gpu_lambda_update_bit_exact():
    S = 33
    step = obca_synth.sequence(S, 1, seed=5)[0]
    o = _solver(S)
    _load(o.bar_state, step)
    lb = np.random.default_rng(1).normal(size=(S, 2, 7, 9))
    o.bar_state.lamb_bar.copy_(_t(lb))
    o.lambda_update(1.0)
    want = oc.lambda_update_batched(step["local_x"], step["lamb_ij"], step["Z_bar"], lb, 1.0)
    assert np.array_equal(o.bar_state.lamb_bar.cpu().numpy(), want)
    o.lambda_update(0.37)
    want = oc.lambda_update_batched(step["local_x"], step["lamb_ij"], step["Z_bar"], want, 0.37)
    assert np.array_equal(o.bar_state.lamb_bar.cpu().numpy(), want)
    o.lambda_update(0.0)                                         # rho = 0 leaves the multipliers alone
    assert np.array_equal(o.bar_state.lamb_bar.cpu().numpy(), want)


@pytest.mark.gpu
def test_gpu_iterate_next_state_bit_exact_and_saturates():
    S = 9
    step = obca_synth.sequence(S, 1, seed=6)[0]
    o = _solver(S)
    _load(o.bar_state, step)
    before = {f: getattr(o.bar_state, f).cpu().numpy() for f in FIELDS}
    ret = o.iterate_next_state(o.bar_state)
    assert ret is o.bar_state
    for f in FIELDS:
        assert np.array_equal(getattr(o.bar_state, f).cpu().numpy(), oc.shift_batched(before[f])), f
    for _ in range(7):
        o.iterate_next_state(o.bar_state)
    for f in FIELDS:                                             # after >= H-1 shifts every row is the last one
        a = getattr(o.bar_state, f).cpu().numpy()
        assert np.array_equal(a, np.broadcast_to(before[f][:, :, -1:], a.shape)), f


@pytest.mark.gpu
def test_gpu_residuals_and_check_converge():
    S = 17
    seq = obca_synth.sequence(S, 30, seed=7)
    o = _solver(S)
    prev = o.create_bar_state()
    for t in (0, 29):
        _load(o.bar_state, seq[t])
        lb = np.random.default_rng(t).normal(size=(S, 2, 7, 9))
        o.bar_state.lamb_bar.copy_(_t(lb))
        c1, c0 = seq[t]["ctrl"], seq[max(t - 1, 0)]["ctrl"] * 0.5
        pr, du = o.residuals(_t(c1), _t(c0), prev)
        ok = o.check_converge(0.1, 0.1).cpu().numpy()
        for s in range(S):
            p, d = oc.residuals(c1[s], c0[s], lb[s], np.full((2, 7, 9), 1e-3))
            assert abs(pr[s].item() - p) <= 1e-12 * p and abs(du[s].item() - d) <= 1e-12 * d
            st = oc.MidState()
            st.A, st.b, st.lamb_ij = seq[t]["A"][s], seq[t]["b"][s], seq[t]["lamb_ij"][s]
            assert bool(ok[s]) == oc.check_converge(st, 0.1, 0.1), (t, s)
    assert ok.any() and not ok.all()


def _run_gpu_loop(S, seq, max_iter=50):
    import torch
    o = _solver(S)
    o.max_iter = max_iter
    prev = o.create_bar_state()
    o.bar_state = prev.clone()                                   # optimizer.bar_state = deepcopy(bar_state_prev)
    ctrl_prev = torch.zeros((S, 2, 14), dtype=torch.float64, device="cuda")
    o.begin_time_step()
    primal, dual, check = [[] for _ in range(S)], [[] for _ in range(S)], [[] for _ in range(S)]
    for step in seq:
        active = (o.stop == 0).cpu().numpy()
        if not active.any():
            break
        _load(o.bar_state, step)
        ck = o.check_converge(0.1, 0.1).cpu().numpy()
        o.consensus_tail(prev, _t(step["ctrl"]), ctrl_prev, rho=1.0)
        p, d = o.primal_res.cpu().numpy(), o.dual_res.cpu().numpy()
        for s in np.flatnonzero(active):
            primal[s].append(float(p[s])); dual[s].append(float(d[s])); check[s].append(bool(ck[s]))
    return o, prev, primal, dual, check


@pytest.mark.gpu
def test_gpu_consensus_loop_matches_golden_and_oracle():
    g = golden()
    S = g["S"]
    seq = obca_synth.sequence(S, g["n_iter_max"], seed=g["seed"])
    o, prev, primal, dual, check = _run_gpu_loop(S, seq)
    assert o.n_iter.cpu().tolist() == g["n_iter"] and o.stop.cpu().tolist() == g["stop"]
    assert check == g["check"]
    for s in range(S):
        assert np.allclose(primal[s], g["primal"][s], rtol=1e-12, atol=0)
        assert np.allclose(dual[s], g["dual"][s], rtol=1e-12, atol=0)
    # a stopped scenario keeps its own lamb_bar while the batch goes on: bit-exact against the oracle
    assert np.array_equal(prev.lamb_bar.cpu().numpy(), np.array(g["prev_lamb_bar"]))
    assert np.array_equal(prev.Z_bar.cpu().numpy(), np.array(g["prev_Z_bar"]))
    assert np.array_equal(o.bar_state.lamb_bar.cpu().numpy(), np.array(g["cur_lamb_bar"]))


@pytest.mark.gpu
def test_gpu_consensus_loop_large_batch_properties():
    S = 4096
    seq = obca_synth.sequence(S, 12, seed=9)
    o, prev, primal, dual, _ = _run_gpu_loop(S, seq, max_iter=10)
    n_iter = o.n_iter.cpu().numpy()
    stop = o.stop.cpu().numpy()
    assert n_iter.max() == 11 and (n_iter[stop == 1] <= 11).all()      # `i_iter > max_iter` exit
    assert stop.all()
    ref = obca_synth.run_oracle([{k: v[:64] for k, v in st.items()} for st in seq], 64, max_iter=10)
    assert np.array_equal(n_iter[:64], ref["n_iter"])
    assert np.array_equal(prev.lamb_bar.cpu().numpy()[:64], ref["prev"]["lamb_bar"])
    # the last two rows of every member of a shifted state are equal
    for f in FIELDS:
        a = getattr(prev, f).cpu().numpy()
        assert np.array_equal(a[:, :, -1], a[:, :, -2]), f
