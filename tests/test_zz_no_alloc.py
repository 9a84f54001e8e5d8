"""piadmm_edge_list_static (include/piadmm.h; csrc/piadmm_static.cu): the allocation-free collision-edge test, whose
scratch is a file-scope __device__ array, against the existing entry point piadmm_detect_edges (handle-owned,
cudaMalloc'ed scratch) on the same trajectories.  Integer output: edge counts and pair lists must be IDENTICAL.

Collected last on purpose: a failure here hides no other test.  Runs alone in a fresh
process and after the other tests alike -- it keeps no state besides the library's static array, which every call
rewrites before it reads."""
import ctypes as C

import numpy as np
import pytest
import torch

import bench
from pi_admm_b200 import _abi
from pi_admm_b200.scenarios import c3_intersection, c4_scenarios, c5_grid
from pi_admm_b200.solver import PIADMMSolver
from util import make_params

pytestmark = pytest.mark.gpu


def _existing(p, xt, spd, refwin, D):
    """pos_old after the first x-update and the edge list piadmm_detect_edges makes of it (phase path)."""
    sol = PIADMMSolver(p, S=xt.shape[0], V=xt.shape[1], D=D)
    sol.setup(xt, spd, refwin)
    sol.begin_step()
    sol.x_update()
    sol.detect_edges()
    torch.cuda.synchronize()
    pos = sol.buf["pos"].clone()
    edge_n = sol.buf["edge_n"].cpu().numpy().copy()
    edge_ij = sol.buf["edge_ij"].cpu().numpy().copy()
    assert not (sol.buf["status"].cpu().numpy() & 4).any(), "scene overflows its neighbour slots: not a fair comparison"
    sol.close()
    return pos, edge_n, edge_ij


def _static(lib, p, pos, e_cap):
    S, V = pos.shape[0], pos.shape[1]
    edge_n = torch.full((S,), -7, dtype=torch.int32, device=pos.device)
    edge_ij = torch.full((S, max(e_cap, 1), 2), -7, dtype=torch.int32, device=pos.device)
    rc = lib.piadmm_edge_list_static(C.c_void_p(pos.data_ptr()), S, V, int(p.num_ho), float(p.dis_thres),
                                     int(p.thr_squared), e_cap, C.c_void_p(edge_n.data_ptr()),
                                     C.c_void_p(edge_ij.data_ptr()),
                                     C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.piadmm_static_last_error().decode()
    torch.cuda.synchronize()
    return edge_n.cpu().numpy(), edge_ij.cpu().numpy()


def _same(got_n, got_ij, ref_n, ref_ij, what):
    assert np.array_equal(got_n, ref_n), f"{what}: edge counts differ"
    for s in range(len(ref_n)):
        n = int(ref_n[s])
        assert np.array_equal(got_ij[s, :n], ref_ij[s, :n]), f"{what}: edge list of scenario {s} differs"


def test_static_edge_list_equals_detect_edges_twice_in_a_row():
    lib = _abi.load_library()
    # scene A: the 4096-vehicle grid of config 5 (fills the static bit matrix to its limit: 4096 * 128 words)
    pa_ = bench.coupled_params(30)
    xa = c5_grid(N=30)
    pos_a, n_a, ij_a = _existing(pa_, *xa, D=8)
    # scene B: 512 two-vehicle scenarios (config-4 shape, other threshold rule, N = 15) -- much smaller, so most of what
    # call A left in the array is still there when B runs
    pb = make_params("python_current")
    xb = c4_scenarios(S=512, N=pb.num_ho)
    pos_b, n_b, ij_b = _existing(pb, *xb, D=1)
    # scene C: the 64-vehicle intersection of config 3
    xc = c3_intersection(V=64, N=30)
    pos_c, n_c, ij_c = _existing(pa_, *xc, D=8)
    assert n_a[0] > 100 and n_b.sum() > 100 and n_c[0] > 10      # the comparison is not about empty lists

    free0 = None
    for rep in range(2):                                          # two rounds of back-to-back calls on the shared array
        g = _static(lib, pa_, pos_a, ij_a.shape[1]); _same(*g, n_a, ij_a, f"c5 grid, round {rep}")
        if free0 is None:
            free0 = torch.cuda.mem_get_info()[0]                  # after the one-off symbol lookup of the first call
        g = _static(lib, pb, pos_b, ij_b.shape[1]); _same(*g, n_b, ij_b, f"c4 batch, round {rep}")
        g = _static(lib, pa_, pos_c, ij_c.shape[1]); _same(*g, n_c, ij_c, f"c3 intersection, round {rep}")
    # (the entry point allocates nothing; the margin is for torch's caching allocator, which hands this test's own output
    # tensors out of 2 MiB blocks and may fetch one more)
    assert free0 - torch.cuda.mem_get_info()[0] < (16 << 20), "device memory shrank across allocation-free calls"

    # a list longer than e_cap: the count is the full one, the stored prefix is the sorted prefix, the rest untouched
    cap = 5
    got_n, got_ij = _static(lib, pa_, pos_c, cap)
    assert got_n[0] == n_c[0] and np.array_equal(got_ij[0, :cap], ij_c[0, :cap])


def test_static_edge_list_rejects_scenes_beyond_its_scratch():
    lib = _abi.load_library()
    dev = torch.device("cuda")
    n = torch.zeros(1, dtype=torch.int32, device=dev)
    ij = torch.zeros(1, 8, 2, dtype=torch.int32, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # 4128 vehicles in one scene: 4128 * 129 words > 524288 (pos is never touched: the size check comes first)
    V = 4128
    pos = torch.zeros(1, V, 2, 31, dtype=torch.float64, device=dev)
    rc = lib.piadmm_edge_list_static(C.c_void_p(pos.data_ptr()), 1, V, 30, 2.0, 1, 8, C.c_void_p(n.data_ptr()),
                                     C.c_void_p(ij.data_ptr()), st)
    assert rc == -4                                               # PIADMM_E_UNSUPPORTED
    assert b"static scratch" in lib.piadmm_static_last_error()
    # more than 8192 vehicles in a batch of small scenarios
    rc = lib.piadmm_edge_list_static(C.c_void_p(pos.data_ptr()), 4097, 2, 15, 2.0, 1, 1, C.c_void_p(n.data_ptr()),
                                     C.c_void_p(ij.data_ptr()), st)
    assert rc == -4
    rc = lib.piadmm_edge_list_static(None, 1, 2, 15, 2.0, 1, 1, C.c_void_p(n.data_ptr()), C.c_void_p(ij.data_ptr()), st)
    assert rc == -1


@pytest.mark.parametrize("S,V,N", [(3, 1, 15), (2, 5, 1), (1, 33, 31), (7, 40, 8)])
def test_static_edge_list_small_and_odd_shapes_against_numpy(S, V, N):
    """Shapes the solver-based comparison does not reach (a single vehicle, N = 1, N = 31, V just over a word), against the
    statement itself in numpy: edge_mat[i, j] = any_k(d^2 < thr) for i < j, pairs in np.where order (casadi/main.py:110-113,121)."""
    lib = _abi.load_library()
    rng = np.random.default_rng(S * 100 + V)
    pos = rng.normal(0, 2.5, (S, V, 2, N + 1))
    thr = 1.7
    t = torch.from_numpy(pos).cuda()
    cap = max(1, V * (V - 1) // 2)
    en = torch.full((S,), -1, dtype=torch.int32, device="cuda"); eij = torch.full((S, cap, 2), -1, dtype=torch.int32, device="cuda")
    rc = lib.piadmm_edge_list_static(C.c_void_p(t.data_ptr()), S, V, N, thr, 0, cap, C.c_void_p(en.data_ptr()), C.c_void_p(eij.data_ptr()),
                                     C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    en, eij = en.cpu().numpy(), eij.cpu().numpy()
    for s in range(S):
        mat = np.zeros((V, V), int)
        for i in range(V):
            for j in range(i + 1, V):
                d2 = (pos[s, i, 0] - pos[s, j, 0]) ** 2 + (pos[s, i, 1] - pos[s, j, 1]) ** 2
                mat[i, j] = int(np.any(d2 < thr))
        r, c = np.where(mat == 1)
        assert en[s] == len(r) and np.array_equal(eij[s, :len(r)], np.stack((r, c), 1))
