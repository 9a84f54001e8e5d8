"""The JOINT flow: the outer loop of casadi_old_PI_ADMM/main.py:60-164 /
matlab_old_files/ADMM_fmincon_multi_veh_intesection_static.m:57-164 with the edge side solved as ONE problem over all
vehicles of a scene (``cost_function_edge``, ...static.m:250-291) instead of one problem per colliding pair.

Every vehicle keeps ONE consensus copy z-hat_i and ONE multiplier lambda_i (neighbour slot 0 of a D = 1 solver).  Per outer
iteration:  vehicle QPs (piadmm_x_update)  ->  collision test (piadmm_edge_list_static, casadi/main.py:110-113)  ->  the
joint edge problem over every pair that collided so far in this step (piadmm_solve_joint: one thread-block cluster per
scene)  ->  z-hat = non-linear roll-out of the joint controls (casadi/main.py:156-158), lambda += kI (pos - z-hat)
(...static.m:131-132; kI = params.kI, or rho when that is negative), global Frobenius residuals (old main.py:154-155) over the vehicles that are in a pair.

The kernels do the arithmetic of the two solves and the collision test; the pair bookkeeping between them (union of the
edge lists, at most 8 pairs per vehicle) and the elementwise multiplier update are a few torch statements per outer
iteration with one host read-back of the edge list -- this flow is the alternative z-step for dense scenes, not the
latency path (that is the pairwise persistent kernel).  No CPU fallback: needs the CUDA library like everything else here.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _abi
from .solver import PIADMMSolver

PAIR_CAP = 8        # pairs per vehicle the cluster kernel holds (JT_D in csrc/piadmm_joint.cuh)


def nonlinear_rollout(xt, spd, u, dt, L):
    """dynamic_update_edge (casadi/PI_ADMM_class.py:77-105), numeric: xt (S,V,3), spd (S,V), u (S,V,N) -> (S,V,2,N+1)."""
    z = torch.zeros_like(u[..., :1])
    th = xt[..., 2:3] + (spd[..., None] / L * dt) * torch.cat((z, torch.cumsum(u, -1)), -1)
    step = spd[..., None] * dt
    x = xt[..., 0:1] + step * torch.cat((z, torch.cumsum(torch.cos(th[..., :-1]), -1)), -1)
    y = xt[..., 1:2] + step * torch.cat((z, torch.cumsum(torch.sin(th[..., :-1]), -1)), -1)
    return torch.stack((x, y), 2)


class JointRSUSolver:
    def __init__(self, params: _abi.Params, S: int, V: int, device=None):
        p = params.copy().update(al_all=0, x_memo=0, dual_mode=_abi.DUAL_PLAIN)
        self.sol = PIADMMSolver(p, S=S, V=V, D=1, device=device)
        self.p, self.S, self.V, self.N = self.sol.params, S, V, self.sol.N
        self.e_cap = max(1, V * (V - 1) // 2)
        dev = self.sol.device
        self.edge_n = torch.zeros(S, dtype=torch.int32, device=dev)
        self.edge_ij = torch.zeros((S, self.e_cap, 2), dtype=torch.int32, device=dev)

    def setup(self, xt, spd, refwin):
        self.sol.setup(xt, spd, refwin)

    def close(self):
        self.sol.close()

    def _detect(self):
        sol = self.sol
        rc = sol.lib.piadmm_edge_list_static(C.c_void_p(sol.buf["pos"].data_ptr()), self.S, self.V, self.N,
                                             float(self.p.dis_thres), int(self.p.thr_squared), self.e_cap,
                                             C.c_void_p(self.edge_n.data_ptr()), C.c_void_p(self.edge_ij.data_ptr()), sol._stream())
        if rc != 0:
            raise _abi.PIADMMError(sol.lib.piadmm_static_last_error().decode())
        n = self.edge_n.cpu().numpy()
        ij = self.edge_ij.cpu().numpy()
        return [set(map(tuple, ij[s, :n[s]].tolist())) for s in range(self.S)]

    def iterate(self):
        """All outer iterations of one MPC step.  Returns a dict of numpy results per scenario."""
        sol, p, S, V, N = self.sol, self.p, self.S, self.V, self.N
        b = sol.buf
        sol.begin_step()
        rho = float(p.rho)
        kI = float(p.kI) if float(p.kI) >= 0.0 else rho
        ever = [set() for _ in range(S)]
        active = np.ones(S, bool)
        n_outer = np.zeros(S, np.int32); converged = np.zeros(S, bool); dropped = np.zeros(S, np.int32)
        err_rk = [[] for _ in range(S)]; err_sk = [[] for _ in range(S)]; joint_iters = [[] for _ in range(S)]
        pairs = [[] for _ in range(S)]
        in_pair = torch.zeros((S, V), dtype=torch.bool, device=sol.device)
        last_hat = torch.zeros_like(b["hat"][:, :, 0])
        status = np.zeros(S, np.int32)
        for it in range(int(p.iter_num)):
            b["done"].copy_(torch.from_numpy((~active).astype(np.int32)))     # finished scenarios skip their vehicle QPs
            sol.x_update()
            hits = self._detect()
            for s in range(S):
                if not active[s]:
                    continue
                n_outer[s] += 1
                ever[s] |= hits[s]
                if not ever[s]:
                    active[s] = False; converged[s] = True
            if not active.any():
                break
            P = 1
            for s in range(S):
                if active[s]:
                    deg = [0] * V
                    pl = []
                    dropped[s] = 0
                    for (i, j) in sorted(ever[s]):
                        if deg[i] < PAIR_CAP and deg[j] < PAIR_CAP:
                            pl.append((i, j)); deg[i] += 1; deg[j] += 1
                        else:
                            dropped[s] += 1
                    pairs[s] = pl
                P = max(P, len(pairs[s]))
            pij = np.zeros((S, P, 2), np.int32); pn = np.zeros(S, np.int32)
            ip = np.zeros((S, V), bool)
            for s in range(S):
                if active[s] and pairs[s]:
                    pij[s, :len(pairs[s])] = np.array(pairs[s], np.int32); pn[s] = len(pairs[s])
                for (i, j) in pairs[s]:
                    ip[s, i] = ip[s, j] = True
            in_pair = torch.from_numpy(ip).to(sol.device)
            uh, its, st = sol.solve_joint(pn, pij)
            new = nonlinear_rollout(b["xt"], b["spd"], uh, float(p.dt), float(p.L))
            act = (in_pair & torch.from_numpy(active).to(sol.device)[:, None])[:, :, None, None]
            hat, dual = b["hat"][:, :, 0], b["dual"][:, :, 0]
            hat.copy_(torch.where(act, new, hat))
            dual.copy_(torch.where(act, dual + kI * (b["pos"] - hat), dual))
            b["nbr"][:, :, 0].copy_(torch.where(in_pair, 0, -1).to(torch.int32))   # AL term of the vehicle QP: on once in a pair
            zero = torch.zeros((), dtype=torch.float64, device=sol.device)
            rk = torch.sqrt(torch.where(act, (b["pos"] - hat) ** 2, zero).sum((1, 2, 3))).cpu().numpy()
            sk = torch.sqrt(torch.where(act, (rho * (last_hat - hat)) ** 2, zero).sum((1, 2, 3))).cpu().numpy()
            its, st = its.cpu().numpy(), st.cpu().numpy()
            last_hat = hat.clone()
            for s in range(S):
                if not active[s]:
                    continue
                err_rk[s].append(float(rk[s])); err_sk[s].append(float(sk[s])); joint_iters[s].append(int(its[s]))
                status[s] |= int(st[s])
                if rk[s] <= p.eps_pri and sk[s] <= p.eps_dual:
                    active[s] = False; converged[s] = True
            if not active.any():
                break
        torch.cuda.synchronize(sol.device)
        return dict(u=b["u"].cpu().numpy(), pos=b["pos"].cpu().numpy(), hat=b["hat"][:, :, 0].cpu().numpy(),
                    dual=b["dual"][:, :, 0].cpu().numpy(), n_outer=n_outer, converged=converged, err_rk=err_rk, err_sk=err_sk,
                    joint_iters=joint_iters, pairs=pairs, dropped=dropped, status=status)
