"""Outer-loop convergence of the coupled scenes (BASELINE configs 3 and 5) under different dual-update gains, on the
oracle's C twin (CPU; a diagnostic for profiles/r2_convergence.md, never a product path).

  python tools/coupled_gain_scan.py            # c3 and c5: the reference's gains, PI=0, and a (kI, kP) grid

Exit kinds: "stop" = the reference's stop test passed with a non-empty edge list (casadi/main.py:174-178), "empty" =
the edge list emptied (residuals exactly 0, :115-118 / :165-178), "CAP" = 100 outer iterations (PI_ADMM_class.py:25).
A "stop" with max|dual| at the clamp and max|u| at the steering bound is the false collapse described in the profile.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from oracle import ctwin  # noqa: E402
from pi_admm_b200.scenarios import c5_grid  # noqa: E402


def one(scene, N, D, cap=100, **kw):
    xt, spd, refwin = scene
    p = bench.coupled_params(N)
    p.update(iter_num=cap, kI=-1.0, theta1=5.0, theta2=3.0)          # the reference's gains unless overridden
    p.update(**kw)
    r = ctwin.step(p, xt, spd, refwin, D=D, nthreads=len(os.sched_getaffinity(0)))
    n, en = int(r["n_outer"][0]), int(r["edge_n"][0])
    rk, sk = r["err_rk"][0, :n], r["err_sk"][0, :n]
    kind = "CAP" if int(r["status"][0]) & 8 else ("empty" if en == 0 else "stop")
    return (f"n_outer={n:3d} {kind:5s} rk_last={rk[-1]:9.3g} sk_last={sk[-1]:9.3g} max|dual|={np.abs(r['dual']).max():6.3g} "
            f"max|u|={np.abs(r['u']).max():.3f} edge_solves={int(r['n_edge_solves'][0]):5d} rk[:6]={np.array2string(rk[:6], precision=3)}")


def main():
    scenes = {"c3 (64-vehicle intersection)": (bench.coupled_scene("c3")[0], 30, 8),
              "c5 (4096-vehicle grid, rejection rule clear=3 m)": (bench.coupled_scene("c5")[0], 30, 8),
              "c5 round-1 generator (clear=0: 15 pairs start inside dis_thres)": (c5_grid(N=30, clear=0.0), 30, 8)}
    for name, (scene, N, D) in scenes.items():
        print(f"## {name}")
        print("reference gains (kI=rho=2, kP=5-3/(1+exp(-dmin))): ", one(scene, N, D), flush=True)
        print("PI=0 (plain lambda += rho r, clamp +-30):            ", one(scene, N, D, PI=0), flush=True)
        if "round-1" in name:
            continue
        for kI in (1.0, 1.5, 2.0):
            for kP in (0.0, 0.25, 0.5, 1.0, 2.0):
                print(f"kI={kI:3.1f} kP={kP:4.2f}:                                      ", one(scene, N, D, kI=kI, theta1=kP, theta2=0.0), flush=True)
        print()


if __name__ == "__main__":
    main()
