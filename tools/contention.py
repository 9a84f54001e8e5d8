"""Contention curve: one config-4 scenario replicated n times (n warps busy with identical work) -> fused-kernel time."""
import os, sys
import numpy as np, torch
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tools"))
from single_scenario_latency import run
from pi_admm_b200.scenarios import c4_scenarios
xt, spd, rw = c4_scenarios(S=4096, N=15)
s = 1220
for n in (1, 2, 148, 296, 592, 1184, 1776, 2368):
    rep = lambda a: np.repeat(a[s:s + 1], n, axis=0)
    r = run(rep(xt), rep(spd), rep(rw), 15)
    print(n, 'copies:', round(r[0], 1), 'us', flush=True)
