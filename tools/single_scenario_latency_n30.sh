#!/bin/bash
# per-iteration latency of one N=30 scenario (LPQ=32: a 64-thread CTA, halves exchange through shared memory)
python - <<'PY'
import sys, os
sys.path.insert(0, 'tools'); sys.path.insert(0, '.')
import numpy as np
from single_scenario_latency import run
from pi_admm_b200.scenarios import c4_scenarios
xt, spd, rw = c4_scenarios(S=256, N=30)
for s in range(12):
    r = run(xt[s:s+1], spd[s:s+1], rw[s:s+1], 30)
    print(f"N=30 scenario {s}: alone {r[0]:.1f} us veh_inner {r[1]} edge_inner {r[2]} outer {r[3]}")
PY
