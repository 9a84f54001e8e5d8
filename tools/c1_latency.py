"""Config 1 (the reference's own case: ONE two-vehicle scenario, 35 MPC steps) on the device: ms per MPC step."""
import os, sys, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import torch
from pi_admm_b200 import main_intersection
for sem in ("python_current", "matlab_pi"):
    for _ in range(2):
        out = main_intersection.run(semantics=sem, device_loop=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out = main_intersection.run(semantics=sem, device_loop=True)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    n = len(out["u_vec"])
    print(f"FEW={os.environ.get('PIADMM_FEW', '1')} config 1 {sem}: {n} MPC steps on the device in {dt * 1e3:.2f} ms = {dt * 1e3 / n:.3f} ms per step "
          f"(outer iterations {int(sum(out['n_outer']))}, launches {out['launches']})", flush=True)
