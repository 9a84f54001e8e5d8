"""Generates tests/golden/c2_consensus.json: the consensus loop of config 2
(Distributed_planner/decentralized_overtaking_ADMM.py:40-97) run by the numpy oracle
(oracle/obca_consensus.py) on the seeded synthetic sequences of tests/obca_synth.py (the IPOPT solves in
between are out of scope).  "Parity unpinned": the reference has no stored outputs.
    python tests/golden/make_golden_c2.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
import obca_synth  # noqa: E402

S, N_ITER, SEED = 8, 56, 1


def main():
    seq = obca_synth.sequence(S, N_ITER, seed=SEED)
    r = obca_synth.run_oracle(seq, S)
    out = dict(S=S, n_iter_max=N_ITER, seed=SEED, n_iter=r["n_iter"].tolist(), stop=r["stop"].tolist(),
               primal=r["primal"], dual=r["dual"], check=r["check"],
               prev_lamb_bar=r["prev"]["lamb_bar"].tolist(), prev_Z_bar=r["prev"]["Z_bar"].tolist(),
               cur_lamb_bar=r["cur_lamb_bar"].tolist())
    with open(os.path.join(HERE, "c2_consensus.json"), "w") as f:
        json.dump(out, f)
    print("n_iter", out["n_iter"], "stop", out["stop"])


if __name__ == "__main__":
    main()
