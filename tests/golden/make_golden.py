"""Generates tests/golden/oracle_golden.npz from the CPU oracle (oracle/deepctr_oracle.py).

The reference cannot run here (Python 2 + Theano / TF 0.x) and has no fixtures of its own, so these vectors
are ORACLE outputs on fixed seeded inputs: they guard the oracle and the CUDA path against drift; they do
not pin parity with the original engines.   Run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
import deepctr_oracle as O  # noqa: E402


def inputs():
    rng = np.random.default_rng(20160524)
    X, F, d, N = 96, 16, 11, 24
    table = rng.uniform(-0.2, 0.2, (X, d))
    ids = np.stack([rng.integers(0, X // F, N) + f * (X // F) for f in range(F)], 1)
    ids[:, 0] = 2
    ids[5] = ids[4]
    wts = rng.uniform(0.2, 1.2, (N, F))
    y = rng.integers(0, 2, N).astype(np.float64)
    return rng, X, F, d, N, table, ids, wts, y


def compute():
    rng, X, F, d, N, table, ids, wts, y = inputs()
    out = {"ids": ids, "wts": wts, "y": y, "table0": table}
    # Theano-form FNN, 3 steps (FNN_wnzh.py semantics)
    Ws, bs = O.theano_fnn_init(1 + F * d, 20, 10)
    Ws[2] = rng.uniform(-0.2, 0.2, Ws[2].shape)
    out["th_w3"] = Ws[2].copy()
    net = O.OracleNet(table, -0.1, Ws, bs, F, d, "none", True, "none", "tanh")
    masks = [None, rng.binomial(1, 0.5, (1, 20)).astype(np.float64), rng.binomial(1, 0.5, (1, 10)).astype(np.float64)]
    out["th_r1"], out["th_r2"] = masks[1], masks[2]
    losses = []
    for _ in range(3):
        l, p = net.train_step_theano(ids, y, 0.01, 0.1, masks=masks)
        losses.append(l)
    out["th_losses"], out["th_p"], out["th_table"], out["th_w1"] = np.array(losses), p, net.table, net.Ws[0]
    # FM, 2 steps with lambda
    fm = O.OracleFM(table, 0.05)
    for _ in range(2):
        l, p = fm.train_step(ids, wts, y, 0.05, 0.01)
    out["fm_loss"], out["fm_p"], out["fm_table"], out["fm_b"] = np.array([l]), p, fm.table, np.array([fm.b])
    # TF-form forward of the three embedding layers
    for cross in ("none", "ip", "op"):
        e = O.gather_fields(table, ids, wts)
        out["z_" + cross] = O.build_z(e, cross, 0.25, False)
    return out


if __name__ == "__main__":
    np.savez_compressed(os.path.join(HERE, "oracle_golden.npz"), **compute())
    print("wrote oracle_golden.npz")
