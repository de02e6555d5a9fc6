"""torchrun script: the N-GPU data-parallel step must equal the 1-GPU step on the same GLOBAL batch.
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 tests/dp_check.py
Every rank builds (a) a DP model over its slice and (b) a single-GPU model over the whole global batch, runs
3 steps of each and compares tables / weights / predictions (rtol 2e-5: the dense all-reduce changes the fp32
summation order; the embedding update is order-identical by construction)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import deepctr_oracle as O  # noqa: E402  (the checker: every rank also runs the float64 oracle on the GLOBAL batch)


def main():
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    from deepctr_b200 import dist as ddist
    from deepctr_b200.theano_fnn import FNN_wnzh, init_mlp
    from deepctr_b200.models import FM
    rng = np.random.default_rng(11)
    X, F, k, N = 640, 16, 11, 64
    GN = N * world
    feat_field = np.repeat(np.arange(F), X // F)
    table = rng.uniform(-0.2, 0.2, (X, k))
    Ws, bs = init_mlp(1 + F * k, 48, 24)
    Ws[2] = rng.uniform(-0.2, 0.2, Ws[2].shape)
    fmm = dict(w_0=-0.3, table=table, feat_field=feat_field)
    dp = FNN_wnzh(fmm, batch_size=N, lr=0.01, dropout=1.0, hidden1=48, hidden2=24, lambda_fm=0.1, weights=(Ws, bs))
    ddist.attach(dp, dist)
    ref = FNN_wnzh(fmm, batch_size=GN, lr=0.01, dropout=1.0, hidden1=48, hidden2=24, lambda_fm=0.1, weights=(Ws, bs))
    ref.use_graph = False
    fm_dp = FM(N, [X, F, k - 1], ["uniform", -0.05, 0.05, [1, 2], None], ["sgd", 0.05, "mean"], [0.01], "train", N)
    ddist.attach(fm_dp, dist)
    fm_ref = FM(GN, [X, F, k - 1], ["uniform", -0.05, 0.05, [1, 2], None], ["sgd", 0.05, "mean"], [0.01], "train", GN)
    fm_ref.use_graph = False
    onet = O.OracleNet(np.concatenate([table, np.zeros((1, k))], 0), -0.3, Ws, bs, F, k, "none", True, "none", "tanh")
    ofm = O.OracleFM(fm_ref.engine.get_table().astype(np.float64), 0.0)
    for step in range(6):            # 2 eager steps, the capture, then graph replays of kernels + NCCL collectives
        feats = np.stack([rng.integers(0, X // F, GN) + f * (X // F) for f in range(F)], 1)
        feats[:, 0] = 3
        y = rng.integers(0, 2, GN)
        wts = rng.uniform(0.5, 1.5, (GN, F)).astype(np.float32)
        sl = slice(rank * N, (rank + 1) * N)
        l1, p1 = dp.train(feats[sl], y[sl])
        l0, p0 = ref.train(feats, y)
        np.testing.assert_allclose(l1, l0, rtol=2e-5)
        np.testing.assert_allclose(p1, p0[sl], rtol=2e-5, atol=1e-6)
        # feature 3 is hit by every sample: its row takes GN sequential-decay updates per step, so the fp32 noise between two
        # summation orders of the dense gradients (all-reduce vs split-K) grows with the global batch
        np.testing.assert_allclose(dp.get_table(), ref.get_table(), rtol=2e-5, atol=max(2e-6, 5e-8 * GN))
        np.testing.assert_allclose(dp.get_weights()[0], ref.get_weights()[0], rtol=2e-5, atol=2e-6)
        # ... and the ORACLE on the global batch: 1e-5 relative per step (north_star), absolute floor for the hot row as above
        ol, op = onet.train_step_theano(feats, y, 0.01, 0.1)
        tol = 1e-5 * (step + 1)
        np.testing.assert_allclose(l1, ol, rtol=tol)
        np.testing.assert_allclose(p1, op[sl], rtol=tol, atol=1e-6)
        np.testing.assert_allclose(dp.get_table(), onet.table[:-1], rtol=tol, atol=max(2e-6, 5e-8 * GN))
        np.testing.assert_allclose(dp.get_weights()[0], onet.Ws[0], rtol=tol, atol=2e-6)
        f1 = fm_dp.train_step({fm_dp.sp_id_hldr: feats[sl].reshape(-1), fm_dp.sp_wt_hldr: wts[sl].reshape(-1), fm_dp.lbl_hldr: y[sl]})
        f0 = fm_ref.train_step({fm_ref.sp_id_hldr: feats.reshape(-1), fm_ref.sp_wt_hldr: wts.reshape(-1), fm_ref.lbl_hldr: y})
        np.testing.assert_allclose(f1[1], f0[1][sl], rtol=2e-5, atol=1e-6)
        np.testing.assert_allclose(fm_dp.engine.get_table(), fm_ref.engine.get_table(), rtol=2e-5, atol=2e-6)
        fl, fp = ofm.train_step(feats, wts, y, 0.05, 0.01, "mean")
        np.testing.assert_allclose(f1[0], fl, rtol=tol)
        np.testing.assert_allclose(f1[1], fp[sl], rtol=tol, atol=1e-6)
        np.testing.assert_allclose(fm_dp.engine.get_table(), ofm.table, rtol=tol, atol=2e-6)
    # Adam (dense, TF form) under data parallelism: FM with L2, against the single-GPU Adam step on the global batch
    ad_dp = FM(N, [X, F, k - 1], ["uniform", -0.05, 0.05, [1, 2], None], ["adam", 1e-2, 1e-8, "mean"], [0.01], "train", N)
    ddist.attach(ad_dp, dist)
    ad_ref = FM(GN, [X, F, k - 1], ["uniform", -0.05, 0.05, [1, 2], None], ["adam", 1e-2, 1e-8, "mean"], [0.01], "train", GN)
    ad_ref.use_graph = False
    for step in range(5):
        feats = np.stack([rng.integers(0, X // F, GN) + f * (X // F) for f in range(F)], 1)
        y = rng.integers(0, 2, GN)
        wts = rng.uniform(0.5, 1.5, (GN, F)).astype(np.float32)
        sl = slice(rank * N, (rank + 1) * N)
        a1 = ad_dp.train_step({ad_dp.sp_id_hldr: feats[sl].reshape(-1), ad_dp.sp_wt_hldr: wts[sl].reshape(-1), ad_dp.lbl_hldr: y[sl]})
        a0 = ad_ref.train_step({ad_ref.sp_id_hldr: feats.reshape(-1), ad_ref.sp_wt_hldr: wts.reshape(-1), ad_ref.lbl_hldr: y})
        np.testing.assert_allclose(a1[1], a0[1][sl], rtol=5e-5, atol=1e-6)
        np.testing.assert_allclose(ad_dp.engine.get_table(), ad_ref.engine.get_table(), rtol=2e-4, atol=2e-6)
    # replicas are bit-identical to each other
    t = dp.engine.table.clone()
    ts = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(ts, t)
    assert all(torch.equal(ts[0], x) for x in ts)
    if rank == 0:
        print("dp_check ok: world=%d" % world)
    ddist.shutdown(dist, [dp, fm_dp, ad_dp])


if __name__ == "__main__":
    main()
