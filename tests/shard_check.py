"""torchrun script: the row-sharded step (table rows on rank row % world, id / embedding / gradient all-to-all) must equal
the 1-GPU step on the same GLOBAL batch.
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29556 tests/shard_check.py
Every rank builds (a) a sharded model over its slice of the batch and its rows of the table and (b) a single-GPU model over
the whole batch and table, runs 3 steps of each and compares loss, predictions, MLP weights and the table rows it owns
(rtol 2e-5: the dense all-reduce and the per-owner fold change the fp32 summation order)."""
import os
import pickle
import sys
import tempfile

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
    import deepctr_b200.models as M
    from deepctr_b200 import dist as ddist
    rng = np.random.default_rng(5)
    n_num, cat_sizes, r = 3, np.array([50, 7, 201, 3, 90]), 4
    offsets = [n_num + int(cat_sizes[:i].sum()) for i in range(len(cat_sizes))]
    X_dim, F, d = n_num + int(cat_sizes.sum()), n_num + len(cat_sizes), r + 1
    N = 48
    GN = N * world
    for cls_name, hidden, ptm in [("FNN", [24, 12], ["sgd", 0.05, "sum"]), ("FNN_IP_L3", [32, 16, 8], ["sgd", 0.05, "sum"]),
                                  ("FNN_OP_L3", [32, 16, 8], ["sgd", 0.05, "sum"]), ("FNN", [24, 12], ["adam", 1e-2, 1e-8, "sum"])]:
        tol = 2e-5 if ptm[0] == "sgd" else 2e-4      # Adam divides by sqrt(v): summation-order noise is amplified early on
        cls = getattr(M, cls_name)
        cw = {"none": 0, "ip": F * (F - 1) // 2, "op": d * d}[cls.cross]
        dims = [F * d + cw + 1] + hidden + [1]
        vm = {"W": rng.uniform(-0.3, 0.3, (X_dim, 1)), "V": rng.uniform(-0.3, 0.3, (X_dim, r)), "b": np.array([0.1])}
        for i in range(len(dims) - 1):
            vm["h%d_w" % (i + 1)] = rng.uniform(-1, 1, (dims[i], dims[i + 1])) / np.sqrt(dims[i])
            vm["h%d_b" % (i + 1)] = rng.uniform(-0.1, 0.1, dims[i + 1])
        vm = {k: v.astype(np.float32) for k, v in vm.items()}
        path = os.path.join(tempfile.gettempdir(), "shard_check_%s_%d.pickle" % (cls_name, rank))
        with open(path, "wb") as f:
            pickle.dump(vm, f, protocol=2)
        rch = [X_dim, F, r] + hidden + ["tanh"]
        init = ["uniform", -0.01, 0.01, list(range(1, 12)), path]
        sh = cls(cat_sizes, offsets, N, rch, init, ptm, [1.0], "train", N, shard=(rank, world))
        ddist.attach_sharded(sh, dist, X_dim)
        ref = cls(cat_sizes, offsets, GN, rch, init, ptm, [1.0], "train", GN)
        ref.use_graph = False
        nl = len(dims) - 1
        onet = O.OracleNet(np.concatenate([vm["W"], vm["V"]], 1), vm["b"][0], [vm["h%d_w" % (i + 1)] for i in range(nl)],
                           [vm["h%d_b" % (i + 1)] for i in range(nl)], F, d, cls.cross, False, "tanh", "tanh")
        for step in range(3):
            c_ids = np.stack([rng.integers(0, s, GN) for s in cat_sizes], 1)
            c_ids[GN // 2] = c_ids[0]
            v_wts = rng.uniform(0, 1, (GN, n_num)).astype(np.float32)
            c_wts = np.ones((GN, len(cat_sizes)), np.float32)
            y = rng.integers(0, 2, GN)
            sl = slice(rank * N, (rank + 1) * N)
            l1, p1 = sh.train_step({sh.v_wt_hldr: v_wts[sl], sh.c_id_hldr: c_ids[sl], sh.c_wt_hldr: c_wts[sl], sh.lbl_hldr: y[sl]})
            l0, p0 = ref.train_step({ref.v_wt_hldr: v_wts, ref.c_id_hldr: c_ids, ref.c_wt_hldr: c_wts, ref.lbl_hldr: y})
            np.testing.assert_allclose(l1, l0, rtol=2e-5)
            np.testing.assert_allclose(p1, p0[sl], rtol=2e-5, atol=1e-6)
            mine = ref.engine.get_table()[rank::world]
            np.testing.assert_allclose(sh.engine.get_table()[:mine.shape[0]], mine, rtol=tol, atol=2e-6)
            np.testing.assert_allclose(sh.engine.get_layer(0)[0], ref.engine.get_layer(0)[0], rtol=tol, atol=2e-6)
            if ptm[0] == "sgd":       # ... and the ORACLE on the global batch, 1e-5 relative per step (north_star)
                g_ids = np.concatenate([np.broadcast_to(np.arange(n_num), (GN, n_num)), c_ids + np.array(offsets)], 1)
                ol, op = onet.train_step_tf(g_ids, np.concatenate([v_wts, c_wts], 1), y, 0.05)
                t2 = 1e-5 * (step + 1)
                np.testing.assert_allclose(l1, ol, rtol=t2)
                np.testing.assert_allclose(p1, op[sl], rtol=t2, atol=1e-6)
                np.testing.assert_allclose(sh.engine.get_table()[:mine.shape[0]], onet.table[rank::world], rtol=t2, atol=2e-6)
        # checkpoint of the sharded model = the reference's [X_dim, ...] format: gathered, written by rank 0, reloadable unsharded
        ck = os.path.join(tempfile.gettempdir(), "shard_check_dump_%s.pickle" % cls_name)
        sh.dump(ck)
        dist.barrier()
        with open(ck, "rb") as f:
            got = pickle.load(f)
        np.testing.assert_array_equal(np.concatenate([got["W"], got["V"]], 1), np.concatenate([sh.var_map()["W"], sh.var_map()["V"]], 1))
        np.testing.assert_allclose(np.concatenate([got["W"], got["V"]], 1), ref.engine.get_table(), rtol=tol, atol=2e-6)
        assert got["W"].shape == (X_dim, 1) and got["h1_w"].shape == (dims[0], dims[1])
        re = cls(cat_sizes, offsets, GN, rch, ["uniform", -0.01, 0.01, list(range(1, 12)), ck], ptm, [1.0], "train", GN)
        np.testing.assert_array_equal(re.engine.get_table(), np.concatenate([got["W"], got["V"]], 1))
        # collective inference through the sharded lookup
        ids_eval = np.concatenate([np.broadcast_to(np.arange(n_num), (GN, n_num)), c_ids + np.array(offsets)], 1)
        wts_eval = np.concatenate([v_wts, c_wts], 1)
        pe = sh.eval_preds.eval({sh.eval_id_hldr: ids_eval[sl], sh.eval_wts_hldr: wts_eval[sl]})
        pr = ref.eval_preds.eval({ref.eval_id_hldr: ids_eval, ref.eval_wts_hldr: wts_eval})
        np.testing.assert_allclose(pe, pr[sl], rtol=2e-5, atol=1e-6)
        if rank == 0:
            print("shard_check %s %s ok: world=%d loss=%.6f" % (cls_name, ptm[0], world, l1))
    ddist.shutdown(dist)


if __name__ == "__main__":
    main()
