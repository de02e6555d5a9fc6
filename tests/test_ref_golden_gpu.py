"""CUDA path (through the reference-facing model classes and the C-ABI) against vectors produced by EXECUTING THE REFERENCE
(tests/golden/ref_golden.pkl, written by tests/golden/make_golden_ref.py from oracle/_ref).  Nothing here reads
/root/reference or calls the oracle.  Tolerance: 1e-5 relative per step as north_star states (absolute floor 2e-6 on values
of order 1); after k SGD steps in fp32 the trajectory is allowed k * 1e-5."""
import os
import pickle

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_golden.pkl")
RTOL = 1e-5


def close(a, b, rtol=RTOL, atol=2e-6):
    np.testing.assert_allclose(np.asarray(a, np.float64), np.asarray(b, np.float64), rtol=rtol, atol=atol)


@pytest.fixture(scope="module")
def gold():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    with open(GOLD, "rb") as f:
        return pickle.load(f)


@pytest.fixture(params=["simt", "tc", "mlp2", "mlp2tc"])
def gemm_path(request, monkeypatch):
    monkeypatch.setenv("DCTR_GEMM", {"mlp2": "simt", "mlp2tc": "tc"}.get(request.param, request.param))
    monkeypatch.setenv("DCTR_MLP2", {"mlp2": "on", "mlp2tc": "tc"}.get(request.param, "off"))
    return request.param


def _pad(lists, J):
    out = np.full((len(lists), J), -1, dtype=np.int64)
    for i, r in enumerate(lists):
        out[i, :len(r)] = r
    return out


def test_fnn_wnzh_whole_script(gold, gemm_path):
    """The reference script's whole run (23 epochs x 3 steps on the S1 demo set, with a duplicate-field, a missing-field and a
    repeated-feature line) replayed on the GPU with the dropout draws it logged: gx and the weights of the first steps, the
    predictions of the train and test files after every epoch, AUC to 3 decimals, and the final feat_weights table."""
    from sklearn.metrics import roc_auc_score
    from deepctr_b200.theano_fnn import FNN_wnzh
    g = gold["wnzh"]
    h = g["hyper"]
    N = h["batch_size"]
    J = max(len(ft) for _, ft in g["train"] + g["test"])
    m = FNN_wnzh(dict(w_0=g["w_0"], table=g["table"], feat_field=g["feat_field"]), batch_size=N, lr=h["lr"], dropout=h["dropout"],
                 hidden1=h["hidden1"], hidden2=h["hidden2"], lambda1=h["lambda1"], lambda_fm=h["lambda_fm"], acti_type=h["acti_type"],
                 max_feats=J)
    train_f, train_y = _pad([ft for _, ft in g["train"]], J), np.array([y for y, _ in g["train"]])
    test_f, test_y = _pad([ft for _, ft in g["test"]], J), np.array([y for y, _ in g["test"]])
    n_batch = len(train_y) // N
    k = 0
    worst = 0.0
    for ep, erec in enumerate(g["epochs"]):
        for j in range(n_batch):
            st = g["steps"][k]
            if "w1" in st:            # train() returns the PRE-update weights: the state before this step
                close(m.get_weights()[0], st["w1"], (k + 1) * RTOL, 2e-6)
                close(m.get_weights()[1], st["w2"], (k + 1) * RTOL, 2e-6)
            m.set_masks(st["r1"], st["r2"])
            m.train(train_f[j * N:(j + 1) * N], train_y[j * N:(j + 1) * N])
            if "gx" in st:
                gx = m.engine.dz0[:, :st["gx"].shape[1]].cpu().numpy()
                # the fused tcgen05 kernel keeps ONE accumulator per output (tensor memory is full) and evaluates tanh through
                # ex2 / rcp: its elements carry an absolute error of a few 1e-6 of the LARGEST gradient of the batch
                atol = 1e-7 if gemm_path != "mlp2tc" else 4e-6 * float(np.abs(st["gx"]).max())
                close(gx, st["gx"], (k + 1) * RTOL, atol)
            k += 1
        for feats, yy, want in ((train_f, train_y, erec["train_p"]), (test_f, test_y, erec["test_p"])):
            p = m.predict(feats)
            worst = max(worst, float(np.max(np.abs(p - want) / np.maximum(np.abs(want), 1e-3))))
            close(p, want, k * RTOL, 2e-6)
            assert round(roc_auc_score(yy, p), 3) == round(roc_auc_score(yy, want), 3)
    # the script stops one epoch later than the last recorded evaluation pair: run the remaining steps
    while k < len(g["steps"]):
        j = k % n_batch
        m.set_masks(g["steps"][k]["r1"], g["steps"][k]["r2"])
        m.train(train_f[j * N:(j + 1) * N], train_y[j * N:(j + 1) * N])
        k += 1
    close(m.get_table(), g["final_table"], k * RTOL, 2e-6)
    print("FNN_wnzh %d steps on %s: worst relative prediction error %.2e" % (k, gemm_path, worst))


def test_fnn_wnzh_evaluate_on_device(gold):
    from sklearn.metrics import roc_auc_score
    from deepctr_b200.theano_fnn import FNN_wnzh
    g = gold["wnzh"]
    J = max(len(ft) for _, ft in g["test"])
    m = FNN_wnzh(dict(w_0=g["w_0"], table=g["table"], feat_field=g["feat_field"]), batch_size=64, dropout=0.5, max_feats=J)
    m.engine.set_layer(2, np.linspace(-0.3, 0.3, 100).reshape(100, 1), [0.05])
    f, y = _pad([ft for _, ft in g["test"]], J), np.array([yy for yy, _ in g["test"]])
    ev = m.evaluate(f, y)
    p = m.predict(f)
    assert abs(ev["auc"] - roc_auc_score(y, p)) < 1e-9
    assert abs(ev["rmse"] - np.sqrt(np.mean((p.astype(np.float64) - y) ** 2))) < 1e-7


@pytest.mark.parametrize("key", ["FM/sgd/sum", "FM/adam/mean", "LR/sgd/sum", "LR/ftrl/sum"])
def test_fm_lr_cases(gold, key, tmp_path):
    import deepctr_b200.models as M
    r = gold["tf"][key]
    name = key.split("/")[0]
    path = str(tmp_path / "init.pickle")
    with open(path, "wb") as f:
        pickle.dump(r["init"], f, protocol=2)
    m = getattr(M, name)(r["N"], r["rch"], ["uniform", -1, 1, [1, 2], path], r["ptmzr"], [r["lam"]], "train", r["E"])
    for k, st in enumerate(r["steps"]):
        l, p = m.train_step({m.sp_id_hldr: st["ids"].flatten(), m.sp_wt_hldr: st["wts"].flatten(), m.lbl_hldr: st["y"]})
        close(l, st["loss"], (k + 1) * RTOL)
        close(p, st["preds"], (k + 1) * RTOL)
    ep = m.eval_preds.eval({m.eval_id_hldr: r["eval_ids"].flatten(), m.eval_wts_hldr: r["eval_wts"].flatten()})
    close(ep, r["eval_preds"], 4 * RTOL)
    m.dump(str(tmp_path / "out.pickle"))
    with open(str(tmp_path / "out.pickle"), "rb") as f:
        out = pickle.load(f)
    for kname in r["final"]:
        close(out[kname], r["final"][kname], 4 * RTOL)


FNN_KEYS = ["FNN/sgd/0.5/sum/tanh", "FNN/adam/0.5/sum/relu", "FNN/ftrl/1/mean/sigmoid", "FNN_IP_L3/sgd/0.5/sum/tanh",
            "FNN_IP_L5/sgd/1/mean/relu", "FNN_IP_L7/adam/0.5/sum/tanh", "FNN_OP_L3/sgd/0.5/sum/tanh"]


@pytest.mark.parametrize("key", FNN_KEYS)
def test_fnn_family_cases(gold, key, tmp_path, gemm_path):
    import deepctr_b200.models as M
    r = gold["tf"][key]
    name = key.split("/")[0]
    path = str(tmp_path / "init.pickle")
    with open(path, "wb") as f:
        pickle.dump(r["init"], f, protocol=2)
    N = r["steps"][0]["y"].shape[0]
    E = r["eval_ids"].shape[0]
    m = getattr(M, name)(r["cat_sizes"], list(r["offsets"]), N, r["rch"], ["uniform", -1, 1, list(range(1, 20)), path], r["ptmzr"],
                         [r["keep"]], "train", E)
    # Adam's first steps divide by sqrt(v) ~ |g|: a relative error of the gradient passes straight into the update
    tol = 4 * RTOL if r["ptmzr"][0] != "sgd" else RTOL
    for k, st in enumerate(r["steps"]):
        if st["masks"] is not None:
            m.set_dropout_masks(st["masks"])
        l, p = m.train_step({m.v_wt_hldr: st["v_wts"], m.c_id_hldr: st["c_ids"], m.c_wt_hldr: st["c_wts"], m.lbl_hldr: st["y"]})
        close(l, st["loss"], (k + 1) * tol)
        close(p, st["preds"], (k + 1) * tol)
    ep = m.eval_preds.eval({m.eval_id_hldr: r["eval_ids"], m.eval_wts_hldr: r["eval_wts"]})
    close(ep, r["eval_preds"], 4 * tol)
    vm = m.var_map()
    for kname in r["final"]:
        close(vm[kname], r["final"][kname], 4 * tol, 4e-6)


def test_out_of_range_id_raises(gold, tmp_path):
    """TF raises InvalidArgument for an id outside the table (asserted on the reference in tests/test_oracle_vs_ref.py); the
    build reads such an id as row 0, never writes through it, and raises after the step."""
    import deepctr_b200.models as M
    r = gold["tf"]["FM/sgd/sum"]
    path = str(tmp_path / "init.pickle")
    with open(path, "wb") as f:
        pickle.dump(r["init"], f, protocol=2)
    m = M.FM(r["N"], r["rch"], ["uniform", -1, 1, [1, 2], path], r["ptmzr"], [r["lam"]], "train", r["E"])
    st = r["steps"][0]
    before = m.engine.get_table().copy()
    for bad in (r["rch"][0], -1, 2 ** 31 - 5):
        ids = st["ids"].copy()
        ids[2, 1] = bad
        with pytest.raises(IndexError):
            m.train_step({m.sp_id_hldr: ids.flatten(), m.sp_wt_hldr: st["wts"].flatten(), m.lbl_hldr: st["y"]})
    assert np.isfinite(m.engine.get_table()).all() and m.engine.get_table().shape == before.shape
    l, p = m.train_step({m.sp_id_hldr: st["ids"].flatten(), m.sp_wt_hldr: st["wts"].flatten(), m.lbl_hldr: st["y"]})   # status was reset
    assert np.isfinite(l)


def test_fnn_wnzh_epoch_protocol(gold):
    """FNN_wnzh.fit = the script's epoch loop with its patience rule and best-weight snapshot (FNN_wnzh.py:286-343): fed the
    dropout draws the reference run logged, it stops after the same epoch, reports the same best epoch and test AUC, and
    get_pred's predictions with the snapshot (best MLP weights on the FINAL table, :257-263, 346-347) agree."""
    from deepctr_b200.theano_fnn import FNN_wnzh
    g = gold["wnzh"]
    h = g["hyper"]
    N = h["batch_size"]
    J = max(len(ft) for _, ft in g["train"] + g["test"])
    m = FNN_wnzh(dict(w_0=g["w_0"], table=g["table"], feat_field=g["feat_field"]), batch_size=N, lr=h["lr"], dropout=h["dropout"],
                 hidden1=h["hidden1"], hidden2=h["hidden2"], lambda1=h["lambda1"], lambda_fm=h["lambda_fm"], acti_type=h["acti_type"],
                 max_feats=J)
    train_f, train_y = _pad([ft for _, ft in g["train"]], J), np.array([y for y, _ in g["train"]])
    test_f, test_y = _pad([ft for _, ft in g["test"]], J), np.array([y for y, _ in g["test"]])
    res = m.fit(train_f, train_y, test_f, test_y, epochs=100, masks=iter((s["r1"], s["r2"]) for s in g["steps"]))
    assert len(res["history"]) == len(g["epochs"]) and res["min_err_epoch"] == g["min_err_epoch"]
    assert abs(res["min_err"] - g["min_err"]) < 5e-4
    m.set_weights(res["best"])
    close(m.predict(test_f), g["best_test_pred"], len(g["steps"]) * RTOL, 2e-6)
