"""oracle/deepctr_oracle.py == the reference's OWN code (oracle/_ref: /root/reference/python/*.py after mechanical
Py2->3 fixes, executed on the theano / tensorflow shims of oracle/shims/).  CPU only.  Skipped when neither
/root/reference nor a prebuilt oracle/_ref is present."""
import os
import pickle

import numpy as np
import pytest

import ref_run
import ref_cases as C
import deepctr_oracle as O

pytestmark = pytest.mark.skipif(not ref_run.available(), reason="oracle/_ref not built and /root/reference absent")


def close(a, b, rtol=1e-9, atol=1e-12):
    np.testing.assert_allclose(np.asarray(a, np.float64), np.asarray(b, np.float64), rtol=rtol, atol=atol)


def read_fm_lines(path):
    out = []
    with open(path) as f:
        for line in f:
            s = line.replace(":", " ").split()
            if s:
                out.append((int(s[0]), [int(s[j]) for j in range(1, len(s), 2)]))
    return out


# ======================================================================================================================
# Theano generation: the whole FNN_wnzh.py script
# ======================================================================================================================
@pytest.fixture(scope="module", params=[False, True], ids=["one_feature_per_field", "duplicate_and_missing_fields"])
def wnzh_run(request, tmp_path_factory):
    wd = str(tmp_path_factory.mktemp("wnzh"))
    case = C.write_s1_demo(wd, dup_field=request.param)
    g, calls = ref_run.run_fnn_wnzh(wd)
    return wd, case, g, calls


def test_fnn_wnzh_script_equals_oracle(wnzh_run):
    """Every train() / predict() call the reference script makes, replayed on the oracle from the same files: the dense
    input x it gathered (i.e. the table state after all earlier updates), gx, the weights it returned, every prediction,
    the final feat_weights dict."""
    wd, case, g, calls = wnzh_run
    F, d, X = case["F"], case["d"], case["X"]
    assert g["k"] == d and g["xdim"] == 1 + F * d and g["batch_size"] == 100
    assert g["lr"] == 0.001 and g["dropout"] == 0.5 and g["lambda_fm"] == 0.1 and g["lambda1"] == 0.0   # advertiser 2997 branch
    # init (FNN_wnzh.py:16-17, 106-130): numpy legacy global RNG seeded 1234
    Ws, bs = O.theano_fnn_init(g["xdim"], 300, 100, "tanh")
    train_calls = [c for c in calls if c["fn"] == "function1"]
    close(train_calls[0]["outputs"][1], Ws[0], 0, 0)
    close(train_calls[0]["outputs"][2], Ws[1], 0, 0)
    # only r1 and r2 are drawn: the input mask r0 is cut out of the compiled function
    assert sorted(train_calls[0]["random"]) == ["binomial1", "binomial2"]
    table = np.concatenate([case["table"], np.zeros((1, d))], 0)
    net = O.OracleNet(table, case["w_0"], Ws, bs, F, d, "none", True, "none", "tanh")
    feat_field = {f: int(case["feat_field"][f]) for f in range(X)}
    assert g["feat_field"] == feat_field
    train = read_fm_lines(os.path.join(wd, "data", "train.fm.txt"))
    test = read_fm_lines(os.path.join(wd, "data", "test.fm.txt"))
    n_batch = len(train) // 100
    it = iter(calls)
    n_epochs = len(train_calls) // n_batch
    for ep in range(n_epochs):
        for j in range(n_batch):
            c = next(it)
            assert c["fn"] == "function1"
            batch = train[j * 100:(j + 1) * 100]
            feats, y = [b[1] for b in batch], np.array([b[0] for b in batch])
            masks = [None, c["random"]["binomial1"].astype(np.float64), c["random"]["binomial2"].astype(np.float64)]
            pre = [w.copy() for w in net.Ws] + [b.copy() for b in net.bs]
            loss, p, aux = net.train_step_theano_feats(feats, case["feat_field"], y, g["lr"], g["lambda_fm"], masks, g["lambda1"])
            close(c["inputs"][0], aux["x"])                      # feats_to_layer_one_array on the current feat_weights
            close(c["inputs"][1], y)
            gx, w1, w2, w3, b1, b2, b3 = c["outputs"]
            close(gx, aux["gx"], 1e-8, 1e-13)
            for got, want in zip((w1, w2, w3, b1, b2, b3), (pre[0], pre[1], pre[2].reshape(-1), pre[3], pre[4], pre[5][0])):
                close(got, want, 1e-8, 1e-13)                    # train() returns the PRE-update weights
        for which in (train, test):                              # print_err(train_file), get_err_bat(test_file): one chunk each
            c = next(it)
            assert c["fn"] == "function2"
            ids = np.full((len(which), F), X, dtype=np.int64)
            for t, (_, ft) in enumerate(which):
                for feat in ft:
                    ids[t, feat_field[feat]] = feat
            close(c["outputs"][0], net.predict(ids), 1e-8, 1e-13)
    fw = g["feat_weights"]
    got = np.array([fw[f] for f in range(X)])
    close(got, net.table[:X], 1e-8, 1e-13)
    assert not net.table[X].any()


def test_fnn_wnzh_vectorised_scatter_equals_script(wnzh_run):
    """the sort + ordered-fold formulation the CUDA path uses (stable sort by row, per-occurrence affine recurrence) gives
    the reference's table on the update list of ALL features (slot = field of the feature)."""
    wd, case, g, calls = wnzh_run
    F, d, X = case["F"], case["d"], case["X"]
    train = read_fm_lines(os.path.join(wd, "data", "train.fm.txt"))
    c = [c for c in calls if c["fn"] == "function1"][0]
    batch = train[:100]
    gx = c["outputs"][0]
    rows = np.array([feat for _, ft in batch for feat in ft])
    slots = np.array([t * F + case["feat_field"][feat] for t, (_, ft) in enumerate(batch) for feat in ft])
    ge = gx[:, 1:].reshape(100 * F, d)[slots]
    table = case["table"].copy()
    O.seq_decay_scatter(table, rows, ge, g["lr"], 1 - 2. * g["lambda_fm"] * g["lr"] / 100)
    table2 = case["table"].copy()
    O.seq_decay_scatter_closed_form(table2, rows, ge, g["lr"], 1 - 2. * g["lambda_fm"] * g["lr"] / 100)
    # reference state after its first update loop = the x it gathers for ... rebuild from the second call is indirect;
    # replay the literal loop instead (it is the script's own statement, asserted above)
    want = case["table"].copy()
    for t, (_, ft) in enumerate(batch):
        for feat in ft:
            for l in range(d):
                want[feat, l] = want[feat, l] * (1 - 2. * g["lambda_fm"] * g["lr"] / 100) - g["lr"] * gx[t, g["feat_layer_one_index"](feat, l)]
    close(table, want, 1e-12, 1e-15)
    close(table2, want, 1e-10, 1e-14)
