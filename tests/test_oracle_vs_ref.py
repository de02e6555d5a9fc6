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


# ======================================================================================================================
# TF generation: FM / LR / FNN / FNN_IP_L3 / L5 / L7 / FNN_OP_L3 graph builders on the tensorflow shim
# ======================================================================================================================
@pytest.fixture(scope="module")
def tfm():
    return ref_run.tf_modules()


def _binary_masks(rng, keep, shapes):
    return [np.floor(keep + rng.random(s)) for s in shapes]


PTMZR = {"sgd": lambda r: ["sgd", 0.05, r], "adam": lambda r: ["adam", 0.01, 1e-8, r], "ftrl": lambda r: ["ftrl", 0.05, r]}


@pytest.mark.parametrize("ptmzr", ["sgd", "adam", "ftrl"])
@pytest.mark.parametrize("name", ["FM", "LR"])
@pytest.mark.parametrize("reduce", ["sum", "mean"])
def test_fm_lr_steps_equal_oracle(tfm, tmp_path, name, ptmzr, reduce):
    run_fm_case(tfm, tmp_path, name, ptmzr, reduce)


def run_fm_case(tfm, tmp_path, name, ptmzr, reduce, record=None):
    tf = tfm["tf"]
    rec = record if record is not None else {}
    c = C.tf_case(name)
    rng, N, E, X, F, r, lam = c["rng"], c["N"], c["E"], c["X_dim"], c["X_feas"], c["rank"], c["lam"]
    vm = {"W": rng.uniform(-0.3, 0.3, (X, 1)).astype(np.float32), "b": np.array([0.05], np.float32)}
    if name == "FM":
        vm["V"] = rng.uniform(-0.3, 0.3, (X, r)).astype(np.float32)
    path = C.dump_pickle(vm, str(tmp_path / "init.pickle"))
    pt = PTMZR[ptmzr](reduce)
    rch = [X, F, r] if name == "FM" else [X, F]
    model = getattr(tfm[name], name)(N, rch, ["uniform", -1, 1, [1, 2], path], pt, [lam], "train", E)
    table = np.concatenate([vm["W"], vm["V"]], 1) if name == "FM" else vm["W"]
    ofm = O.OracleFM(table, vm["b"][0])
    rec.update(init=vm, rch=rch, ptmzr=pt, lam=lam, N=N, E=E, steps=[])
    with tf.Session(graph=model.graph) as sess:
        tf.initialize_all_variables().run()
        for step in range(3):
            ids = rng.integers(0, X, (N, F))
            ids[1] = ids[0]
            ids[:, -1] = X - 1                                    # the padded slot of ipinyou.py:58-59: id = X_dim - 1, weight 0
            wts = rng.uniform(0.5, 1.5, (N, F)).astype(np.float32)
            wts[:, -1] = 0
            y = rng.integers(0, 2, N)
            _, l, p = sess.run([model.ptmzr, model.loss, model.train_preds],
                               feed_dict={model.sp_id_hldr: ids.flatten(), model.sp_wt_hldr: wts.flatten(), model.lbl_hldr: y})
            if ptmzr == "sgd":
                ol, op = ofm.train_step(ids, wts, y, pt[1], lam, reduce)
            elif ptmzr == "adam":
                ol, op = ofm.train_step_adam(ids, wts, y, pt[1], pt[2], lam, reduce)
            else:
                ol, op = ofm.train_step_ftrl(ids, wts, y, pt[1], lam, reduce)
            close(l, ol, 1e-10)
            close(p, op, 1e-10)
            rec["steps"].append(dict(ids=ids, wts=wts, y=y, loss=float(l), preds=np.asarray(p)))
        eids, ewts = rng.integers(0, X, (E, F)), rng.uniform(0.5, 1.5, (E, F)).astype(np.float32)
        wt_h = model.eval_wts_hldr if name == "FM" else model.eval_wt_hldr          # FM.py:45 vs LR.py:44
        ep = model.eval_preds.eval(feed_dict={model.eval_id_hldr: eids.flatten(), wt_h: ewts.flatten()})
        close(ep, ofm.predict(eids, ewts), 1e-10)
        rec.update(eval_ids=eids, eval_wts=ewts, eval_preds=np.asarray(ep))
        model.dump(str(tmp_path / "out.pickle"))
        with pytest.raises(IndexError):                           # TF: InvalidArgument for an id outside the table
            bad = eids.copy()
            bad[0, 0] = X
            model.eval_preds.eval(feed_dict={model.eval_id_hldr: bad.flatten(), wt_h: ewts.flatten()})
    with open(str(tmp_path / "out.pickle"), "rb") as f:
        out = pickle.load(f)
    rec["final"] = out
    got = np.concatenate([out["W"], out["V"]], 1) if name == "FM" else out["W"]
    close(got, ofm.table, 1e-9)
    close(out["b"], [ofm.b], 1e-9)


def run_fnn_case(tfm, tmp_path, name, ptmzr="sgd", keep=1.0, reduce="sum", act="tanh", record=None):
    """3 optimiser steps + eval + dump of one FNN-family class of the reference against the oracle; `record` (a dict)
    receives every input and output (tests/golden/make_golden_ref.py)."""
    tf = tfm["tf"]
    c = C.tf_case(name)
    c["act"] = act
    rng, N, E = c["rng"], c["N"], c["E"]
    cat_sizes, offsets, X_dim, X_feas, d, cw = C.fnn_geometry(c)
    vm, dims = C.fnn_var_map(c)
    path = C.dump_pickle(vm, str(tmp_path / "init.pickle"))
    pt = PTMZR[ptmzr](reduce)
    rch = [X_dim, X_feas, c["rank"]] + c["hidden"] + [act]
    cls = getattr(tfm[name], name)
    model = cls(cat_sizes, offsets, N, rch, ["uniform", -1, 1, list(range(1, 20)), path], pt, [keep], "train", E)
    nl = len(dims) - 1
    net = O.OracleNet(np.concatenate([vm["W"], vm["V"]], 1), vm["b"][0], [vm["h%d_w" % (i + 1)] for i in range(nl)],
                      [vm["h%d_b" % (i + 1)] for i in range(nl)], X_feas, d, c["cross"], False, act, act)
    mrng = np.random.default_rng(99)
    shapes = [(N, w) for w in dims[:-1]]
    rec = record if record is not None else {}
    rec.update(init=vm, cat_sizes=cat_sizes, offsets=np.asarray(offsets), rch=rch, ptmzr=pt, keep=keep, steps=[])
    with tf.Session(graph=model.graph) as sess:
        tf.initialize_all_variables().run()
        for step in range(c["steps"]):
            v_wts, c_ids, c_wts, y = C.fnn_batch(c, N)
            binm = _binary_masks(mrng, keep, shapes) if keep < 1 else None
            tf.DROPOUT_HOOK = (lambda shape, kp, k: binm[k]) if keep < 1 else None
            _, l, p = sess.run([model.ptmzr, model.loss, model.train_preds],
                               feed_dict={model.v_wt_hldr: v_wts, model.c_id_hldr: c_ids, model.c_wt_hldr: c_wts, model.lbl_hldr: y})
            tf.DROPOUT_HOOK = None
            g_ids = np.concatenate([np.broadcast_to(np.arange(13), (N, 13)), c_ids + np.asarray(offsets)], 1)
            g_wts = np.concatenate([v_wts, c_wts], 1)
            masks = [m / keep for m in binm] if keep < 1 else None
            if ptmzr == "sgd":
                ol, op = net.train_step_tf(g_ids, g_wts, y, pt[1], masks, reduce)
            elif ptmzr == "adam":
                ol, op = net.train_step_tf_adam(g_ids, g_wts, y, pt[1], pt[2], masks, reduce)
            else:
                ol, op = net.train_step_tf_ftrl(g_ids, g_wts, y, pt[1], masks, reduce)
            close(l, ol, 1e-9)
            close(p, op, 1e-9)
            rec["steps"].append(dict(v_wts=v_wts, c_ids=c_ids, c_wts=c_wts, y=y, masks=masks, loss=float(l), preds=np.asarray(p)))
        # eval graph: GLOBAL ids [E, X_feas], offsets subtracted in-graph (FNN.py:65-69), no dropout
        v_wts, c_ids, c_wts, _ = C.fnn_batch(c, E)
        g_ids = np.concatenate([np.broadcast_to(np.arange(13), (E, 13)), c_ids + np.asarray(offsets)], 1)
        g_wts = np.concatenate([v_wts, c_wts], 1)
        ep = model.eval_preds.eval(feed_dict={model.eval_id_hldr: g_ids, model.eval_wts_hldr: g_wts})
        close(ep, net.predict(g_ids, g_wts), 1e-9)
        rec.update(eval_ids=g_ids, eval_wts=g_wts, eval_preds=np.asarray(ep))
        model.dump(str(tmp_path / "out.pickle"))
    with open(str(tmp_path / "out.pickle"), "rb") as f:
        out = pickle.load(f)
    rec["final"] = out
    close(np.concatenate([out["W"], out["V"]], 1), net.table, 1e-8, 1e-12)
    close(out["b"], [net.bias], 1e-8, 1e-12)
    for i in range(nl):
        close(out["h%d_w" % (i + 1)], net.Ws[i], 1e-8, 1e-12)
        close(out["h%d_b" % (i + 1)], net.bs[i], 1e-8, 1e-12)
    if c["cross"] != "op":
        assert model.log.startswith("input dim: %d, features: %d, rank: %d, embedding: %d" % (X_dim, X_feas, c["rank"], dims[0]))
    return rec


@pytest.mark.parametrize("name", ["FNN", "FNN_IP_L3", "FNN_IP_L5", "FNN_IP_L7"])
@pytest.mark.parametrize("ptmzr,keep,reduce,act", [("sgd", 1.0, "sum", "tanh"), ("sgd", 0.5, "mean", "relu"), ("adam", 0.5, "sum", "sigmoid"),
                                                    ("ftrl", 1.0, "sum", "tanh")])
def test_fnn_family_steps_equal_oracle(tfm, tmp_path, name, ptmzr, keep, reduce, act):
    """the reference's own graph-building code (pair order, concat order, bias column last, activation applied to the
    layer INPUT incl. z1 (Q7), dropout on every layer input) against the oracle, 3 optimiser steps + eval + dump."""
    run_fnn_case(tfm, tmp_path, name, ptmzr, keep, reduce, act)


def test_fnn_op_l3_declared_width_defect(tfm, tmp_path):
    """FNN_OP_L3.py:18,24 declares h1_w with P = F(F-1)/2 cross columns while forward() produces d*d (:93-100).  The
    placeholders fix F = 39 (P = 741, never a square), so with the DECLARED shape the reference's own matmul fails on
    every configuration ..."""
    tf = tfm["tf"]
    c = C.tf_case("FNN_OP_L3")
    cat_sizes, offsets, X_dim, X_feas, d, cw = C.fnn_geometry(c)
    P = X_feas * (X_feas - 1) // 2
    assert d * d != P
    vm, dims = C.fnn_var_map(c, declared_width=X_feas * d + P + 1)
    path = C.dump_pickle(vm, str(tmp_path / "init.pickle"))
    model = tfm["FNN_OP_L3"].FNN_OP_L3(cat_sizes, offsets, c["N"], [X_dim, X_feas, c["rank"]] + c["hidden"] + ["tanh"],
                                      ["uniform", -1, 1, list(range(1, 20)), path], ["sgd", 0.05, "sum"], [1.0], "train", c["E"])
    assert ("embedding: %d" % (X_feas * d + P + 1)) in model.log
    v_wts, c_ids, c_wts, y = C.fnn_batch(c, c["N"])
    with tf.Session(graph=model.graph) as sess:
        tf.initialize_all_variables().run()
        with pytest.raises(ValueError, match="matmul"):
            sess.run([model.ptmzr, model.loss], feed_dict={model.v_wt_hldr: v_wts, model.c_id_hldr: c_ids, model.c_wt_hldr: c_wts,
                                                           model.lbl_hldr: y})


@pytest.mark.parametrize("ptmzr,keep", [("sgd", 1.0), ("adam", 0.5)])
def test_fnn_op_l3_true_width_equals_oracle(tfm, tmp_path, ptmzr, keep):
    """... but init_var_map takes a checkpointed h1_w as it is (tf_util.py:43-45), so a pickle whose h1_w has the
    forward's TRUE width F*d + d*d + 1 makes the reference's own OP graph run: that is the build's sizing (Q2), and there
    the reference equals the oracle (outer products summed over pairs i<j, flattened row-major)."""
    run_fnn_case(tfm, tmp_path, "FNN_OP_L3", ptmzr, keep, "sum", "tanh")


def test_build_inds(tfm):
    """tf_util.py:7-12: COO indices, sample-major -- the flat [N*F] order of sp_id_hldr."""
    inds = tfm["tf_util"].build_inds(3, 4)
    assert inds.shape == (12, 2)
    assert (inds[:, 0] * 4 + inds[:, 1] == np.arange(12)).all()


# ======================================================================================================================
# host logic either side of the step: baseline.py / ipinyou.py functions, executed as they are
# ======================================================================================================================
def _tiny_criteo(workdir, n_train=47, n_eval=30, seed=3):
    from deepctr_b200 import data as D
    c = C.tf_case("FNN")
    rng = np.random.default_rng(seed)
    cat_sizes = c["cat_sizes"]
    paths = {}
    for name, n in (("train", n_train), ("eval", n_eval)):
        y = rng.integers(0, 2, n)
        num = np.stack([rng.integers(0, int(m) + 1, n) for m in D.CRITEO_MAX_VALS], 1)
        cat = np.stack([rng.integers(0, s, n) for s in cat_sizes], 1)
        paths[name] = os.path.join(workdir, "data", name + ".tsv")
        os.makedirs(os.path.dirname(paths[name]), exist_ok=True)
        D.write_criteo_tsv(paths[name], y, num, cat)
    return c, paths


@pytest.fixture()
def baseline_mod(tmp_path):
    wd = str(tmp_path)
    c, paths = _tiny_criteo(wd)
    b = ref_run.load_baseline(wd)
    # the module body hard-codes the Criteo geometry (baseline.py:30-43); point its globals at the tiny set instead
    b.cat_sizes = np.array(c["cat_sizes"])
    b.offsets = [13 + sum(b.cat_sizes[:i]) for i in range(len(b.cat_sizes))]        # :40, same expression
    b.X_dim, b.X_feas = 13 + np.sum(b.cat_sizes), 13 + len(b.cat_sizes)
    b.train_path, b.eval_path, b.test_path, b.data_set = paths["train"], paths["eval"], paths["eval"], None
    return wd, c, paths, b


def test_get_fxy_and_get_batch_xy(baseline_mod):
    """get_fxy (baseline.py:190-198) == the text oracle and data.parse_criteo_line; get_batch_xy (:233-243: at end of file
    reopen and top up) == ingest.LineCursor(wrap=True) over the parsed lines."""
    import text_oracle as T
    from deepctr_b200 import data as D
    from deepctr_b200.ingest import LineCursor
    wd, c, paths, b = baseline_mod
    lines = [ln for ln in open(paths["train"]).read().split("\n") if ln]
    for ln in lines[:10]:
        y, f, x = b.get_fxy(ln)
        y2, f2, x2 = T.criteo_get_fxy(ln, b.offsets, b.max_vals)
        y3, f3, x3 = D.parse_criteo_line(ln, b.offsets, b.max_vals)
        assert (y, list(f)) == (y2, list(f2)) == (y3, list(f3))
        assert list(x) == list(x2) == list(x3)
    ids, wts, ys = T.criteo_block(lines, b.offsets, b.max_vals)
    cur = LineCursor(len(lines), wrap=True)
    for size in (20, 20, 20, 5, 47, 13):
        labels, rows, cols, vals = b.get_batch_xy(size)
        idx = cur.next(size)
        assert (cols == ids[idx]).all() and (labels == ys[idx]).all()
        np.testing.assert_array_equal(np.float32(vals), wts[idx])


def test_early_stop_and_watch_train(baseline_mod):
    from deepctr_b200.metrics import early_stop, recalibrate
    from sklearn.metrics import roc_auc_score
    wd, c, paths, b = baseline_mod
    rng = np.random.default_rng(0)
    for trial in range(300):
        b.least_step, b.skip_window = int(rng.integers(0, 50)), int(rng.integers(1, 4))
        b.smooth_window, b.stop_window = int(rng.integers(1, 6)), int(rng.integers(2, 6))
        step = int(rng.integers(0, 100))
        errs = list(rng.random(int(rng.integers(1, 40))))
        for metric in ("auc", "rmse"):
            import contextlib, io
            with contextlib.redirect_stdout(io.StringIO()):
                want = b.early_stop(step, list(errs), metric)
            got = early_stop(step, errs, metric, b.least_step, b.skip_window, b.smooth_window, b.stop_window)
            assert got == want == O.early_stop_literal(step, errs, metric, b.least_step, b.skip_window, b.smooth_window, b.stop_window)
    with ref_run.in_reference_cwd(wd):
        m = b.watch_train(7, [0, 1, 1, 0], [0.1, 0.8, 0.4, 0.3], np.array([0.2, 0.9, 0.5]), [0, 1, 1])
        assert m["batch_auc"] == np.float32(roc_auc_score([0, 1, 1, 0], [0.1, 0.8, 0.4, 0.3]))
        # (single-class input: the 2016 sklearn raised ValueError -> -1, :249-254; sklearn 1.9 here warns and returns nan, so
        #  that branch of the reference cannot be exercised; the build's device metric returns -1 there, test_kernels_gpu)
    p = rng.random(50)
    q = p.copy()
    q /= q + (1 - q) / b.nds_rate                                 # baseline.py:369-370, the statement itself
    close(recalibrate(p, b.nds_rate), q, 1e-15)


def test_baseline_train_loop_runs_the_reference_protocol(baseline_mod):
    """baseline.train() itself (baseline.py:284-379) on the tiny set: the loop feeds c_ids minus offsets, evaluates every
    `epoch` steps with recalibration, dumps a checkpoint and stops through early_stop.  The oracle driven by the same lines
    reproduces every logged loss; the build's Session / Fetch / dump protocol mirrors these calls (models.py)."""
    import text_oracle as T
    wd, c, paths, b = baseline_mod
    vm, dims = C.fnn_var_map(c)
    init = C.dump_pickle(vm, os.path.join(wd, "init.pickle"))
    b.algo, b.batch_size, b.eval_size, b.eval_buf_size, b.buffer_size = "FNN", 4, 10, 30, 20
    b.epoch, b.ckpt, b.least_step, b.skip_window, b.smooth_window, b.stop_window = 3, 6, 6, 1, 2, 2
    b._rch_argv = [b.X_dim, b.X_feas, c["rank"]] + c["hidden"] + ["tanh"]
    b._init_argv, b._ptmzr_argv, b._reg_argv = ["uniform", -1, 1, list(range(9)), init], ["sgd", 0.05, "sum"], [1.0]
    b.log_path, b.model_path = os.path.join(wd, "log", "run"), os.path.join(wd, "model", "m.pickle")
    np.random.seed(5)
    import contextlib, io
    out = io.StringIO()
    with ref_run.in_reference_cwd(wd), contextlib.redirect_stdout(out):
        b.train()
    text = out.getvalue()
    assert "early stop at step" in text
    losses = [float(ln.split("loss: ")[1].split("\t")[0]) for ln in text.split("\n") if ln.startswith("step: ")]
    stop_step = int(text.split("early stop at step ")[1].split()[0])
    assert stop_step % b.epoch == 0 and len(losses) == stop_step // b.epoch
    # replay on the oracle: same ring of lines, batches of 4 out of buffers of 20 (ragged buffer tail dropped, :338)
    lines = [ln for ln in open(paths["train"]).read().split("\n") if ln]
    ids, wts, ys = T.criteo_block(lines, b.offsets, b.max_vals)
    nl = len(dims) - 1
    net = O.OracleNet(np.concatenate([vm["W"], vm["V"]], 1), vm["b"][0], [vm["h%d_w" % (i + 1)] for i in range(nl)],
                      [vm["h%d_b" % (i + 1)] for i in range(nl)], b.X_feas, c["rank"] + 1, "none", False, "tanh", "tanh")
    from deepctr_b200.ingest import LineCursor
    cur, step, got = LineCursor(len(lines), wrap=True), 0, []
    while step < stop_step:
        idx = cur.next(20)
        for i in range(len(idx) // 4):
            sel = idx[i * 4:(i + 1) * 4]
            l, _ = net.train_step_tf(ids[sel], wts[sel].astype(np.float64), ys[sel], 0.05)
            step += 1
            if step % b.epoch == 0:
                got.append(l)
            if step == stop_step:
                break
    close(losses, got, 1e-5)                                       # '%g' prints 6 significant digits
    with open(b.model_path + "_%d" % 6, "rb") as f:                # ckpt every 6 steps (:376-377)
        assert sorted(pickle.load(f)) == sorted(vm)


def test_ipinyou_readers(tmp_path):
    """collect / stat / load_ipinyou_data (ipinyou.py:11-65) == ingest.chunk_shuffle_perm + the yzx oracle + data.stat_yzx."""
    import text_oracle as T
    from deepctr_b200 import data as D
    from deepctr_b200.ingest import chunk_shuffle_perm
    ip = ref_run.load_ipinyou()
    rng = np.random.default_rng(2)
    n = 257
    ids = [sorted(set(rng.integers(0, 500, int(rng.integers(3, 17))).tolist())) for _ in range(n)]
    y = rng.integers(0, 2, n)
    path = str(tmp_path / "train.yzx.txt")
    with open(path, "w") as f:
        for t in range(n):
            f.write("%d %d %s\n" % (y[t], 0, " ".join("%d:1" % i for i in ids[t])))
    np.random.seed(11)
    max_dim, max_fea = ip.stat(path)
    assert (max_dim, max_fea) == D.stat_yzx(path)
    X_dim = max_dim + 2                                           # ipinyou.py:119
    lines = open(path).read().split("\n")[:-1]
    np.random.seed(42)
    perm = chunk_shuffle_perm(n, 100)
    np.random.seed(42)
    fin = open(path)
    got_ids, got_vals, got_y = [], [], []
    while True:
        X_ind, X_val, yy = ip.load_ipinyou_data(fin, 100, X_dim - 1, max_fea)
        if X_ind is None:
            break
        got_ids.append(X_ind), got_vals.append(X_val), got_y.append(yy)
    feats, wts, yo, nnz = T.sparse_block([lines[k] for k in perm], max_fea, X_dim - 1, 1)
    assert (np.concatenate(got_ids) == feats).all() and (np.concatenate(got_vals) == wts).all() and (np.concatenate(got_y) == yo).all()


def test_bench_times_the_reference_script_itself():
    """bench.py's B0 leg: the converted FNN_wnzh.py runs under a time budget (the Theano stand-in stops it at the first call after the
    deadline) and the loop iterations are timed from the call log; the whole-script run of the other tests is unaffected by the
    budget (it is reset)."""
    import bench
    b0 = bench.reference_script_baseline(budget_s=1.5, n_train=400, n_test=100)
    assert b0 is not None and "unavailable" not in b0 and b0["kind"] == "reference" and b0["cores"] == 1
    assert b0["value"] > 0 and b0["value_without_train_call"] > b0["value"] and 0 < b0["train_call_ms"] < b0["ms_per_step"]
    import sys
    assert sys.modules["theano"].BUDGET_S is None


@pytest.mark.parametrize("script,lr", [("SNN_DAE.py", 0.0005), ("SNN_RBM.py", 0.001)])
def test_snn_dae_script_equals_oracle(tmp_path, script, lr):
    """the reference's own SNN_DAE.py / SNN_RBM.py (the same graph and loop; lr 0.0005 / 0.001, dropout 0.99 / 0.98 for
    advertiser 2997, SNN_DAE.py:41-49, SNN_RBM.py:52-58) (converted, run top to bottom from the weight pickle it loads, SNN_DAE.py:83-86)
    against OracleSNN, call by call: the dense inputs x = sigmoid(sum ww0[feat] + bb0) of every train() call (get_fi_h1_y :234-260 --
    they carry the ww0 / bb0 updates of all earlier steps, :282-288), gx and the weights train() returns (:113-163, dropout draws
    taken from the call log), and the per-epoch predict() passes over the train and test files (:164-196)."""
    rng = np.random.default_rng(3)
    X_DIM, H0, H1, H2, N = 133465, 200, 300, 100, 1000          # x_dim, hidden0/1/2 of advertiser 2997 and batch_size (:19,41-45)
    lambda1 = 0.0                                                 # :46-49
    wd = str(tmp_path)
    os.makedirs(os.path.join(wd, "data"))
    n_ids = 4000

    def write(name, n):
        feats = np.stack([rng.integers(0, n_ids // 16, n) + f * (n_ids // 16) for f in range(16)], 1)
        feats[5] = feats[4]                                        # a repeated sample: its rows are updated twice, in order
        y = rng.integers(0, 2, n)
        with open(os.path.join(wd, "data", name), "w") as f:
            for t in range(n):
                f.write("%d %s\n" % (y[t], " ".join("%d:1" % i for i in feats[t])))
        return feats, y
    (tr_f, tr_y), (te_f, te_y) = write("train.fm.txt", 2 * N), write("test.fm.txt", 300)
    ww0, bb0 = rng.uniform(-0.05, 0.05, (X_DIM, H0)), rng.uniform(-0.1, 0.1, H0)
    ww1, bb1 = rng.uniform(-0.1, 0.1, (H0, H1)), rng.uniform(-0.1, 0.1, H1)
    ww2, bb2 = rng.uniform(-0.1, 0.1, (H1, H2)), rng.uniform(-0.1, 0.1, H2)
    _, calls = ref_run.run_snn_dae(wd, (ww0, bb0, ww1, bb1, ww2, bb2), budget_s=4.0, script=script)
    net = O.OracleSNN(ww0, bb0, [ww1, ww2, np.zeros((H2, 1))], [bb1, bb2, np.zeros(1)], act="tanh")      # ww3 = 0, bb3 = 0 (:93-94)
    steps = epochs = 0
    k = 0
    close = lambda a, b, tol=1e-9: np.testing.assert_allclose(np.asarray(a, np.float64), np.asarray(b, np.float64), rtol=tol, atol=1e-12)
    while k < len(calls):
        c = calls[k]
        if c["fn"] == "function1":                                # train(x, y) of batch j = steps % 2
            j = steps % 2
            feats, y = tr_f[j * N:(j + 1) * N], tr_y[j * N:(j + 1) * N]
            x = net.dense(feats)
            close(c["inputs"][0], x)                              # the bag forward on the CURRENT ww0 / bb0
            close(c["inputs"][1], y)
            r1, r2 = c["random"]["binomial0"], c["random"]["binomial1"]
            assert r1.shape == (1, H1) and r2.shape == (1, H2)
            _, _, gx, _, _ = net.loss_and_grads(x, y.astype(np.float64), r1, r2, lambda1)
            out = c["outputs"]
            close(out[0], gx, 1e-8)
            close(out[1], net.Ws[0]); close(out[2], net.Ws[1]); close(out[3], net.Ws[2].reshape(-1))      # pre-update values
            close(out[4], net.bs[0]); close(out[5], net.bs[1]); close(out[6], net.bs[2][0])
            net.train_step(feats, y, lr, r1, r2, lambda1, faithful=True)
            steps += 1
            k += 1
        else:                                                     # the epoch's two predict() passes: train file, then test file
            close(c["outputs"][0], net.predict(tr_f), 1e-8)
            close(calls[k + 1]["outputs"][0], net.predict(te_f), 1e-8) if k + 1 < len(calls) else None
            epochs += 1
            k += 2
    assert steps >= 4 and epochs >= 2, (steps, epochs)
    # the vectorised update of the oracle (what the GPU tests compare the CUDA bag step with) equals the literal loop
    a = O.OracleSNN(ww0[:n_ids], bb0, [ww1, ww2, rng.uniform(-0.1, 0.1, (H2, 1))], [bb1, bb2, np.zeros(1)])
    b = O.OracleSNN(ww0[:n_ids], bb0, a.Ws, a.bs)
    a.train_step(tr_f[:64], tr_y[:64], 0.05, faithful=True)
    b.train_step(tr_f[:64], tr_y[:64], 0.05, faithful=False)
    close(a.ww0, b.ww0, 1e-9); close(a.bb0, b.bb0, 1e-9)


def test_data_fm_class_equals_the_loaders(tmp_path):
    """data_fm.DataFM (data_fm.py:14-84, the class form of FNN_wnzh.py:62-96, 224-253), run as it is: the parsed model (w_0, k,
    xdim, feat_weights, feat_field), feat_layer_one_index, and get_batch_data -- feature lists, labels and the host-gathered dense
    x of every line, duplicate-field and missing-field lines included -- equal data.load_fm_model / load_fm_samples and the
    oracle's gather of the same ids."""
    import importlib
    import sys
    from deepctr_b200 import data as D
    case = C.write_s1_demo(str(tmp_path), dup_field=True)
    model_path = os.path.join(str(tmp_path), "data", "fm.model.txt")
    train_path = os.path.join(str(tmp_path), "data", "train.fm.txt")
    sys.modules.pop("data_fm", None)
    with ref_run._paths():
        dfm = importlib.import_module("data_fm").DataFM(model_path)
        f_ref, x_ref, y_ref = dfm.get_batch_data(train_path, 1, 60)          # lines 1..60 (linecache is 1-based)
        idx = [dfm.feat_layer_one_index(feat, l) for feat in (0, 5, case["X"] - 1) for l in (0, 3, dfm.k - 1)]
    m = D.load_fm_model(model_path)
    assert dfm.w_0 == m["w_0"] and dfm.k == m["k"] == case["d"] and dfm.xdim == 1 + 16 * m["k"]
    assert np.array_equal(np.array([dfm.feat_weights[f] for f in range(case["X"])]), m["table"])
    assert np.array_equal(np.array([dfm.feat_field[f] for f in range(case["X"])]), m["feat_field"])
    assert idx == [1 + m["feat_field"][feat] * m["k"] + l for feat in (0, 5, case["X"] - 1) for l in (0, 3, m["k"] - 1)]
    feats, y = D.load_fm_samples(train_path)
    assert np.array_equal(y_ref, y[:60]) and [list(r[r >= 0]) for r in feats[:60]] == [list(r) for r in f_ref]
    # the dense rows: slot of a feature = its field, a later feature of the same field overwrites, a missing field stays 0
    x = np.zeros((60, dfm.xdim))
    x[:, 0] = m["w_0"]
    for t in range(60):
        for feat in feats[t][feats[t] >= 0]:
            o = 1 + m["feat_field"][feat] * m["k"]
            x[t, o:o + m["k"]] = m["table"][feat]
    assert np.array_equal(x_ref, x)
    assert any(len(r) == 17 for r in f_ref) and any(len(r) == 15 for r in f_ref)      # the adversarial lines are inside the sample
