"""GPU parity of full train / predict steps through the reference-facing model classes against the oracle:
FM, FNN, FNN_IP_L3/L5/L7, FNN_OP_L3 (TF semantics) and FNN_wnzh (Theano semantics).  Weights are injected
through the reference's own pickle-checkpoint path (tf_util.py:41-49) so both sides start identical."""
import pickle

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

RTOL = 1e-5


@pytest.fixture(scope="module", autouse=True)
def _need_cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


@pytest.fixture(autouse=True, params=["simt", "tc", "mlp2", "mlp2tc"])
def gemm_path(request, monkeypatch):
    """Every model-level parity test runs on the per-layer GEMM paths (fp32 FFMA, tcgen05 3xTF32) and with the fused
    two-hidden-layer kernels of csrc/mlp2.cu (which only apply to models with two hidden layers)."""
    monkeypatch.setenv("DCTR_GEMM", {"mlp2": "simt", "mlp2tc": "tc"}.get(request.param, request.param))
    monkeypatch.setenv("DCTR_MLP2", {"mlp2": "on", "mlp2tc": "tc"}.get(request.param, "off"))
    return request.param


def close(a, b, rtol=RTOL, atol=1e-6):
    np.testing.assert_allclose(np.asarray(a, np.float64), np.asarray(b, np.float64), rtol=rtol, atol=atol)


def criteo_like(rng, n_num, cat_sizes):
    cat_sizes = np.asarray(cat_sizes)
    offsets = [n_num + int(cat_sizes[:i].sum()) for i in range(len(cat_sizes))]
    return cat_sizes, offsets, n_num + int(cat_sizes.sum()), n_num + len(cat_sizes)


def make_fnn(cls_name, tmp_path, rng, N, hidden, act, n_num=3, cat_sizes=(50, 7, 200, 3, 90), rank=4, keep=1.0, lr=0.05,
             reduce="sum", use_graph=False, ptmzr=None):
    import deepctr_b200.models as M
    import deepctr_oracle as O
    cls = getattr(M, cls_name)
    cat_sizes, offsets, X_dim, X_feas = criteo_like(rng, n_num, cat_sizes)
    d = rank + 1
    cw = {"none": 0, "ip": X_feas * (X_feas - 1) // 2, "op": d * d}[cls.cross]
    dims = [X_feas * d + cw + 1] + list(hidden) + [1]
    vm = {"W": rng.uniform(-0.3, 0.3, (X_dim, 1)), "V": rng.uniform(-0.3, 0.3, (X_dim, rank)), "b": np.array([0.1])}
    for i in range(len(dims) - 1):
        vm["h%d_w" % (i + 1)] = rng.uniform(-1, 1, (dims[i], dims[i + 1])) / np.sqrt(dims[i])
        vm["h%d_b" % (i + 1)] = rng.uniform(-0.1, 0.1, dims[i + 1])
    vm = {k: v.astype(np.float32) for k, v in vm.items()}
    path = str(tmp_path / (cls_name + ".pickle"))
    with open(path, "wb") as f:
        pickle.dump(vm, f, protocol=2)
    rch = [X_dim, X_feas, rank] + list(hidden) + [act]
    model = cls(cat_sizes, offsets, N, rch, ["uniform", -0.01, 0.01, [1, 2, 3, 4, 5, 6, 7, 8, 9, 10], path],
                ptmzr or ["sgd", lr, reduce], [keep], "train", 2 * N)
    model.use_graph = use_graph
    table = np.concatenate([vm["W"], vm["V"]], 1)
    net = O.OracleNet(table, vm["b"][0], [vm["h%d_w" % (i + 1)] for i in range(len(dims) - 1)],
                      [vm["h%d_b" % (i + 1)] for i in range(len(dims) - 1)], X_feas, d, cls.cross, False, act, act)
    return model, net, (cat_sizes, offsets, n_num)


def fnn_batch(rng, N, cat_sizes, offsets, n_num):
    c_ids = np.stack([rng.integers(0, s, N) for s in cat_sizes], 1)
    c_ids[N // 2] = c_ids[0]
    v_wts = rng.random((N, n_num)).astype(np.float32)
    c_wts = np.ones((N, len(cat_sizes)), np.float32)
    y = rng.integers(0, 2, N)
    g_ids = np.concatenate([np.broadcast_to(np.arange(n_num), (N, n_num)), c_ids + np.asarray(offsets)], 1)
    g_wts = np.concatenate([v_wts, c_wts], 1)
    return v_wts, c_ids, c_wts, y, g_ids, g_wts


def compare_state(model, net, rtol=RTOL):
    vm = model.var_map()
    close(np.concatenate([vm["W"], vm["V"]], 1), net.table, rtol, 2e-6)
    close(vm["b"], [net.bias], rtol, 2e-6)
    for i in range(len(net.Ws)):
        close(vm["h%d_w" % (i + 1)], net.Ws[i], rtol, 2e-6)
        close(vm["h%d_b" % (i + 1)], net.bs[i], rtol, 2e-6)


@pytest.mark.parametrize("cls_name,hidden,act,reduce", [
    ("FNN", [40, 30], "tanh", "sum"), ("FNN", [33, 17], "relu", "mean"), ("FNN_IP_L3", [40, 30, 20], "tanh", "sum"),
    ("FNN_OP_L3", [40, 30, 20], "sigmoid", "sum"), ("FNN_IP_L5", [40, 36, 30, 20, 10], "relu", "sum"),
    ("FNN_IP_L7", [48, 40, 36, 30, 20, 10, 6], "tanh", "sum")])
def test_tf_family_steps(tmp_path, cls_name, hidden, act, reduce):
    rng = np.random.default_rng(101)
    N = 96
    model, net, (cat_sizes, offsets, n_num) = make_fnn(cls_name, tmp_path, rng, N, hidden, act, reduce=reduce)
    sess = model.session()
    for step in range(3):
        v_wts, c_ids, c_wts, y, g_ids, g_wts = fnn_batch(rng, N, cat_sizes, offsets, n_num)
        _, loss, preds = sess.run([model.ptmzr, model.loss, model.train_preds],
                                  feed_dict={model.v_wt_hldr: v_wts, model.c_id_hldr: c_ids, model.c_wt_hldr: c_wts, model.lbl_hldr: y})
        oloss, opreds = net.train_step_tf(g_ids, g_wts, y, model.lr, reduce=reduce)
        close([loss], [oloss], 2e-5)
        close(preds, opreds, 2e-5)
        compare_state(model, net, 5e-5 if step else 2e-5)
    # eval path: global ids [E, X_feas] (FNN.py:65-69), E not a multiple of the batch
    v_wts, c_ids, c_wts, y, g_ids, g_wts = fnn_batch(rng, 150, cat_sizes, offsets, n_num)
    ev = model.eval_preds.eval(feed_dict={model.eval_id_hldr: g_ids, model.eval_wts_hldr: g_wts})
    close(ev, net.predict(g_ids, g_wts), 5e-5)


@pytest.mark.parametrize("cls_name,hidden,use_graph", [("FNN", [40, 30], False), ("FNN_IP_L3", [40, 30, 20], True)])
def test_tf_family_adam_steps(tmp_path, cls_name, hidden, use_graph):
    """['adam', lr, eps, reduce] (tf_util.py:18-21; what baseline.py:93 runs): TF's dense Adam over table + MLP.  Rows
    untouched by a step still move (their m decays), and the bias correction advances inside the replayed CUDA graph."""
    rng = np.random.default_rng(202)
    N = 64
    model, net, (cat_sizes, offsets, n_num) = make_fnn(cls_name, tmp_path, rng, N, hidden, "tanh", use_graph=use_graph,
                                                       ptmzr=["adam", 1e-2, 1e-8, "sum"])
    assert "epsilon" in model.log
    for step in range(6):
        v_wts, c_ids, c_wts, y, g_ids, g_wts = fnn_batch(rng, N, cat_sizes, offsets, n_num)
        if step >= 3:
            c_ids[:, 0] = 1                     # field 0 now only touches row 1: its other rows coast on their momentum
            g_ids[:, n_num] = 1 + offsets[0]
        loss, preds = model.train_step({model.v_wt_hldr: v_wts, model.c_id_hldr: c_ids, model.c_wt_hldr: c_wts, model.lbl_hldr: y})
        before = net.table.copy()
        oloss, opreds = net.train_step_tf_adam(g_ids, g_wts, y, 1e-2, 1e-8)
        close([loss], [oloss], 2e-5)
        close(preds, opreds, 5e-5)
        # Adam divides by sqrt(v) + eps: early steps move every touched weight by ~lr whatever the gradient's size, so
        # the fp32-vs-fp64 comparison is on the weights, with an absolute floor of a few 1e-6
        compare_state(model, net, 2e-4)
    untouched = np.setdiff1d(np.arange(offsets[0], offsets[0] + cat_sizes[0]), [1 + offsets[0]])
    moved = np.abs(net.table[untouched] - before[untouched]).max()
    assert moved > 0, "dense Adam: rows without a gradient this step must still move"


def test_tf_dropout_masks_injected(tmp_path):
    rng = np.random.default_rng(202)
    N = 64
    model, net, (cat_sizes, offsets, n_num) = make_fnn("FNN_IP_L3", tmp_path, rng, N, [24, 16, 8], "tanh", keep=0.5)
    eng = model.engine
    masks = [((rng.random((N, w)) < 0.5) / 0.5).astype(np.float32) for w in eng.widths]
    model.set_dropout_masks(masks)
    v_wts, c_ids, c_wts, y, g_ids, g_wts = fnn_batch(rng, N, cat_sizes, offsets, n_num)
    loss, preds = model.train_step({model.v_wt_hldr: v_wts, model.c_id_hldr: c_ids, model.c_wt_hldr: c_wts, model.lbl_hldr: y})
    oloss, opreds = net.train_step_tf(g_ids, g_wts, y, model.lr, masks=[m.astype(np.float64) for m in masks])
    close([loss], [oloss], 2e-5)
    close(preds, opreds, 2e-5)
    compare_state(model, net, 2e-5)


def test_fm_steps_and_checkpoint(tmp_path):
    import deepctr_oracle as O
    from deepctr_b200.models import FM
    rng = np.random.default_rng(303)
    N, X, F, rank = 128, 3000, 16, 10
    for lam, reduce in [(0.0, "sum"), (1e-2, "mean")]:
        fm = FM(N, [X, F, rank], ["uniform", -0.05, 0.05, [0x3210, 0x7654], None], ["sgd", 0.05, reduce], [lam], "train", 100)
        fm.use_graph = False
        ofm = O.OracleFM(fm.engine.get_table(), 0.0)
        for step in range(3):
            ids = rng.integers(0, X - 1, (N, F))
            wts = np.ones((N, F), np.float32)
            ids[:, -2:] = X - 1
            wts[:, -2:] = 0           # rows padded to X_feas with (X_dim-1, 0) (ipinyou.py:58-59)
            y = rng.integers(0, 2, N)
            loss, preds = fm.train_step({fm.sp_id_hldr: ids.reshape(-1), fm.sp_wt_hldr: wts.reshape(-1), fm.lbl_hldr: y})
            oloss, opreds = ofm.train_step(ids, wts, y, 0.05, lam, reduce)
            if lam == 0.0:     # with lambda != 0 the reported loss leaves out the l2 term of FM.py:39 (its gradient IS applied)
                close([loss], [oloss], 2e-5)
            close(preds, opreds, 2e-5)
            close(fm.engine.get_table(), ofm.table, 2e-5, 2e-6)
            close(fm.engine.bias().cpu(), [ofm.b], 2e-5, 2e-6)
        ev = fm.eval_preds.eval(feed_dict={fm.eval_wt_hldr: wts.reshape(-1), fm.eval_id_hldr: ids.reshape(-1)})
        close(ev, ofm.predict(ids, wts), 2e-5)
        p = str(tmp_path / "fm.pickle")
        fm.dump(p)
        fm2 = FM(N, [X, F, rank], ["uniform", -0.05, 0.05, [1, 2], p], None, None, "test", None)
        close(fm2.test_preds.eval(feed_dict={fm2.sp_id_hldr: ids.reshape(-1), fm2.sp_wt_hldr: wts.reshape(-1)}), ev, 1e-6)


@pytest.mark.parametrize("kind,lam,reduce,use_graph", [("fm", 0.0, "sum", False), ("fm", 1e-2, "mean", True), ("lr", 1e-2, "sum", True)])
def test_fm_lr_adam_steps(kind, lam, reduce, use_graph):
    """['adam', lr, eps, reduce] for FM / LR (what baseline.py runs on Criteo): dense TF Adam over W, V, b with the L2
    gradient lam * theta inside g (FM.py:39, LR.py:37); the fold only produces the per-row gradient (dctr_scatter_grad_fm)"""
    import deepctr_oracle as O
    from deepctr_b200.models import FM, LR
    rng = np.random.default_rng(505)
    N, X, F, rank = 64, 400, 8, 6
    init = ["uniform", -0.05, 0.05, [11, 12], None]
    if kind == "fm":
        m = FM(N, [X, F, rank], init, ["adam", 1e-2, 1e-8, reduce], [lam], "train", 50)
    else:
        m = LR(N, [X, F], init, ["adam", 1e-2, 1e-8, reduce], [lam], "train", 50)
    m.use_graph = use_graph
    o = O.OracleFM(m.engine.get_table(), 0.0)
    for step in range(6):
        ids = rng.integers(0, X // 2 if step >= 3 else X, (N, F))       # from step 3 on half of the rows only coast
        ids[:, 0] = 7
        wts = rng.uniform(0.5, 1.5, (N, F)).astype(np.float32)
        y = rng.integers(0, 2, N)
        loss, preds = m.train_step({m.sp_id_hldr: ids.reshape(-1), m.sp_wt_hldr: wts.reshape(-1), m.lbl_hldr: y})
        oloss, opreds = o.train_step_adam(ids, wts, y, 1e-2, 1e-8, lam, reduce)
        close(preds, opreds, 5e-5)
        close(m.engine.get_table(), o.table, 2e-4, 2e-6)
        close(m.engine.bias().cpu(), [o.b], 2e-4, 2e-6)
    if lam == 0.0:
        close([loss], [oloss], 2e-5)


def test_evaluate_on_device_matches_host_metrics():
    """model.evaluate: chunked predict + AUC / log-loss / RMSE without copying the predictions back == sklearn on the
    host-side predictions of the same model"""
    from sklearn.metrics import roc_auc_score
    from deepctr_b200.models import FM
    rng = np.random.default_rng(606)
    N, X, F, rank = 64, 300, 8, 4
    fm = FM(N, [X, F, rank], ["uniform", -0.5, 0.5, [3, 4], None], ["sgd", 0.05, "sum"], [0.0], "train", 100)
    E = 3 * N + 17
    ids = rng.integers(0, X, (E, F))
    wts = rng.uniform(0.5, 1.5, (E, F)).astype(np.float32)
    y = rng.integers(0, 2, E)
    host = fm.eval_preds.eval(feed_dict={fm.eval_id_hldr: ids.reshape(-1), fm.eval_wts_hldr: wts.reshape(-1)})
    got = fm.evaluate(ids, wts, y)
    np.testing.assert_allclose(got["auc"], roc_auc_score(y, host), rtol=1e-12)
    np.testing.assert_allclose(got["rmse"], np.sqrt(np.mean((host.astype(np.float64) - y) ** 2)), rtol=1e-9)
    assert got["n_pos"] == int(y.sum())


def test_lr_steps_and_checkpoint(tmp_path):
    """LR.py = the FM path with rank 0: logit = sum_f x_f W[id_f] + b, dense L2 on W and b (LR.py:37, 52-53)"""
    import deepctr_oracle as O
    from deepctr_b200.LR import LR
    rng = np.random.default_rng(404)
    N, X, F = 96, 500, 16
    for lam, reduce in [(0.0, "sum"), (1e-2, "mean")]:
        lr_m = LR(N, [X, F], ["uniform", -0.05, 0.05, [7, 8], None], ["sgd", 0.05, reduce], [lam], "train", 100)
        lr_m.use_graph = False
        assert lr_m.engine.get_table().shape == (X, 1)
        olr = O.OracleFM(lr_m.engine.get_table(), 0.0)
        for step in range(3):
            ids = rng.integers(0, X, (N, F))
            ids[:, 0] = 5                       # a row hit by every sample
            wts = rng.uniform(0.5, 1.5, (N, F)).astype(np.float32)
            y = rng.integers(0, 2, N)
            loss, preds = lr_m.train_step({lr_m.sp_id_hldr: ids.reshape(-1), lr_m.sp_wt_hldr: wts.reshape(-1), lr_m.lbl_hldr: y})
            oloss, opreds = olr.train_step(ids, wts, y, 0.05, lam, reduce)
            want = opreds
            if lam == 0.0:
                close([loss], [oloss], 2e-5)
            close(preds, want, 2e-5)
            close(lr_m.engine.get_table(), olr.table, 2e-5, 2e-6)
            close(lr_m.engine.bias().cpu(), [olr.b], 2e-5, 2e-6)
        # the logit really is the plain weighted sum (no pairwise term at rank 0)
        t, b = lr_m.engine.get_table().astype(np.float64), float(lr_m.engine.bias().cpu()[0])
        ev = lr_m.eval_preds.eval(feed_dict={lr_m.eval_wt_hldr: wts.reshape(-1), lr_m.eval_id_hldr: ids.reshape(-1)})
        close(ev, 1.0 / (1.0 + np.exp(-((t[ids, 0] * wts).sum(1) + b))), 2e-5)
        p = str(tmp_path / "lr.pickle")
        lr_m.dump(p)
        lr2 = LR(N, [X, F], ["uniform", -0.05, 0.05, [1, 2], p], None, None, "test", None)
        close(lr2.test_preds.eval(feed_dict={lr2.sp_id_hldr: ids.reshape(-1), lr2.sp_wt_hldr: wts.reshape(-1)}), ev, 1e-6)


def make_theano(rng, N, X=640, F=16, k=11, h1=48, h2=24, dropout=1.0, lambda1=0.0, use_graph=False):
    import deepctr_oracle as O
    from deepctr_b200.theano_fnn import FNN_wnzh, init_mlp
    feat_field = np.repeat(np.arange(F), X // F)
    table = rng.uniform(-0.2, 0.2, (X, k))
    Ws, bs = init_mlp(1 + F * k, h1, h2)
    oWs, obs = O.theano_fnn_init(1 + F * k, h1, h2)
    assert all(np.array_equal(a, b) for a, b in zip(Ws, oWs))
    Ws[2] = rng.uniform(-0.2, 0.2, Ws[2].shape)       # w3 = 0 would zero every hidden gradient on step 1
    m = FNN_wnzh(dict(w_0=-0.3, table=table, feat_field=feat_field), batch_size=N, lr=0.01, dropout=dropout, hidden1=h1,
                 hidden2=h2, lambda1=lambda1, lambda_fm=0.1, weights=(Ws, bs))
    m.use_graph = use_graph
    net = O.OracleNet(table, -0.3, Ws, bs, F, k, "none", True, "none", "tanh")
    return m, net, X, F


def theano_batch(rng, N, X, F):
    feats = np.stack([rng.integers(0, X // F, N) + f * (X // F) for f in range(F)], 1)
    feats[:, 0] = 3                       # a hot feature shared by the whole batch: sequential decay matters
    feats[N // 3] = feats[0]
    perm = rng.permutation(F)
    return feats[:, perm], rng.integers(0, 2, N)      # features arrive in arbitrary order; slot = field


@pytest.mark.parametrize("dropout,lambda1", [(1.0, 0.0), (0.5, 0.0), (0.5, 0.01)])
def test_theano_fnn_steps(dropout, lambda1):
    rng = np.random.default_rng(404)
    N = 100
    m, net, X, F = make_theano(rng, N, dropout=dropout, lambda1=lambda1)
    for step in range(4):
        feats, y = theano_batch(rng, N, X, F)
        masks = None
        if dropout < 1.0:
            r1 = rng.binomial(1, dropout, (1, 48)).astype(np.float64)
            r2 = rng.binomial(1, dropout, (1, 24)).astype(np.float64)
            m.set_masks(r1, r2)
            masks = [None, r1, r2]
        loss, p = m.train(feats, y)
        ids, _ = m.arrange(feats)
        oloss, op = net.train_step_theano(ids, y, 0.01, 0.1, masks=masks, lambda1=lambda1, faithful=(step == 0))
        if lambda1 == 0:       # the reported loss is the cross-entropy sum; the lambda1 term only enters the gradient
            close([loss], [oloss], 3e-5)
        close(p, op, 3e-5)
        close(m.get_table(), net.table, 5e-5, 2e-6)
        w1, w2, w3, b1, b2, b3 = m.get_weights()
        close(w1, net.Ws[0], 5e-5, 2e-6)
        close(w3, net.Ws[2].reshape(-1), 5e-5, 2e-6)
        close([b3], net.bs[2], 5e-5, 2e-6)
    feats, _ = theano_batch(rng, 250, X, F)
    close(m.predict(feats), net.predict(m.arrange(feats)[0]), 5e-5)


def test_missing_field_and_unknown_slot():
    rng = np.random.default_rng(505)
    N = 32
    m, net, X, F = make_theano(rng, N)
    feats, y = theano_batch(rng, N, X, F)
    feats = feats.copy()
    feats[::3, 5] = -1                    # some samples lack one field: that slot of x stays 0 (FNN_wnzh.py:91-95)
    ids, wts = m.arrange(feats)
    assert (ids[::3] == m.pad_row).sum() == len(ids[::3])
    before = m.get_table().copy()
    loss, p = m.train(feats, y)
    # oracle with the same zero slot: give the pad a zero row
    net.table = np.concatenate([net.table, np.zeros((1, net.d))], 0)
    _, op = net.train_step_theano(ids, y, 0.01, 0.1)
    close(p, op, 3e-5)
    close(m.get_table(), net.table[:-1], 5e-5, 2e-6)
    assert np.abs(m.engine.get_table()[m.pad_row]).max() == 0.0
    untouched = np.setdiff1d(np.arange(X), feats[feats >= 0])
    assert np.array_equal(m.get_table()[untouched], before[untouched])


def test_graph_replay_equals_eager_and_is_deterministic(tmp_path):
    rng = np.random.default_rng(606)
    N = 128
    runs = []
    for use_graph in (False, True, True):
        r = np.random.default_rng(1)
        m, _, X, F = make_theano(np.random.default_rng(2), N, use_graph=use_graph)
        losses = []
        for step in range(6):
            feats, y = theano_batch(r, N, X, F)
            losses.append(m.train(feats, y)[0])
        runs.append((losses, m.get_table(), m.get_weights()[0]))
    for other in runs[1:]:
        assert runs[0][0] == other[0]                     # bit-identical losses: eager == graph == graph
        assert np.array_equal(runs[0][1], other[1])
        assert np.array_equal(runs[0][2], other[2])


def test_auc_matches_oracle_after_epochs():
    """AUC identical to 3 decimals after a fixed number of epochs on the synthetic demo set (north_star)."""
    from sklearn.metrics import roc_auc_score
    from deepctr_b200 import data
    rng = np.random.default_rng(1234)
    N, X, F, k = 100, 2048, 16, 11
    sizes = np.full(F, X // F)
    planted = rng.normal(0, 0.4, (X, k))
    tr = data.make_field_batch(rng, 2000, sizes, 1.1)
    te = data.make_field_batch(rng, 500, sizes, 1.1)
    ytr, yte = data.planted_labels(rng, tr, planted), data.planted_labels(rng, te, planted)
    m, net, _, _ = make_theano(np.random.default_rng(5), N, X=X, h1=64, h2=32)
    for epoch in range(3):
        for j in range(len(ytr) // N):
            sl = slice(j * N, (j + 1) * N)
            m.train(tr[sl], ytr[sl])
            net.train_step_theano(tr[sl], ytr[sl], 0.01, 0.1)
    a_gpu = roc_auc_score(yte, m.predict(te))
    a_cpu = roc_auc_score(yte, net.predict(te))
    assert round(a_gpu, 3) == round(a_cpu, 3), (a_gpu, a_cpu)
    close(m.predict(te), net.predict(te), 1e-3, 1e-4)
